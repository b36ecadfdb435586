# fairness check for bench.py's cpu_baseline: the REAL reference (imported from /root/reference) against the oracle
# port (oracle/float_ref.py) on this container's CPU cores, same 768x512 HOP iteration (forward, backward, clip, Adam)
import os, sys, time
sys.path.insert(0, "tests/golden"); sys.path.insert(0, ".")
import torch
from refimport import import_reference
import_reference()
from make_golden import build_reference
from coolchic.enc.training.loss import loss_function
from oracle import float_ref as O
import bench

torch.set_num_threads(os.cpu_count())
H, W = bench.H, bench.W
fe, arch, p0 = build_reference("hop", H, W, 1, 16, seed=0)
target = O.synthetic_image(0, H, W)
fe.set_to_train()
opt = torch.optim.Adam(fe.parameters(), lr=1e-2)
def ref_step():
    out = fe.forward(reference_frames=[], quantizer_noise_type="kumaraswamy", quantizer_type="softround",
                     soft_round_temperature=torch.tensor(0.3), noise_parameter=torch.tensor(2.0), AC_MAX_VAL=-1,
                     flag_additional_outputs=False)
    lo = loss_function(out.decoded_image, out.rate, target, lmbda=1e-3, rate_mlp_bit=0.0, compute_logs=False)
    opt.zero_grad(); lo.loss.backward()
    torch.nn.utils.clip_grad_norm_(fe.parameters(), 1e-1, norm_type=2.0, error_if_nonfinite=False)
    opt.step()
ts = []
for i in range(6):
    t = time.time(); ref_step(); ts.append(time.time() - t)
ts = sorted(ts[2:]); print("reference  :", f"{1 / ts[len(ts) // 2]:.2f} it/s on {os.cpu_count()} cores")
its, cores, secs = bench.cpu_port_iters_per_s(4, 2)
print("oracle port:", f"{its:.2f} it/s on {cores} cores")
