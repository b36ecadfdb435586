# replay rate of the captured iteration for the quantiser / noise combinations of the c3x phases (no validations)
import sys, time, types
sys.path.insert(0, ".")
import torch
import bench
from coolchic_b200.enc.training import train as T
from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
from coolchic_b200.enc.component.frame import FrameEncoder
from oracle import float_ref as O
dev = torch.device("cuda:0")
H, W, HOP = bench.H, bench.W, bench.HOP
prm = CoolChicEncoderParameter(layers_synthesis=list(HOP["syn"]), n_ft_per_res=[1] * 7, dim_arm=HOP["dim_arm"],
                               n_hidden_layers_arm=HOP["nh"], ups_k_size=8, ups_preconcat_k_size=7, encoder_gain=16)
prm.set_image_size((H, W))
fe = FrameEncoder(prm); fe.to_device(dev)
target = O.synthetic_image(100, H, W).to(dev)
for q, n, mask in [("softround", "kumaraswamy", T.MODULE_MASK["all"]), ("softround", "gaussian", T.MODULE_MASK["all"]), ("ste", "none", T.MODULE_MASK["all"]), ("ste", "none", T.MODULE_MASK["latent"])]:
    tr = T._FusedTrainer(fe, target, 1e-3, q, n, mask, 0)
    tr.st.lr.fill_(1e-3); tr.st.T.fill_(0.3); tr.st.sigma.fill_(0.25 if n == "gaussian" else 2.0)
    for _ in range(50): tr.graph.replay()
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(2000): tr.graph.replay()
    torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"{q:10s} {n:12s} mask {mask}: {dt / 2000 * 1e6:7.1f} us/it  {2000 / dt:7.0f} it/s", flush=True)
