"""Where do the full-size latent-gradient mismatches sit, and who is closer to float64?  (run on the GPU box)
python scratch/diag_fullsize.py fs_hop_768x512"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import common as Cm
from test_gpu_fullsize import run_ours

name = sys.argv[1]
inp = Cm.fullsize_inputs(name)
z = np.load(os.path.join(Cm.GOLDEN_DIR, name + ".npz"))
dec, rate, grads = run_ours(inp)
# float64 oracle on the CPU (same graph as the reference, double precision)
O = Cm.O
p64 = {k: v.double() for k, v in inp["p"].items()}
leaves = {k: v.clone().requires_grad_(True) for k, v in p64.items()}
lats = [leaves[f"latent_grids.{i}.data"] for i in range(7)]
raw, r64, _ = O.coolchic_forward(leaves, inp["arch"], True, inp["noise_type"], inp["q_type"], torch.tensor(inp["T"]).double(),
                                 torch.tensor(inp["sigma"]).double(), inp["noise"].double(), latents=lats)
loss, _, _ = O.loss_ref(O.frame_output(raw, True), r64, inp["target"].double(), inp["lmbda"])
loss.backward()
for i in range(7):
    k = f"grad/latent_grids.{i}.data"
    idx = z["idx/" + k]; ref = z["val/" + k].astype(np.float64); mx = float(z["max/" + k])
    g = grads[f"latent_grids.{i}.data"].detach().reshape(-1).cpu().double().numpy()[idx]
    t = lats[i].grad.reshape(-1).numpy()[idx]
    h, w = lats[i].shape[-2:]
    e_our_ref = np.abs(g - ref) / mx; e_our_t = np.abs(g - t) / mx; e_ref_t = np.abs(ref - t) / mx
    print(f"level {i} ({h}x{w}) max|g| {mx:.3e}: ours-ref {e_our_ref.max():.2e}  ours-f64 {e_our_t.max():.2e}  ref-f64 {e_ref_t.max():.2e}"
          f"   n>3e-4: ours-ref {int((e_our_ref > 3e-4).sum())} ours-f64 {int((e_our_t > 3e-4).sum())} ref-f64 {int((e_ref_t > 3e-4).sum())} of {idx.size}")
    for j in np.argsort(-e_our_ref)[:4]:
        if e_our_ref[j] > 3e-4:
            y, x = divmod(int(idx[j]) % (h * w), w)
            print(f"      (y {y}, x {x}) ours {g[j]:+.6e} ref {ref[j]:+.6e} f64 {t[j]:+.6e}")
