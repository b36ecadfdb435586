"""Where do the full-size latent-gradient mismatches sit, and who is closer to the reference's float64 run?
(run on the GPU box)   python scratch/diag_fullsize.py fs_hop_1080p"""
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
for i in range(7):
    k = f"grad/latent_grids.{i}.data"
    idx = z["idx/" + k].astype(np.int64); ref = z["val/" + k].astype(np.float64); t = z["val64/" + k]; mx = float(z["max/" + k])
    gt = grads[f"latent_grids.{i}.data"].detach()
    g = gt.reshape(-1).cpu().double().numpy()[idx]
    h, w = gt.shape[-2:]
    e_or, e_ot, e_rt = np.abs(g - ref) / mx, np.abs(g - t) / mx, np.abs(ref - t) / mx
    print(f"level {i} ({h}x{w}) max|g| {mx:.3e}: ours-ref {e_or.max():.2e}  ours-f64 {e_ot.max():.2e}  ref-f64 {e_rt.max():.2e}"
          f"   n>3e-4: ours-ref {int((e_or > 3e-4).sum())} ours-f64 {int((e_ot > 3e-4).sum())} ref-f64 {int((e_rt > 3e-4).sum())} of {idx.size}")
    for j in np.argsort(-np.maximum(e_or, e_ot))[:6]:
        if max(e_or[j], e_ot[j]) > 1e-4:
            y, x = divmod(int(idx[j]) % (h * w), w)
            print(f"      (y {y}, x {x}) ours {g[j]:+.6e} ref32 {ref[j]:+.6e} ref64 {t[j]:+.6e}   ours-f64 {e_ot[j]:.1e} ref-f64 {e_rt[j]:.1e}")
