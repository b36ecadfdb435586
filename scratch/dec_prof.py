"""Driver for profiling the fixed-point decoder kernels: three decodes of the kodim01 fixture (ncu captures the last)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from dec_common import load_dec  # noqa: E402
from coolchic_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
z, grids, prm = load_dec("kodim01-lmbda-0001")
H, W = (int(v) for v in z["img_hw"])
lat = torch.from_numpy(np.concatenate([g.reshape(-1) for g in grids]).astype(np.int32)).to(dev)
layers = [dict(w=torch.from_numpy(l["w"].astype(np.int32)).to(dev), b=torch.from_numpy(l["b"].astype(np.int32)).to(dev),
               ks=l["ks"], residual=l["residual"], relu=l["relu"]) for l in prm["syn"]]
for _ in range(3):
    r = ops.decode_int(lat, (H, W), len(grids), prm["ups_w"], prm["pre_w"], layers, want_ups=False)
torch.cuda.synchronize()
print("ok", int(r["out"].abs().max()))
