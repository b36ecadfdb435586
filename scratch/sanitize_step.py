"""A few fused iterations + module forward/backward at small shapes, for compute-sanitizer:
    compute-sanitizer --tool memcheck  python scratch/sanitize_step.py
    compute-sanitizer --tool racecheck python scratch/sanitize_step.py
"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import common as Cm
from coolchic_b200 import ops

dev = torch.device("cuda:0")
HOP = Cm.ARCHS["hop"]
for (H, W) in ((64, 96), (135, 241), (256, 384)):
    spec = ops.ArchSpec(H, W, HOP["dim_arm"], HOP["nh"], list(HOP["syn"]), 7, 8, 7, 16.0)
    st = ops.FusedState(spec, 1, dev)
    st.load(Cm.O.init_params(Cm.make_arch("hop", H, W), seed=3, B=1, random=True))
    st.lr.fill_(1e-2); st.T.fill_(0.3); st.sigma.fill_(2.0)
    tg = Cm.O.synthetic_image(2, H, W).to(dev)
    for _ in range(2):
        st.train_step(tg, 1e-3, "softround", "kumaraswamy", seed=5)
    torch.cuda.synchronize()
    print(H, W, st.out4.cpu().tolist())
    st.close()
print("done")
