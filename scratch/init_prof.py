# where does _FusedTrainer.__init__ spend its (variable) time?  Re-runs its body step by step with timers.
import sys, time, types, gc
sys.path.insert(0, ".")
import torch
import bench
from coolchic_b200 import ops
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
def tick(lbl, t0):
    torch.cuda.synchronize(); t = time.perf_counter(); print(f"  {lbl:14s} {1e3*(t-t0):8.2f} ms"); return time.perf_counter()
old = None
for rep in range(5):
    print("rep", rep)
    t = time.perf_counter()
    self = types.SimpleNamespace()
    cc = fe.coolchic_encoder; self.cc = cc
    self.spec = cc._spec()
    t = tick("spec", t)
    self.st = ops.FusedState(self.spec, 1, dev)
    t = tick("FusedState", t)
    self.target = target
    T._FusedTrainer.push(self)
    t = tick("push", t)
    st = self.st
    step = lambda: st.train_step(self.target, 1e-3, "ste", "none", optimize_mask=7, seed=rep)
    step()
    t = tick("warm step", t)
    T._FusedTrainer.push(self); st.arena[1:].zero_(); st.step.zero_(); st.philox_offset.zero_()
    t = tick("push+zero", t)
    g = torch.cuda.CUDAGraph()
    t = tick("CUDAGraph()", t)
    with torch.cuda.graph(g):
        step()
    t = tick("capture", t)
    for _ in range(300): g.replay()
    t = tick("300 replays", t)
    old = None
    t = tick("drop previous", t)
    gc.collect()
    t = tick("gc", t)
    old = (self, g)
