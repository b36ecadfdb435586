# does the replay rate drift with the iteration count (data-dependent kernels? clocks?)  chunks of 200 replays, CUDA events
import sys, time
sys.path.insert(0, ".")
import torch, pynvml
import bench
from coolchic_b200.enc.training import train as T
from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
from coolchic_b200.enc.component.frame import FrameEncoder
from oracle import float_ref as O
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
dev = torch.device("cuda:0")
H, W, HOP = bench.H, bench.W, bench.HOP
prm = CoolChicEncoderParameter(layers_synthesis=list(HOP["syn"]), n_ft_per_res=[1] * 7, dim_arm=HOP["dim_arm"],
                               n_hidden_layers_arm=HOP["nh"], ups_k_size=8, ups_preconcat_k_size=7, encoder_gain=16)
prm.set_image_size((H, W))
fe = FrameEncoder(prm); fe.to_device(dev)
target = O.synthetic_image(100, H, W).to(dev)
tr = T._FusedTrainer(fe, target, 1e-3, "softround", "kumaraswamy", T.MODULE_MASK["all"], 0)
tr.st.lr.fill_(1e-2); tr.st.T.fill_(0.3); tr.st.sigma.fill_(2.0)
for chunk in range(12):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); t = time.perf_counter()
    e0.record()
    for _ in range(200): tr.graph.replay()
    t_launch = time.perf_counter() - t
    clk = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
    e1.record(); torch.cuda.synchronize()
    wall = time.perf_counter() - t
    lat = tr.st.params["latent_grids.0.data"]
    print(f"chunk {chunk:2d}: {e0.elapsed_time(e1) / 200 * 1e3:6.1f} us/it (events)  wall {wall / 200 * 1e6:6.1f}  host launch {t_launch / 200 * 1e6:5.1f} us/it  "
          f"sm {clk} MHz  loss {float(tr.st.out4[0]):.5f}  nonzero latents {float((lat.abs() > 0.03).float().mean()):.3f}", flush=True)
