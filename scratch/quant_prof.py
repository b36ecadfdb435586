# cProfile of quantize_model on a trained-ish 768x512 HOP encoder (host overhead per candidate)
import sys, time, types, cProfile, pstats, io, contextlib
sys.path.insert(0, ".")
import torch
import bench
from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
from coolchic_b200.enc.component.frame import FrameEncoder
from coolchic_b200.enc.training.quantizemodel import quantize_model
from coolchic_b200.enc.training.train import train
from oracle import float_ref as O
dev = torch.device("cuda:0")
H, W, HOP = bench.H, bench.W, bench.HOP
prm = CoolChicEncoderParameter(layers_synthesis=list(HOP["syn"]), n_ft_per_res=[1] * 7, dim_arm=HOP["dim_arm"],
                               n_hidden_layers_arm=HOP["nh"], ups_k_size=8, ups_preconcat_k_size=7, encoder_gain=16)
prm.set_image_size((H, W))
fe = FrameEncoder(prm); fe.to_device(dev)
target = O.synthetic_image(100, H, W).to(dev)
mgr = types.SimpleNamespace(lmbda=1e-3, iterations_counter=0, total_training_time_sec=0.0)
train(fe, target, mgr, max_iterations=600, patience=600)
sd = {k: v.clone() for k, v in fe.coolchic_encoder.state_dict().items()}
for rep in range(2):
    fe.coolchic_encoder.load_state_dict(sd)
    pr = cProfile.Profile()
    torch.cuda.synchronize(); t = time.perf_counter()
    pr.enable()
    with contextlib.redirect_stdout(io.StringIO()):
        quantize_model(fe, target, mgr)
    pr.disable()
    torch.cuda.synchronize(); print("quantize_model", round(time.perf_counter() - t, 3), "s")
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45); print(s.getvalue()[:9000])
