"""Where do the 3 ms of the per-sample-weight forward + backward go?  (configs[4], 8 x 256x256 crops, vmap path)

    python scratch/ps_prof.py

Prints wall time per call, CUDA kernel time per call (torch.profiler) and the top kernels / CPU ops.
"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    from torch.func import functional_call

    from coolchic_b200.enc.component.nlcoolchic import LatentFreeCoolChicEncoder
    from coolchic_b200.enc.training.loss import loss_function

    dev = torch.device("cuda:0")
    B, Hh, Ww = 8, 256, 256
    torch.manual_seed(3)
    lf = LatentFreeCoolChicEncoder(bench.hop_param(Hh, Ww, gain=1)).to(dev)
    lf.train()
    leaves = {}
    for k, v in lf.named_parameters():
        if k.endswith(".bias") and "upsampling" in k:
            leaves[k] = v.detach().unsqueeze(0).expand(B, *v.shape)
        else:
            leaves[k] = (v.detach().unsqueeze(0) + 0.01 * torch.randn(B, *v.shape, device=dev)).requires_grad_(True)
    hw = [(-(-Hh // 2 ** i), -(-Ww // 2 ** i)) for i in range(7)]
    lats = [(3.0 * torch.randn(B, 1, h, w, device=dev)).requires_grad_(True) for h, w in hw]
    target = bench.synthetic_image(12, Hh, Ww, B).to(dev)
    kw = dict(quantizer_noise_type="gaussian", quantizer_type="softround",
              soft_round_temperature=torch.tensor(0.3, device=dev), noise_parameter=torch.tensor(0.25, device=dev),
              AC_MAX_VAL=-1, flag_additional_outputs=False)

    def fwd(lat_list, params):
        return functional_call(lf, params, (lat_list,), kwargs=kw)

    def one():
        for t in lats + [v for v in leaves.values() if v.requires_grad]:
            t.grad = None
        raw, rate, _ = torch.vmap(fwd, randomness="different")([t.unsqueeze(1) for t in lats], leaves)
        loss_function(raw.squeeze(1), rate.squeeze(1), target, lmbda=1e-3).loss.backward()

    for _ in range(5):
        one()
    torch.cuda.synchronize()
    n = 30
    t0 = time.perf_counter()
    for _ in range(n):
        one()
    torch.cuda.synchronize()
    print(f"wall per call: {(time.perf_counter() - t0) / n * 1e3:.3f} ms")
    from torch.profiler import ProfilerActivity, profile

    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for _ in range(10):
            one()
        torch.cuda.synchronize()
    ev = prof.key_averages()
    cuda_total = sum(e.self_device_time_total for e in ev) / 10 / 1e3
    print(f"CUDA kernel time per call: {cuda_total:.3f} ms")
    print(ev.table(sort_by="self_device_time_total", row_limit=22, max_name_column_width=60))
    print(ev.table(sort_by="self_cpu_time_total", row_limit=25, max_name_column_width=60))


if __name__ == "__main__":
    main()
