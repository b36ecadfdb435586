#!/usr/bin/env python
"""bench.py -- overfitting iterations/s of one Kodak-shaped 768x512 image per B200 (BASELINE.json).

A "step" is one iteration of the reference's encode loop (enc/training/train.py:244-268: forward in
train mode, loss, backward, clip_grad_norm_(0.1), Adam) on a synthetic 768x512 RGB image with the HOP
decoder (arm 16,2 / synthesis 48-1,3-1,3-3,3-3), softround + kumaraswamy noise as phase 1 of
preset_cfg/c3x.yaml.  N > 1: one process / image per GPU (image-sharded, no collective on the path;
one all_gather of the per-image results at the end), weak scaling.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Prints ONE JSON line (rank 0).  Keys are documented in DESIGN.md ("Measurement").
"""

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H, W = 512, 768
HOP = dict(dim_arm=16, nh=2, syn=["48-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"])
LMBDA, LR = 1e-3, 1e-2
Q_TYPE, NOISE_TYPE, T0, SIGMA0 = "softround", "kumaraswamy", 0.3, 2.0
METRIC = "overfit_iters_per_s_768x512"
UNIT = "it/s"
# the SAME dict in both arms (ours / --impl reference): what is computed, not how it is timed
CONFIG = {"workload": "768x512 RGB (configs[1]), HOP decoder (arm 16,2 / syn 48-1,3-1,3-3,3-3), one iteration of "
                      "train.py:244-268 (forward in train mode, loss, backward, clip_grad_norm_(0.1), Adam), "
                      "softround + kumaraswamy noise, one image per GPU",
          "lambda": LMBDA, "lr": LR, "M": 524256, "N": H * W}


def synthetic_image(idx: int, Hh: int, Ww: int, B: int = 1):
    """Seeded smooth-plus-texture test image, [B,3,H,W] in [0,1] on 8-bit levels (SURVEY.md 8d "synthetic inputs":
    a sum of six random sinusoids per channel plus box-blurred noise, what read_png would yield)."""
    import math

    import torch
    import torch.nn.functional as F

    g = torch.Generator().manual_seed(1000 + idx)
    v, u = torch.meshgrid(torch.linspace(0, 1, Hh), torch.linspace(0, 1, Ww), indexing="ij")
    out = torch.empty(B, 3, Hh, Ww)
    for b in range(B):
        for c in range(3):
            x = torch.full((Hh, Ww), 0.5)
            for k in range(1, 7):
                a = torch.rand((), generator=g) / k
                fx, fy = 8 * torch.rand((), generator=g), 8 * torch.rand((), generator=g)
                ph = 2 * math.pi * torch.rand((), generator=g)
                x = x + 0.25 * a * torch.sin(2 * math.pi * (fx * u + fy * v) + ph)
            n = 0.03 * torch.randn(1, 1, Hh, Ww, generator=g)
            n = F.avg_pool2d(F.pad(n, (2, 2, 2, 2), mode="replicate"), 5, stride=1)
            out[b, c] = x + n[0, 0]
    return torch.round(out.clamp(0, 1) * 255) / 255


def hop_param(Hh: int, Ww: int, gain: int = 16):
    from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter

    prm = CoolChicEncoderParameter(layers_synthesis=list(HOP["syn"]), n_ft_per_res=[1] * 7, dim_arm=HOP["dim_arm"],
                                   n_hidden_layers_arm=HOP["nh"], ups_k_size=8, ups_preconcat_k_size=7, encoder_gain=gain)
    prm.set_image_size((Hh, Ww))
    return prm


def initial_params(Hh: int, Ww: int, B: int, seed: int):
    """The reference's own initialisation (zero latents, bicubic / Dirac filters, zero residual layers, random ARM output
    and synthesis head: coolchic.py:513-531, arm.py:66-84, synthesis.py:86-116), drawn by the module mirror."""
    import torch

    from coolchic_b200.enc.component.coolchic import CoolChicEncoder

    torch.manual_seed(seed)
    sd = {k: v.detach().clone() for k, v in CoolChicEncoder(hop_param(Hh, Ww)).state_dict().items()}
    if B > 1:
        for k in list(sd):
            if k.startswith("latent_grids."):
                sd[k] = sd[k].repeat(B, 1, 1, 1)
    return sd


def pin_to_gpu_numa(gpu_index: int):
    """Bind this rank to the host cores next to its GPU: /sys/bus/pci/devices/<bus id>/local_cpulist."""
    try:
        bus = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(gpu_index)],
                             capture_output=True, text=True, timeout=10).stdout.strip().lower()
        bus = bus[-12:] if len(bus) > 12 else bus  # 00000000:1b:00.0 -> 0000:1b:00.0
        cpus = set()
        for part in open(f"/sys/bus/pci/devices/{bus}/local_cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return {"cores": len(cpus), "first": min(cpus), "last": max(cpus)}
    except Exception as e:  # no sysfs entry, container without the device tree: run unpinned
        return {"unpinned": str(e)[:80]}
    return {"unpinned": "empty local_cpulist"}


def arm_macs(C, nh):
    return nh * C * C + 2 * C


def algorithmic_flops(M, N, C, nh, S):
    """2 FLOP per MAC forward, 4 backward (input + weight gradients)."""
    arm = M * arm_macs(C, nh)
    syn = N * (7 * S + S * 3 + 81 + 81)
    ups = N * 64  # separable polyphase + refine, all stages (~64 MAC / pixel)
    return dict(arm_fwd=2 * arm, arm_bwd=4 * arm, syn_fwd=2 * syn, syn_bwd=4 * syn, ups_fwd=2 * ups, ups_bwd=4 * ups,
                total=6 * (arm + syn + ups))


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        super().__init__(daemon=True)
        self.gpu, self.rows, self._stop_evt = gpu_index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                      str(self.gpu)], capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        reasons = []
        for i, name in enumerate(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")):
            if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def cpu_port_iters_per_s(n_timed: int, n_warm: int = 2, threads=None, device="cpu", Hh=H, Ww=W):
    """The reference's PyTorch path for this step, restated by oracle/float_ref.py with the same torch ops
    (F.conv2d / conv_transpose2d / unfold / index_select / linear + autograd + clip + Adam).  device "cpu": on
    `threads` host cores (default: all).  device "cuda": the same eager ops on the GPU (the bar SURVEY.md 2.2 names:
    the reference's own PyTorch path on the same B200).  Returns (it/s, cores, seconds spent).
    This baseline leg is the only place bench.py runs anything of oracle/."""
    import torch

    from oracle import float_ref as O

    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    arch = O.Arch(img_size=(Hh, Ww), dim_arm=HOP["dim_arm"], n_hidden_layers_arm=HOP["nh"], layers_synthesis=HOP["syn"])
    p = O.init_params(arch, seed=0, random=False)
    g = torch.Generator().manual_seed(0)
    p["arm.mlp.4.weight"] = torch.randn(2, 16, generator=g) / 4.0
    target = O.synthetic_image(0, Hh, Ww)
    T, S = torch.tensor(T0), torch.tensor(SIGMA0)
    if device != "cpu":
        p = {k: v.to(device) for k, v in p.items()}
        target, T, S = target.to(device), T.to(device), S.to(device)
        torch.backends.cuda.matmul.allow_tf32 = False  # fp32 like the CPU path ("highest")
        torch.backends.cudnn.allow_tf32 = False
    state = {}
    t_all = time.time()
    times = []
    for it in range(n_warm + n_timed):
        if device != "cpu":
            torch.cuda.synchronize()
        t0 = time.time()
        O.train_step_ref(p, arch, target, state, LMBDA, LR, NOISE_TYPE, Q_TYPE, T, S)
        if device != "cpu":
            torch.cuda.synchronize()
        if it >= n_warm:
            times.append(time.time() - t0)
    times.sort()
    med = times[len(times) // 2]
    return 1.0 / med, cores, time.time() - t_all


def cpu_debug_preset_wall(Hh=H, Ww=W):
    """wall time of warmup() + the three train() phases of preset_cfg/debug.yaml on the host cores
    (warm-up 3 x 10 + 2 x 10 iterations, phases 50 + 10 + 10 iterations, freq_valid 400 / 100 / 5 => one validation at
    the end of each train() call plus two in the last phase): 120 training iterations and 10 eval-mode validations of
    the oracle port, all host cores."""
    import torch

    from oracle import float_ref as O

    torch.set_num_threads(os.cpu_count() or 1)
    arch = O.Arch(img_size=(Hh, Ww), dim_arm=HOP["dim_arm"], n_hidden_layers_arm=HOP["nh"], layers_synthesis=HOP["syn"])
    target = O.synthetic_image(0, Hh, Ww)
    T, S = torch.tensor(T0), torch.tensor(SIGMA0)

    def validate(p):
        with torch.no_grad():
            raw, rate, _ = O.coolchic_forward(p, arch, False)
            O.loss_ref(O.frame_output(raw, False), rate, target, LMBDA)

    t0 = time.time()
    n_it = n_val = 0
    cands = [O.init_params(arch, seed=i, random=False) for i in range(3)]
    for p in cands:
        g = torch.Generator().manual_seed(1)
        p["arm.mlp.4.weight"] = torch.randn(2, 16, generator=g) / 4.0
    for n_cand in (3, 2):  # debug.yaml warm-up: 3 candidates x 10 iterations, best 2 x 10 more
        for p in cands[:n_cand]:
            state = {}
            validate(p)  # train() opens with test()
            for _ in range(10):
                O.train_step_ref(p, arch, target, state, LMBDA, LR, NOISE_TYPE, Q_TYPE, T, S)
            validate(p)
            validate(p)  # warmup() scores the candidate
            n_it, n_val = n_it + 10, n_val + 3
    p = cands[0]
    for iters, freq in ((50, 400), (10, 100), (10, 5)):
        state = {}
        validate(p)
        n_val += 1
        for i in range(iters):
            O.train_step_ref(p, arch, target, state, LMBDA, LR, NOISE_TYPE, Q_TYPE, T, S)
            if (i + 1) % freq == 0 or i + 1 == iters:
                validate(p)
                n_val += 1
        n_it += iters
    return {"seconds": time.time() - t0, "iterations": n_it, "validations": n_val,
            "preset": "preset_cfg/debug.yaml (warm-up 3x10 + 2x10, phases 50 + 10 + 10), 768x512, HOP"}


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the step on the box's host cores.
    /root/reference does not exist on the GPU box, so this times the oracle port (same torch ops)."""
    if rank != 0:
        return
    n = max(3, min(args.steps, 30))
    its, cores, secs = cpu_port_iters_per_s(n_timed=n, n_warm=max(1, min(args.warmup, 3)))
    line = {
        "impl": "reference", "metric": METRIC, "value": its, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 / its, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": CONFIG,
        "timing": "CPU PyTorch path (oracle/float_ref.py, the reference's torch ops) on all host cores, wall clock",
        "cpu_baseline": {"value": its, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"median of {n} iterations of the same 768x512 workload ({secs:.0f} s of CPU work)"},
        "e2e": {"value": its, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


_INIT_LOCK = threading.Lock()

C3X_PHASES = [  # /root/reference/preset_cfg/c3x.yaml `all_phases` (the warm-up search before them: C3X_WARMUP below)
    dict(start_lr=1e-2, max_iterations=10600, patience=5000, optimized_module=("all",), cosine_scheduling_lr=True,
         quantizer_type="softround", quantizer_noise_type="gaussian", softround_temperature=(0.3, 0.1),
         noise_parameter=(0.25, 0.1)),
    dict(start_lr=1e-4, max_iterations=1500, patience=1500, optimized_module=("all",), cosine_scheduling_lr=True,
         quantizer_type="ste", quantizer_noise_type="none", softround_temperature=(1e-4, 1e-4),
         noise_parameter=(1.0, 1.0), quantize_model=True),
    dict(start_lr=1e-4, max_iterations=1000, patience=50, optimized_module=("latent",), cosine_scheduling_lr=False,
         quantizer_type="ste", quantizer_noise_type="none", softround_temperature=(1e-4, 1e-4),
         noise_parameter=(1.0, 1.0), frequency_validation=10),
]


_C3X_WARMUP_TP = dict(lr=1e-2, max_itr=400, freq_valid=400, patience=100000, schedule_lr=False,
                      softround_temperature=[0.3, 0.3], noise_parameter=[2.0, 2.0], quantizer_noise_type="kumaraswamy",
                      quantizer_type="softround", optimized_module=["all"])
C3X_WARMUP = {"phases": [{"candidates": 5, "training_phase": _C3X_WARMUP_TP},  # preset_cfg/c3x.yaml `warmup`
                         {"candidates": 2, "training_phase": _C3X_WARMUP_TP}]}


def run_c3x_encode(dev, image_index, extras=True):
    """One whole image through the reference-facing API: warmup() over five random FrameEncoders, then train() over
    the three training phases of preset_cfg/c3x.yaml (with their validations, best-model bookkeeping and schedules)
    and quantize_model().  Runs on the CURRENT stream (several may be in flight, one per host thread)."""
    import types

    import torch

    from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
    from coolchic_b200.enc.component.frame import FrameEncoder
    from coolchic_b200.enc.training.quantizemodel import quantize_model
    from coolchic_b200.enc.training.train import test as evaluate
    from coolchic_b200.enc.training.train import train
    from coolchic_b200.enc.training.warmup import warmup

    rank = image_index
    with _INIT_LOCK:  # the candidates' random initialisation draws from torch's global generator
        torch.manual_seed(4321 + image_index)
        prm = hop_param(H, W)
        candidates = [FrameEncoder(prm) for _ in range(C3X_WARMUP["phases"][0]["candidates"])]  # 5 random initialisations
    target = synthetic_image(1000 + image_index, H, W).to(dev)  # seeds 1000 .. 1023 + 1000
    sync = lambda: torch.cuda.current_stream(dev).synchronize()  # noqa: E731
    mgr = types.SimpleNamespace(lmbda=LMBDA, iterations_counter=0, total_training_time_sec=0.0,
                                preset={"warmup": C3X_WARMUP})
    sync()
    t0 = time.perf_counter()
    fe = warmup(mgr, candidates, target, dev)
    sync()
    phases = [{"warmup_seconds": round(time.perf_counter() - t0, 3), "iterations": mgr.iterations_counter,
               "candidates": [ph["candidates"] for ph in C3X_WARMUP["phases"]]}]
    for i, ph in enumerate(C3X_PHASES):
        t1 = time.perf_counter()
        n0 = mgr.iterations_counter
        ph = dict(ph)
        do_quantize = ph.pop("quantize_model", False)
        train(fe, target, mgr, seed=1000 * rank + i, **ph)
        sync()
        dt = time.perf_counter() - t1
        phases.append({"iterations": mgr.iterations_counter - n0, "seconds": round(dt, 3),
                       "q": ph["quantizer_type"], "optimized": list(ph["optimized_module"])})
        if do_quantize:  # the NN weight quantisation search of the preset (about 490 full-size eval forwards)
            t2 = time.perf_counter()
            quantize_model(fe, target, mgr, verbose=False)
            sync()
            phases.append({"quantize_model_seconds": round(time.perf_counter() - t2, 3),
                           "q_step": {m: {k: float(v) for k, v in d.items()}
                                      for m, d in fe.coolchic_encoder.nn_q_step.items()}})
    secs = time.perf_counter() - t0
    logs = evaluate(fe, target, mgr)
    row = {"seconds": secs, "iterations": mgr.iterations_counter, "loss": float(logs.loss), "mse": float(logs.mse),
           "psnr_db": float(logs.psnr_db), "rate_latent_bpp": float(logs.rate_latent_bpp),
           "rate_nn_bpp": float(logs.rate_nn_bpp)}
    if not extras:
        return row
    from coolchic_b200.enc.utils.decoder_int import decoder_image

    dec = decoder_image(fe, target)  # what the C++ decoder would reconstruct (fixed-point emulation, not timed)
    from coolchic_b200.enc.bitstream.encode import latent_layers_for_bac

    latent_layers_for_bac(fe)
    torch.cuda.synchronize(dev)
    t3 = time.perf_counter()
    layers = latent_layers_for_bac(fe)  # (symbols, mu, log-scale) of every latent for the CABAC coder: integer ARM kernel
    handoff_ms = (time.perf_counter() - t3) * 1e3
    row.update({"phases": phases, "decoder_psnr_db": float(dec["psnr_db"]), "bitstream_handoff_ms": round(handoff_ms, 3),
                "latent_symbols": int(sum(l["y"].numel() for l in layers))})
    return row


def run_config3(dev, n_images, in_flight):
    """BASELINE.json configs[2]: n_images Kodak-shaped synthetic images, complete c3x preset each, image i on rank
    i mod world (enc/training/sharding.py), `in_flight` images at a time per GPU, ONE all_gather of the result rows
    at the end (NCCL when world > 1).  -> (rows of all images, this rank's wall seconds incl. the gather)."""
    import torch

    from coolchic_b200.enc.training import sharding

    def encode_one(i):
        return run_c3x_encode(dev, i, extras=False)

    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    rows = sharding.encode_images(n_images, encode_one, device=dev, concurrency=in_flight)
    torch.cuda.synchronize(dev)
    return rows, time.perf_counter() - t0


def time_other_workload(dev, Hh, Ww, B, peak_fp32, flush, n_iter=60):
    """The same fused step on another BASELINE.json shape: iterations/s with L2 flushed before every iteration (CUDA
    events around each graph replay) and back to back, plus the whole step against the fp32 peak."""
    import torch

    from coolchic_b200 import ops

    spec = ops.ArchSpec(Hh, Ww, HOP["dim_arm"], HOP["nh"], HOP["syn"], 7, 8, 7, 16.0)
    st = ops.FusedState(spec, B, dev)
    st.load(initial_params(Hh, Ww, B, seed=7))
    st.lr.fill_(LR)
    st.T.fill_(T0)
    st.sigma.fill_(SIGMA0)
    target = synthetic_image(11, Hh, Ww, B).to(dev)
    for _ in range(3):
        st.train_step(target, LMBDA, Q_TYPE, NOISE_TYPE, seed=5)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        st.train_step(target, LMBDA, Q_TYPE, NOISE_TYPE, seed=5)
    for _ in range(3):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n_iter):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    ms_warm = e0.elapsed_time(e1) / n_iter
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n_iter)]
    for a, b in ev:
        flush.zero_()
        a.record()
        g.replay()
        b.record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev) / n_iter
    M, Npix = st.plan.M, Hh * Ww
    flops = B * algorithmic_flops(M, Npix, HOP["dim_arm"], HOP["nh"], 48)["total"]
    alg_bytes = B * (28 * M + 12 * Npix)
    return {"it_per_s": 1e3 / ms, "ms_per_step": ms, "it_per_s_warm_l2": 1e3 / ms_warm, "H": Hh, "W": Ww, "B": B,
            "M_per_image": M, "l2": "flushed before every timed iteration",
            "step_fp32": {"achieved": flops / (ms * 1e-3) / 1e12, "peak": peak_fp32, "unit": "TFLOP/s",
                          "frac": flops / (ms * 1e-3) / 1e12 / peak_fp32},
            "algorithmic_bytes_per_step": alg_bytes, "state_bytes": int(4 * st.arena.numel() + st.ws.numel())}


def time_per_sample_crops(dev, B=8, Hh=256, Ww=256, n_iter=30):
    """BASELINE.json configs[4], the DeltaWholeNet shape: LatentFreeCoolChicEncoder under
    torch.vmap(functional_call(...), randomness="different") with ONE WEIGHT SET PER CROP (hypernet/hypernet.py:1266-1288),
    forward + loss + backward through the module API (the per-sample-weight kernels, one launch set for the batch)."""
    import torch
    from torch.func import functional_call

    from coolchic_b200.enc.component.nlcoolchic import LatentFreeCoolChicEncoder
    from coolchic_b200.enc.training.loss import loss_function

    torch.manual_seed(3)
    lf = LatentFreeCoolChicEncoder(hop_param(Hh, Ww, gain=1)).to(dev)
    lf.train()
    leaves = {}
    for k, v in lf.named_parameters():
        if k.endswith(".bias") and "upsampling" in k:  # never used
            leaves[k] = v.detach().unsqueeze(0).expand(B, *v.shape)
        else:
            leaves[k] = (v.detach().unsqueeze(0) + 0.01 * torch.randn(B, *v.shape, device=dev)).requires_grad_(True)
    hw = [(-(-Hh // 2 ** i), -(-Ww // 2 ** i)) for i in range(7)]
    lats = [(3.0 * torch.randn(B, 1, h, w, device=dev)).requires_grad_(True) for h, w in hw]
    target = synthetic_image(12, Hh, Ww, B).to(dev)
    kw = dict(quantizer_noise_type="gaussian", quantizer_type="softround",
              soft_round_temperature=torch.tensor(0.3, device=dev), noise_parameter=torch.tensor(0.25, device=dev),
              AC_MAX_VAL=-1, flag_additional_outputs=False)

    def fwd(lat_list, params):
        return functional_call(lf, params, (lat_list,), kwargs=kw)

    def one():
        for t in lats + [v for v in leaves.values() if v.requires_grad]:
            t.grad = None
        raw, rate, _ = torch.vmap(fwd, randomness="different")([t.unsqueeze(1) for t in lats], leaves)
        loss_function(raw.squeeze(1), rate.squeeze(1), target, lmbda=LMBDA).loss.backward()

    for _ in range(3):
        one()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n_iter):
        one()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n_iter
    M = sum(h * w for h, w in hw)
    flops = B * algorithmic_flops(M, Hh * Ww, HOP["dim_arm"], HOP["nh"], 48)["total"]
    return {"fwd_bwd_per_s": 1e3 / ms, "ms_per_fwd_bwd": ms, "crops_per_s": B * 1e3 / ms, "B": B, "H": Hh, "W": Ww,
            "weights": "one set per crop (vmap over functional_call)", "algorithmic_gflop": flops / 1e9,
            "note": "module API: includes the Python / autograd / vmap dispatch of every call (no CUDA graph)"}


def time_arm_int(dev, do_cpu):
    """The fixed-point ARM over every latent position of one 768x512 image: mu,
    log-scale and the CABAC context indices in one launch, against the C restatement (one host core, sequential like
    the decoder's own arm_cpu.cpp)."""
    import numpy as np
    import torch

    from coolchic_b200 import ops

    spec = ops.ArchSpec(H, W, HOP["dim_arm"], HOP["nh"], HOP["syn"], 7, 8, 7, 16.0)
    rng = np.random.default_rng(0)
    shapes = [(-(-H // 2 ** i), -(-W // 2 ** i)) for i in range(7)]
    grids = [rng.integers(-12, 13, size=s).astype(np.int32) for s in shapes]
    C, nh = HOP["dim_arm"], HOP["nh"]
    wq = [rng.integers(-96, 97, size=(C, C)).astype(np.int32) for _ in range(nh)] + [rng.integers(-96, 97, size=(2, C)).astype(np.int32)]
    bq = [rng.integers(-4096, 4097, size=(C,)).astype(np.int32) for _ in range(nh)] + [rng.integers(-4096, 4097, size=(2,)).astype(np.int32)]
    y = torch.from_numpy(np.concatenate([g.reshape(-1) for g in grids]))[None].to(dev)
    wd, bd = [torch.from_numpy(w).to(dev) for w in wq], [torch.from_numpy(b).to(dev) for b in bq]
    for _ in range(3):
        outs = ops.arm_int(spec, y, wd, bd)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_rep = 30
    e0.record()
    for _ in range(n_rep):
        outs = ops.arm_int(spec, y, wd, bd)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n_rep
    M = int(y.numel())
    macs = M * (nh * C * C + 2 * C)
    out = {"positions": M, "ms_per_image": ms, "positions_per_s": M / (ms * 1e-3), "launches": 1,
           "int32_mac_per_s": macs / (ms * 1e-3), "algorithmic_bytes": M * 4 * 6,  # y in; mu, log-scale, 3 indices out
           "achieved_gbs": M * 4 * 6 / (ms * 1e-3) / 1e9}
    if do_cpu:
        from oracle import int_ref as I

        t0 = time.perf_counter()
        mu, ls = I.arm_int_pyramid(grids, wq, bq)
        out["cpu_c_port_ms"] = (time.perf_counter() - t0) * 1e3
        out["bit_exact_vs_port"] = bool(np.array_equal(outs[0][0].cpu().numpy(), mu) and
                                        np.array_equal(outs[1][0].cpu().numpy(), ls))
    return out


def time_dec_int(dev, do_cpu):
    """The fixed-point decoder's upsampling + synthesis on a shipped Kodak bitstream's integers:
    device time per image (CUDA events, back-to-back launches) against the numpy port on the host."""
    import numpy as np
    import torch

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from dec_common import load_dec
    from coolchic_b200 import ops

    name = "kodim01-lmbda-0001"
    z, grids, prm = load_dec(name)
    Hh, Ww = (int(v) for v in z["img_hw"])
    lat = torch.from_numpy(np.concatenate([g.reshape(-1) for g in grids]).astype(np.int32)).to(dev)
    layers = [dict(w=torch.from_numpy(l["w"].astype(np.int32)).to(dev), b=torch.from_numpy(l["b"].astype(np.int32)).to(dev),
                   ks=l["ks"], residual=l["residual"], relu=l["relu"]) for l in prm["syn"]]

    def call():
        return ops.decode_int(lat, (Hh, Ww), len(grids), prm["ups_w"], prm["pre_w"], layers, want_ups=False)

    for _ in range(3):
        r = call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_rep = 30
    e0.record()
    for _ in range(n_rep):
        r = call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n_rep
    n, npix, M = len(grids), Hh * Ww, int(lat.numel())
    # algorithmic bytes: latents read once; every level stack written once and read once; the synthesis reads the n
    # planes, each layer writes / the next reads 3 planes; pixels written
    stacks = sum((n - t) * g.size for t, g in enumerate(grids) if t < n - 1)
    alg_bytes = 4 * M + 2 * 4 * stacks + 4 * 3 * npix * (2 * len(layers) - 3) + 3 * npix
    out = {"image": name, "H": Hh, "W": Ww, "ms_per_image": ms, "launches": (n - 1) + len(layers) - 1,  # one per pyramid stage; fused 1x1 pair; pixels by the last 3x3 layer
           "algorithmic_bytes": alg_bytes, "achieved_gbs": alg_bytes / (ms * 1e-3) / 1e9}
    if do_cpu:
        from oracle import dec_int_ref as D

        t0 = time.perf_counter()
        ref = D.decode(grids, prm)
        out["cpu_port_ms"] = (time.perf_counter() - t0) * 1e3
        out["bit_exact_vs_port"] = bool(np.array_equal(r["out"].cpu().numpy(), ref["out"]))
    return out


def time_in_flight(dev, spec, n, iters):
    """Aggregate iterations/s of n independent 768x512 images trained at the same time on one GPU: one FusedState
    (arena, workspace, weight slot), one captured iteration graph and one stream per image; device timed."""
    import torch

    from coolchic_b200 import ops

    states, graphs, streams, targets = [], [], [], []
    for i in range(n):
        st = ops.FusedState(spec, 1, dev)
        st.load(initial_params(H, W, 1, seed=50 + i))
        st.lr.fill_(LR)
        st.T.fill_(T0)
        st.sigma.fill_(SIGMA0)
        tg = synthetic_image(60 + i, H, W).to(dev)
        s = ops.new_stream(dev)
        with torch.cuda.stream(s):
            for _ in range(3):
                st.train_step(tg, LMBDA, Q_TYPE, NOISE_TYPE, seed=70 + i)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            st.train_step(tg, LMBDA, Q_TYPE, NOISE_TYPE, seed=70 + i)
        states.append(st); graphs.append(g); streams.append(s); targets.append(tg)
    torch.cuda.synchronize(dev)
    cur = torch.cuda.current_stream(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for rep_ in range(2):  # first pass = warm-up
        e0.record(cur)
        for s in streams:
            s.wait_stream(cur)
        for _ in range(iters):
            for g, s in zip(graphs, streams):
                with torch.cuda.stream(s):
                    g.replay()
        for s in streams:
            cur.wait_stream(s)
        e1.record(cur)
        torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1)
    return {"images": n, "aggregate_it_per_s": n * iters / (ms * 1e-3), "per_image_it_per_s": iters / (ms * 1e-3),
            "slots": [st.slot.index for st in states]}


def run_ours(args, rank, world, local_rank):
    import ctypes as C

    import torch

    from coolchic_b200 import _native as N
    from coolchic_b200 import ops

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    gpu_index = int(os.environ.get("CUDA_VISIBLE_DEVICES", "0,1,2,3,4,5,6,7").split(",")[local_rank]) \
        if os.environ.get("CUDA_VISIBLE_DEVICES") else local_rank
    affinity = pin_to_gpu_numa(gpu_index)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)

    spec = ops.ArchSpec(H, W, HOP["dim_arm"], HOP["nh"], HOP["syn"], 7, 8, 7, 16.0)
    st = ops.FusedState(spec, 1, dev)
    st.load(initial_params(H, W, 1, seed=rank))  # the reference's initialisation (zero latents, ...)
    st.lr.fill_(LR)
    st.T.fill_(T0)
    st.sigma.fill_(SIGMA0)
    target_host = synthetic_image(rank, H, W).pin_memory()
    target = target_host.to(dev, non_blocking=True)
    M, Npix = st.plan.M, H * W

    def step():
        st.train_step(target, LMBDA, Q_TYPE, NOISE_TYPE, seed=1234 + rank)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    # ---- one iteration captured in a CUDA graph (the Python loop then costs one launch per iteration)
    graph = torch.cuda.CUDAGraph()
    cap_stream = torch.cuda.Stream(device=dev, priority=-1) if os.environ.get("CCB_BENCH_HIPRI") else None
    with torch.cuda.graph(graph, stream=cap_stream):
        step()
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    sampler = ClockSampler(gpu_index)
    sampler.start()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- timed region 1 (the reported value): L2 flushed before every iteration
    K = args.steps
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier()
    for a, b in ev:
        flush.zero_()
        a.record()
        graph.replay()
        b.record()
    barrier()
    ms_cold = sum(a.elapsed_time(b) for a, b in ev)
    # ---- timed region 2: back to back (state resident in L2, as in a real encode)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(K):
        graph.replay()
    e1.record()
    barrier()
    ms_warm = e0.elapsed_time(e1)
    clocks = sampler.stop()
    loss_now = st.out4.cpu().tolist()

    # ---- per-stage device time of the same iteration (CUDA events between stages, L2 flushed)
    stage = {k: 0.0 for k in N.STAGES}
    n_prof = 5
    for _ in range(n_prof):
        flush.zero_()
        torch.cuda.synchronize()
        for k, v in st.train_step_timed(target, LMBDA, Q_TYPE, NOISE_TYPE, seed=1234 + rank).items():
            stage[k] += v / n_prof

    # the dominant kernel once more in the form it takes when nothing has to co-reside with it (two CTAs per SM)
    os.environ["CCB_TIMED_ARM_ALONE"] = "1"
    arm_alone_ms = 0.0
    for _ in range(n_prof):
        flush.zero_()
        torch.cuda.synchronize()
        arm_alone_ms += st.train_step_timed(target, LMBDA, Q_TYPE, NOISE_TYPE, seed=1234 + rank)["arm_bwd"] / n_prof
    del os.environ["CCB_TIMED_ARM_ALONE"]

    # ---- e2e: the C-ABI call on HOST buffers (upload image + parameters, K iterations, download)
    n_par = st.plan.n_param_floats
    params_host = torch.empty(n_par, dtype=torch.float32).pin_memory()
    params_host.copy_(st.arena[0].cpu())
    out4_host = torch.zeros(4, dtype=torch.float32).pin_memory()
    q = ops.make_quant_cfg(spec, True, Q_TYPE, NOISE_TYPE, T0, SIGMA0, seed=99 + rank)

    def e2e_call(n_iters):
        N.check(N.lib().ccb_overfit_host(st.plan.handle, params_host.data_ptr(), target_host.data_ptr(), C.byref(q),
                                         LMBDA, LR, N.OPT_ALL, n_iters, out4_host.data_ptr()), "ccb_overfit_host")

    e2e_call(3)
    barrier()
    t0 = time.perf_counter()
    e2e_call(K)
    t_e2e = time.perf_counter() - t0
    barrier()

    # ---- several independent images in flight on this GPU (one stream, one weight slot, one graph each): the
    # iteration of ONE image uses well under half of the issue slots, so a dataset job keeps more than one going
    in_flight = {}
    for n_if in (2, 3, 4):
        in_flight[str(n_if)] = time_in_flight(dev, spec, n_if, min(K, 100))

    enc, rows, cfg3_secs = None, None, 0.0
    if not args.no_encode:
        # one image alone first: it pays the one-off costs of a process (allocator growth, lazy module loads) and
        # gives the latency of a single encode; then BASELINE.json configs[2] through enc/training/sharding.py
        barrier()
        run_c3x_encode(dev, 100 + rank, extras=False)
        barrier()
        enc = run_c3x_encode(dev, 200 + rank)
        barrier()
        rows, cfg3_secs = run_config3(dev, args.images, args.in_flight)
        barrier()

    times = torch.tensor([ms_cold, ms_warm, t_e2e * 1e3, enc["seconds"] if enc else 0.0, cfg3_secs], device=dev,
                         dtype=torch.float64)
    all_cfg3 = [torch.zeros_like(times[4:5]) for _ in range(world)] if dist is not None else None
    if dist is not None:
        dist.all_gather(all_cfg3, times[4:5].clone())
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_cold, ms_warm, ms_e2e, enc_secs, cfg3_wall = times.tolist()
    cfg3_per_rank = [float(t.item()) for t in all_cfg3] if all_cfg3 is not None else [cfg3_secs]
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel
    flops = algorithmic_flops(M, Npix, HOP["dim_arm"], HOP["nh"], 48)
    peak_fp32 = ops.fp32_peak_tflops(local_rank)
    # the dominant single kernel: k_arm_bwd (the "arm_bwd" stage of the timed, serial schedule is exactly that
    # one launch); measured live here with CUDA events, not taken from a profile
    dom = "arm_bwd"
    achieved = flops[dom] / (stage[dom] * 1e-3) / 1e12
    traffic = None
    try:  # dram__bytes_read.sum + dram__bytes_write.sum of that kernel from the committed ncu --set full capture
        traffic = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))["k_arm_bwd"]["dram_bytes"]
    except Exception:
        pass
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    alg_bytes = 28 * M + 12 * Npix
    L, n_tail = 7, len(HOP["syn"]) - 2
    launches = int(N.lib().ccb_plan_train_step_launches(st.plan.handle))  # kernels of one captured iteration
    line = {
        "metric": METRIC, "value": world * K / (ms_cold * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
        "warmup": args.warmup, "ms_per_step": ms_cold / K, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": CONFIG,
        "timing": "one CUDA-graph replay per iteration, CUDA events, L2 flushed before every timed iteration (256 MiB "
                  "memset, untimed), noise from the in-kernel Philox generator, max over ranks",
        "value_warm_l2": world * K / (ms_warm * 1e-3),
        "images_in_flight": in_flight,
        "cpu_affinity": affinity,
        "e2e": {"value": world * K / (ms_e2e * 1e-3), "unit": UNIT,
                "h2d_bytes_per_step": (4 * n_par + 12 * Npix) / K, "d2h_bytes_per_step": (4 * n_par + 16) / K,
                "call": "ccb_overfit_host(params_host, image_host, K iterations) on pinned host buffers: upload of the "
                        "image and the parameter vector, graph capture + instantiation, K iterations, download of "
                        "parameters and loss; wall clock around the call.  The device buffers of the plan were "
                        "allocated by an earlier (untimed) call"},
        "gpu_launches": launches * K,
        "clocks": clocks,
        "roofline": {"bound": "fp32_fma", "kernel": "k_" + dom, "achieved": achieved, "peak": peak_fp32,
                     "unit": "TFLOP/s", "frac": achieved / peak_fp32, "traffic": traffic,
                     "note": "k_arm_bwd is timed as the step launches it: ONE working CTA per SM (half the register "
                             "file) so that the main chain keeps co-residing. `alone` = the same launch with two CTAs "
                             "per SM (the module path's form), measured live as well",
                     "alone": {"launch_ms": arm_alone_ms, "achieved": flops[dom] / (arm_alone_ms * 1e-3) / 1e12,
                               "frac": flops[dom] / (arm_alone_ms * 1e-3) / 1e12 / peak_fp32, "unit": "TFLOP/s"},
                     "critical_path_stage": {"stage": "syn_bwd (k_syn_conv3x2_bwd + k_syn_head_bwd)",
                                             "achieved": flops["syn_bwd"] / (stage["syn_bwd"] * 1e-3) / 1e12,
                                             "frac": flops["syn_bwd"] / (stage["syn_bwd"] * 1e-3) / 1e12 / peak_fp32,
                                             "unit": "TFLOP/s"},
                     "peak_source": "ccb_fp32_peak FFMA micro-benchmark run in this process (MEASURED_PEAKS.json has "
                                    "no fp32 figure)",
                     "algorithmic_flops_per_launch": flops[dom], "launch_ms": stage[dom],
                     "step_hbm": {"achieved": alg_bytes / (ms_cold / K * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                  "frac": alg_bytes / (ms_cold / K * 1e-3) / 1e9 / hbm_peak,
                                  "algorithmic_bytes_per_step": alg_bytes},
                     "step_fp32": {"achieved": flops["total"] / (ms_cold / K * 1e-3) / 1e12, "peak": peak_fp32,
                                   "unit": "TFLOP/s", "frac": flops["total"] / (ms_cold / K * 1e-3) / 1e12 / peak_fp32}},
        "stage_ms": {k: round(v, 5) for k, v in stage.items()},
        "last_out4": loss_now,
    }
    if world == 1 and not args.no_encode:
        # the other BASELINE.json shapes (parity-tested against the reference in tests/test_gpu_fullsize.py)
        line["other_workloads"] = {
            "1080x1920 (configs[3])": time_other_workload(dev, 1080, 1920, 1, peak_fp32, flush),
            "8 x 256x256 crops, shared weights (configs[4], NOWholeNet shape)":
                time_other_workload(dev, 256, 256, 8, peak_fp32, flush),
            "8 x 256x256 crops, per-sample weights (configs[4], DeltaWholeNet shape)": time_per_sample_crops(dev)}
    if world == 1:
        # the fixed-point decoder's upsampling + synthesis, bit-exact, on a shipped bitstream's integers
        line["dec_int"] = time_dec_int(dev, not args.no_cpu)
        line["arm_int"] = time_arm_int(dev, not args.no_cpu)  # rows I1 / I2: the fixed-point ARM of the bitstream writer
    if enc:
        # the second half of BASELINE.json's metric: whole images through FrameEncoder + train() (max over ranks)
        psnr = [r["psnr_db"] for r in rows]
        line["encode"] = {
            "images_per_hour": args.images * 3600.0 / cfg3_wall,
            "images": args.images, "images_in_flight_per_gpu": args.in_flight, "wall_seconds": cfg3_wall,
            "wall_seconds_per_rank": [round(t, 3) for t in cfg3_per_rank],
            "single_image_seconds": enc_secs, "images_per_hour_one_at_a_time": world * 3600.0 / enc_secs,
            "schedule": "BASELINE.json configs[2]: seeded 768x512 synthetic images, image i on rank i mod N "
                        "(enc/training/sharding.py), complete preset_cfg/c3x.yaml each: warm-up (5 candidates x 400 "
                        "iterations trained side by side, best 2 x 400 more) + all_phases (10600 softround/gaussian + "
                        "1500 ste iterations, the NN weight quantisation search, up to 1000 ste latent-only "
                        "iterations), validations included; wall = max over ranks incl. the one all_gather of the "
                        "result rows; single_image_seconds = one image alone on the GPU (rank0 holds its details)",
            "psnr_db_min_mean_max": [min(psnr), sum(psnr) / len(psnr), max(psnr)],
            "rate_latent_bpp_mean": sum(r["rate_latent_bpp"] for r in rows) / len(rows),
            "iterations_total": int(sum(r["iterations"] for r in rows)),
            "rank0": enc}
    if world == 1 and not args.no_cpu:
        its, cores, secs = cpu_port_iters_per_s(n_timed=args.cpu_iters, n_warm=2)
        its1, _, secs1 = cpu_port_iters_per_s(n_timed=3, n_warm=1, threads=1)
        line["cpu_baseline"] = {"value": its, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"median of {args.cpu_iters} iterations of the same 768x512 workload on the "
                                          f"host ({secs:.0f} s of CPU work), oracle/float_ref.py",
                                "value_1_thread": its1, "sample_1_thread": f"median of 3 iterations ({secs1:.0f} s)",
                                "debug_preset": cpu_debug_preset_wall()}
        try:  # the same torch ops, eager, on this B200: how much of the ratio is "GPU vs CPU" and how much the kernels
            gits, _, gsecs = cpu_port_iters_per_s(n_timed=20, n_warm=5, device=str(dev))
            line["torch_gpu_baseline"] = {"value": gits, "unit": UNIT, "kind": "port, eager PyTorch on the same GPU, "
                                          "fp32 (TF32 off)", "sample": f"median of 20 iterations ({gsecs:.1f} s)",
                                          "ours_over_it": line["value"] / gits}
        except Exception as e:  # the port is written for the CPU: report instead of failing the bench
            line["torch_gpu_baseline"] = {"unavailable": f"{type(e).__name__}: {str(e)[:160]}"}
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-iters", type=int, default=12)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-encode", action="store_true", help="skip the whole-image c3x encodes (configs[2])")
    ap.add_argument("--images", type=int, default=24, help="images of configs[2] (sharded over the GPUs)")
    ap.add_argument("--in-flight", type=int, default=4, help="images encoded at the same time per GPU")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
