"""Shared helpers for the test-suite: golden loading, architectures, error metrics."""

import glob
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import float_ref as O  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

ARCHS = {
    "hop": dict(dim_arm=16, nh=2, syn=["48-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"]),
    "default": dict(dim_arm=24, nh=2, syn=["40-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"]),
    "mop": dict(dim_arm=16, nh=2, syn=["16-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"]),
    "lop": dict(dim_arm=8, nh=2, syn=["16-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"]),
    "vlop": dict(dim_arm=8, nh=1, syn=["8-1-linear-relu", "3-1-linear-none", "3-3-residual-none"]),
    "arm32": dict(dim_arm=32, nh=2, syn=["16-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"]),
}

# Float-path tolerance (BASELINE.json north_star: "within a stated relative tolerance
# (e.g. 1e-4 fp32)"): norm-wise,  max|a-b| <= RTOL * max|b| (+ATOL).
RTOL = 1e-4
ATOL = 1e-7


def make_arch(name: str, H: int, W: int, gain: float = 16.0) -> O.Arch:
    a = ARCHS[name]
    return O.Arch(img_size=(H, W), dim_arm=a["dim_arm"], n_hidden_layers_arm=a["nh"],
                  layers_synthesis=list(a["syn"]), encoder_gain=float(gain))


def golden_cases():
    """Float-path cases (the int_*.npz files belong to the fixed-point ARM tests)."""
    names = sorted(os.path.basename(f)[:-4] for f in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
    return [n for n in names if not n.startswith(("int_", "quant_", "cool_", "fs_", "ps_", "conv_"))]


def int_golden_cases(prefix: str):
    return sorted(os.path.basename(f)[:-4] for f in glob.glob(os.path.join(GOLDEN_DIR, prefix + "*.npz")))


class Golden:
    def __init__(self, name: str):
        self.name = name
        self.z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False)
        g = self.z
        self.H, self.W, self.B = int(g["H"]), int(g["W"]), int(g["B"])
        self.arch = make_arch(str(g["arch"]), self.H, self.W, float(g["gain"]))
        self.q_type, self.noise_type = str(g["q_type"]), str(g["noise_type"])
        self.T, self.sigma = torch.tensor(float(g["T"])), torch.tensor(float(g["sigma"]))
        self.lmbda, self.lr, self.n_steps = float(g["lmbda"]), float(g["lr"]), int(g["n_steps"])
        self.latent_free = "latent_free" in g.files
        self.target = torch.from_numpy(g["target"])

    def tensors(self, prefix: str):
        return {k[len(prefix):]: torch.from_numpy(self.z[k]).clone() for k in self.z.files if k.startswith(prefix)}

    def get(self, key: str):
        return torch.from_numpy(self.z[key]) if key in self.z.files else None


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    """max|a-b| / max(max|b|, tiny) on CPU float64."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.numel() == 0:
        return 0.0
    return float((a - b).abs().max() / max(float(b.abs().max()), 1e-30))


def assert_close(a, b, what="", rtol=RTOL, atol=ATOL):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (what, a.shape, b.shape)
    if a.numel() == 0:
        return
    err = float((a - b).abs().max())
    ref = float(b.abs().max())
    assert err <= rtol * ref + atol, f"{what}: max|a-b|={err:.3e} vs max|b|={ref:.3e} (rel {err / max(ref, 1e-30):.3e})"


# ---------------------------------------------------------------------------------------------
# How tight can fp32 parity be?  The reference's OWN fp32 results sit 1e-4 .. 3e-4 (norm-wise) away
# from the exact (float64) value on the ARM gradients and the rate (cancellation in
# F(y+.5) - F(y-.5), expm1 backward = result + 1: arm.py:355-370), so a second fp32 implementation
# cannot be asked to be closer to the reference than the reference is to the truth.  Tolerance per
# tensor = RTOL + 2 * err(reference fp32, float64 oracle), both measured on the same inputs.
# ---------------------------------------------------------------------------------------------
def fp64_truth(g: "Golden"):
    """Step-0 forward/backward of the oracle in float64 on the golden inputs.
    -> dict(rate=, decoded=, grads={name: tensor}, loss=)."""
    p = {k: v.double() for k, v in g.tensors("p0/").items()}
    leaves = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    lats = [leaves[f"latent_grids.{i}.data"] for i in range(7)]
    noise = g.get("s0/noise")
    noise = None if noise is None else noise.double()
    raw, rate, _ = O.coolchic_forward(leaves, g.arch, True, g.noise_type, g.q_type, g.T.double(), g.sigma.double(),
                                      noise, latents=lats)
    dec = raw if g.latent_free else O.frame_output(raw, True)
    loss, _, _ = O.loss_ref(dec, rate, g.target.double(), g.lmbda)
    loss.backward()
    grads = {k: v.grad for k, v in leaves.items() if v.grad is not None}
    return dict(rate=rate.detach(), decoded=dec.detach(), grads=grads, loss=float(loss.detach()))


def parity_tol(ref32: torch.Tensor, truth64: torch.Tensor, rtol: float = RTOL) -> float:
    # capped: a ReLU / clamp branch that flips between fp32 and fp64 is not a precision statement
    return rtol + 2.0 * min(rel_err(ref32, truth64), 2.5e-4)


# ---------------------------------------------------------------------------------------------
# Full-size / per-sample-weight goldens (tests/golden/make_fullsize_golden.py): the INPUTS are regenerated from
# seeds (CPU torch generators are deterministic), the golden file keeps the reference's outputs on a fixed sample.
# ---------------------------------------------------------------------------------------------
FULLSIZE_CASES = {
    # name: (arch, H, W, B, per_sample, q_type, noise_type, T, sigma, gain)
    "fs_hop_768x512": ("hop", 512, 768, 1, False, "softround", "kumaraswamy", 0.3, 2.0, 16),
    "fs_debug_768x512": ("debug", 512, 768, 1, False, "softround", "gaussian", 0.3, 0.25, 16),
    "fs_hop_1080p": ("hop", 1080, 1920, 1, False, "softround", "kumaraswamy", 0.3, 2.0, 16),
    "fs_hop_crops_b8": ("hop", 256, 256, 8, False, "softround", "gaussian", 0.3, 0.25, 1),
    "ps_hop_crops_b8": ("hop", 256, 256, 8, True, "softround", "gaussian", 0.3, 0.25, 1),
    "ps_hop_b3_40x56": ("hop", 40, 56, 3, True, "softround", "gaussian", 0.3, 0.25, 1),
    "ps_lop_b2_33x47": ("lop", 33, 47, 2, True, "softround", "kumaraswamy", 0.3, 2.0, 16),
}
ARCHS["debug"] = dict(dim_arm=8, nh=2, syn=["16-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"])


def fullsize_inputs(name: str):
    """-> dict(arch, p (one weight set + latents [B,1,h,w]), target, noise [B,M], per-sample weights or None)."""
    import zlib

    arch_name, H, W, B, per_sample, q_type, noise_type, T, sigma, gain = FULLSIZE_CASES[name]
    arch = make_arch(arch_name, H, W, gain)
    seed = zlib.crc32(name.encode()) % 1000
    p = O.init_params(arch, seed=seed, B=B, random=True)
    if gain == 1:
        for i in range(7):
            p[f"latent_grids.{i}.data"] = p[f"latent_grids.{i}.data"] * 12.0
    # a well-conditioned rate: with init_params' log-scale bias (3.5 -> scale 0.6) the random latents sit ~11 scales from
    # the random mu, i.e. p ~ 2^-16, right ON the clamp_min threshold of coolchic.py:419-423 for a large share of the
    # positions -- there two fp32 evaluations take different branches and whole weight gradients move by 1e-2.  A trained
    # encoder spends 0..6 bits per latent; bias 6.5 (scale ~ 12) puts the test there.
    nh = arch.n_hidden_layers_arm
    p[f"arm.mlp.{2 * nh}.bias"] = p[f"arm.mlp.{2 * nh}.bias"] + torch.tensor([0.0, 3.0])
    p[f"arm.mlp.{2 * nh}.weight"] = p[f"arm.mlp.{2 * nh}.weight"] * 0.2  # mu / log-scale spread ~0.7, not ~3.6
    target = O.synthetic_image(11, H, W, B)
    g = torch.Generator().manual_seed(seed + 77)
    M = arch.n_latent_values
    noise = torch.randn(B, M, generator=g) if noise_type == "gaussian" else torch.rand(B, M, generator=g)
    ps = None
    if per_sample:
        ps = {}
        for k, v in p.items():
            if k.startswith("latent_grids"):
                continue
            scale = 0.1 * float(v.abs().mean()) + 1e-3
            ps[k] = v.unsqueeze(0) + scale * torch.randn(B, *v.shape, generator=g)
    return dict(arch=arch, p=p, target=target, noise=noise, per_sample=ps, B=B, H=H, W=W, q_type=q_type,
                noise_type=noise_type, T=T, sigma=sigma, lmbda=1e-3, gain=gain)


def sample_indices(shape, seed: int, n_random: int = 4096):
    """Flat indices of a fixed sample of a tensor: n_random random positions plus, for [..., h, w] tensors with
    h, w >= 4, the two outermost rows / columns of every plane (where a mis-indexed border would show)."""
    n = int(np.prod(shape))
    if n <= 2 * n_random:
        return np.arange(n, dtype=np.int64)
    rng = np.random.default_rng(seed)
    idx = [rng.integers(0, n, size=n_random, dtype=np.int64)]
    if len(shape) >= 2 and shape[-1] >= 4 and shape[-2] >= 4:
        h, w = shape[-2], shape[-1]
        planes = n // (h * w)
        rows = np.array([0, 1, h - 2, h - 1])
        cols = np.array([0, 1, w - 2, w - 1])
        border = np.concatenate([(rows[:, None] * w + np.arange(w)[None, :]).reshape(-1),
                                 (np.arange(h)[:, None] * w + cols[None, :]).reshape(-1)])
        idx.append((np.arange(planes)[:, None] * (h * w) + border[None, :]).reshape(-1))
    return np.unique(np.concatenate(idx))


# ---------------------------------------------------------------------------------------------
# Convergence goldens (tests/golden/make_convergence_golden.py, tests/test_gpu_convergence.py): the same schedule dicts
# feed the reference's PresetConfig and this repo's warmup() / train().
# ---------------------------------------------------------------------------------------------
CONVERGENCE_HW = (64, 96)
CONVERGENCE_IMAGE = 21
CONVERGENCE_LMBDA = 1e-3


def _tp(**kw):
    d = dict(lr=1e-2, max_itr=10, freq_valid=100, patience=100000, quantize_model=False, schedule_lr=False,
             softround_temperature=[0.3, 0.3], noise_parameter=[2.0, 2.0], quantizer_noise_type="kumaraswamy",
             quantizer_type="softround", optimized_module=["all"])
    d.update(kw)
    return d


CONVERGENCE_PRESETS = {
    # preset_cfg/debug.yaml, verbatim
    "debug": dict(
        preset_name="debug",
        warmup=dict(phases=[dict(candidates=3, training_phase=_tp(max_itr=10, freq_valid=400)),
                            dict(candidates=2, training_phase=_tp(max_itr=10, freq_valid=400))]),
        all_phases=[
            _tp(max_itr=50, schedule_lr=True, quantizer_noise_type="gaussian", softround_temperature=[0.3, 0.1],
                noise_parameter=[0.25, 0.1]),
            _tp(lr=1e-4, max_itr=10, patience=10, quantizer_type="ste", quantizer_noise_type="none", quantize_model=True,
                softround_temperature=[1e-4, 1e-4], noise_parameter=[1.0, 1.0]),
            _tp(lr=1e-4, max_itr=10, patience=50, optimized_module=["latent"], freq_valid=5, quantizer_type="ste",
                quantizer_noise_type="none", softround_temperature=[1e-4, 1e-4], noise_parameter=[1.0, 1.0]),
        ]),
    # the shape of preset_cfg/c3x.yaml, shortened to what a 64x96 image needs to converge
    "mini": dict(
        preset_name="mini",
        warmup=dict(phases=[dict(candidates=3, training_phase=_tp(max_itr=100, freq_valid=100)),
                            dict(candidates=2, training_phase=_tp(max_itr=100, freq_valid=100))]),
        all_phases=[
            _tp(max_itr=1500, patience=500, schedule_lr=True, quantizer_noise_type="gaussian",
                softround_temperature=[0.3, 0.1], noise_parameter=[0.25, 0.1]),
            _tp(lr=1e-4, max_itr=200, patience=200, schedule_lr=True, quantizer_type="ste", quantizer_noise_type="none",
                quantize_model=True, softround_temperature=[1e-4, 1e-4], noise_parameter=[1.0, 1.0]),
            _tp(lr=1e-4, max_itr=100, patience=50, optimized_module=["latent"], freq_valid=10, quantizer_type="ste",
                quantizer_noise_type="none", softround_temperature=[1e-4, 1e-4], noise_parameter=[1.0, 1.0]),
        ]),
}
