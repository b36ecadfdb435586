"""ORACLE (test infrastructure, not product code) -- float path.

CPU fp32 restatement, in plain functional PyTorch, of the reference's per-image
overfitting step (pepborrell/cool-chic).  Only `tests/`, `__graft_entry__.smoke()`
and `bench.py`'s cpu_baseline / `--impl reference` legs may import this module; the
product (`coolchic_b200`) never does.

Third-party arithmetic: the reference's float path *is* PyTorch library calls
(`pyproject.toml:9-10` pins torch 2.5.1), so the restatement uses the
same primitives (F.pad / unfold / index_select / F.linear / F.conv2d /
F.conv_transpose2d) in the same places.  That keeps (a) the rounding behaviour and
(b) the CPU cost representative of the reference when this port is timed as the
CPU baseline.  Gradients come from autograd, exactly as `loss.backward()` does in
the reference (enc/training/train.py:264).

Parity pin: `tests/golden/make_golden.py` runs the *real* reference modules
(imported from /root/reference in the build container) and commits their
inputs/outputs under tests/golden/*.npz; `tests/test_oracle_golden.py` checks this
file against them.  The float path has no reference-owned known-answer tests
(SURVEY.md §4), so those differential fixtures are the pin.

Each function cites the reference lines it restates (paths relative to
/root/reference/coolchic).
"""

from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor

# --------------------------------------------------------------------------- #
# Architecture description
# --------------------------------------------------------------------------- #


@dataclass
class Arch:
    """What `CoolChicEncoderParameter` carries (enc/component/coolchic.py:55-124)."""

    img_size: Tuple[int, int]
    dim_arm: int = 16
    n_hidden_layers_arm: int = 2
    layers_synthesis: Sequence[str] = (
        "48-1-linear-relu",
        "3-1-linear-none",
        "3-3-residual-relu",
        "3-3-residual-none",
    )
    n_latents: int = 7
    encoder_gain: float = 16.0
    ups_k_size: int = 8
    ups_preconcat_k_size: int = 7
    latent_sizes: List[Tuple[int, int]] = field(init=False)

    def __post_init__(self):
        H, W = self.img_size
        # ceil pyramid: enc/component/coolchic.py:184-189
        self.latent_sizes = [
            (int(math.ceil(H / 2**i)), int(math.ceil(W / 2**i)))
            for i in range(self.n_latents)
        ]

    @property
    def n_latent_values(self) -> int:
        return sum(h * w for h, w in self.latent_sizes)


# 9x9 causal window, flat indices of the selected taps
# (enc/component/core/arm.py:373-506; same sets as cpp/cc-frame-decoder.cpp:111-154)
_CTX_ROWS = {
    8: {1: [4], 2: [4], 3: [3, 4, 5], 4: [1, 2, 3]},
    16: {1: [4, 5], 2: [2, 3, 4, 5, 6], 3: [1, 2, 3, 4, 5, 6], 4: [1, 2, 3]},
    24: {
        0: [4],
        1: [2, 3, 4, 5, 6],
        2: [1, 2, 3, 4, 5, 6, 7],
        3: [1, 2, 3, 4, 5, 6, 7],
        4: [0, 1, 2, 3],
    },
    32: {
        0: [2, 3, 4, 5],
        1: [1, 2, 3, 4, 5, 6, 7],
        2: [1, 2, 3, 4, 5, 6, 7, 8],
        3: [0, 1, 2, 3, 4, 5, 6, 7, 8],
        4: [0, 1, 2, 3],
    },
}


def ctx_flat_indices(dim_arm: int) -> List[int]:
    """Flat indices into the 9x9 window, row-major, ascending (arm.py:398-506)."""
    if dim_arm not in _CTX_ROWS:
        raise ValueError(f"ARM context size must be 8, 16, 24 or 32. Found {dim_arm}.")
    out = []
    for r in sorted(_CTX_ROWS[dim_arm]):
        out += [9 * r + c for c in _CTX_ROWS[dim_arm][r]]
    assert len(out) == dim_arm
    return out


def ctx_offsets(dim_arm: int) -> List[Tuple[int, int]]:
    """(dy, dx) of every context tap relative to the coded latent (dy<=0)."""
    return [(i // 9 - 4, i % 9 - 4) for i in ctx_flat_indices(dim_arm)]


# --------------------------------------------------------------------------- #
# Parameter container helpers (names = the reference's state_dict keys, §8b)
# --------------------------------------------------------------------------- #


def parse_synthesis(layers: Sequence[str], in_ft: int):
    """-> list of (cin, cout, k, residual, relu)  (core/synthesis.py:238-260)."""
    out = []
    for spec in layers:
        o, k, mode, nl = spec.split("-")
        o = 3 if o == "X" else int(o)
        assert mode in ("linear", "residual") and nl in ("none", "relu")
        out.append((in_ft, o, int(k), mode == "residual", nl == "relu"))
        in_ft = o
    return out


def init_params(arch: Arch, seed: int = 0, B: int = 1, random: bool = True) -> Dict[str, Tensor]:
    """Random (non-degenerate) parameter set with the reference's names and shapes.

    `random=False` gives the reference's own initialisation values for the
    deterministic parts (zeros latents, bicubic / Dirac filters, zero residual
    layers: arm.py:66-84, upsampling.py:135-166,255-295, synthesis.py:86-116).
    """
    g = torch.Generator().manual_seed(seed)
    p: Dict[str, Tensor] = {}
    C, nh = arch.dim_arm, arch.n_hidden_layers_arm
    for i, (h, w) in enumerate(arch.latent_sizes):
        p[f"latent_grids.{i}.data"] = (
            0.3 * torch.randn(B, 1, h, w, generator=g) if random else torch.zeros(B, 1, h, w)
        )
    for k in range(nh):
        p[f"arm.mlp.{2 * k}.weight"] = (
            0.15 * torch.randn(C, C, generator=g) if random else torch.zeros(C, C)
        )
        p[f"arm.mlp.{2 * k}.bias"] = 0.1 * torch.randn(C, generator=g) if random else torch.zeros(C)
    p[f"arm.mlp.{2 * nh}.weight"] = torch.randn(2, C, generator=g) / 4.0
    p[f"arm.mlp.{2 * nh}.bias"] = (
        torch.tensor([0.05, 3.5]) + 0.1 * torch.randn(2, generator=g) if random else torch.zeros(2)
    )
    n_ups = arch.n_latents - 1
    k_up, k_pre = arch.ups_k_size, arch.ups_preconcat_k_size
    if k_up < 8:
        core = torch.tensor([0.25, 0.75])
    else:
        core = torch.tensor([0.0351562, 0.1054687, -0.2617187, -0.8789063])
    for i in range(n_ups):
        w_up = torch.zeros(k_up // 2)
        w_up[k_up // 2 - core.numel():] = core
        w_pre = torch.zeros((k_pre + 1) // 2)
        w_pre[-1] = 1.0
        if random:
            w_up = w_up + 0.05 * torch.randn(w_up.shape, generator=g)
            w_pre = w_pre + 0.05 * torch.randn(w_pre.shape, generator=g)
        p[f"upsampling.conv_transpose2ds.{i}.parametrizations.weight.original"] = w_up
        p[f"upsampling.conv_transpose2ds.{i}.bias"] = torch.zeros(1)
        p[f"upsampling.conv2ds.{i}.parametrizations.weight.original"] = w_pre
        p[f"upsampling.conv2ds.{i}.bias"] = torch.zeros(1)
    for li, (cin, cout, k, res, relu) in enumerate(parse_synthesis(arch.layers_synthesis, arch.n_latents)):
        a = math.sqrt(1.0 / (cin * k * k))
        if res and not random:
            wgt = torch.zeros(cout, cin, k, k)
        else:
            scale = (0.3 if res else 1.0) * a
            wgt = (torch.rand(cout, cin, k, k, generator=g) - 0.5) * 2 * scale
        p[f"synthesis.layers.{2 * li}.weight"] = wgt
        p[f"synthesis.layers.{2 * li}.bias"] = (
            0.05 * torch.randn(cout, generator=g) if random else torch.zeros(cout)
        )
    return p


# --------------------------------------------------------------------------- #
# Q: quantiser  (core/quantizer.py:16-41, 62-107, 116-232)
# --------------------------------------------------------------------------- #


def softround(x: Tensor, t: Tensor) -> Tensor:
    fl = torch.floor(x)
    delta = x - fl - 0.5
    return fl + 0.5 * torch.tanh(delta / t) / torch.tanh(1 / (2 * t)) + 0.5


def kumaraswamy_noise(u: Tensor, a: Tensor) -> Tensor:
    b = (2**a * (a - 1) + 1) / a
    return (1 - (1 - u) ** (1 / b)) ** (1 / a) - 0.5


def quantize_ref(
    x: Tensor,
    noise_type: str,
    q_type: str,
    T: Optional[Tensor],
    sigma: Optional[Tensor],
    raw_noise: Optional[Tensor] = None,
) -> Tensor:
    """`raw_noise` is what the reference would have drawn: N(0,1) samples for
    "gaussian" (quantizer.py:192), U(0,1) samples for "kumaraswamy" (:195-197).
    If None it is drawn here with the same torch call."""
    noise = None
    if noise_type == "gaussian":
        n0 = torch.randn_like(x) if raw_noise is None else raw_noise
        noise = n0 * sigma
    elif noise_type == "kumaraswamy":
        u = torch.rand_like(x) if raw_noise is None else raw_noise
        noise = kumaraswamy_noise(u, sigma)
    elif noise_type != "none":
        raise ValueError(noise_type)

    if q_type == "none":
        return x + noise
    if q_type == "softround_alone":
        return softround(x, T)
    if q_type == "softround":
        return softround(softround(x, T) + noise, T)
    if q_type == "ste":
        # quantizer.py:222-225 rebinds y INSIDE torch.no_grad(), so the returned tensor
        # carries no autograd history: value (s - s) + round(x) == round(x), and the
        # latents receive grad=None in this mode (verified against the live reference;
        # ARM / upsampling / synthesis weights still get gradients).
        y = softround(x, T)
        with torch.no_grad():
            y = y - softround(x, T) + torch.round(x)
        return y
    if q_type == "true_ste":
        return x + (torch.round(x) - x).detach()
    if q_type == "hardround":
        return torch.round(x)
    raise ValueError(q_type)


# --------------------------------------------------------------------------- #
# G + A + R: context gather, ARM MLP, Laplace rate
# --------------------------------------------------------------------------- #


def gather_contexts(yhat: Tensor, ctx_idx: Tensor, mask_size: int = 9) -> Tensor:
    """[B,1,h,w] -> [B,h*w,C]   (core/arm.py:308-352)."""
    B, C1, h, w = yhat.shape
    assert C1 == 1
    pad = (mask_size - 1) // 2
    xp = F.pad(yhat, (pad, pad, pad, pad), mode="constant", value=0.0)
    win = xp.unfold(2, mask_size, 1).unfold(3, mask_size, 1).reshape(B, h * w, mask_size * mask_size)
    return torch.index_select(win, 2, ctx_idx.to(win.device))


def arm_mlp(ctx: Tensor, p: Dict[str, Tensor], nh: int) -> Tuple[Tensor, Tensor, Tensor]:
    """(core/arm.py:86-101, 227-268): residual+ReLU hidden layers, linear head."""
    h = ctx
    for k in range(nh):
        h = torch.relu(F.linear(h, p[f"arm.mlp.{2 * k}.weight"], p[f"arm.mlp.{2 * k}.bias"]) + h)
    out = F.linear(h, p[f"arm.mlp.{2 * nh}.weight"], p[f"arm.mlp.{2 * nh}.bias"])
    mu, log_scale = out[..., 0], out[..., 1]
    scale = torch.exp(torch.clamp(log_scale - 4, min=-4.6, max=5.0))
    return mu, scale, log_scale


def laplace_cdf(x: Tensor, mu: Tensor, scale: Tensor) -> Tensor:
    """(core/arm.py:355-370)"""
    d = x - mu
    return 0.5 - 0.5 * d.sign() * torch.expm1(-d.abs() / scale)


def rate_bits(y: Tensor, mu: Tensor, scale: Tensor) -> Tensor:
    """(enc/component/coolchic.py:419-424)"""
    proba = torch.clamp_min(laplace_cdf(y + 0.5, mu, scale) - laplace_cdf(y - 0.5, mu, scale), 2**-16)
    return -torch.log2(proba)


# --------------------------------------------------------------------------- #
# U: upsampling pyramid  (core/upsampling.py)
# --------------------------------------------------------------------------- #


def symmetric_kernel(orig: Tensor, k: int) -> Tensor:
    """a b c d -> a b c d d c b a (even k) / a b c d c b a (odd k)   (upsampling.py:46-68)."""
    return torch.cat([orig, torch.flip(orig, dims=[0])[k % 2:]])


def refine_conv(x: Tensor, orig: Tensor, k: int, train: bool) -> Tensor:
    """Separable symmetric odd conv, zero padded, + residual (upsampling.py:168-209)."""
    w = symmetric_kernel(orig, k).view(1, -1)
    pad = k // 2
    if train:
        k2 = torch.kron(w, w.T).view(1, 1, k, k)
        return F.conv2d(x, k2, None, 1, pad) + x
    yw = F.conv2d(x, w.view(1, 1, 1, k), padding=(0, pad))
    return F.conv2d(yw, w.view(1, 1, k, 1), padding=(pad, 0)) + x


def upsample2x(x: Tensor, orig: Tensor, k: int, train: bool) -> Tensor:
    """Separable symmetric even transposed conv, stride 2, replicate padded, cropped
    to exactly twice the size (upsampling.py:297-355)."""
    w = symmetric_kernel(orig, k).view(1, -1)
    p0 = k // 2
    crop = 2 * p0 - 1 + k // 2
    if train:
        k2 = torch.kron(w, w.T).view(1, 1, k, k)
        xp = F.pad(x, (p0, p0, p0, p0), mode="replicate")
        yc = F.conv_transpose2d(xp, k2, stride=2)
        Hc, Wc = yc.shape[-2:]
        return yc[:, :, crop:Hc - crop, crop:Wc - crop]
    xp = F.pad(x, (p0, p0, 0, 0), mode="replicate")
    yc = F.conv_transpose2d(xp, w.view(1, 1, 1, k), stride=(1, 2))
    y = yc[..., crop:yc.shape[-1] - crop]
    xp = F.pad(y, (0, 0, p0, p0), mode="replicate")
    yc = F.conv_transpose2d(xp, w.view(1, 1, k, 1), stride=(2, 1))
    return yc[:, :, crop:yc.shape[-2] - crop, :]


def upsample_pyramid(lat: List[Tensor], p: Dict[str, Tensor], arch: Arch, train: bool) -> Tensor:
    """(upsampling.py:476-506) -> [B, L, H, W], channel 0 = refined full-res latent."""
    rev = list(reversed(lat))
    up = rev[0]
    n = arch.n_latents - 1
    for idx, target in enumerate(rev[1:]):
        B, c, h, w = up.shape
        x = upsample2x(
            up.reshape(B * c, 1, h, w),
            p[f"upsampling.conv_transpose2ds.{idx % n}.parametrizations.weight.original"],
            arch.ups_k_size,
            train,
        )
        x = x.reshape(B, c, x.shape[-2], x.shape[-1])[:, :, : target.shape[-2], : target.shape[-1]]
        hi = refine_conv(
            target,
            p[f"upsampling.conv2ds.{idx % n}.parametrizations.weight.original"],
            arch.ups_preconcat_k_size,
            train,
        )
        up = torch.cat((hi, x), dim=1)
    return up


# --------------------------------------------------------------------------- #
# S: synthesis  (core/synthesis.py:69-84, 264-277)
# --------------------------------------------------------------------------- #


def synthesis(z: Tensor, p: Dict[str, Tensor], arch: Arch) -> Tensor:
    x = z
    for li, (cin, cout, k, res, relu) in enumerate(parse_synthesis(arch.layers_synthesis, arch.n_latents)):
        pad = (k - 1) // 2
        xp = F.pad(x, (pad, pad, pad, pad), mode="replicate")
        y = F.conv2d(xp, p[f"synthesis.layers.{2 * li}.weight"], p[f"synthesis.layers.{2 * li}.bias"])
        if res:
            y = y + x
        x = torch.relu(y) if relu else y
    return x


# --------------------------------------------------------------------------- #
# F1 / FE / LO: forward, frame clamp, loss
# --------------------------------------------------------------------------- #


def coolchic_forward(
    p: Dict[str, Tensor],
    arch: Arch,
    training: bool,
    noise_type: str = "kumaraswamy",
    q_type: str = "softround",
    T: Optional[Tensor] = None,
    sigma: Optional[Tensor] = None,
    raw_noise: Optional[Tensor] = None,
    ac_max_val: int = -1,
    latents: Optional[List[Tensor]] = None,
    detailed: bool = False,
):
    """(enc/component/coolchic.py:291-479; nlcoolchic.py:151-336 when `latents` given)

    Returns raw_out [B,3,H,W], rate [B,M] and, if detailed, a dict with the flat
    mu / scale / log_scale / quantised latent tensors.
    """
    lat = latents if latents is not None else [p[f"latent_grids.{i}.data"] for i in range(arch.n_latents)]
    B = lat[0].shape[0]
    flat = torch.cat([l.reshape(B, -1) for l in lat], dim=1)
    yhat_flat = quantize_ref(
        flat * arch.encoder_gain,
        noise_type if training else "none",
        q_type if training else "hardround",
        T,
        sigma,
        raw_noise,
    )
    if ac_max_val != -1:
        yhat_flat = torch.clamp(yhat_flat, -ac_max_val, ac_max_val + 1)
    yhat, cnt = [], 0
    for l in lat:
        n = l[0].numel()
        yhat.append(yhat_flat[:, cnt:cnt + n].reshape(l.shape))
        cnt += n
    ctx_idx = torch.tensor(ctx_flat_indices(arch.dim_arm))
    ctx = torch.cat([gather_contexts(y, ctx_idx) for y in yhat], dim=1)
    mu, scale, log_scale = arm_mlp(ctx, p, arch.n_hidden_layers_arm)
    rate = rate_bits(yhat_flat, mu, scale)
    raw = synthesis(upsample_pyramid(yhat, p, arch, training), p, arch)
    extra = {}
    if detailed:
        extra = {"mu": mu, "scale": scale, "log_scale": log_scale, "yhat": yhat_flat}
    return raw, rate, extra


def frame_output(raw: Tensor, training: bool, bitdepth: int = 8) -> Tensor:
    """(enc/component/frame.py:172-183), rgb / yuv444."""
    x = raw
    if not training:
        m = 2**bitdepth - 1
        x = torch.round(x * m) / m
    return torch.clamp(x, 0.0, 1.0)


def loss_ref(decoded: Tensor, rate: Tensor, target: Tensor, lmbda: float, rate_mlp_bit: float = 0.0):
    """(enc/training/loss.py:135-150) -> loss, mse, rate_bpp"""
    mse = ((decoded - target) ** 2).mean()
    n_pix = decoded.shape[-2] * decoded.shape[-1]
    bpp = (rate.sum(dim=1).mean(dim=0) + rate_mlp_bit) / n_pix
    return mse + lmbda * bpp, mse, bpp


# --------------------------------------------------------------------------- #
# OPT: clip_grad_norm_ + Adam  (enc/training/train.py:184, 244-266)
# --------------------------------------------------------------------------- #


def clip_and_adam_ref(
    params: Dict[str, Tensor],
    grads: Dict[str, Optional[Tensor]],
    state: Dict[str, Dict[str, Tensor]],
    lr: float,
    optimized: Optional[Sequence[str]] = None,
    max_norm: float = 1e-1,
    betas=(0.9, 0.999),
    eps: float = 1e-8,
) -> float:
    """In-place restatement of `torch.nn.utils.clip_grad_norm_(all, 0.1, 2.0)` followed
    by `torch.optim.Adam.step()` with PyTorch defaults (no amsgrad / weight decay).
    The norm spans every parameter that has a gradient (train.py:220,265), the update
    only `optimized` (train.py:165-184).  Returns the total norm (python float)."""
    with torch.no_grad():
        gs = [g for g in grads.values() if g is not None]
        total = torch.linalg.vector_norm(torch.stack([torch.linalg.vector_norm(g, 2.0) for g in gs]), 2.0)
        coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
        for g in gs:
            g.mul_(coef)
        for name, prm in params.items():
            g = grads.get(name)
            if g is None or (optimized is not None and name not in optimized):
                continue
            st = state.setdefault(
                name, {"step": torch.zeros(()), "exp_avg": torch.zeros_like(prm), "exp_avg_sq": torch.zeros_like(prm)}
            )
            st["step"] += 1
            t = float(st["step"])
            st["exp_avg"].lerp_(g, 1 - betas[0])
            st["exp_avg_sq"].mul_(betas[1]).addcmul_(g, g, value=1 - betas[1])
            bc1, bc2 = 1 - betas[0] ** t, 1 - betas[1] ** t
            denom = (st["exp_avg_sq"].sqrt() / math.sqrt(bc2)).add_(eps)
            prm.addcdiv_(st["exp_avg"], denom, value=-(lr / bc1))
    return float(total)


def train_step_ref(
    p: Dict[str, Tensor],
    arch: Arch,
    target: Tensor,
    state: Dict[str, Dict[str, Tensor]],
    lmbda: float,
    lr: float,
    noise_type: str,
    q_type: str,
    T: Tensor,
    sigma: Tensor,
    raw_noise: Optional[Tensor] = None,
    optimized: Optional[Sequence[str]] = None,
):
    """One iteration of enc/training/train.py:244-268 on leaf tensors `p` (updated in place).
    Returns (loss, mse, bpp, grads-before-clipping, total_norm)."""
    leaves = {k: v.detach().requires_grad_(True) for k, v in p.items()}
    raw, rate, _ = coolchic_forward(leaves, arch, True, noise_type, q_type, T, sigma, raw_noise)
    loss, mse, bpp = loss_ref(frame_output(raw, True), rate, target, lmbda)
    loss.backward()
    grads = {k: (v.grad.clone() if v.grad is not None else None) for k, v in leaves.items()}
    clipped = {k: (g.clone() if g is not None else None) for k, g in grads.items()}
    norm = clip_and_adam_ref(p, clipped, state, lr, optimized)
    return float(loss.detach()), float(mse.detach()), float(bpp.detach()), grads, norm


# --------------------------------------------------------------------------- #
# Synthetic inputs
# --------------------------------------------------------------------------- #


def synthetic_image(idx: int, H: int, W: int, B: int = 1) -> Tensor:
    """Seeded smooth-plus-texture test image, [B,3,H,W] in [0,1], 8-bit levels."""
    g = torch.Generator().manual_seed(1000 + idx)
    v, u = torch.meshgrid(torch.linspace(0, 1, H), torch.linspace(0, 1, W), indexing="ij")
    out = torch.empty(B, 3, H, W)
    for b in range(B):
        for c in range(3):
            x = torch.full((H, W), 0.5)
            for k in range(1, 7):
                a = torch.rand((), generator=g) / k
                fx, fy = 8 * torch.rand((), generator=g), 8 * torch.rand((), generator=g)
                ph = 2 * math.pi * torch.rand((), generator=g)
                x = x + 0.25 * a * torch.sin(2 * math.pi * (fx * u + fy * v) + ph)
            n = 0.03 * torch.randn(1, 1, H, W, generator=g)
            n = F.avg_pool2d(F.pad(n, (2, 2, 2, 2), mode="replicate"), 5, stride=1)
            out[b, c] = x + n[0, 0]
    return torch.round(out.clamp(0, 1) * 255) / 255
