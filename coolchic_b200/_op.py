"""The CoolChicEncoder forward / backward as `torch.library` custom operators over the C ABI.

Why operators and not a `torch.autograd.Function`: the reference wraps its frame encoder in
`torch.compile(frame_encoder, dynamic=False, mode="reduce-overhead", fullgraph=True)` for every non-debug preset
(enc/component/video.py:271-279, encode_simpler.py:244-252) and runs the latent-free decoder under
`torch.vmap(functional_call(...), randomness="different")` with per-sample weights (hypernet/hypernet.py:1266-1288).
An operator registered with a fake (shape) implementation, an autograd formula and a vmap rule is opaque to
Dynamo (no graph break), differentiable, and batched by ONE launch of the per-sample-weight kernels
(`ccb_tensors.weight_batch_stride`, include/ccb.h).

  coolchic_b200::forward   (latents[L], weights[n], noise?, T?, sigma?, scalars...) -> 8 tensors
  coolchic_b200::backward  (d_raw, d_rate?, ws, noise?, latents[L], weights[n], T?, sigma?, scalars...) -> L + n tensors

No tensor memory is owned by the library: the workspace is a returned uint8 tensor that autograd saves for the
backward.  CUDA only; there is no CPU implementation (the operators raise on non-CUDA tensors).
"""

from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch
from torch import Tensor

from . import _native as N

_LAYER_SEP = ","


def _require_cuda(t: Tensor, what: str) -> None:
    if not t.is_cuda:
        raise N.CcbError(
            f"{what} is on {t.device}: coolchic_b200 runs on CUDA (sm_100a) only and has no CPU fallback"
        )


def _ptr(t: Optional[Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _latent_hw(H: int, W: int, n_latents: int) -> List[Tuple[int, int]]:
    """ceil(H / 2^i) x ceil(W / 2^i) (coolchic.py:184-189)."""
    return [((H + (1 << i) - 1) >> i, (W + (1 << i) - 1) >> i) for i in range(n_latents)]


def _weight_shapes(dim_arm: int, nh: int, layers: Sequence[str], n_latents: int, ups_k: int, ups_pre_k: int):
    """Shapes of the weight tensors in ArchSpec.weight_names() order (one weight set)."""
    out = []
    for k in range(nh + 1):
        o = dim_arm if k < nh else 2
        out += [(o, dim_arm), (o,)]
    out += [(ups_k // 2,)] * (n_latents - 1)
    out += [((ups_pre_k + 1) // 2,)] * (n_latents - 1)
    cin = n_latents
    for (o, k, _, _) in N.parse_synthesis_layers(layers):
        out += [(o, cin, k, k), (o,)]
        cin = o
    return out


def _numel(shape) -> int:
    n = 1
    for s in shape:
        n *= int(s)
    return n


def _fill_tensors(t: N.Tensors, nh: int, n_latents: int, n_syn: int, latents: Optional[Sequence[Optional[Tensor]]],
                  weight_ptrs: Sequence[Optional[int]], stride: int) -> None:
    if latents is not None:
        for i, l in enumerate(latents):
            t.latent[i] = _ptr(l)
    it = iter(weight_ptrs)
    for k in range(nh + 1):
        t.arm_w[k] = next(it)
        t.arm_b[k] = next(it)
    for i in range(n_latents - 1):
        t.ups_w[i] = next(it)
    for i in range(n_latents - 1):
        t.pre_w[i] = next(it)
    for l in range(n_syn):
        t.syn_w[l] = next(it)
        t.syn_b[l] = next(it)
    t.weight_batch_stride = int(stride)


def _pack_per_sample(weights: Sequence[Tensor], B: int) -> Tensor:
    """Per-sample weights [B, ...] each -> one contiguous [B, n_weights] buffer (what weight_batch_stride addresses)."""
    return torch.cat([w.detach().reshape(B, -1).float() for w in weights], dim=1).contiguous()


def _packed_ptrs(pack: Tensor, shapes) -> List[int]:
    base, off, out = pack.data_ptr(), 0, []
    for shp in shapes:
        out.append(base + 4 * off)
        off += _numel(shp)
    return out


def _scalar(q: N.QuantCfg, name: str, t: Optional[Tensor], f: float, keep: list) -> None:
    """soft_round_temperature / noise_parameter arrive as CUDA 0-d tensors (train.py:210,215), python floats
    (test.py:405-411) or CPU 0-d tensors (utils/nn.py:37-39): a CUDA tensor is read on the device (no sync)."""
    if t is not None and t.is_cuda:
        v = t.detach().reshape(()).to(torch.float32)
        keep.append(v)
        setattr(q, name + "_dev", v.data_ptr())
        setattr(q, name + "_host", 0.0)
    else:
        setattr(q, name + "_dev", None)
        setattr(q, name + "_host", float(t) if t is not None else float(f))


def _quant_cfg(training: bool, q_type: str, noise_type: str, ac_max_val: int, gain: float, T_t, sigma_t, T_f, sigma_f,
               noise: Optional[Tensor], keep: list) -> N.QuantCfg:
    if q_type not in N.Q_TYPES:
        raise ValueError(f"Unknown quantizer_type {q_type}")
    if noise_type not in N.NOISE_TYPES:
        noise_type = "none"  # the reference's match statement ignores unknown noise names (quantizer.py:186-199)
    q = N.QuantCfg()
    q.training = int(training)
    q.q_type = N.Q_TYPES[q_type]
    q.noise_type = N.NOISE_TYPES[noise_type]
    q.ac_max_val = int(ac_max_val)
    q.gain = float(gain)
    _scalar(q, "T", T_t, T_f, keep)
    _scalar(q, "sigma", sigma_t, sigma_f, keep)
    q.noise = _ptr(noise)
    q.seed = 0
    q.offset_host = 0
    q.offset_dev = None
    return q


def _noise_needed(training: bool, q_type: str, noise_type: str) -> bool:
    return training and noise_type in ("gaussian", "kumaraswamy") and q_type in ("none", "softround")


def _get_plan(B: int, H: int, W: int, n_latents: int, dim_arm: int, nh: int, layers: str, ups_k: int, ups_pre_k: int,
              device: torch.device) -> N.Plan:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    return N.get_plan(B, H, W, n_latents, dim_arm, nh, tuple(layers.split(_LAYER_SEP)), ups_k, ups_pre_k, idx)


def _empty(dev) -> Tensor:
    return torch.empty(0, device=dev, dtype=torch.float32)


# cudagraph_unsafe: under torch.compile(mode="reduce-overhead") Inductor partitions its CUDA graphs around the operator
# (the call creates / looks up the plan and launches ~12 kernels through ctypes; the surrounding pointwise code is
# still captured).  The fused trainer captures its own graph of ccb_train_step instead (enc/training/train.py).
_TAGS = (torch._C.Tag.cudagraph_unsafe,)


@torch.library.custom_op("coolchic_b200::forward", mutates_args=(), tags=_TAGS)
def cc_forward(latents: List[Tensor], weights: List[Tensor], noise: Optional[Tensor], T_t: Optional[Tensor],
               sigma_t: Optional[Tensor], T_f: float, sigma_f: float, H: int, W: int, dim_arm: int, nh: int,
               layers: str, ups_k: int, ups_pre_k: int, gain: float, training: bool, q_type: str, noise_type: str,
               ac_max_val: int, detailed: bool, per_sample: bool) -> List[Tensor]:
    """-> [raw_out [B,3,H,W], rate [B,M], workspace (uint8), noise drawn here (or empty), mu, scale, log_scale,
    yhat ([B,M] each when `detailed`, else empty)].  per_sample: every weight tensor carries a leading batch
    dimension (one weight set per batch element)."""
    L = len(latents)
    lat = [t.detach().contiguous().float() for t in latents]
    for t in lat + list(weights):
        _require_cuda(t, "CoolChicEncoder parameter")
    dev = lat[0].device
    B = int(lat[0].shape[0])
    plan = _get_plan(B, H, W, L, dim_arm, nh, layers, ups_k, ups_pre_k, dev)
    for i, (h, w) in enumerate(plan.latent_hw):
        if tuple(lat[i].shape) != (B, 1, h, w):
            raise N.CcbError(f"latent {i} has shape {tuple(lat[i].shape)}, expected {(B, 1, h, w)}")
    layer_list = layers.split(_LAYER_SEP)
    shapes = _weight_shapes(dim_arm, nh, layer_list, L, ups_k, ups_pre_k)
    if len(weights) != len(shapes):
        raise N.CcbError(f"expected {len(shapes)} weight tensors, got {len(weights)}")
    with torch.cuda.device(dev):
        keep: list = []
        drawn = None
        if noise is None and _noise_needed(training, q_type, noise_type):
            # the draw the reference makes (quantizer.py:190-197), same generator / same stream
            drawn = (torch.randn if noise_type == "gaussian" else torch.rand)(B, plan.M, device=dev,
                                                                                dtype=torch.float32)
        nz = drawn if drawn is not None else noise
        if nz is not None:
            nz = nz.detach().contiguous().float().reshape(B, plan.M)
        q = _quant_cfg(training, q_type, noise_type, ac_max_val, gain, T_t, sigma_t, T_f, sigma_f, nz, keep)
        params = N.Tensors()
        if per_sample:
            for w, shp in zip(weights, shapes):
                if tuple(w.shape) != (B,) + tuple(shp):
                    raise N.CcbError(f"per-sample weight has shape {tuple(w.shape)}, expected {(B,) + tuple(shp)}")
            pack = _pack_per_sample(weights, B)
            keep.append(pack)
            _fill_tensors(params, nh, L, len(layer_list), lat, _packed_ptrs(pack, shapes), pack.shape[1])
        else:
            wts = [t.detach().contiguous().float() for t in weights]
            keep.append(wts)
            _fill_tensors(params, nh, L, len(layer_list), lat, [w.data_ptr() for w in wts], 0)
        ws = torch.empty(plan.ws_bytes, dtype=torch.uint8, device=dev)
        raw = torch.empty(B, 3, H, W, device=dev, dtype=torch.float32)
        rate = torch.empty(B, plan.M, device=dev, dtype=torch.float32)
        if detailed:
            mu, scale, ls, yh = (torch.empty_like(rate) for _ in range(4))
        else:
            mu = scale = ls = yh = None
        N.check(N.lib().ccb_forward(plan.handle, C.byref(params), C.byref(q), ws.data_ptr(), raw.data_ptr(),
                                    rate.data_ptr(), _ptr(mu), _ptr(scale), _ptr(ls), _ptr(yh), _stream()),
                "ccb_forward")
    e = lambda t: t if t is not None else _empty(dev)  # noqa: E731
    return [raw, rate, ws, e(drawn if drawn is None else nz), e(mu), e(scale), e(ls), e(yh)]


@cc_forward.register_fake
def _cc_forward_fake(latents, weights, noise, T_t, sigma_t, T_f, sigma_f, H, W, dim_arm, nh, layers, ups_k, ups_pre_k,
                     gain, training, q_type, noise_type, ac_max_val, detailed, per_sample):
    L = len(latents)
    B = latents[0].shape[0]
    dev = latents[0].device
    M = sum(h * w for h, w in _latent_hw(H, W, L))
    plan = _get_plan(int(B), H, W, L, dim_arm, nh, layers, ups_k, ups_pre_k, dev)
    f = lambda *s: torch.empty(*s, device=dev, dtype=torch.float32)  # noqa: E731
    drawn = noise is None and _noise_needed(training, q_type, noise_type)
    det = (lambda: f(B, M)) if detailed else (lambda: f(0))
    return [f(B, 3, H, W), f(B, M), torch.empty(plan.ws_bytes, device=dev, dtype=torch.uint8),
            f(B, M) if drawn else f(0), det(), det(), det(), det()]


@torch.library.custom_op("coolchic_b200::backward", mutates_args=(), tags=_TAGS)
def cc_backward(d_raw: Tensor, d_rate: Optional[Tensor], ws: Tensor, noise: Optional[Tensor], latents: List[Tensor],
                weights: List[Tensor], T_t: Optional[Tensor], sigma_t: Optional[Tensor], T_f: float, sigma_f: float,
                H: int, W: int, dim_arm: int, nh: int, layers: str, ups_k: int, ups_pre_k: int, gain: float,
                training: bool, q_type: str, noise_type: str, ac_max_val: int, per_sample: bool,
                want_latent_grads: bool) -> List[Tensor]:
    """What loss.backward() does through coolchic_b200::forward (train.py:264): -> gradients of the L latents (empty
    tensors when not wanted: the reference yields None in "ste" / eval mode, quantizer.py:222-225) followed by the
    gradients of the weights, shaped like the inputs (per-sample weights: per-sample gradients)."""
    L = len(latents)
    lat = [t.detach().contiguous().float() for t in latents]
    dev = lat[0].device
    B = int(lat[0].shape[0])
    plan = _get_plan(B, H, W, L, dim_arm, nh, layers, ups_k, ups_pre_k, dev)
    layer_list = layers.split(_LAYER_SEP)
    shapes = _weight_shapes(dim_arm, nh, layer_list, L, ups_k, ups_pre_k)
    n_syn = len(layer_list)
    with torch.cuda.device(dev):
        keep: list = []
        nz = None if noise is None else noise.detach().contiguous().float().reshape(B, plan.M)
        q = _quant_cfg(training, q_type, noise_type, ac_max_val, gain, T_t, sigma_t, T_f, sigma_f, nz, keep)
        d_raw = d_raw.contiguous().float()
        d_rate_t = None if d_rate is None else d_rate.contiguous().float()
        g_lat = [torch.empty_like(t) if want_latent_grads else None for t in lat]
        params, grads = N.Tensors(), N.Tensors()
        if per_sample:
            pack = _pack_per_sample(weights, B)
            gpack = torch.empty_like(pack)
            _fill_tensors(params, nh, L, n_syn, lat, _packed_ptrs(pack, shapes), pack.shape[1])
            _fill_tensors(grads, nh, L, n_syn, g_lat, _packed_ptrs(gpack, shapes), gpack.shape[1])
        else:
            wts = [t.detach().contiguous().float() for t in weights]
            g_wts = [torch.empty_like(t) for t in wts]
            _fill_tensors(params, nh, L, n_syn, lat, [w.data_ptr() for w in wts], 0)
            _fill_tensors(grads, nh, L, n_syn, g_lat, [g.data_ptr() for g in g_wts], 0)
        N.check(N.lib().ccb_backward(plan.handle, C.byref(params), C.byref(q), ws.data_ptr(), d_raw.data_ptr(),
                                     _ptr(d_rate_t), 0.0, C.byref(grads), _stream()), "ccb_backward")
        if per_sample:
            g_wts, off = [], 0
            for w, shp in zip(weights, shapes):
                n = _numel(shp)
                g_wts.append(gpack[:, off:off + n].reshape(w.shape).contiguous())
                off += n
    return [g if g is not None else _empty(dev) for g in g_lat] + list(g_wts)


@cc_backward.register_fake
def _cc_backward_fake(d_raw, d_rate, ws, noise, latents, weights, T_t, sigma_t, T_f, sigma_f, H, W, dim_arm, nh, layers,
                      ups_k, ups_pre_k, gain, training, q_type, noise_type, ac_max_val, per_sample, want_latent_grads):
    dev = latents[0].device
    gl = [torch.empty_like(t, dtype=torch.float32) if want_latent_grads else torch.empty(0, device=dev) for t in latents]
    return gl + [torch.empty_like(w, dtype=torch.float32) for w in weights]


def _setup_context(ctx, inputs, output):
    (latents, weights, noise, T_t, sigma_t, T_f, sigma_f, H, W, dim_arm, nh, layers, ups_k, ups_pre_k, gain, training,
     q_type, noise_type, ac_max_val, detailed, per_sample) = inputs
    ws, drawn = output[2], output[3]
    nz = noise if noise is not None else (drawn if drawn.numel() > 0 else None)
    ctx.n_lat, ctx.n_w = len(latents), len(weights)
    ctx.has = (nz is not None, T_t is not None, sigma_t is not None)
    ctx.save_for_backward(ws, *latents, *weights, *[t for t in (nz, T_t, sigma_t) if t is not None])
    ctx.scalars = (T_f, sigma_f, H, W, dim_arm, nh, layers, ups_k, ups_pre_k, gain, training, q_type, noise_type,
                   ac_max_val, per_sample)
    ctx.no_latent_grad = (not training) or q_type in ("ste", "hardround")
    ctx.lat_needs = [t.requires_grad for t in latents]
    ctx.w_needs = [t.requires_grad for t in weights]


def _backward(ctx, grads):
    d_raw, d_rate = grads[0], grads[1]
    saved = list(ctx.saved_tensors)
    ws = saved[0]
    latents = saved[1:1 + ctx.n_lat]
    weights = saved[1 + ctx.n_lat:1 + ctx.n_lat + ctx.n_w]
    rest = saved[1 + ctx.n_lat + ctx.n_w:]
    opt = []
    for present in ctx.has:
        opt.append(rest.pop(0) if present else None)
    nz, T_t, sigma_t = opt
    (T_f, sigma_f, H, W, dim_arm, nh, layers, ups_k, ups_pre_k, gain, training, q_type, noise_type, ac_max_val,
     per_sample) = ctx.scalars
    if d_raw is None:
        d_raw = torch.zeros(latents[0].shape[0], 3, H, W, device=latents[0].device, dtype=torch.float32)
    if d_rate is None:
        d_rate = torch.zeros(latents[0].shape[0], sum(h * w for h, w in _latent_hw(H, W, ctx.n_lat)),
                             device=latents[0].device, dtype=torch.float32)
    want_lat = any(ctx.lat_needs) and not ctx.no_latent_grad
    out = torch.ops.coolchic_b200.backward(d_raw, d_rate, ws, nz, latents, weights, T_t, sigma_t, T_f, sigma_f, H, W,
                                           dim_arm, nh, layers, ups_k, ups_pre_k, gain, training, q_type, noise_type,
                                           ac_max_val, per_sample, want_lat)
    g_lat = [g if (want_lat and need) else None for g, need in zip(out[:ctx.n_lat], ctx.lat_needs)]
    g_w = [g if need else None for g, need in zip(out[ctx.n_lat:], ctx.w_needs)]
    return (g_lat, g_w) + (None,) * 19


torch.library.register_autograd("coolchic_b200::forward", _backward, setup_context=_setup_context)


# ------------------------------------------------------------------------------------------------
# vmap: torch.vmap(functional_call(decoder, per_sample_params, (latents,)), randomness="different")
# (hypernet/hypernet.py:1266-1288) becomes ONE call of the per-sample-weight kernels.
# ------------------------------------------------------------------------------------------------
def _front(t: Tensor, dim: Optional[int], n: int) -> Tensor:
    """Batched tensor with its vmap dimension moved to the front; an unbatched one is expanded."""
    if dim is None:
        return t.unsqueeze(0).expand(n, *t.shape)
    return t.movedim(dim, 0)


@torch.library.register_vmap("coolchic_b200::forward")
def _cc_forward_vmap(info, in_dims, latents, weights, noise, T_t, sigma_t, T_f, sigma_f, H, W, dim_arm, nh, layers,
                     ups_k, ups_pre_k, gain, training, q_type, noise_type, ac_max_val, detailed, per_sample):
    Bv = info.batch_size
    lat_dims, w_dims, noise_dim, T_dim, sigma_dim = in_dims[0], in_dims[1], in_dims[2], in_dims[3], in_dims[4]
    if T_dim is not None or sigma_dim is not None:
        raise NotImplementedError("coolchic_b200: per-sample soft_round_temperature / noise_parameter under vmap")
    if detailed:
        raise NotImplementedError("Batching is not yet supported for additional outputs.")
    lats = [_front(t, d, Bv) for t, d in zip(latents, lat_dims)]  # [Bv, Bi, 1, h, w]
    Bi = lats[0].shape[1]
    lats = [t.reshape(Bv * Bi, *t.shape[2:]) for t in lats]
    wts = []
    for t, d in zip(weights, w_dims):
        t = _front(t, d, Bv)  # [Bv, (Bi,) ...]
        if per_sample:  # already one set per inner batch element
            t = t.reshape(Bv * Bi, *t.shape[2:])
        elif Bi > 1:
            t = t.repeat_interleave(Bi, dim=0)
        wts.append(t)
    if noise is not None:
        noise = _front(noise, noise_dim, Bv).reshape(Bv * Bi, -1)
    out = torch.ops.coolchic_b200.forward(lats, wts, noise, T_t, sigma_t, T_f, sigma_f, H, W, dim_arm, nh, layers,
                                          ups_k, ups_pre_k, gain, training, q_type, noise_type, ac_max_val, False, True)
    raw, rate = out[0], out[1]
    res = [raw.reshape(Bv, Bi, *raw.shape[1:]), rate.reshape(Bv, Bi, rate.shape[1])] + list(out[2:])
    return res, [0, 0] + [None] * (len(out) - 2)


def coolchic_forward(H: int, W: int, dim_arm: int, nh: int, layers: Sequence[str], ups_k: int, ups_pre_k: int,
                     gain: float, latents: Sequence[Tensor], weights: Sequence[Tensor], training: bool, q_type: str,
                     noise_type: str, T, sigma, ac_max_val: int = -1, noise: Optional[Tensor] = None,
                     detailed: bool = False, per_sample: Optional[bool] = None):
    """Differentiable forward through the custom operator.  Returns (raw_out, rate) or, if `detailed`,
    (raw_out, rate, mu, scale, log_scale, yhat).  T / sigma: python floats or 0-d tensors (CUDA or CPU)."""
    if per_sample is None:
        per_sample = weights[0].dim() == 3  # arm.mlp.0.weight is [out, C]; one more dimension = one set per sample
    T_t = T if isinstance(T, Tensor) else None
    s_t = sigma if isinstance(sigma, Tensor) else None
    T_f = 0.0 if T_t is not None else (1.0 if T is None else float(T))
    s_f = 0.0 if s_t is not None else (1.0 if sigma is None else float(sigma))
    out = torch.ops.coolchic_b200.forward(list(latents), list(weights), noise, T_t, s_t, T_f, s_f, int(H), int(W),
                                          int(dim_arm), int(nh), _LAYER_SEP.join(layers), int(ups_k), int(ups_pre_k),
                                          float(gain), bool(training), q_type, noise_type, int(ac_max_val),
                                          bool(detailed), bool(per_sample))
    if detailed:
        return out[0], out[1], out[4], out[5], out[6], out[7]
    return out[0], out[1]
