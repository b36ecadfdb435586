"""Torch-facing wrappers over the C ABI: one autograd.Function for the whole
CoolChicEncoder forward/backward, plus the fused loss / clip+Adam / train-step /
integer-ARM entry points.  PyTorch is only the owner of device memory and streams here.
"""

from __future__ import annotations

import ctypes as C
import threading
from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import Tensor

from . import _native as N
from . import _op


def _require_cuda(t: Tensor, what: str) -> None:
    if not t.is_cuda:
        raise N.CcbError(
            f"{what} is on {t.device}: coolchic_b200 runs on CUDA (sm_100a) only and has no CPU fallback"
        )


def _ptr(t: Optional[Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


# torch.cuda.CUDAGraph.capture_begin()/capture_end() of two host threads at the same time crash inside torch (segfault
# in capture_end, seen one run in four with threaded warm-ups): captures are serialised process-wide.  Replays and
# ordinary launches of other threads go on meanwhile (captures here use capture_error_mode="thread_local").
CAPTURE_LOCK = threading.Lock()


class _OwnedStream:
    """Keeps a ccb_stream_create'd stream alive for as long as its torch wrapper is referenced."""

    def __init__(self, device: int):
        h = C.c_void_p()
        N.check(N.lib().ccb_stream_create(device, C.byref(h)), "ccb_stream_create")
        self.device, self.handle = device, h.value

    def __del__(self):
        import sys

        try:
            # not at interpreter exit (the CUDA runtime may already be gone; the process frees everything anyway)
            if getattr(self, "handle", None) and not sys.is_finalizing():
                N.lib().ccb_stream_destroy(self.device, self.handle)
                self.handle = None
        except Exception:
            pass


def new_stream(device) -> "torch.cuda.Stream":
    """A PRIVATE non-blocking stream on `device`.  torch.cuda.Stream() draws from a pool of 32 streams per device in
    round-robin order: with several encoders training side by side (each with a main, an auxiliary and a capture
    stream) two host threads end up on the same cudaStream_t, and one's graph capture breaks the other's replays."""
    device = torch.device(device)
    idx = device.index if device.index is not None else torch.cuda.current_device()
    owner = _OwnedStream(idx)
    s = torch.cuda.ExternalStream(owner.handle, device=torch.device("cuda", idx))
    s._ccb_owner = owner
    return s


class ArchSpec:
    """Hashable description of a decoder architecture (what CoolChicEncoderParameter holds)."""

    def __init__(self, H, W, dim_arm, n_hidden_layers_arm, layers_synthesis, n_latents=7, ups_k_size=8,
                 ups_preconcat_k_size=7, encoder_gain=16.0):
        self.H, self.W = int(H), int(W)
        self.dim_arm, self.nh = int(dim_arm), int(n_hidden_layers_arm)
        self.layers_synthesis = tuple(layers_synthesis)
        self.n_latents = int(n_latents)
        self.ups_k, self.ups_pre_k = int(ups_k_size), int(ups_preconcat_k_size)
        self.gain = float(encoder_gain)
        self.n_syn = len(self.layers_synthesis)

    def plan(self, B: int, device: torch.device) -> N.Plan:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        return N.get_plan(B, self.H, self.W, self.n_latents, self.dim_arm, self.nh, self.layers_synthesis,
                          self.ups_k, self.ups_pre_k, idx)

    # ---- canonical ordering of the weight tensors (matches ccb_tensors and the reference's names)
    def weight_names(self) -> List[str]:
        names = []
        for k in range(self.nh + 1):
            names += [f"arm.mlp.{2 * k}.weight", f"arm.mlp.{2 * k}.bias"]
        for i in range(self.n_latents - 1):
            names.append(f"upsampling.conv_transpose2ds.{i}.parametrizations.weight.original")
        for i in range(self.n_latents - 1):
            names.append(f"upsampling.conv2ds.{i}.parametrizations.weight.original")
        for l in range(self.n_syn):
            names += [f"synthesis.layers.{2 * l}.weight", f"synthesis.layers.{2 * l}.bias"]
        return names

    def latent_names(self) -> List[str]:
        return [f"latent_grids.{i}.data" for i in range(self.n_latents)]


def pack_tensors(spec: ArchSpec, latents: Optional[Sequence[Optional[Tensor]]],
                 weights: Sequence[Optional[Tensor]]) -> N.Tensors:
    """latents: n_latents tensors (or None entries); weights: in `spec.weight_names()` order."""
    t = N.Tensors()
    if latents is not None:
        for i, l in enumerate(latents):
            t.latent[i] = _ptr(l)
    it = iter(weights)
    for k in range(spec.nh + 1):
        t.arm_w[k] = _ptr(next(it))
        t.arm_b[k] = _ptr(next(it))
    for i in range(spec.n_latents - 1):
        t.ups_w[i] = _ptr(next(it))
    for i in range(spec.n_latents - 1):
        t.pre_w[i] = _ptr(next(it))
    for l in range(spec.n_syn):
        t.syn_w[l] = _ptr(next(it))
        t.syn_b[l] = _ptr(next(it))
    return t


def make_quant_cfg(spec: ArchSpec, training: bool, q_type: str, noise_type: str, T, sigma, ac_max_val: int = -1,
                   noise: Optional[Tensor] = None, seed: int = 0, offset: int = 0,
                   offset_dev: Optional[Tensor] = None, keep: Optional[list] = None) -> N.QuantCfg:
    if q_type not in N.Q_TYPES:
        raise ValueError(f"Unknown quantizer_type {q_type}")
    if noise_type not in N.NOISE_TYPES:
        # the reference's match statement silently ignores unknown noise names (quantizer.py:186-199)
        noise_type = "none"
    q = N.QuantCfg()
    q.training = int(training)
    q.q_type = N.Q_TYPES[q_type]
    q.noise_type = N.NOISE_TYPES[noise_type]
    q.ac_max_val = int(ac_max_val)
    q.gain = spec.gain
    for name, val in (("T", T), ("sigma", sigma)):
        if isinstance(val, Tensor) and val.is_cuda:
            v = val.detach().reshape(()).to(torch.float32)
            if keep is not None:
                keep.append(v)
            setattr(q, name + "_dev", v.data_ptr())
            setattr(q, name + "_host", 0.0)
        else:
            setattr(q, name + "_dev", None)
            setattr(q, name + "_host", float(val) if val is not None else 1.0)
    q.noise = _ptr(noise)
    q.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    q.offset_host = int(offset)
    q.offset_dev = _ptr(offset_dev)
    return q


def _noise_needed(training: bool, q_type: str, noise_type: str) -> bool:
    return training and noise_type in ("gaussian", "kumaraswamy") and q_type in ("none", "softround")


def coolchic_forward(spec: ArchSpec, latents: Sequence[Tensor], weights: Sequence[Tensor], training: bool,
                     q_type: str = "softround", noise_type: str = "kumaraswamy", T=0.3, sigma=1.0,
                     ac_max_val: int = -1, noise: Optional[Tensor] = None, detailed: bool = False):
    """Differentiable forward (the `coolchic_b200::forward` custom operator, _op.py: Dynamo-traceable, with an
    autograd formula and a vmap rule).  `weights` in `spec.weight_names()` order; tensors with one extra leading
    dimension are per-sample weights.  Returns (raw_out [B,3,H,W], rate [B,M]) or, if detailed, additionally
    (mu, scale, log_scale, yhat) as [B,M] tensors."""
    return _op.coolchic_forward(spec.H, spec.W, spec.dim_arm, spec.nh, spec.layers_synthesis, spec.ups_k,
                                spec.ups_pre_k, spec.gain, latents, weights, training, q_type, noise_type, T, sigma,
                                ac_max_val, noise, detailed)


# --------------------------------------------------------------------------------------------
# Fused trainer state: one flat arena per image for params / grads / Adam moments
# --------------------------------------------------------------------------------------------
class FusedState:
    """Device-resident state of one image's overfit (train.py:157-266 minus the Python loop):
    a flat parameter arena [latents | weights] with views named like the reference's
    state_dict, matching gradient / exp_avg / exp_avg_sq arenas, Adam step counter, learning
    rate, Philox offset and the 4 result scalars (loss, mse, bpp, grad norm)."""

    def __init__(self, spec: ArchSpec, B: int, device: torch.device, slot_wait_s: float = 0.0):
        self.spec, self.B, self.device = spec, B, device
        self.plan = spec.plan(B, device)
        n = self.plan.n_param_floats
        # params, grads, m, v: rows padded to a multiple of 4 floats so that each starts 16-byte aligned (128-bit Adam)
        self._arena_rows = torch.zeros(4, (n + 3) // 4 * 4, device=device, dtype=torch.float32)
        self.arena = self._arena_rows[:, :n]
        self.views: List[Dict[str, Tensor]] = [dict() for _ in range(4)]
        shapes = self._shapes()
        off = 0
        for name, shp in shapes:
            cnt = 1
            for s in shp:
                cnt *= s
            for k in range(4):
                self.views[k][name] = self.arena[k, off:off + cnt].view(shp)
            off += cnt
        assert off == n, (off, n)
        self.ws = torch.empty(self.plan.ws_bytes, dtype=torch.uint8, device=device)
        self.out4 = torch.zeros(4, device=device, dtype=torch.float32)
        self.lr = torch.zeros((), device=device, dtype=torch.float32)
        self.step = torch.zeros((), device=device, dtype=torch.int32)
        self.philox_offset = torch.zeros((), device=device, dtype=torch.int64)
        self.T = torch.full((), 0.3, device=device, dtype=torch.float32)
        self.sigma = torch.full((), 1.0, device=device, dtype=torch.float32)
        self._structs = [self._pack(k) for k in range(4)]
        self.aux_stream = new_stream(device)  # ARM branch of the fused step (private: see new_stream)
        # this encoder's constant-bank weight slot (released with the object): concurrent FusedStates never share one
        self.slot = N.WeightSlot(device.index if device.index is not None else torch.cuda.current_device(),
                                 wait_s=slot_wait_s)

    def close(self) -> None:
        """Give the weight slot back now (the caller guarantees that no work of this state is still queued)."""
        self.slot.release()

    def _shapes(self) -> List[Tuple[str, Tuple[int, ...]]]:
        sp, B = self.spec, self.B
        C = sp.dim_arm
        out = [(f"latent_grids.{i}.data", (B, 1, h, w)) for i, (h, w) in enumerate(self.plan.latent_hw)]
        for k in range(sp.nh + 1):
            o = C if k < sp.nh else 2
            out += [(f"arm.mlp.{2 * k}.weight", (o, C)), (f"arm.mlp.{2 * k}.bias", (o,))]
        for i in range(sp.n_latents - 1):
            out.append((f"upsampling.conv_transpose2ds.{i}.parametrizations.weight.original", (sp.ups_k // 2,)))
        for i in range(sp.n_latents - 1):
            out.append((f"upsampling.conv2ds.{i}.parametrizations.weight.original", ((sp.ups_pre_k + 1) // 2,)))
        cin = sp.n_latents
        for l, (o, k, _, _) in enumerate(N.parse_synthesis_layers(sp.layers_synthesis)):
            out += [(f"synthesis.layers.{2 * l}.weight", (o, cin, k, k)), (f"synthesis.layers.{2 * l}.bias", (o,))]
            cin = o
        return out

    def _pack(self, k: int) -> N.Tensors:
        v = self.views[k]
        return pack_tensors(self.spec, [v[n] for n in self.spec.latent_names()],
                            [v[n] for n in self.spec.weight_names()])

    @property
    def params(self) -> Dict[str, Tensor]:
        return self.views[0]

    @property
    def grads(self) -> Dict[str, Tensor]:
        return self.views[1]

    def load(self, params: Dict[str, Tensor]) -> None:
        for k, v in self.views[0].items():
            v.copy_(params[k].to(self.device).reshape(v.shape))

    def train_step(self, target: Tensor, lmbda: float, q_type: str, noise_type: str, noise: Optional[Tensor] = None,
                   optimize_mask: int = N.OPT_ALL, max_norm: float = 1e-1, seed: int = 0,
                   betas=(0.9, 0.999), eps: float = 1e-8, two_streams: bool = True) -> None:
        """One iteration of train.py:244-268; results land in self.out4 (no host sync)."""
        keep: list = []
        q = make_quant_cfg(self.spec, True, q_type, noise_type, self.T, self.sigma, -1, noise, seed=seed, keep=keep)
        with torch.cuda.device(self.device):
            N.check(N.lib().ccb_train_step(
                self.plan.handle, C.byref(self._structs[0]), C.byref(self._structs[1]), C.byref(self._structs[2]),
                C.byref(self._structs[3]), C.byref(q), self.ws.data_ptr(), target.data_ptr(), float(lmbda),
                int(optimize_mask), float(max_norm), self.lr.data_ptr(), float(betas[0]), float(betas[1]),
                float(eps), self.step.data_ptr(), self.philox_offset.data_ptr() if noise is None else None,
                self.out4.data_ptr(), _stream(), self.aux_stream.cuda_stream if two_streams else None,
                self.slot.index), "ccb_train_step")


def _fs_train_step_timed(self, target: Tensor, lmbda: float, q_type: str, noise_type: str,
                         optimize_mask: int = N.OPT_ALL, max_norm: float = 1e-1, seed: int = 0,
                         betas=(0.9, 0.999), eps: float = 1e-8) -> Dict[str, float]:
    """train_step with per-stage CUDA-event timing (synchronises).  -> {stage: ms}."""
    keep: list = []
    q = make_quant_cfg(self.spec, True, q_type, noise_type, self.T, self.sigma, -1, None, seed=seed, keep=keep)
    ms = (C.c_float * len(N.STAGES))()
    with torch.cuda.device(self.device):
        N.check(N.lib().ccb_train_step_timed(
            self.plan.handle, C.byref(self._structs[0]), C.byref(self._structs[1]), C.byref(self._structs[2]),
            C.byref(self._structs[3]), C.byref(q), self.ws.data_ptr(), target.data_ptr(), float(lmbda),
            int(optimize_mask), float(max_norm), self.lr.data_ptr(), float(betas[0]), float(betas[1]),
            float(eps), self.step.data_ptr(), self.philox_offset.data_ptr(), self.out4.data_ptr(), _stream(), ms,
            self.slot.index), "ccb_train_step_timed")
    return {n: float(ms[i]) for i, n in enumerate(N.STAGES)}


FusedState.train_step_timed = _fs_train_step_timed


def _fs_clip_adam_step(self, optimize_mask: int = N.OPT_ALL, max_norm: float = 1e-1, betas=(0.9, 0.999),
                       eps: float = 1e-8, skip_latents: bool = False) -> Tensor:
    """clip_grad_norm_ + Adam on the gradient arena (train.py:265-266).  Returns the norm (device)."""
    norm = torch.zeros((), device=self.device, dtype=torch.float32)
    g = self._structs[1]
    if skip_latents:  # latents without gradient (quantizer "ste"): excluded from norm and update
        g = pack_tensors(self.spec, [None] * self.spec.n_latents, [self.views[1][n] for n in self.spec.weight_names()])
    with torch.cuda.device(self.device):
        N.check(N.lib().ccb_clip_adam_step(
            self.plan.handle, self.ws.data_ptr(), C.byref(self._structs[0]), C.byref(g), C.byref(self._structs[2]),
            C.byref(self._structs[3]), int(optimize_mask), float(max_norm), self.lr.data_ptr(), float(betas[0]),
            float(betas[1]), float(eps), self.step.data_ptr(), norm.data_ptr(), _stream()), "ccb_clip_adam_step")
    return norm


FusedState.clip_adam_step = _fs_clip_adam_step


def fused_loss(spec: ArchSpec, plan: N.Plan, ws: Tensor, raw: Tensor, rate: Tensor, target: Tensor, lmbda: float):
    """-> (out4 = [loss, mse, bpp, 0], d_raw_out)."""
    out4 = torch.zeros(4, device=raw.device, dtype=torch.float32)
    d_raw = torch.empty_like(raw)
    with torch.cuda.device(raw.device):
        N.check(N.lib().ccb_loss(plan.handle, ws.data_ptr(), raw.data_ptr(), rate.data_ptr(), target.data_ptr(),
                                 float(lmbda), d_raw.data_ptr(), out4.data_ptr(), _stream()), "ccb_loss")
    return out4, d_raw


def arm_int(spec: ArchSpec, yhat_q: Tensor, w_q: Sequence[Tensor], b_q: Sequence[Tensor], indices: bool = True):
    """Fixed-point ARM (cpp/arm_cpu.cpp:18-106) for every position.  yhat_q: int32 [B,M];
    w_q / b_q: int32 per layer.  -> mu_q8, log_scale_q8 [, mu_rounded, mu_index, sig_index]."""
    _require_cuda(yhat_q, "yhat_q")
    B = yhat_q.shape[0]
    plan = spec.plan(B, yhat_q.device)
    assert yhat_q.dtype == torch.int32 and tuple(yhat_q.shape) == (B, plan.M)
    yhat_q = yhat_q.contiguous()
    w_q = [w.contiguous() for w in w_q]
    b_q = [b.contiguous() for b in b_q]
    assert all(t.dtype == torch.int32 and t.is_cuda for t in w_q + b_q)
    wp = (C.c_void_p * len(w_q))(*[w.data_ptr() for w in w_q])
    bp = (C.c_void_p * len(b_q))(*[b.data_ptr() for b in b_q])
    outs = [torch.empty_like(yhat_q) for _ in range(5 if indices else 2)]
    ptrs = [o.data_ptr() for o in outs] + [None] * (5 - len(outs))
    with torch.cuda.device(yhat_q.device):
        N.check(N.lib().ccb_arm_int(plan.handle, yhat_q.data_ptr(), wp, bp, *ptrs, _stream()), "ccb_arm_int")
    return tuple(outs)


def decode_int(latents_q: Tensor, img_size: Tuple[int, int], n_latents: int, ups_w, pre_w, syn_layers,
               want_ups: bool = False, want_pixels: bool = True):
    """Fixed-point upsampling + synthesis of the DECODER (cpp/cc-frame-decoder.cpp run_ups / run_syn), bit-exact, on
    the GPU.  latents_q: int32 [M] integer latents, finest level first (what ccb_arm_int takes); ups_w [n_ups, k],
    pre_w [n_pre, kp]: integer filters (12 fractional bits, host arrays / CPU tensors); syn_layers: list of dicts
    {w: int32 [n_out, n_in, ks, ks], b: int32 [n_out], ks, residual, relu} with CUDA tensors for w / b.
    -> dict(out=int32 [3, H, W] at 12 bits [, ups=int32 [n, H, W]] [, pixels=uint8 [H, W, 3]])."""
    _require_cuda(latents_q, "latents_q")
    H, W = int(img_size[0]), int(img_size[1])
    dev = latents_q.device
    assert latents_q.dtype == torch.int32
    latents_q = latents_q.contiguous().reshape(-1)
    ups_h = torch.as_tensor(ups_w, dtype=torch.int32).cpu().contiguous()
    pre_h = torch.as_tensor(pre_w, dtype=torch.int32).cpu().contiguous()
    keep = []
    arr = (N.DecLayer * len(syn_layers))()
    for i, l in enumerate(syn_layers):
        w, b = l["w"].contiguous(), l["b"].contiguous()
        assert w.dtype == torch.int32 and b.dtype == torch.int32 and w.is_cuda and b.is_cuda
        keep += [w, b]
        arr[i] = N.DecLayer(int(w.shape[0]), int(l["ks"]), int(bool(l["residual"])), int(bool(l["relu"])), w.data_ptr(),
                            b.data_ptr())
    L = N.lib()
    nbytes = int(L.ccb_dec_int_workspace_bytes(H, W, n_latents, len(syn_layers), arr))
    if nbytes == 0:
        raise N.CcbError("ccb_dec_int_workspace_bytes: invalid geometry")
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    out = torch.empty((3, H, W), dtype=torch.int32, device=dev)
    ups = torch.empty((n_latents, H, W), dtype=torch.int32, device=dev) if want_ups else None
    pix = torch.empty((H, W, 3), dtype=torch.uint8, device=dev) if want_pixels else None
    with torch.cuda.device(dev):
        N.check(L.ccb_dec_int(H, W, n_latents, latents_q.data_ptr(), int(ups_h.shape[1]), int(ups_h.shape[0]),
                              ups_h.data_ptr(), int(pre_h.shape[1]), int(pre_h.shape[0]), pre_h.data_ptr(),
                              len(syn_layers), arr, ws.data_ptr(), nbytes, _ptr(ups), out.data_ptr(), _ptr(pix),
                              _stream()), "ccb_dec_int")
    res = {"out": out}
    if want_ups:
        res["ups"] = ups
    if want_pixels:
        res["pixels"] = pix
    res["_keepalive"] = (ws, keep, ups_h, pre_h)
    return res


def fp32_peak_tflops(device: int = 0, n_iters: int = 20000) -> float:
    v = C.c_float(0)
    N.check(N.lib().ccb_fp32_peak(device, n_iters, C.byref(v)), "ccb_fp32_peak")
    return float(v.value)
