"""ctypes binding of libccb.so (include/ccb.h).  No torch types cross this boundary: only raw
device pointers, sizes and the CUDA stream handle.

The product path has NO CPU fallback: if the shared library is missing, or an entry point
returns an error, this module raises.
"""

from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libccb.so")

CCB_MAX_LATENTS = 8
CCB_MAX_ARM_LAYERS = 5
CCB_MAX_SYN_LAYERS = 8

Q_TYPES = {"none": 0, "softround_alone": 1, "softround": 2, "hardround": 3, "ste": 4, "true_ste": 5}
NOISE_TYPES = {"none": 0, "gaussian": 1, "kumaraswamy": 2}
STAGES = ["quantise", "arm_fwd", "ups_fwd", "syn_fwd", "loss", "syn_bwd", "ups_bwd", "arm_bwd", "reduce",
          "quant_bwd", "clip_adam"]
OPT_LATENT, OPT_ARM, OPT_UPS, OPT_SYN, OPT_ALL = 1, 2, 4, 8, 15

# every symbol include/ccb.h declares
EXPORTS = [
    "ccb_strerror", "ccb_last_cuda_error", "ccb_version", "ccb_plan_create", "ccb_plan_destroy",
    "ccb_plan_n_latent_values", "ccb_plan_latent_geometry", "ccb_plan_workspace_bytes",
    "ccb_plan_n_weight_params", "ccb_plan_n_param_floats", "ccb_plan_train_step_launches", "ccb_forward", "ccb_backward", "ccb_loss",
    "ccb_clip_adam_step", "ccb_train_step", "ccb_train_step_timed", "ccb_overfit_host", "ccb_arm_int", "ccb_fp32_peak",
    "ccb_dec_int", "ccb_dec_int_workspace_bytes", "ccb_weight_slot_acquire", "ccb_weight_slot_release",
    "ccb_loss_noalloc", "ccb_stream_create", "ccb_stream_destroy",
]
NOALLOC_MAX_ELEMENTS = 4096 * 4096  # CCB_NOALLOC_MAX_BLOCKS * 4096 image elements (include/ccb.h)
N_WEIGHT_SLOTS = 8


class DecLayer(C.Structure):
    """ccb_dec_layer: one synthesis layer of the fixed-point decoder (w, b: device int32 pointers)."""

    _fields_ = [("n_out", C.c_int32), ("ks", C.c_int32), ("residual", C.c_int32), ("relu", C.c_int32),
                ("w", C.c_void_p), ("b", C.c_void_p)]


class SynLayer(C.Structure):
    _fields_ = [("out_ch", C.c_int32), ("k", C.c_int32), ("residual", C.c_int32), ("relu", C.c_int32)]


class Arch(C.Structure):
    _fields_ = [
        ("B", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("n_latents", C.c_int32),
        ("dim_arm", C.c_int32), ("n_hidden_arm", C.c_int32), ("ups_k", C.c_int32), ("ups_pre_k", C.c_int32),
        ("n_syn_layers", C.c_int32), ("syn", SynLayer * CCB_MAX_SYN_LAYERS),
    ]


class Tensors(C.Structure):
    _fields_ = [
        ("latent", C.c_void_p * CCB_MAX_LATENTS),
        ("arm_w", C.c_void_p * CCB_MAX_ARM_LAYERS),
        ("arm_b", C.c_void_p * CCB_MAX_ARM_LAYERS),
        ("ups_w", C.c_void_p * CCB_MAX_LATENTS),
        ("pre_w", C.c_void_p * CCB_MAX_LATENTS),
        ("syn_w", C.c_void_p * CCB_MAX_SYN_LAYERS),
        ("syn_b", C.c_void_p * CCB_MAX_SYN_LAYERS),
        ("weight_batch_stride", C.c_int64),  # floats; 0: one weight set for the batch, > 0: per-sample weights
    ]


class QuantCfg(C.Structure):
    _fields_ = [
        ("training", C.c_int32), ("q_type", C.c_int32), ("noise_type", C.c_int32), ("ac_max_val", C.c_int32),
        ("gain", C.c_float), ("T_host", C.c_float), ("sigma_host", C.c_float),
        ("T_dev", C.c_void_p), ("sigma_dev", C.c_void_p), ("noise", C.c_void_p),
        ("seed", C.c_uint64), ("offset_dev", C.c_void_p), ("offset_host", C.c_uint64),
    ]


class CcbError(RuntimeError):
    pass


_lib: Optional[C.CDLL] = None


def lib() -> C.CDLL:
    """Load libccb.so (built by __graft_entry__.build() / csrc/build.sh).  Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CcbError(
            f"{LIB_PATH} not found: build the sm_100a kernels first (python -c 'import __graft_entry__ as g; "
            f"g.build()').  coolchic_b200 has no CPU fallback."
        )
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, f32, f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_double
    L.ccb_strerror.restype = C.c_char_p
    L.ccb_strerror.argtypes = [C.c_int]
    L.ccb_last_cuda_error.restype = C.c_char_p
    L.ccb_version.restype = C.c_int
    L.ccb_plan_create.argtypes = [C.POINTER(vp), C.POINTER(Arch), C.c_int]
    L.ccb_plan_destroy.argtypes = [vp]
    L.ccb_plan_destroy.restype = None
    L.ccb_plan_n_latent_values.argtypes = [vp]
    L.ccb_plan_n_latent_values.restype = i64
    L.ccb_plan_latent_geometry.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i64)]
    L.ccb_plan_workspace_bytes.argtypes = [vp]
    L.ccb_plan_workspace_bytes.restype = C.c_size_t
    L.ccb_plan_n_weight_params.argtypes = [vp]
    L.ccb_plan_n_weight_params.restype = i64
    L.ccb_plan_train_step_launches.argtypes = [vp]
    L.ccb_plan_train_step_launches.restype = C.c_int
    L.ccb_plan_n_param_floats.argtypes = [vp]
    L.ccb_plan_n_param_floats.restype = i64
    L.ccb_forward.argtypes = [vp, C.POINTER(Tensors), C.POINTER(QuantCfg), vp, vp, vp, vp, vp, vp, vp, vp]
    L.ccb_backward.argtypes = [vp, C.POINTER(Tensors), C.POINTER(QuantCfg), vp, vp, vp, f32, C.POINTER(Tensors), vp]
    L.ccb_loss.argtypes = [vp, vp, vp, vp, vp, f32, vp, vp, vp]
    L.ccb_clip_adam_step.argtypes = [vp, vp, C.POINTER(Tensors), C.POINTER(Tensors), C.POINTER(Tensors),
                                     C.POINTER(Tensors), C.c_int, f32, vp, f64, f64, f64, vp, vp, vp]
    L.ccb_train_step.argtypes = [vp, C.POINTER(Tensors), C.POINTER(Tensors), C.POINTER(Tensors), C.POINTER(Tensors),
                                 C.POINTER(QuantCfg), vp, vp, f32, C.c_int, f32, vp, f64, f64, f64, vp, vp, vp, vp, vp,
                                 C.c_int]
    L.ccb_train_step_timed.argtypes = L.ccb_train_step.argtypes[:-2] + [C.POINTER(f32), C.c_int]
    L.ccb_weight_slot_acquire.argtypes = [C.c_int, C.POINTER(C.c_int)]
    L.ccb_weight_slot_release.argtypes = [C.c_int, C.c_int]
    L.ccb_loss_noalloc.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, i64, vp, vp, vp, f32, vp, vp, vp]
    L.ccb_loss_noalloc.restype = C.c_int
    L.ccb_stream_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.ccb_stream_create.restype = C.c_int
    L.ccb_stream_destroy.argtypes = [C.c_int, vp]
    L.ccb_stream_destroy.restype = C.c_int
    L.ccb_train_step_timed.restype = C.c_int
    L.ccb_overfit_host.argtypes = [vp, vp, vp, C.POINTER(QuantCfg), f32, f32, C.c_int, C.c_int, vp]
    L.ccb_arm_int.argtypes = [vp, vp, C.POINTER(vp), C.POINTER(vp), vp, vp, vp, vp, vp, vp]
    L.ccb_fp32_peak.argtypes = [C.c_int, C.c_int, C.POINTER(f32)]
    L.ccb_dec_int_workspace_bytes.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(DecLayer)]
    L.ccb_dec_int_workspace_bytes.restype = C.c_size_t
    L.ccb_dec_int.argtypes = [C.c_int, C.c_int, C.c_int, vp, C.c_int, C.c_int, vp, C.c_int, C.c_int, vp, C.c_int,
                              C.POINTER(DecLayer), vp, C.c_size_t, vp, vp, vp, vp]
    L.ccb_dec_int.restype = C.c_int
    for name in ("ccb_plan_create", "ccb_plan_latent_geometry", "ccb_forward", "ccb_backward", "ccb_loss",
                 "ccb_clip_adam_step", "ccb_train_step", "ccb_train_step_timed", "ccb_overfit_host", "ccb_arm_int", "ccb_fp32_peak",
                 "ccb_weight_slot_acquire", "ccb_weight_slot_release"):
        getattr(L, name).restype = C.c_int
    _lib = L
    return L


def check(rc: int, what: str) -> None:
    if rc != 0:
        L = lib()
        msg = L.ccb_strerror(rc).decode()
        if rc == -3:
            msg += ": " + L.ccb_last_cuda_error().decode()
        raise CcbError(f"{what} failed ({rc}): {msg}")


def parse_synthesis_layers(layers, n_out_default: int = 3):
    """'48-1-linear-relu' -> (out, k, residual, relu); 'X' = number of output channels
    (core/synthesis.py:238-260)."""
    out = []
    for spec in layers:
        o, k, mode, nl = spec.split("-")
        if mode not in ("linear", "residual") or nl not in ("none", "relu"):
            raise ValueError(f"Unknown synthesis layer description: {spec}")
        out.append((n_out_default if o == "X" else int(o), int(k), int(mode == "residual"), int(nl == "relu")))
    return out


def make_arch(B, H, W, n_latents, dim_arm, n_hidden_arm, layers_synthesis, ups_k, ups_pre_k) -> Arch:
    a = Arch()
    a.B, a.H, a.W, a.n_latents = int(B), int(H), int(W), int(n_latents)
    a.dim_arm, a.n_hidden_arm, a.ups_k, a.ups_pre_k = int(dim_arm), int(n_hidden_arm), int(ups_k), int(ups_pre_k)
    parsed = parse_synthesis_layers(layers_synthesis)
    if len(parsed) > CCB_MAX_SYN_LAYERS:
        raise CcbError("too many synthesis layers")
    a.n_syn_layers = len(parsed)
    for i, (o, k, r, nl) in enumerate(parsed):
        a.syn[i] = SynLayer(o, k, r, nl)
    return a


class Plan:
    """Owns one ccb_plan handle.  Cached per (arch, device) by `get_plan`."""

    def __init__(self, arch: Arch, device: int):
        self.arch = arch
        self.device = device
        h = C.c_void_p()
        check(lib().ccb_plan_create(C.byref(h), C.byref(arch), device), "ccb_plan_create")
        self.handle = h
        L = arch.n_latents
        hh, ww, oo = (C.c_int32 * CCB_MAX_LATENTS)(), (C.c_int32 * CCB_MAX_LATENTS)(), (C.c_int64 * CCB_MAX_LATENTS)()
        check(lib().ccb_plan_latent_geometry(h, hh, ww, oo), "ccb_plan_latent_geometry")
        self.latent_hw = [(hh[i], ww[i]) for i in range(L)]
        self.latent_off = [oo[i] for i in range(L)]
        self.M = int(lib().ccb_plan_n_latent_values(h))
        self.ws_bytes = int(lib().ccb_plan_workspace_bytes(h))
        self.n_weight_params = int(lib().ccb_plan_n_weight_params(h))
        self.n_param_floats = int(lib().ccb_plan_n_param_floats(h))

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                lib().ccb_plan_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


_plans = {}


def get_plan(B, H, W, n_latents, dim_arm, n_hidden_arm, layers_synthesis, ups_k, ups_pre_k, device: int) -> Plan:
    key = (B, H, W, n_latents, dim_arm, n_hidden_arm, tuple(layers_synthesis), ups_k, ups_pre_k, device)
    p = _plans.get(key)
    if p is None:
        p = Plan(make_arch(B, H, W, n_latents, dim_arm, n_hidden_arm, layers_synthesis, ups_k, ups_pre_k), device)
        _plans[key] = p
    return p


class WeightSlot:
    """One constant-bank weight slot of a device (include/ccb.h: ccb_weight_slot_acquire), held for the lifetime of
    the object.  `index` is -1 when every slot of the device is taken: the fused iteration then runs the kernels that
    read the weights from the tensors (same results)."""

    def __init__(self, device: int, wait_s: float = 0.0):
        """wait_s > 0: when every slot is taken, retry for up to that long before settling for -1 (train() does:
        with more trainers alive than slots, a trainer then starts when another one finishes instead of running the
        other kernels, whose results differ in the last bits)."""
        import time

        self.device = int(device)
        slot = C.c_int(-1)
        deadline = time.monotonic() + max(0.0, wait_s)
        while True:
            rc = lib().ccb_weight_slot_acquire(self.device, C.byref(slot))
            if rc == 0 or time.monotonic() >= deadline:
                break
            time.sleep(0.002)
        self.index = int(slot.value) if rc == 0 else -1

    def release(self) -> None:
        if getattr(self, "index", -1) >= 0:
            try:
                lib().ccb_weight_slot_release(self.device, self.index)
            except Exception:
                pass
            self.index = -1

    def __del__(self):
        import sys

        if not sys.is_finalizing():  # at interpreter exit the CUDA runtime may already be gone
            self.release()
