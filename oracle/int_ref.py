"""TEST INFRASTRUCTURE ONLY -- never imported by coolchic_b200.

ctypes front-ends for
  * oracle/_lib/libarmint_ref.so   (oracle/armint_ref.c: the C restatement of the fixed-point ARM, rows I1/I2)
  * oracle/_ref/libccref.so        (the reference's own decoder sources compiled in place + oracle/ref_driver.cpp)
  * oracle/_ref/ccencapi*.so       (the reference's pybind11 entropy encoder, unmodified)
plus the float->fixed-point weight conversion of ArmInt.set_param_from_float
(coolchic/enc/bitstream/armint.py:190-209).

The `_ref` libraries exist only where oracle/build.sh found the reference checkout; `have_ref()` says so.
"""

from __future__ import annotations

import ctypes as C
import glob
import os
import sys
from typing import List, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_lib", "libarmint_ref.so")
_REF = os.path.join(_HERE, "_ref", "libccref.so")

ARM_PRECISION = 8
ARM_SCALE = 1 << ARM_PRECISION

_lib = None
_ref = None

_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            raise RuntimeError(f"{_LIB} missing: run `bash oracle/build.sh`")
        L = C.CDLL(_LIB)
        L.arm_int_grid.argtypes = [_i32p, C.c_int, C.c_int, C.c_int, C.c_int, _i32p, _i32p, _i32p, _i32p, _i32p, _i32p]
        L.arm_int_grid.restype = C.c_int
        L.mu_sig_indices.argtypes = [_i32p, _i32p, C.c_long, _i32p, _i32p, _i32p]
        L.mu_sig_indices.restype = None
        _lib = L
    return _lib


def have_ref() -> bool:
    return os.path.exists(_REF)


def ref() -> C.CDLL:
    global _ref
    if _ref is None:
        if not have_ref():
            raise RuntimeError(f"{_REF} missing: oracle/build.sh builds it only where /root/reference exists")
        L = C.CDLL(_REF)
        L.ccref_open.argtypes = [C.c_char_p]
        L.ccref_open.restype = C.c_void_p
        L.ccref_close.argtypes = [C.c_void_p]
        L.ccref_close.restype = None
        L.ccref_header.argtypes = [C.c_void_p, _i32p]
        L.ccref_layer_dims.argtypes = [C.c_void_p, C.c_int, _i32p, C.POINTER(C.c_int)]
        L.ccref_latent.argtypes = [C.c_void_p, C.c_int, _i32p]
        L.ccref_layer_bytes.argtypes = [C.c_void_p, C.c_int, _u8p]
        L.ccref_arm_weights.argtypes = [C.c_void_p, _i32p, _i32p, _i32p, _i32p]
        L.ccref_decode_layer.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _i32p, _i32p, _i32p, _i32p,
                                         C.c_int, _i32p]
        L.ccref_mu_indices.argtypes = [_i32p, _i32p, C.c_int, _i32p, _i32p, _i32p]
        L.ccref_dec_header.argtypes = [C.c_void_p, _i32p]
        L.ccref_ups_weights.argtypes = [C.c_void_p, _i32p, _i32p]
        L.ccref_syn_weights.argtypes = [C.c_void_p, C.c_int, C.c_int, _i32p, C.c_int, _i32p]
        L.ccref_run_ups.argtypes = [C.c_void_p, _i32p]
        L.ccref_run_syn.argtypes = [C.c_void_p, _i32p]
        _ref = L
    return _ref


def ccencapi():
    """The reference's pybind11 encoder module (cpp/ccencapi.cpp), built unmodified into oracle/_ref."""
    d = os.path.join(_HERE, "_ref")
    if not glob.glob(os.path.join(d, "ccencapi*.so")):
        raise RuntimeError("oracle/_ref/ccencapi*.so missing")
    if d not in sys.path:
        sys.path.insert(0, d)
    import ccencapi as m  # type: ignore

    return m


# ---------------------------------------------------------------------------------------------
# restatement front-end
# ---------------------------------------------------------------------------------------------
def quantise_arm_weights(ws: Sequence[np.ndarray], bs: Sequence[np.ndarray]) -> Tuple[List[np.ndarray], List[np.ndarray]]:
    """Float ARM weights -> fixed point, as ArmInt.set_param_from_float (armint.py:190-209):
    W*256 and b*65536, rounded half away from zero by `trunc(v + sign(v)*0.5)`, stored as int32."""
    def fx(v, scale):
        v = np.asarray(v, dtype=np.float32) * np.float32(scale)
        v = v + np.sign(v).astype(np.float32) * np.float32(0.5)
        return np.trunc(v).astype(np.int32)

    return [fx(w, ARM_SCALE) for w in ws], [fx(b, ARM_SCALE * ARM_SCALE) for b in bs]


def arm_int_grid(y: np.ndarray, w_q: Sequence[np.ndarray], b_q: Sequence[np.ndarray]) -> Tuple[np.ndarray, np.ndarray]:
    """y: [H,W] int32 latents; w_q: n_hidden x [C,C] (out,in) then [2,C]; b_q likewise.  -> (mu_q8, ls_q8) [H,W]."""
    y = np.ascontiguousarray(y, dtype=np.int32)
    H, W = y.shape
    nh = len(w_q) - 1
    Cc = w_q[-1].shape[1]
    wh = np.ascontiguousarray(np.stack([np.asarray(w, np.int32) for w in w_q[:-1]]) if nh else np.zeros((1,), np.int32))
    bh = np.ascontiguousarray(np.stack([np.asarray(b, np.int32) for b in b_q[:-1]]) if nh else np.zeros((1,), np.int32))
    wo = np.ascontiguousarray(w_q[-1], dtype=np.int32)
    bo = np.ascontiguousarray(b_q[-1], dtype=np.int32)
    mu = np.empty((H, W), np.int32)
    ls = np.empty((H, W), np.int32)
    rc = lib().arm_int_grid(y, H, W, Cc, nh, wh.reshape(-1), bh.reshape(-1), wo.reshape(-1), bo.reshape(-1),
                            mu.reshape(-1), ls.reshape(-1))
    if rc != 0:
        raise ValueError(f"arm_int_grid: unsupported configuration C={Cc} nh={nh}")
    return mu, ls


def mu_sig_indices(mu: np.ndarray, ls: np.ndarray):
    mu = np.ascontiguousarray(mu, dtype=np.int32).reshape(-1)
    ls = np.ascontiguousarray(ls, dtype=np.int32).reshape(-1)
    a, b, c = (np.empty_like(mu) for _ in range(3))
    lib().mu_sig_indices(mu, ls, mu.size, a, b, c)
    return a, b, c


def arm_int_pyramid(grids: Sequence[np.ndarray], w_q, b_q):
    """All grids of one image, flattened in the plan's [M] order (finest first)."""
    mus, lss = [], []
    for g in grids:
        m, l = arm_int_grid(g, w_q, b_q)
        mus.append(m.reshape(-1))
        lss.append(l.reshape(-1))
    return np.concatenate(mus), np.concatenate(lss)


# ---------------------------------------------------------------------------------------------
# compiled-reference front-end
# ---------------------------------------------------------------------------------------------
class CoolFile:
    """A shipped .cool bitstream decoded by the reference's own C++ up to and including the ARM stage."""

    def __init__(self, path: str):
        L = ref()
        self.h = L.ccref_open(path.encode())
        if not self.h:
            raise RuntimeError(f"reference decoder could not open {path}")
        hd = np.zeros(8, np.int32)
        L.ccref_header(self.h, hd)
        (self.img_h, self.img_w, self.n_latents, self.dim_arm, self.n_hidden, self.hls_sig_blksize,
         self.ac_max_val_latent, self.n_2d_grid) = (int(v) for v in hd)
        self.grids, self.layer_bytes, self.zero_layer = [], [], []
        for l in range(self.n_latents):
            hw = np.zeros(2, np.int32)
            nb = C.c_int(0)
            L.ccref_layer_dims(self.h, l, hw, C.byref(nb))
            g = np.zeros((int(hw[0]), int(hw[1])), np.int32)
            z = L.ccref_latent(self.h, l, g.reshape(-1))
            buf = np.zeros(max(nb.value, 1), np.uint8)
            n = L.ccref_layer_bytes(self.h, l, buf)
            self.grids.append(g)
            self.layer_bytes.append(bytes(buf[:n]))
            self.zero_layer.append(bool(z))
        Cc, nh = self.dim_arm, self.n_hidden
        wt = np.zeros((max(nh, 1), Cc, Cc), np.int32)
        bh = np.zeros((max(nh, 1), Cc), np.int32)
        wo = np.zeros((2, Cc), np.int32)
        bo = np.zeros(2, np.int32)
        L.ccref_arm_weights(self.h, wt.reshape(-1), bh.reshape(-1), wo.reshape(-1), bo)
        self.w_hidden_t = wt[:nh]  # [in][out], as the C++ kernel consumes it
        # (out,in) order, as ArmIntLinear / our kernels take it
        self.w_q = [np.ascontiguousarray(wt[l].T) for l in range(nh)] + [wo]
        self.b_q = [bh[l].copy() for l in range(nh)] + [bo]

    def decoder_params(self) -> dict:
        """Integer upsampling filters and synthesis layers as the reference decoder holds them after
        read_ups / read_syn (cc-frame-decoder.cpp:254-357)."""
        L = ref()
        hd = np.zeros(6 + 4 * 8, np.int32)
        L.ccref_dec_header(self.h, hd)
        k, kp, n_ups, n_pre, n_syn, n_br = (int(v) for v in hd[:6])
        ups = np.zeros((n_ups, k), np.int32)
        pre = np.zeros((n_pre, kp), np.int32)
        L.ccref_ups_weights(self.h, ups.reshape(-1), pre.reshape(-1))
        layers, n_in = [], self.n_latents
        for i in range(n_syn):
            n_out, ks, res, relu = (int(v) for v in hd[6 + 4 * i:10 + 4 * i])
            w = np.zeros((n_out, n_in, ks, ks), np.int32)
            b = np.zeros(n_out, np.int32)
            L.ccref_syn_weights(self.h, i, w.size, w.reshape(-1), b.size, b)
            layers.append(dict(w=w, b=b, ks=ks, residual=bool(res), relu=bool(relu)))
            n_in = n_out
        return dict(ups_w=ups, pre_w=pre, syn=layers, n_branches=n_br)

    def run_ups_syn(self):
        """The reference's own run_ups + run_syn on this frame -> (ups planes [n,H,W], output [3,H,W]) int32."""
        L = ref()
        u = np.zeros((self.n_latents, self.img_h, self.img_w), np.int32)
        L.ccref_run_ups(self.h, u.reshape(-1))
        o = np.zeros((3, self.img_h, self.img_w), np.int32)
        L.ccref_run_syn(self.h, o.reshape(-1))
        return u, o

    def close(self):
        if self.h:
            ref().ccref_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def ref_decode_layer(data: bytes, H: int, W: int, w_q, b_q, hls_sig_blksize: int) -> np.ndarray:
    """Reference sequential ARM + CABAC decode (arm_cpu.cpp) of one sub-stream."""
    nh = len(w_q) - 1
    Cc = w_q[-1].shape[1]
    wt = np.ascontiguousarray(np.stack([np.asarray(w, np.int32).T for w in w_q[:-1]]) if nh else np.zeros((1,), np.int32))
    bh = np.ascontiguousarray(np.stack([np.asarray(b, np.int32) for b in b_q[:-1]]) if nh else np.zeros((1,), np.int32))
    out = np.zeros((H, W), np.int32)
    buf = np.frombuffer(data, dtype=np.uint8).copy()
    ref().ccref_decode_layer(buf, buf.size, H, W, Cc, nh, wt.reshape(-1), bh.reshape(-1),
                             np.ascontiguousarray(w_q[-1], np.int32).reshape(-1), np.ascontiguousarray(b_q[-1], np.int32),
                             hls_sig_blksize, out.reshape(-1))
    return out


def ref_encode_layer(x: np.ndarray, mu: np.ndarray, ls: np.ndarray, hls_sig_blksize: int, tmp_path: str) -> bytes:
    """Reference entropy encoder (ccencapi.cpp: cc_code_latent_layer_bac) -> sub-stream bytes."""
    H, W = x.shape
    m = ccencapi()
    m.cc_code_latent_layer_bac(tmp_path, [int(v) for v in x.reshape(-1)], [int(v) for v in np.asarray(mu).reshape(-1)],
                               [int(v) for v in np.asarray(ls).reshape(-1)], H, W, hls_sig_blksize)
    with open(tmp_path, "rb") as f:
        return f.read()
