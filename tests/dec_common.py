"""Shared loaders for the fixed-point decoder tests (tests/golden/int_dec_*.npz, made by make_dec_golden.py)."""

import glob
import hashlib
import os

import numpy as np

from common import GOLDEN_DIR


def dec_cases():
    return sorted(os.path.basename(f)[len("int_dec_"):-4] for f in glob.glob(os.path.join(GOLDEN_DIR, "int_dec_*.npz")))


def sha(a, dtype=np.int32):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=dtype).tobytes()).hexdigest()


def load_dec(name):
    """-> (z, grids [n] int32 arrays, params dict as oracle/dec_int_ref.py takes it)"""
    z = np.load(os.path.join(GOLDEN_DIR, f"int_dec_{name}.npz"), allow_pickle=False)
    n = int(z["n_latents"])
    src = z if "grid0" in z.files else np.load(os.path.join(GOLDEN_DIR, f"int_cool_{name}.npz"), allow_pickle=False)
    grids = [src[f"grid{k}"].astype(np.int32) for k in range(n)]
    layers = []
    for i, (n_out, ks, res, relu) in enumerate(z["syn_desc"].tolist()):
        layers.append(dict(w=z[f"syn{i}_w"], b=z[f"syn{i}_b"], ks=ks, residual=bool(res), relu=bool(relu)))
    return z, grids, dict(ups_w=z["ups_w"], pre_w=z["pre_w"], syn=layers)


def random_case(seed, H, W, n, ups_k=8, pre_k=7, syn=((16, 1, 0, 1), (3, 1, 0, 0), (3, 3, 1, 1), (3, 3, 1, 0))):
    """Random latents / filters / synthesis weights in the ranges real bitstreams show."""
    rng = np.random.default_rng(seed)
    grids, h, w = [], H, W
    for _ in range(n):
        grids.append(rng.integers(-6, 7, size=(h, w)).astype(np.int32))
        h, w = (h + 1) // 2, (w + 1) // 2

    def filt(count, taps, phases, gain):
        # random signs / shapes with a bounded absolute sum per phase (real filters have gain ~1..1.5 per pass),
        # so that the int32 sums of the pipeline cannot wrap
        k = rng.integers(-1000, 1001, size=(count, taps)).astype(np.float64)
        for ph in range(phases):
            k[:, ph::phases] *= gain * 4096.0 / np.maximum(np.abs(k[:, ph::phases]).sum(axis=1, keepdims=True), 1.0)
        return np.trunc(k).astype(np.int32)

    ups_w = filt(n - 1, ups_k, 2, 1.1)
    pre_w = filt(n - 1, pre_k, 1, 0.6)
    layers, n_in = [], n
    for n_out, ks, res, relu in syn:
        bound = max(2, 3000 // (n_in * ks * ks))
        layers.append(dict(w=rng.integers(-bound, bound + 1, size=(n_out, n_in, ks, ks)).astype(np.int32),
                           b=rng.integers(-(1 << 22), 1 << 22, size=(n_out,)).astype(np.int32), ks=ks,
                           residual=bool(res), relu=bool(relu)))
        n_in = n_out
    return grids, dict(ups_w=ups_w, pre_w=pre_w, syn=layers)
