"""GPU parity of the fixed-point ARM kernel (ccb_arm_int): BIT-EXACT against the
oracle (oracle/armint_ref.c), against ArmInt(pure_int) outputs generated from the reference, against the digests of
shipped .cool bitstreams, and -- where oracle/_ref travelled with the repo -- by feeding the kernel's (mu, log-scale)
to the reference's own entropy encoder and requiring the shipped sub-streams byte for byte."""

import hashlib
import os

import numpy as np
import pytest
import torch

from common import GOLDEN_DIR, int_golden_cases
from coolchic_b200 import ops
from oracle import int_ref as I

pytestmark = pytest.mark.gpu
SYN = ["16-1-linear-relu", "3-1-linear-none"]


def _sha(*arrs):
    h = hashlib.sha256()
    for a in arrs:
        h.update(np.ascontiguousarray(a, dtype=np.int32).tobytes())
    return h.hexdigest()


def pyramid_shapes(H, W, n=7):
    return [(-(-H // 2**i), -(-W // 2**i)) for i in range(n)]


def run_kernel(H, W, C, nh, grids_per_image, wq, bq):
    """grids_per_image: B lists of 7 int32 grids.  -> 5 numpy arrays [B, M]."""
    dev = torch.device("cuda:0")
    spec = ops.ArchSpec(H, W, C, nh, SYN)
    flat = np.stack([np.concatenate([g.reshape(-1) for g in grids]) for grids in grids_per_image]).astype(np.int32)
    y = torch.from_numpy(flat).to(dev)
    outs = ops.arm_int(spec, y, [torch.from_numpy(np.ascontiguousarray(w, np.int32)).to(dev) for w in wq],
                       [torch.from_numpy(np.ascontiguousarray(b, np.int32)).to(dev) for b in bq])
    torch.cuda.synchronize()
    return [o.cpu().numpy() for o in outs]


def oracle_all(grids, wq, bq):
    mu, ls = I.arm_int_pyramid(grids, wq, bq)
    return (mu, ls) + tuple(I.mu_sig_indices(mu, ls))


def random_grids(rng, H, W, lo=-12, hi=13):
    return [rng.integers(lo, hi, size=s).astype(np.int32) for s in pyramid_shapes(H, W)]


def weights_of(z):
    nh = int(z["nh"])
    return [z[f"wq{i}"] for i in range(nh + 1)], [z[f"bq{i}"] for i in range(nh + 1)]


@pytest.mark.parametrize("name", int_golden_cases("int_armint_"))
def test_armint_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    C, nh = int(z["C"]), int(z["nh"])
    wq, bq = weights_of(z)
    g0 = z["grid"].astype(np.int32)
    H, W = g0.shape
    rng = np.random.default_rng(3)
    grids = random_grids(rng, H, W)
    grids[0] = g0
    got = run_kernel(H, W, C, nh, [grids], wq, bq)
    n0 = H * W
    assert np.array_equal(got[0][0, :n0], z["mu_q8"])
    assert np.array_equal(got[1][0, :n0], z["ls_q8"])
    for a, b in zip(got, oracle_all(grids, wq, bq)):
        assert np.array_equal(a[0], b)


@pytest.mark.parametrize("name", int_golden_cases("int_cool_"))
def test_cool_fixture(name, tmp_path):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    C, nh = int(z["C"]), int(z["nh"])
    H, W = (int(v) for v in z["img_hw"])
    wq, bq = weights_of(z)
    grids = [z[f"grid{l}"].astype(np.int32) for l in range(int(z["n_latents"]))]
    mu, ls, mr, mi, si = (a[0] for a in run_kernel(H, W, C, nh, [grids], wq, bq))
    assert np.array_equal(mu[:4096], z["head_mu"]) and np.array_equal(ls[:4096], z["head_ls"])
    assert _sha(mu, ls) == str(z["sha_mu_ls"])
    assert _sha(mr, mi, si) == str(z["sha_indices"])
    if I.have_ref():
        # the kernel's outputs drive the reference's entropy encoder: shipped bytes must come back
        blk = int(z["hls_sig_blksize"])
        off = 0
        for l, g in enumerate(grids):
            n = g.size
            if n <= 128 * 192 or name == "int_cool_kodim02-lmbda-002":
                data = I.ref_encode_layer(g, mu[off:off + n].reshape(g.shape), ls[off:off + n].reshape(g.shape), blk,
                                          str(tmp_path / "l.bin"))
                assert data == z[f"bytes{l}"].tobytes(), (name, l)
            off += n


@pytest.mark.parametrize("C,nh,H,W,B", [(8, 0, 17, 19, 1), (8, 1, 33, 47, 2), (16, 2, 64, 96, 1), (16, 3, 37, 53, 3),
                                        (24, 2, 40, 56, 1), (32, 2, 135, 241, 1), (32, 3, 8, 8, 2), (16, 2, 1, 1, 1)])
def test_random_vs_oracle(C, nh, H, W, B):
    rng = np.random.default_rng(C * 100 + nh * 10 + B)
    ws = [np.eye(C, dtype=np.float32) * 0.2 + rng.normal(0, 0.1, (C, C)).astype(np.float32) for _ in range(nh)]
    ws.append(rng.normal(0, 0.08, (2, C)).astype(np.float32))
    bs = [rng.normal(0, 1.0, C).astype(np.float32) for _ in range(nh)] + [np.array([-0.3, 3.0], np.float32)]
    wq, bq = I.quantise_arm_weights(ws, bs)
    images = [random_grids(rng, H, W, -20, 21) for _ in range(B)]
    got = run_kernel(H, W, C, nh, images, wq, bq)
    for b in range(B):
        for a, e in zip(got, oracle_all(images[b], wq, bq)):
            assert np.array_equal(a[b], e)


def test_int32_wraparound():
    """Huge weights overflow int32 in every layer; the kernel must wrap exactly as the C code does."""
    rng = np.random.default_rng(9)
    C, nh, H, W = 16, 2, 24, 40
    wq = [rng.integers(-2**20, 2**20, size=(C, C)).astype(np.int32) for _ in range(nh)] + \
         [rng.integers(-2**20, 2**20, size=(2, C)).astype(np.int32)]
    bq = [rng.integers(-2**30, 2**30, size=C).astype(np.int32) for _ in range(nh)] + \
         [rng.integers(-2**30, 2**30, size=2).astype(np.int32)]
    grids = random_grids(rng, H, W, -100, 100)
    got = run_kernel(H, W, C, nh, [grids], wq, bq)
    exp = oracle_all(grids, wq, bq)
    assert np.array_equal(got[0][0], exp[0]) and np.array_equal(got[1][0], exp[1])


def test_full_size_properties():
    """768x512 and 1080p pyramids: shift invariance away from borders (a translated grid gives translated outputs)
    and agreement with the oracle on the finest grid."""
    rng = np.random.default_rng(11)
    C, nh = 16, 2
    ws = [np.eye(C, dtype=np.float32) * 0.2 + rng.normal(0, 0.1, (C, C)).astype(np.float32) for _ in range(nh)]
    ws.append(rng.normal(0, 0.08, (2, C)).astype(np.float32))
    bs = [rng.normal(0, 1.0, C).astype(np.float32) for _ in range(nh)] + [np.array([0.0, 3.0], np.float32)]
    wq, bq = I.quantise_arm_weights(ws, bs)
    for H, W in ((512, 768), (1080, 1920)):
        grids = random_grids(rng, H, W, -8, 9)
        shifted = [g.copy() for g in grids]
        shifted[0] = np.roll(grids[0], (5, 7), axis=(0, 1))
        a = run_kernel(H, W, C, nh, [grids], wq, bq)
        b = run_kernel(H, W, C, nh, [shifted], wq, bq)
        n0 = H * W
        for k in range(5):
            x = a[k][0, :n0].reshape(H, W)[4:-12, 4:-12]
            y = b[k][0, :n0].reshape(H, W)[9:-7, 11:-5]
            assert np.array_equal(x, y)
        if H == 512:
            mu, ls = I.arm_int_grid(grids[0], wq, bq)
            assert np.array_equal(a[0][0, :n0], mu.reshape(-1)) and np.array_equal(a[1][0, :n0], ls.reshape(-1))
