"""GPU parity of the fixed-point decoder's upsampling + synthesis (ccb_dec_int, through the C ABI): bit-exact against
the reference's own C++ on shipped bitstreams (digests in tests/golden/int_dec_*.npz: three Kodak images and two CLIC
images with odd height / odd width) and against the numpy oracle on random small, ragged cases (every supported filter
length, fused and unfused synthesis stacks, 2..7 latent levels)."""

import numpy as np
import pytest
import torch

from dec_common import dec_cases, load_dec, random_case, sha
from coolchic_b200 import ops
from oracle import dec_int_ref as D

pytestmark = pytest.mark.gpu


def run_gpu(grids, prm, H, W):
    dev = torch.device("cuda:0")
    lat = torch.from_numpy(np.concatenate([g.reshape(-1) for g in grids]).astype(np.int32)).to(dev)
    layers = [dict(w=torch.from_numpy(l["w"].astype(np.int32)).to(dev), b=torch.from_numpy(l["b"].astype(np.int32)).to(dev),
                   ks=l["ks"], residual=l["residual"], relu=l["relu"]) for l in prm["syn"]]
    r = ops.decode_int(lat, (H, W), len(grids), prm["ups_w"], prm["pre_w"], layers, want_ups=True,
                       want_pixels=prm["syn"][-1]["w"].shape[0] == 3)
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in r.items() if not k.startswith("_")}


@pytest.mark.parametrize("name", dec_cases())
def test_shipped_bitstreams_bit_exact(name):
    z, grids, prm = load_dec(name)
    H, W = (int(v) for v in z["img_hw"])
    r = run_gpu(grids, prm, H, W)
    assert np.array_equal(r["ups"][:, ::37, ::41], z["sample_ups"])
    assert sha(r["ups"]) == str(z["sha_ups"])
    assert np.array_equal(r["out"][:, ::37, ::41], z["sample_out"])
    assert sha(r["out"]) == str(z["sha_out"])
    assert sha(r["pixels"], np.uint8) == str(z["sha_pix"])


UNFUSED = ((6, 1, 0, 1), (3, 3, 0, 0), (3, 3, 1, 1), (3, 1, 1, 0))
CASES = [
    # seed, H, W, n_latents, ups_k, pre_k, synthesis
    (1, 37, 53, 5, 8, 7, None),
    (2, 64, 96, 7, 8, 7, None),
    (3, 19, 35, 3, 4, 3, None),
    (4, 41, 29, 4, 4, 5, UNFUSED),
    (5, 8, 12, 2, 8, 7, ((8, 1, 0, 1), (3, 1, 0, 0))),
    (6, 135, 241, 7, 8, 7, UNFUSED),
    (7, 1, 1, 2, 8, 7, None),
    (8, 33, 65, 6, 8, 1, ((40, 1, 0, 1), (3, 1, 0, 0), (3, 3, 1, 0))),
]


@pytest.mark.parametrize("seed,H,W,n,uk,pk,syn", CASES)
def test_random_cases_match_the_oracle(seed, H, W, n, uk, pk, syn):
    kw = {} if syn is None else {"syn": syn}
    grids, prm = random_case(seed, H, W, n, ups_k=uk, pre_k=pk, **kw)
    ref = D.decode(grids, prm)
    r = run_gpu(grids, prm, H, W)
    assert np.array_equal(r["ups"], ref["ups"])
    assert np.array_equal(r["out"], ref["out"])
    assert np.array_equal(r["pixels"], ref["pixels"])


def test_unsupported_filter_lengths_are_refused():
    grids, prm = random_case(9, 16, 16, 3)
    prm["ups_w"] = prm["ups_w"][:, :6]
    from coolchic_b200._native import CcbError

    with pytest.raises(CcbError):
        run_gpu(grids, prm, 16, 16)
