"""The numpy restatement of the fixed-point decoder's upsampling + synthesis (oracle/dec_int_ref.py) against the
reference's own C++: committed vectors (tests/golden/int_dec_*.npz: SHA-256 of the upsampled planes, of the synthesis
output and of the 8-bit pixels of shipped bitstreams) and, where the reference checkout exists, the live decoder."""

import os

import numpy as np
import pytest

from dec_common import dec_cases, load_dec, sha
from oracle import dec_int_ref as D
from oracle import int_ref as I

# the two largest images cost ~10 s each in numpy: the CPU suite keeps the Kodak-shaped ones and one ragged image
CPU_CASES = [c for c in dec_cases() if "davide" not in c]


@pytest.mark.parametrize("name", CPU_CASES)
def test_oracle_reproduces_the_reference_decoder(name):
    z, grids, prm = load_dec(name)
    r = D.decode(grids, prm)
    assert r["ups"].shape == (len(grids),) + tuple(z["img_hw"])
    assert np.array_equal(r["ups"][:, ::37, ::41], z["sample_ups"])
    assert np.array_equal(r["out"][:, ::37, ::41], z["sample_out"])
    assert sha(r["ups"]) == str(z["sha_ups"])
    assert sha(r["out"]) == str(z["sha_out"])
    assert sha(r["pixels"], np.uint8) == str(z["sha_pix"])


def test_rounding_is_toward_zero():
    v = np.array([-4097, -4096, -4095, -1, 0, 1, 4095, 4096, 8191], np.int64)
    assert D.trunc_shift(v, 12).tolist() == [-1, -1, 0, 0, 0, 0, 0, 1, 1]


def test_pixels_round_and_clamp():
    x = np.array([[[-5000, 0, 8]], [[2048, 4095, 4096]], [[4104, 9000, 2056]]], np.int32)  # [3,1,3]
    p = D.to_pixels(x)
    assert p.shape == (1, 3, 3) and p.dtype == np.uint8
    assert p[0, :, 0].tolist() == [0, 0, 0] and p[0, :, 1].tolist() == [128, 255, 255] and p[0, :, 2].tolist() == [255, 255, 128]


def test_zero_level_stays_zero_and_shapes_are_ragged():
    from dec_common import random_case

    grids, prm = random_case(3, 37, 53, 5)
    grids[2][:] = 0
    ups = D.run_ups(grids, prm["ups_w"], prm["pre_w"])
    assert ups.shape == (5, 37, 53) and not ups[2].any() and ups[1].any()


@pytest.mark.skipif(not I.have_ref(), reason="needs oracle/_ref/libccref.so (build container only)")
def test_oracle_against_the_live_reference():
    path = "/root/reference/results/image/kodak/bitstreams/kodim05-lmbda-0004.cool"
    if not os.path.exists(path):
        pytest.skip("reference bitstreams absent")
    f = I.CoolFile(path)
    ups, out = f.run_ups_syn()
    r = D.decode(f.grids, f.decoder_params())
    assert np.array_equal(r["ups"], ups) and np.array_equal(r["out"], out)
