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


def test_decoder_emulation_of_a_trained_encoder_matches_its_float_forward():
    """End to end on the modules: overfit a small image, quantise the networks, then the fixed-point decoder image
    (what ccdecapi would output) must sit within 0.1 dB of the encoder's own eval-mode float image against the
    target -- the gate the reference applies between its encoder logs and its decoded files."""
    import math
    import types

    from common import O
    from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
    from coolchic_b200.enc.component.frame import FrameEncoder
    from coolchic_b200.enc.training.quantizemodel import quantize_model
    from coolchic_b200.enc.training.train import train
    from coolchic_b200.enc.utils.decoder_int import decoder_image

    dev = torch.device("cuda:0")
    H, W = 96, 136
    torch.manual_seed(77)
    p = CoolChicEncoderParameter(layers_synthesis=["16-1-linear-relu", "X-1-linear-none", "X-3-residual-relu",
                                                   "X-3-residual-none"], n_ft_per_res=[1] * 7, dim_arm=8,
                                 n_hidden_layers_arm=2, ups_k_size=8, ups_preconcat_k_size=7, encoder_gain=16)
    p.set_image_size((H, W))
    fe = FrameEncoder(p)
    fe.to_device(dev)
    target = O.synthetic_image(21, H, W).to(dev)
    mgr = types.SimpleNamespace(lmbda=1e-3, iterations_counter=0, total_training_time_sec=0.0)
    train(fe, target, mgr, start_lr=1e-2, max_iterations=600, frequency_validation=100, patience=100000, seed=5)
    train(fe, target, mgr, start_lr=1e-4, max_iterations=100, frequency_validation=50, patience=100000,
          quantizer_type="ste", quantizer_noise_type="none", softround_temperature=(1e-4, 1e-4), seed=6)
    quantize_model(fe, target, mgr)
    fe.set_to_eval()
    with torch.no_grad():
        o = fe.forward(quantizer_noise_type="none", quantizer_type="hardround", AC_MAX_VAL=-1)
    psnr_float = -10.0 * math.log10(float(((o.decoded_image - target) ** 2).mean()))
    d = decoder_image(fe, target)
    between = -10.0 * math.log10(max(float(((o.decoded_image - d["image"]) ** 2).mean()), 1e-12))
    assert psnr_float > 25.0, psnr_float
    assert abs(d["psnr_db"] - psnr_float) <= 0.1, (d["psnr_db"], psnr_float)
    assert between > psnr_float + 10.0, (between, psnr_float)
