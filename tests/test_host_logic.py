"""CPU tests of the host-side mirror of the reference's module API (no kernels run here)."""

import io

import numpy as np
import pytest
import torch

from common import Golden
from coolchic_b200.enc.component.coolchic import CoolChicEncoder, CoolChicEncoderParameter
from coolchic_b200.enc.component.frame import FrameEncoder, load_frame_encoder
from coolchic_b200.enc.training.loss import loss_function
from coolchic_b200.enc.training.train import _cosine_lr, _linear_schedule
from coolchic_b200 import _native as N

HOP_SYN = ["48-1-linear-relu", "X-1-linear-none", "X-3-residual-relu", "X-3-residual-none"]


def hop_param(H=40, W=56):
    p = CoolChicEncoderParameter(layers_synthesis=list(HOP_SYN), n_ft_per_res=[1] * 7, dim_arm=16,
                                 n_hidden_layers_arm=2, ups_k_size=8, ups_preconcat_k_size=7, encoder_gain=16)
    p.set_image_size((H, W))
    return p


def test_state_dict_names_and_shapes_match_the_reference():
    g = Golden("hop_sr_gauss_40x56")
    ref = g.tensors("p0/")
    enc = CoolChicEncoder(hop_param(g.H, g.W))
    sd = enc.state_dict()
    ours = {k: tuple(v.shape) for k, v in sd.items() if "non_zero_pixel_ctx_index" not in k}
    theirs = {k: tuple(v.shape) for k, v in ref.items()}
    assert set(ours) == set(theirs), set(ours) ^ set(theirs)
    for k in ours:
        assert ours[k] == theirs[k], (k, ours[k], theirs[k])
    enc.set_param(type(sd)((k, ref.get(k, v)) for k, v in sd.items()))
    got = enc.get_param()
    for k, v in ref.items():
        assert torch.equal(got[k], v), k


def test_latent_pyramid_uses_ceil():
    enc = CoolChicEncoder(hop_param(135, 241))
    shapes = [tuple(g.data.shape[-2:]) for g in enc.latent_grids]
    assert shapes == [(135, 241), (68, 121), (34, 61), (17, 31), (9, 16), (5, 8), (3, 4)]


def test_frame_encoder_save_load_round_trip():
    fe = FrameEncoder(hop_param(24, 40))
    with torch.no_grad():
        for p in fe.parameters():
            p.add_(torch.randn_like(p) * 0.01)
    blob = fe.save()
    blob.seek(0)
    fe2 = load_frame_encoder(io.BytesIO(blob.read()))
    a, b = fe.get_param(), fe2.get_param()
    assert list(a) == list(b)
    for k in a:
        assert torch.equal(a[k], b[k]), k


def test_forward_without_cuda_fails_loudly():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    fe = FrameEncoder(hop_param(24, 40))
    with pytest.raises((N.CcbError, RuntimeError, AssertionError)):
        fe.forward()


def test_yuv420_is_refused():
    with pytest.raises(NotImplementedError):
        FrameEncoder(hop_param(24, 40), frame_data_type="yuv420")


def test_loss_function_matches_definition():
    torch.manual_seed(0)
    x, y = torch.rand(1, 3, 8, 12), torch.rand(1, 3, 8, 12)
    rate = torch.rand(1, 50) * 4
    out = loss_function(x, rate, y, lmbda=1e-2, compute_logs=True)
    mse = ((x - y) ** 2).mean().item()
    bpp = rate.sum().item() / (8 * 12)
    assert abs(out.loss.item() - (mse + 1e-2 * bpp)) < 1e-6
    assert abs(out.psnr_db - (-10 * np.log10(mse))) < 1e-4 and abs(out.rate_latent_bpp - bpp) < 1e-5
    with pytest.raises(AssertionError):
        loss_function(x, rate[0], y)


def test_schedules_match_torch():
    assert _linear_schedule(0.3, 0.1, 0, 100) == 0.3 and abs(_linear_schedule(0.3, 0.1, 100, 100) - 0.1) < 1e-12
    with pytest.raises(AssertionError):
        _linear_schedule(0.3, 0.1, 101, 100)
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.Adam([p], lr=1e-2)
    sch = torch.optim.lr_scheduler.CosineAnnealingLR(opt, T_max=10000 / 100, eta_min=1e-5)
    for n in range(1, 101):
        opt.step()
        sch.step()
        assert abs(sch.get_last_lr()[0] - _cosine_lr(1e-2, 1e-5, n, 100.0)) < 1e-9
