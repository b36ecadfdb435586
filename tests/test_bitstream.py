"""Bitstream hand-off: the latent sub-streams written from the integer ARM kernel's outputs
(coolchic_b200/enc/bitstream/encode.py) must decode, with the REFERENCE's sequential ARM + CABAC decoder
(cpp/arm_cpu.cpp through oracle/_ref/libccref.so), to exactly the symbols that were sent -- any one-bit difference
between the kernel's (mu, log-scale) and the decoder's own desynchronises the arithmetic decoder.  The coder is the
reference's unmodified `cc_code_latent_layer_bac` (oracle/_ref/ccencapi*.so)."""

import os
import types

import numpy as np
import pytest
import torch

from coolchic_b200.enc.bitstream.encode import (encode_latents, get_ac_max_val_latent, get_sub_bitstream_path,
                                                integerise_arm_param, latent_layers_for_bac)
from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
from coolchic_b200.enc.component.frame import FrameEncoder
from coolchic_b200.enc.training.quantizemodel import quantize_model
from coolchic_b200.enc.training.train import train
from coolchic_b200.enc.utils.decoder_int import integer_latents
from oracle import float_ref as O
from oracle import int_ref as I

needs_ref = pytest.mark.skipif(not I.have_ref(), reason="oracle/_ref not built (no reference checkout where it was built)")


def test_integerise_arm_param_matches_the_oracle():
    """CPU: ArmInt.set_param_from_float restated twice (oracle/int_ref.py and the product) agree, halves included."""
    g = torch.Generator().manual_seed(0)
    ws = [torch.randn(16, 16, generator=g) * 0.3, torch.randn(2, 16, generator=g)]
    bs = [torch.randn(16, generator=g) * 0.1, torch.randn(2, generator=g)]
    ws[0][0, :4] = torch.tensor([0.5, -0.5, 1.5, -2.5]) / 256.0  # exact halves: away from zero
    param = {"mlp.0.weight": ws[0], "mlp.0.bias": bs[0], "mlp.2.weight": ws[1], "mlp.2.bias": bs[1]}
    wq, bq = integerise_arm_param(param)
    ow, ob = I.quantise_arm_weights([w.numpy() for w in ws], [b.numpy() for b in bs])
    for a, b in zip(wq + bq, ow + ob):
        assert a.dtype == torch.int32 and np.array_equal(a.numpy(), b)
    assert wq[0][0, :4].tolist() == [1, -1, 2, -3]


def _parse_tables(hdr: bytes, n_layers: int):
    """The variable fields of a frame header (layout: header.py:255-467), read back from shipped bytes."""
    p = 2 + 1 + 1 + 1 + 1 + 1 + 1 + 3 * n_layers
    rd = lambda n, signed=False: int.from_bytes(hdr[p:p + n], "big", signed=signed)  # noqa: E731
    flow_gain = rd(1); p += 1
    ac_nn = rd(2); p += 2
    ac_lat = rd(2); p += 2
    blk = rd(1, True); p += 1
    tables = []
    for width in (1, 1, 2):
        t = {}
        for nn_name in ("arm", "upsampling", "synthesis"):
            t[nn_name] = {}
            for kind in ("weight", "bias"):
                t[nn_name][kind] = rd(width); p += width
        tables.append(t)
    n_res = rd(1); p += 1
    n_2d = rd(1); p += 1 + n_res
    n_bytes = []
    for _ in range(n_2d):
        n_bytes.append(rd(3)); p += 3
    assert p == len(hdr)
    return flow_gain, ac_nn, ac_lat, blk, tables, n_bytes


@pytest.mark.parametrize("name", ["kodim01-lmbda-0001", "kodim02-lmbda-002", "kodim13-lmbda-00001",
                                  "amy-zhang-15940-lmbda-0004", "davide-ragusa-716-lmbda-0004"])
def test_header_writers_reproduce_shipped_headers(name):
    """CPU: gop_header_bytes / frame_header_bytes, given the architecture of a shipped bitstream and its tables, write
    the shipped header bytes (tests/golden/cool_headers.npz, copied from the reference's results/ by
    tests/golden/make_header_golden.py)."""
    from common import GOLDEN_DIR
    from coolchic_b200.enc.bitstream.encode import frame_header_bytes, gop_header_bytes

    raw = np.load(os.path.join(GOLDEN_DIR, "cool_headers.npz"))[name].tobytes()
    gop, frame = raw[:9], raw[9:]
    H, W = int.from_bytes(gop[2:4], "big"), int.from_bytes(gop[4:6], "big")
    # the architecture, read off the shipped header (it changes with the rate point: kodim02's ARM is 8 wide)
    dim_arm, nh = (frame[3] >> 4) * 8, frame[3] & 15
    ups_k, pre_k, n_layers = frame[4] & 15, frame[5] & 15, frame[7]
    assert frame[4] >> 4 == frame[5] >> 4 == 6 and frame[6] == 1
    syn = []
    for i in range(n_layers):
        out_ft, ks, flags = frame[8 + 3 * i:11 + 3 * i]
        syn.append(f"{out_ft}-{ks}-{['linear', 'residual'][flags >> 4]}-{['none', 'relu'][flags & 15]}")
    prm = CoolChicEncoderParameter(layers_synthesis=syn, n_ft_per_res=[1] * 7, dim_arm=dim_arm, n_hidden_layers_arm=nh,
                                   ups_k_size=ups_k, ups_preconcat_k_size=pre_k, encoder_gain=16)
    prm.set_image_size((H, W))
    fe = FrameEncoder(prm)  # on the CPU: only its description is read
    assert gop_header_bytes(fe) == gop
    flow_gain, ac_nn, ac_lat, blk, (q_idx, scale, n_bytes_nn), n_bytes_lat = _parse_tables(frame, len(syn))
    assert frame_header_bytes(fe, n_bytes_lat, q_idx, scale, n_bytes_nn, ac_nn, ac_lat, blk, display_order=0,
                              flow_gain=flow_gain) == frame


def _trained_encoder(H, W, dim_arm, nh, syn, iters, seed):
    dev = torch.device("cuda:0")
    torch.manual_seed(seed)
    prm = CoolChicEncoderParameter(layers_synthesis=list(syn), n_ft_per_res=[1] * 7, dim_arm=dim_arm,
                                   n_hidden_layers_arm=nh, ups_k_size=8, ups_preconcat_k_size=7, encoder_gain=16)
    prm.set_image_size((H, W))
    fe = FrameEncoder(prm)
    fe.to_device(dev)
    target = O.synthetic_image(seed, H, W).to(dev)
    mgr = types.SimpleNamespace(lmbda=1e-3, iterations_counter=0, total_training_time_sec=0.0)
    train(fe, target, mgr, max_iterations=iters, patience=iters, frequency_validation=50, seed=seed)
    quantize_model(fe, target, mgr)
    return fe


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("H,W,dim_arm,nh,syn,blk", [
    (72, 104, 16, 2, ["48-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"], 16),
    (72, 104, 16, 2, ["48-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"], -16),
    (57, 83, 8, 1, ["16-1-linear-relu", "3-1-linear-none"], 8),
    (40, 56, 24, 2, ["16-1-linear-relu", "3-1-linear-none"], 0),
])
def test_latent_substreams_decode_with_the_reference(H, W, dim_arm, nh, syn, blk, tmp_path):
    fe = _trained_encoder(H, W, dim_arm, nh, syn, iters=300, seed=H + dim_arm)
    cc = fe.coolchic_encoder
    sent = integer_latents(cc).cpu().numpy()
    assert np.abs(sent).max() > 0, "the training run left every latent at zero: nothing to code"
    root = str(tmp_path / "frame.cool")
    n_bytes = encode_latents(fe, root, blk, cc_code_latent_layer_bac=I.ccencapi().cc_code_latent_layer_bac)
    layers = latent_layers_for_bac(fe)
    assert len(n_bytes) == len(layers) == 7
    wq, bq = integerise_arm_param(cc.arm.get_param())
    wq = [w.cpu().numpy() for w in wq]
    bq = [b.cpu().numpy() for b in bq]
    off = 0
    for k, layer in enumerate(layers):
        h, w = layer["h"], layer["w"]
        y = sent[off:off + h * w].reshape(h, w)
        off += h * w
        assert np.array_equal(layer["y"].numpy().reshape(h, w), y)
        path = get_sub_bitstream_path(root, k)
        assert os.path.getsize(path) == n_bytes[k]
        if layer["all_zero"]:
            assert n_bytes[k] == 0 and not y.any()
            continue
        with open(path, "rb") as f:
            data = f.read()
        dec = I.ref_decode_layer(data, h, w, wq, bq, blk)
        assert np.array_equal(dec, y), f"latent level {k} ({h}x{w}) does not survive the reference's decoder"
        # ... and the kernel's (mu, log-scale) are the sequential oracle's, position by position
        mu, ls = I.arm_int_grid(y, wq, bq)
        assert np.array_equal(layer["mu"].numpy().reshape(h, w), mu)
        assert np.array_equal(layer["log_scale"].numpy().reshape(h, w), ls)
    assert off == sent.size
    assert get_ac_max_val_latent(fe) == int(np.abs(sent).max()) + 2


@pytest.mark.gpu
def test_encode_latents_fails_loudly_without_a_coder(tmp_path):
    """No CABAC coder in this package and no silent fallback: without CCLIB.ccencapi the call raises."""
    fe = _trained_encoder(40, 56, 8, 1, ["16-1-linear-relu", "3-1-linear-none"], iters=20, seed=3)
    try:
        import CCLIB.ccencapi  # noqa: F401
        pytest.skip("the reference's CCLIB is installed here")
    except ImportError:
        pass
    with pytest.raises(ImportError, match="CCLIB.ccencapi"):
        encode_latents(fe, str(tmp_path / "x.cool"), 16)


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("H,W,dim_arm,nh,syn,blk", [
    (72, 104, 16, 2, ["48-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"], 16),
    (57, 83, 8, 1, ["16-1-linear-relu", "3-1-linear-none"], -16),
    (128, 96, 24, 2, ["32-1-linear-relu", "3-1-linear-none", "3-3-residual-none"], 8),
    # BASELINE.json's full size (configs[1]): 524 256 symbols through the reference's CABAC and sequential decoder
    (512, 768, 16, 2, ["48-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"], 16),
])
def test_cool_file_decodes_with_the_reference_to_the_emulated_image(H, W, dim_arm, nh, syn, blk, tmp_path):
    """End to end: GPU-trained encoder -> quantize_model -> encode_image (headers + network + latent sub-streams, the
    reference's CABAC) -> the REFERENCE's decoder (oracle/_ref/libccref.so: header parsing, weight decoding, sequential
    ARM, fixed-point upsampling + synthesis).  What it reconstructs must be, bit for bit, what the encoder said it
    sent and what `decoder_image` (the CUDA emulation) predicted."""
    from coolchic_b200.enc.bitstream.encode import encode_image
    from coolchic_b200.enc.utils.decoder_int import decoder_image, integer_decoder_params
    from oracle import dec_int_ref as D

    fe = _trained_encoder(H, W, dim_arm, nh, syn, iters=300, seed=W + nh)
    cc = fe.coolchic_encoder
    before = decoder_image(fe)
    sent = integer_latents(cc).cpu().numpy()
    m = I.ccencapi()
    path = str(tmp_path / "image.cool")
    info = encode_image(fe, path, blk, cc_code_wb_bac=m.cc_code_wb_bac, cc_code_latent_layer_bac=m.cc_code_latent_layer_bac)
    assert info["n_bytes"] == os.path.getsize(path) and not [f for f in os.listdir(tmp_path) if f != "image.cool"]
    after = decoder_image(fe)  # the networks were replaced by their decoded values: a no-op after quantize_model
    assert torch.equal(before["out"], after["out"])

    f = I.CoolFile(path)
    assert (f.img_h, f.img_w, f.n_latents, f.dim_arm, f.n_hidden) == (H, W, 7, dim_arm, nh)
    assert f.hls_sig_blksize == blk and f.ac_max_val_latent == info["ac_max_val_latent"]
    off = 0
    for l, g in enumerate(f.grids):
        assert np.array_equal(g, sent[off:off + g.size].reshape(g.shape)), f"latent level {l}"
        assert len(f.layer_bytes[l]) == info["n_bytes_per_latent"][l] or f.zero_layer[l]
        off += g.size
    wq, bq = integerise_arm_param(cc.arm.get_param())
    for a, b in zip(wq + bq, f.w_q + f.b_q):
        assert np.array_equal(a.cpu().numpy().reshape(b.shape), b)
    ours, theirs = integer_decoder_params(cc), f.decoder_params()
    assert np.array_equal(ours["ups_w"].numpy(), theirs["ups_w"]) and np.array_equal(ours["pre_w"].numpy(), theirs["pre_w"])
    for a, b in zip(ours["syn"], theirs["syn"]):
        assert np.array_equal(a["w"].cpu().numpy(), b["w"]) and np.array_equal(a["b"].cpu().numpy(), b["b"])
        assert (a["ks"], a["residual"], a["relu"]) == (b["ks"], b["residual"], b["relu"])
    _, out = f.run_ups_syn()
    f.close()
    assert np.array_equal(out, after["out"].cpu().numpy()), "the reference decoder and decoder_image disagree"
    assert np.array_equal(D.to_pixels(out), after["pixels"].cpu().numpy())
