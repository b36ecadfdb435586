"""Encoder-side emulation of the fixed-point DECODER: the exact image `ccdecapi` will reconstruct from
the bitstream this encoder would write, computed on the GPU from the encoder's own parameters.

The reference has no such function: its encoder measures PSNR on the float forward and learns the true decoded image
only by running the C++ decoder on the written file.  The integers are built the way the bitstream carries them
(enc/bitstream/encode.py + cpp/cc-frame-decoder.cpp:157-199,254-357): latents `round(y * encoder_gain)`, filters and
synthesis weights at 12 fractional bits, synthesis biases at 24, symmetric filters stored as their first half and
mirrored.  After quantize_model() every parameter is a multiple of a power-of-two step >= 2^-12 (2^-24 for biases), so
the conversion is exact; on a model that has not been quantised it rounds to the decoder's grid.
"""

from __future__ import annotations

from typing import Any, Dict, Optional

import torch
from torch import Tensor

from ... import ops
from ..component.core.synthesis import SynthesisConv2d
from ..component.core.upsampling import symmetric_kernel_1d

UPS_PRECISION = 12
SYN_WEIGHT_PRECISION = 12


def _fixed(t: Tensor, bits: int) -> Tensor:
    return torch.round(t.detach().double() * float(1 << bits)).to(torch.int32)


@torch.no_grad()
def integer_decoder_params(cc: Any) -> Dict[str, Any]:
    """ups_w [n_ups, k] / pre_w [n_pre, kp] (CPU int32) and the synthesis layers (CUDA int32) of a CoolChicEncoder."""
    up = cc.upsampling
    ups_w = torch.stack([_fixed(symmetric_kernel_1d(m.parametrizations.weight.original, m.target_k_size), UPS_PRECISION)
                         for m in up.conv_transpose2ds]).cpu()
    pre_w = torch.stack([_fixed(symmetric_kernel_1d(m.parametrizations.weight.original, m.target_k_size), UPS_PRECISION)
                         for m in up.conv2ds]).cpu()
    layers = []
    children = list(cc.synthesis.layers.children())
    for i, m in enumerate(children):
        if not isinstance(m, SynthesisConv2d):
            continue
        relu = i + 1 < len(children) and isinstance(children[i + 1], torch.nn.ReLU)
        layers.append(dict(w=_fixed(m.weight, SYN_WEIGHT_PRECISION).contiguous(),
                           b=_fixed(m.bias, 2 * SYN_WEIGHT_PRECISION).contiguous(), ks=int(m.weight.shape[-1]),
                           residual=bool(m.residual), relu=bool(relu)))
    return dict(ups_w=ups_w, pre_w=pre_w, syn=layers)


@torch.no_grad()
def integer_latents(cc: Any) -> Tensor:
    """The symbols the bitstream carries: round(latent * encoder_gain) (torch.round, as the eval-mode quantiser),
    flat, finest level first, int32."""
    flat = torch.cat([g.data.reshape(-1) for g in cc.latent_grids])
    return torch.round(flat.detach() * float(cc.encoder_gains)).to(torch.int32)


@torch.no_grad()
def decoder_image(frame_encoder: Any, target: Optional[Tensor] = None) -> Dict[str, Any]:
    """-> dict(pixels uint8 [H, W, 3], image float [1, 3, H, W] in [0, 1], out int32 [3, H, W] at 12 bits
    [, mse, psnr_db against `target` ([1, 3, H, W] in [0, 1])]) -- what the C++ decoder will output."""
    cc = frame_encoder.coolchic_encoder if hasattr(frame_encoder, "coolchic_encoder") else frame_encoder
    prm = integer_decoder_params(cc)
    lat = integer_latents(cc)
    if int(cc.latent_grids[0].data.shape[0]) != 1:
        raise ValueError("decoder_image(): one image at a time (the bitstream holds one frame)")
    H, W = (int(v) for v in cc.param.img_size)
    r = ops.decode_int(lat, (H, W), len(cc.latent_grids), prm["ups_w"], prm["pre_w"], prm["syn"], want_pixels=True)
    img = r["pixels"].permute(2, 0, 1)[None].float() / 255.0
    out = {"pixels": r["pixels"], "image": img, "out": r["out"]}
    if target is not None:
        mse = float(((img - target.to(img.device)) ** 2).mean())
        out["mse"] = mse
        out["psnr_db"] = float("inf") if mse == 0 else -10.0 * torch.log10(torch.tensor(mse)).item()
    return out
