"""loss_function -- mirror of the reference's enc/training/loss.py:20-162 (rgb / yuv444 tensors).
Differentiable torch expression used by callers of the module API; the fused trainer computes the
same quantity inside csrc/ccb_opt.cu (ccb_loss)."""

from dataclasses import dataclass
from typing import Optional

import torch
from torch import Tensor


@dataclass
class LossFunctionOutput:
    loss: Optional[Tensor] = None
    mse: Optional[float] = None
    rate_nn_bpp: Optional[float] = None
    rate_latent_bpp: Optional[float] = None
    psnr_db: Optional[float] = None
    total_rate_bpp: Optional[float] = None

    def __post_init__(self):
        if self.mse is not None:
            import math

            self.psnr_db = -10.0 * math.log10(self.mse + 1e-10)  # loss.py:47 of the reference
        if self.rate_nn_bpp is not None and self.rate_latent_bpp is not None:
            self.total_rate_bpp = self.rate_nn_bpp + self.rate_latent_bpp


def loss_function(decoded_image: Tensor, rate_latent_bit: Tensor, target_image: Tensor, lmbda: float = 1e-3,
                  rate_mlp_bit: float = 0.0, compute_logs: bool = False) -> LossFunctionOutput:
    mse = ((decoded_image - target_image) ** 2).mean()
    n_pixels = decoded_image.size()[-2] * decoded_image.size()[-1]
    assert rate_latent_bit.ndim == 2, "Rate latent bit should have shape [batch, n_latent]."
    avg_rate_latent_bit = rate_latent_bit.sum(dim=1).mean(dim=0)
    rate_bpp = (avg_rate_latent_bit + rate_mlp_bit) / n_pixels
    loss = mse + lmbda * rate_bpp
    return LossFunctionOutput(
        loss=loss,
        mse=mse.detach().item() if compute_logs else None,
        rate_nn_bpp=float(rate_mlp_bit) / n_pixels if compute_logs else None,
        rate_latent_bpp=avg_rate_latent_bit.detach().item() / n_pixels if compute_logs else None,
    )
