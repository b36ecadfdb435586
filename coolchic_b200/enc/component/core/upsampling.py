"""Upsampling parameter container (reference: enc/component/core/upsampling.py:21-89, 92-355,
358-537).  Names follow torch's parametrize layout the reference uses
(`conv_transpose2ds.{i}.parametrizations.weight.original`, `.bias`); the filters run in
csrc/ccb_ups.cu."""

from typing import OrderedDict

import torch
from torch import Tensor, nn

from ...._hyper import set_hypernet_weights
from .arm import _FUSED_ONLY


def symmetric_kernel_1d(original: Tensor, target_k_size: int) -> Tensor:
    """a b c -> a b c c b a (even) / a b c b a (odd)   (upsampling.py:46-68)."""
    return torch.cat([original, torch.flip(original, dims=[0])[target_k_size % 2:]])


class _Original(nn.Module):
    def __init__(self, w: Tensor):
        super().__init__()
        self.original = nn.Parameter(w, requires_grad=True)


class _Parametrizations(nn.Module):
    def __init__(self, w: Tensor):
        super().__init__()
        self.weight = _Original(w)


class _SymmetricFilter(nn.Module):
    def __init__(self, kernel_size: int):
        super().__init__()
        self.target_k_size = kernel_size
        self.param_size = kernel_size // 2 + kernel_size % 2
        self.bias = nn.Parameter(torch.zeros(1), requires_grad=True)  # present in the state_dict, never used
        self.parametrizations = _Parametrizations(self._init_weight())

    def _init_weight(self) -> Tensor:
        raise NotImplementedError

    @property
    def weight(self) -> Tensor:
        """The full symmetric 1-D kernel (what `.weight` returns in the reference)."""
        return symmetric_kernel_1d(self.parametrizations.weight.original, self.target_k_size)

    def initialize_parameters(self) -> None:
        dev = self.parametrizations.weight.original.device
        self.parametrizations = _Parametrizations(self._init_weight().to(dev))
        self.bias = nn.Parameter(torch.zeros(1, device=dev), requires_grad=True)

    def forward(self, x: Tensor) -> Tensor:
        raise NotImplementedError(_FUSED_ONLY)


class UpsamplingSeparableSymmetricConv2d(_SymmetricFilter):
    """Pre-concatenation filter: y = conv(x, w (x) w, zero pad) + x; Dirac init (upsampling.py:135-166)."""

    def __init__(self, kernel_size: int):
        assert kernel_size % 2 == 1, f"Upsampling kernel size must be odd, found {kernel_size}."
        super().__init__(kernel_size)

    def _init_weight(self) -> Tensor:
        w = torch.zeros(self.param_size)
        w[-1] = 1
        return w


class UpsamplingSeparableSymmetricConvTranspose2d(_SymmetricFilter):
    """x2 upsampling filter; bilinear (k < 8) or bicubic init (upsampling.py:255-295)."""

    def __init__(self, kernel_size: int):
        assert kernel_size >= 4 and not kernel_size % 2, (
            f"Upsampling kernel size shall be even and ≥4. Found {kernel_size}"
        )
        super().__init__(kernel_size)

    def _init_weight(self) -> Tensor:
        if self.target_k_size < 8:
            core = torch.tensor([1.0 / 4.0, 3.0 / 4.0])
        else:
            core = torch.tensor([0.0351562, 0.1054687, -0.2617187, -0.8789063])
        w = torch.zeros(self.param_size)
        w[self.param_size - core.numel():] = core
        return w


class Upsampling(nn.Module):
    def __init__(self, ups_k_size: int, ups_preconcat_k_size: int, n_ups_kernel: int, n_ups_preconcat_kernel: int):
        super().__init__()
        self.n_ups_kernel, self.n_ups_preconcat_kernel = n_ups_kernel, n_ups_preconcat_kernel
        self.conv_transpose2ds = nn.ModuleList(
            [UpsamplingSeparableSymmetricConvTranspose2d(ups_k_size) for _ in range(n_ups_kernel)])
        self.conv2ds = nn.ModuleList(
            [UpsamplingSeparableSymmetricConv2d(ups_preconcat_k_size) for _ in range(n_ups_preconcat_kernel)])

    def forward(self, decoder_side_latent):
        raise NotImplementedError(_FUSED_ONLY)

    def ordered_tensors(self, n_stages: int):
        """[ups_0..ups_{n-1}, pre_0..pre_{n-1}] with the reference's `idx % n_kernel` sharing rule
        (upsampling.py:493-500)."""
        ups = [self.conv_transpose2ds[i % self.n_ups_kernel].parametrizations.weight.original for i in range(n_stages)]
        pre = [self.conv2ds[i % self.n_ups_preconcat_kernel].parametrizations.weight.original for i in range(n_stages)]
        return ups + pre

    def get_param(self) -> OrderedDict[str, Tensor]:
        return OrderedDict({k: v.detach().clone() for k, v in self.named_parameters()})

    def set_param(self, param: OrderedDict[str, Tensor]):
        self.load_state_dict(param)

    def reinitialize_parameters(self) -> None:
        for m in list(self.conv_transpose2ds) + list(self.conv2ds):
            m.initialize_parameters()

    def set_hypernet_weights(self, all_weights) -> None:
        set_hypernet_weights(self, all_weights)
