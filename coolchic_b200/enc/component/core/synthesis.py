"""Synthesis parameter container (reference: enc/component/core/synthesis.py:21-116, 175-314).
The convolutions run in csrc/ccb_syn.cu."""

import math
from typing import List, OrderedDict

import torch
from torch import Tensor, nn

from ...._hyper import set_hypernet_weights
from .arm import _FUSED_ONLY


class SynthesisConv2d(nn.Module):
    """k x k conv, replicate padding, optional residual (synthesis.py:21-116)."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size: int, residual: bool = False):
        super().__init__()
        self.residual, self.in_channels, self.out_channels = residual, in_channels, out_channels
        self.kernel_size, self.pad, self.groups = kernel_size, int((kernel_size - 1) / 2), 1
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, kernel_size, kernel_size), requires_grad=True)
        self.bias = nn.Parameter(torch.empty(out_channels), requires_grad=True)
        self.initialize_parameters()

    def initialize_parameters(self) -> None:
        """bias 0; weight 0 if residual else U(-a, a), a = sqrt(1/(cin k^2)) / cout^2 (synthesis.py:86-116)."""
        self.bias = nn.Parameter(torch.zeros_like(self.bias), requires_grad=True)
        if self.residual:
            self.weight = nn.Parameter(torch.zeros_like(self.weight), requires_grad=True)
        else:
            cout, cin, kh, kw = self.weight.size()
            a = math.sqrt(self.groups / (cin * self.groups * kh * kw))
            self.weight = nn.Parameter((torch.rand_like(self.weight) - 0.5) * 2 * a / (cout**2), requires_grad=True)

    def forward(self, x: Tensor) -> Tensor:
        raise NotImplementedError(_FUSED_ONLY)


class Synthesis(nn.Module):
    possible_non_linearity = {"none": nn.Identity, "relu": nn.ReLU}
    possible_mode = ["linear", "residual"]

    def __init__(self, input_ft: int, layers_dim: List[str]):
        super().__init__()
        self.synth_branches = nn.ModuleList()
        self.input_ft = input_ft
        layers = []
        for spec in layers_dim:
            out_ft, k_size, mode, non_linearity = spec.split("-")
            out_ft, k_size = int(out_ft), int(k_size)
            assert mode in Synthesis.possible_mode, f"Unknown mode. Found {mode}. Should be in {Synthesis.possible_mode}"
            assert non_linearity in Synthesis.possible_non_linearity, (
                f"Unknown non linearity. Found {non_linearity}. Should be in {Synthesis.possible_non_linearity.keys()}"
            )
            layers.append(SynthesisConv2d(input_ft, out_ft, k_size, residual=mode == "residual"))
            layers.append(Synthesis.possible_non_linearity[non_linearity]())
            input_ft = out_ft
        self.layers = nn.Sequential(*layers)

    def forward(self, x: Tensor) -> Tensor:
        raise NotImplementedError(_FUSED_ONLY)

    def ordered_tensors(self):
        out = []
        for layer in self.layers.children():
            if isinstance(layer, SynthesisConv2d):
                out += [layer.weight, layer.bias]
        return out

    def get_param(self) -> OrderedDict[str, Tensor]:
        return OrderedDict({k: v.detach().clone() for k, v in self.named_parameters()})

    def set_param(self, param: OrderedDict[str, Tensor]):
        self.load_state_dict(param)

    def reinitialize_parameters(self) -> None:
        for layer in self.layers.children():
            if isinstance(layer, SynthesisConv2d):
                layer.initialize_parameters()

    def set_hypernet_weights(self, all_weights) -> None:
        set_hypernet_weights(self, all_weights)
