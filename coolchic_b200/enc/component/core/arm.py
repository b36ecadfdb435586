"""ARM parameter container (reference: enc/component/core/arm.py:19-101, 157-306).

The MLP itself is evaluated inside the fused CUDA kernel (csrc/ccb_arm.cu) together with the
context gather and the Laplace rate; this module owns the parameters with the reference's
names (`mlp.{0,2,..}.weight|bias`), initialisation and get/set API.
"""

from typing import OrderedDict

import torch
from torch import Tensor, nn

from ...._hyper import set_hypernet_weights

_FUSED_ONLY = (
    "coolchic_b200 evaluates this sub-module inside the fused sm_100a kernels of CoolChicEncoder.forward; "
    "there is no stand-alone (PyTorch / CPU) forward."
)


class ArmLinear(nn.Module):
    """W x + b (+ x if residual).  Init: bias 0; weight 0 if residual else N(0, 1/out^4) (arm.py:66-84)."""

    def __init__(self, in_channels: int, out_channels: int, residual: bool = False):
        super().__init__()
        self.residual, self.in_channels, self.out_channels = residual, in_channels, out_channels
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels), requires_grad=True)
        self.bias = nn.Parameter(torch.empty(out_channels), requires_grad=True)
        self.initialize_parameters()

    def initialize_parameters(self) -> None:
        self.bias = nn.Parameter(torch.zeros_like(self.bias), requires_grad=True)
        if self.residual:
            self.weight = nn.Parameter(torch.zeros_like(self.weight), requires_grad=True)
        else:
            self.weight = nn.Parameter(torch.randn_like(self.weight) / self.weight.size(0) ** 2, requires_grad=True)

    def forward(self, x: Tensor) -> Tensor:
        raise NotImplementedError(_FUSED_ONLY)


class Arm(nn.Module):
    def __init__(self, dim_arm: int, n_hidden_layers_arm: int):
        super().__init__()
        assert dim_arm % 8 == 0, (
            f"ARM context size and hidden layer dimension must be a multiple of 8. Found {dim_arm}."
        )
        self.dim_arm, self.n_hidden_layers_arm = dim_arm, n_hidden_layers_arm
        layers = []
        for _ in range(n_hidden_layers_arm):
            layers += [ArmLinear(dim_arm, dim_arm, residual=True), nn.ReLU()]
        layers.append(ArmLinear(dim_arm, 2, residual=False))
        self.mlp = nn.Sequential(*layers)

    def forward(self, x: Tensor):
        raise NotImplementedError(_FUSED_ONLY)

    def ordered_tensors(self):
        """[w0, b0, w1, b1, ..., w_out, b_out] read by attribute so that tensors swapped in by
        set_hypernet_weights (plain, non-leaf tensors) are honoured."""
        out = []
        for layer in self.mlp.children():
            if isinstance(layer, ArmLinear):
                out += [layer.weight, layer.bias]
        return out

    def get_param(self) -> OrderedDict[str, Tensor]:
        return OrderedDict({k: v.detach().clone() for k, v in self.named_parameters()})

    def set_param(self, param: OrderedDict[str, Tensor]) -> None:
        self.load_state_dict(param)

    def reinitialize_parameters(self) -> None:
        for layer in self.mlp.children():
            if isinstance(layer, ArmLinear):
                layer.initialize_parameters()

    def set_hypernet_weights(self, all_weights) -> None:
        set_hypernet_weights(self, all_weights)
