"""Small shared definitions (subset of the reference's enc/utils/misc.py that the hot path
touches: device literals, network descriptors, exp-Golomb rate of the NN parameters)."""

from dataclasses import dataclass
from typing import Literal, Optional, TypedDict, Union

import torch
from torch import Tensor, nn

POSSIBLE_DEVICE = Literal["cpu", "cuda:0"]
MAX_ARM_MASK_SIZE = 9  # enc/utils/misc.py: 9x9 causal window


class DescriptorNN(TypedDict, total=False):
    """One value for the weights, one for the biases (misc.py DescriptorNN)."""

    weight: Optional[Union[int, float, str, Tensor]]
    bias: Optional[Union[int, float, str, Tensor]]


class DescriptorCoolChic(TypedDict, total=False):
    """One value per sub-network (misc.py DescriptorCoolChic)."""

    arm: Optional[DescriptorNN]
    upsampling: Optional[DescriptorNN]
    synthesis: Optional[DescriptorNN]


MODULES_TO_SEND = ["arm", "upsampling", "synthesis"]


def get_best_device() -> str:
    return "cuda:0" if torch.cuda.is_available() else "cpu"


def exp_golomb_nbins(symbol: Tensor, count: int = 0) -> Tensor:
    """Bits of an exp-Golomb code of exponent `count` (misc.py:255-268)."""
    nbins = 2 * torch.floor(torch.log2(symbol.abs() / (2**count) + 1)) + count + 1 + (symbol != 0)
    return nbins.sum()


@torch.no_grad()
def measure_expgolomb_rate(q_module: nn.Module, q_step: DescriptorNN, expgol_cnt: DescriptorNN) -> DescriptorNN:
    """Rate of the quantised NN parameters (misc.py:186-252).  Returns floats/0-d tensors."""
    sent = {"bias": [], "weight": []}
    rate: DescriptorNN = {"bias": 0.0, "weight": 0.0}
    for name, value in q_module.get_param().items():
        kind = "weight" if ".weight" in name else ("bias" if ".bias" in name else None)
        if kind is None or q_step.get(kind) is None:
            return rate
        sent[kind].append((value / q_step[kind]).view(-1))
    for kind, vals in sent.items():
        if not vals:
            continue
        if expgol_cnt.get(kind) is None:
            return rate
        rate[kind] = exp_golomb_nbins(torch.cat(vals), count=expgol_cnt[kind])
    return rate


# ---- NN quantisation search space (enc/utils/misc.py:111-152 of the reference) ---------------------------
MAX_AC_MAX_VAL = 65535  # 16-bit field of the bitstream header


def _pow2_range(lo: int, hi: int) -> Tensor:
    return 2.0 ** torch.arange(lo, hi + 1, dtype=torch.float32)


POSSIBLE_Q_STEP = {
    "arm": {"weight": _pow2_range(-8, 0), "bias": _pow2_range(-16, 0)},
    "upsampling": {"weight": _pow2_range(-12, 0), "bias": _pow2_range(0, 0)},
    "synthesis": {"weight": _pow2_range(-12, 0), "bias": _pow2_range(-24, 0)},
}
POSSIBLE_EXP_GOL_COUNT = {m: {"weight": torch.arange(0, 13, dtype=torch.float32),
                              "bias": torch.arange(0, 13, dtype=torch.float32)} for m in MODULES_TO_SEND}


def get_q_step_from_parameter_name(parameter_name: str, q_step: DescriptorNN):
    """Weights and biases have their own step; the kind is read off the state-dict name (misc.py:155-185)."""
    for kind in ("weight", "bias"):
        if "." + kind in parameter_name:
            return q_step.get(kind)
    print(f'Parameter name should include ".weight" or ".bias" Found: {parameter_name}')
    return None
