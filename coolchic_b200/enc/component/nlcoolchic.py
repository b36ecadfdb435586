"""LatentFreeCoolChicEncoder -- drop-in for the reference's enc/component/nlcoolchic.py:33-717:
the same fused forward, with the (batched) latents passed as an argument instead of owned.
This is the entry the hypernetwork wholenets call (hypernet/hypernet.py:1176-1184, 1266-1288).
"""

from __future__ import annotations

from typing import Any, Dict, List, Optional, OrderedDict, Tuple

import torch
from torch import Tensor

from .coolchic import CoolChicEncoderParameter, _FusedCoolChic
from .core.quantizer import POSSIBLE_QUANTIZATION_NOISE_TYPE, POSSIBLE_QUANTIZER_TYPE


class LatentFreeCoolChicEncoder(_FusedCoolChic):
    def __init__(self, param: CoolChicEncoderParameter):
        super().__init__()
        self._build_networks(param)

    def forward(
        self,
        latents: List[Tensor],
        quantizer_noise_type: POSSIBLE_QUANTIZATION_NOISE_TYPE = "kumaraswamy",
        quantizer_type: POSSIBLE_QUANTIZER_TYPE = "softround",
        soft_round_temperature: Optional[Tensor] = torch.tensor(0.3),
        noise_parameter: Optional[Tensor] = torch.tensor(1.0),
        AC_MAX_VAL: int = -1,
        flag_additional_outputs: bool = False,
    ) -> Tuple[Tensor, Tensor, Dict[str, Any]]:
        assert len(latents) == self.param.latent_n_grids, (
            f"Expected {self.param.latent_n_grids} latents, got {len(latents)}"
        )
        return self._fused_forward(list(latents), quantizer_noise_type, quantizer_type, soft_round_temperature,
                                   noise_parameter, AC_MAX_VAL, flag_additional_outputs)

    def get_param(self) -> OrderedDict[str, Tensor]:
        from collections import OrderedDict as OD

        param = OD()
        param.update({f"arm.{k}": v for k, v in self.arm.get_param().items()})
        param.update({f"upsampling.{k}": v for k, v in self.upsampling.get_param().items()})
        param.update({f"synthesis.{k}": v for k, v in self.synthesis.get_param().items()})
        return param

    def set_param(self, param: OrderedDict[str, Tensor]):
        self.load_state_dict(param)

    def reinitialize_parameters(self):
        self.arm.reinitialize_parameters()
        self.upsampling.reinitialize_parameters()
        self.synthesis.reinitialize_parameters()
        self._reset_nn_descriptors()
