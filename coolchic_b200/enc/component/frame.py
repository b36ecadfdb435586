"""FrameEncoder -- drop-in for the reference's enc/component/frame.py:36-320 (rgb / yuv444 frames).

Wraps the fused CoolChicEncoder, applies the eval-mode 8-bit rounding and the [0,1] clamp
(frame.py:172-183).  Inter-frame coding is bypassed in the fork (frame.py:165-172) and not provided.
"""

from __future__ import annotations

from dataclasses import dataclass, field
from io import BytesIO
from typing import Any, Dict, List, Optional, OrderedDict, Union

import torch
from torch import Tensor, nn

from .coolchic import CoolChicEncoder, CoolChicEncoderOutput, CoolChicEncoderParameter
from .core.quantizer import POSSIBLE_QUANTIZATION_NOISE_TYPE, POSSIBLE_QUANTIZER_TYPE


@dataclass
class FrameEncoderOutput:
    """decoded_image [B,3,H,W] in [0,1]; rate [B,M] in bits; additional_data (frame.py:36-52)."""

    decoded_image: Union[Tensor, Dict[str, Tensor]]
    rate: Tensor
    additional_data: Dict[str, Any] = field(default_factory=lambda: {})


class FrameEncoder(nn.Module):
    def __init__(self, coolchic_encoder_param: CoolChicEncoderParameter, frame_type: str = "I",
                 frame_data_type: str = "rgb", bitdepth: int = 8):
        super().__init__()
        if frame_data_type == "yuv420":
            raise NotImplementedError("coolchic_b200 covers rgb / yuv444 frames (SURVEY.md 8a row FE)")
        self.coolchic_encoder_param = coolchic_encoder_param
        self.frame_type = frame_type
        self.frame_data_type = frame_data_type
        self.bitdepth = bitdepth
        self.coolchic_encoder = CoolChicEncoder(self.coolchic_encoder_param)

    def forward(
        self,
        reference_frames: Optional[List[Tensor]] = None,
        quantizer_noise_type: POSSIBLE_QUANTIZATION_NOISE_TYPE = "kumaraswamy",
        quantizer_type: POSSIBLE_QUANTIZER_TYPE = "softround",
        soft_round_temperature: Optional[Tensor] = torch.tensor(0.3),
        noise_parameter: Optional[Tensor] = torch.tensor(1.0),
        AC_MAX_VAL: int = -1,
        flag_additional_outputs: bool = False,
    ) -> FrameEncoderOutput:
        raw_out, rate, add_data = self.coolchic_encoder.forward(
            quantizer_noise_type=quantizer_noise_type,
            quantizer_type=quantizer_type,
            soft_round_temperature=soft_round_temperature,
            noise_parameter=noise_parameter,
            AC_MAX_VAL=AC_MAX_VAL,
            flag_additional_outputs=flag_additional_outputs,
        )
        decoded_image = raw_out
        if not self.training:
            max_dynamic = 2 ** (self.bitdepth) - 1
            decoded_image = torch.round(decoded_image * max_dynamic) / max_dynamic
        decoded_image = torch.clamp(decoded_image, 0.0, 1.0)
        additional_data = {}
        if flag_additional_outputs:
            additional_data.update(add_data)
        return FrameEncoderOutput(decoded_image=decoded_image, rate=rate, additional_data=additional_data)

    # ------- Getter / Setter and Initializer (frame.py:199-254)
    def get_param(self) -> OrderedDict[str, Tensor]:
        from collections import OrderedDict as OD

        return OD({f"coolchic_encoder.{k}": v for k, v in self.coolchic_encoder.get_param().items()})

    def set_param(self, param: OrderedDict[str, Tensor]):
        self.load_state_dict(param)

    def reinitialize_parameters(self) -> None:
        self.coolchic_encoder.reinitialize_parameters()

    def set_to_train(self) -> None:
        self.train()
        self.coolchic_encoder.train()

    def set_to_eval(self) -> None:
        self.eval()
        self.coolchic_encoder.eval()

    def to_device(self, device) -> None:
        assert str(device) == "cpu" or str(device).startswith("cuda"), f"Unknown device {device}, should be 'cpu' or a CUDA device"
        self.to(device)
        self.coolchic_encoder.to_device(device)

    def save(self) -> BytesIO:
        """Same keys as the reference's checkpoint (frame.py:256-280)."""
        buffer = BytesIO()
        data = {
            "bitdepth": self.bitdepth,
            "frame_type": self.frame_type,
            "frame_data_type": self.frame_data_type,
            "coolchic_encoder_param": self.coolchic_encoder_param,
            "coolchic_encoder": self.coolchic_encoder.get_param(),
            "coolchic_nn_q_step": self.coolchic_encoder.get_network_quantization_step(),
            "coolchic_nn_expgol_cnt": self.coolchic_encoder.get_network_expgol_count(),
        }
        if self.coolchic_encoder.full_precision_param is not None:
            data["coolchic_full_precision_param"] = self.coolchic_encoder.full_precision_param
        torch.save(data, buffer)
        return buffer


def load_frame_encoder(raw_bytes: BytesIO) -> FrameEncoder:
    """(frame.py:283-320)"""
    raw_bytes.seek(0)
    d = torch.load(raw_bytes, map_location="cpu", weights_only=False)
    fe = FrameEncoder(coolchic_encoder_param=d["coolchic_encoder_param"], frame_type=d["frame_type"],
                      frame_data_type=d["frame_data_type"], bitdepth=d["bitdepth"])
    fe.coolchic_encoder.set_param(d["coolchic_encoder"])
    fe.coolchic_encoder.nn_q_step = d["coolchic_nn_q_step"]
    if "coolchic_nn_expgol_cnt" in d:
        fe.coolchic_encoder.nn_expgol_cnt = d["coolchic_nn_expgol_cnt"]
    if "coolchic_full_precision_param" in d:
        fe.coolchic_encoder.full_precision_param = d["coolchic_full_precision_param"]
    return fe
