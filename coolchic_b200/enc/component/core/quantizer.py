"""Quantiser modes accepted by the fused forward (reference: enc/component/core/quantizer.py:110-118).
The arithmetic itself lives in csrc/ccb_quant.cu."""

from typing import Literal

POSSIBLE_QUANTIZATION_NOISE_TYPE = Literal["kumaraswamy", "gaussian", "none"]
POSSIBLE_QUANTIZER_TYPE = Literal["softround_alone", "softround", "hardround", "ste", "true_ste", "none"]
