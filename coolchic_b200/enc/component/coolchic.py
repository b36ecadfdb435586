"""CoolChicEncoder -- drop-in for the reference's enc/component/coolchic.py:55-795.

Same constructor (`CoolChicEncoderParameter`), same `forward` signature and return triple, same
`state_dict` names and get/set API; the forward and its backward are ONE call each into the
sm_100a kernels (csrc/, through include/ccb.h).  CUDA only: there is no CPU path.
"""

from __future__ import annotations

import math
from dataclasses import dataclass, field, fields
from typing import Any, Dict, List, Optional, OrderedDict, Tuple

import torch
from torch import Tensor, nn

from ... import ops
from ..utils.misc import (
    MAX_ARM_MASK_SIZE,
    MODULES_TO_SEND,
    DescriptorCoolChic,
    DescriptorNN,
    measure_expgolomb_rate,
)
from .core.arm import Arm
from .core.quantizer import POSSIBLE_QUANTIZATION_NOISE_TYPE, POSSIBLE_QUANTIZER_TYPE
from .core.synthesis import Synthesis
from .core.upsampling import Upsampling


@dataclass
class CoolChicEncoderParameter:
    """Architecture of a CoolChicEncoder (coolchic.py:55-124).  Picklable: it is stored inside
    checkpoints by FrameEncoder.save()."""

    layers_synthesis: List[str]
    n_ft_per_res: List[int]
    dim_arm: int = 24
    n_hidden_layers_arm: int = 2
    encoder_gain: int = 16
    ups_k_size: int = 8
    ups_preconcat_k_size: int = 7

    latent_n_grids: int = field(init=False)
    img_size: Optional[Tuple[int, int]] = field(init=False, default=None)

    def __post_init__(self):
        self.latent_n_grids = len(self.n_ft_per_res)

    def set_image_size(self, img_size: Tuple[int, int]) -> None:
        self.img_size = img_size

    def pretty_string(self) -> str:
        s = "CoolChicEncoderParameter value:\n-------------------------------\n"
        for k in fields(self):
            s += f"{k.name:<25}: {str(getattr(self, k.name)):<80}\n"
        return s + "\n"


@dataclass
class CoolChicEncoderOutput:
    raw_out: Tensor
    rate: Tensor
    additional_data: Dict[str, Any]


class CoolChicLatentGrid(nn.Module):
    """One latent resolution (coolchic.py:144-151)."""

    def __init__(self, data: Tensor):
        super().__init__()
        self.data = nn.Parameter(data, requires_grad=True)


def ctx_flat_indices(dim_arm: int) -> Tensor:
    """Indices (into the flattened 9x9 causal window) of the context pixels
    (core/arm.py:373-506); kept as the `non_zero_pixel_ctx_index` buffer for API parity."""
    rows = {
        8: {1: [4], 2: [4], 3: [3, 4, 5], 4: [1, 2, 3]},
        16: {1: [4, 5], 2: [2, 3, 4, 5, 6], 3: [1, 2, 3, 4, 5, 6], 4: [1, 2, 3]},
        24: {0: [4], 1: [2, 3, 4, 5, 6], 2: [1, 2, 3, 4, 5, 6, 7], 3: [1, 2, 3, 4, 5, 6, 7], 4: [0, 1, 2, 3]},
        32: {0: [2, 3, 4, 5], 1: [1, 2, 3, 4, 5, 6, 7], 2: [1, 2, 3, 4, 5, 6, 7, 8],
             3: [0, 1, 2, 3, 4, 5, 6, 7, 8], 4: [0, 1, 2, 3]},
    }
    if dim_arm not in rows:
        raise ValueError(f"ARM context size must be 8, 16, 24 or 32. Found {dim_arm}.")
    out = []
    for r in sorted(rows[dim_arm]):
        out += [9 * r + c for c in rows[dim_arm][r]]
    return torch.tensor(out, dtype=torch.long)


def resolve_synthesis_layers(layers: List[str], n_out: int = 3) -> List[str]:
    """'X' stands for the number of output channels (utils/types.py DecoderConfig)."""
    return [l.replace("X", str(n_out)) for l in layers]


class _FusedCoolChic(nn.Module):
    """Everything CoolChicEncoder and LatentFreeCoolChicEncoder share: sub-networks, bookkeeping
    attributes and the fused forward."""

    def _build_networks(self, param: CoolChicEncoderParameter) -> None:
        self.param = param
        assert self.param.img_size is not None, (
            "You are trying to instantiate a CoolChicEncoder from a CoolChicEncoderParameter with a field "
            "img_size set to None. Use the function coolchic_encoder_param.set_img_size((H, W)) before "
            "instantiating the CoolChicEncoder."
        )
        assert all(c == 1 for c in param.n_ft_per_res), "Only one feature per latent resolution is supported"
        self.encoder_gains = param.encoder_gain
        self.size_per_latent = []
        for i in range(param.latent_n_grids):
            h_grid, w_grid = [int(math.ceil(x / (2**i))) for x in param.img_size]
            self.size_per_latent.append((1, param.n_ft_per_res[i], h_grid, w_grid))
        self._layers_synthesis = resolve_synthesis_layers(list(param.layers_synthesis))
        self.synthesis = Synthesis(sum(s[1] for s in self.size_per_latent), self._layers_synthesis)
        self.upsampling = Upsampling(
            ups_k_size=param.ups_k_size,
            ups_preconcat_k_size=param.ups_preconcat_k_size,
            n_ups_kernel=param.latent_n_grids - 1,
            n_ups_preconcat_kernel=param.latent_n_grids - 1,
        )
        max_context_pixel = int((MAX_ARM_MASK_SIZE**2 - 1) / 2)
        assert param.dim_arm <= max_context_pixel, (
            f"You can not have more context pixels  than {max_context_pixel}. Found {param.dim_arm}"
        )
        self.mask_size = MAX_ARM_MASK_SIZE
        self.register_buffer("non_zero_pixel_ctx_index", ctx_flat_indices(param.dim_arm), persistent=False)
        self.arm = Arm(param.dim_arm, param.n_hidden_layers_arm)
        self.modules_to_send = list(MODULES_TO_SEND)
        self.flops_str = ""
        self.total_flops = 0.0
        self.flops_per_module = {k: 0 for k in self.modules_to_send}
        self.get_flops()
        self.nn_q_step: Dict[str, DescriptorNN] = {k: {"weight": None, "bias": None} for k in self.modules_to_send}
        self.nn_expgol_cnt: Dict[str, DescriptorNN] = {k: {"weight": None, "bias": None} for k in self.modules_to_send}
        self.full_precision_param = None

    # ------------------------------------------------------------------ fused forward
    def _spec(self) -> ops.ArchSpec:
        p = self.param
        return ops.ArchSpec(p.img_size[0], p.img_size[1], p.dim_arm, p.n_hidden_layers_arm, self._layers_synthesis,
                            p.latent_n_grids, p.ups_k_size, p.ups_preconcat_k_size, float(self.encoder_gains))

    def _weight_tensors(self) -> List[Tensor]:
        """Weights in ops.ArchSpec.weight_names() order, read by attribute at call time (callers may
        have swapped Parameters for non-leaf tensors: hypernet/common.py:161-182)."""
        n_stage = self.param.latent_n_grids - 1
        return self.arm.ordered_tensors() + self.upsampling.ordered_tensors(n_stage) + self.synthesis.ordered_tensors()

    def _fused_forward(self, latents: List[Tensor], quantizer_noise_type, quantizer_type, soft_round_temperature,
                       noise_parameter, AC_MAX_VAL: int, flag_additional_outputs: bool, noise: Optional[Tensor] = None):
        B = latents[0].shape[0]
        if noise is None:
            # parity hook: the raw N(0,1) / U(0,1) draw to use instead of torch.randn / torch.rand
            noise = getattr(self, "_injected_noise", None)
        if flag_additional_outputs and B > 1:
            raise NotImplementedError("Batching is not yet supported for additional outputs.")
        out = ops.coolchic_forward(
            self._spec(), latents, self._weight_tensors(), training=self.training, q_type=quantizer_type,
            noise_type=quantizer_noise_type, T=soft_round_temperature, sigma=noise_parameter,
            ac_max_val=AC_MAX_VAL, noise=noise, detailed=flag_additional_outputs)
        additional_data: Dict[str, Any] = {}
        if not flag_additional_outputs:
            raw, rate = out
            return raw, rate, additional_data
        raw, rate, mu, scale, log_scale, yhat = out
        keys = ("detailed_sent_latent", "detailed_mu", "detailed_scale", "detailed_log_scale", "detailed_rate_bit",
                "detailed_centered_latent")
        for k in keys:
            additional_data[k] = []
        additional_data["hpfilters"] = []
        cnt = 0
        for (_, c_i, h_i, w_i) in self.size_per_latent:
            n = c_i * h_i * w_i
            sl = [t[:, cnt:cnt + n].view((1, c_i, h_i, w_i)) for t in (yhat, mu, scale, log_scale, rate)]
            cnt += n
            additional_data["detailed_sent_latent"].append(sl[0])
            additional_data["detailed_mu"].append(sl[1])
            additional_data["detailed_scale"].append(sl[2])
            additional_data["detailed_log_scale"].append(sl[3])
            additional_data["detailed_rate_bit"].append(sl[4])
            additional_data["detailed_centered_latent"].append(sl[0] - sl[1])
        return raw, rate, additional_data

    # ------------------------------------------------------------------ parameters
    def _reset_nn_descriptors(self) -> None:
        self.nn_q_step = {k: {"weight": None, "bias": None} for k in self.modules_to_send}
        self.nn_expgol_cnt = {k: {"weight": None, "bias": None} for k in self.modules_to_send}

    def _store_full_precision_param(self) -> None:
        """(coolchic.py:551-586)"""
        if self.full_precision_param is not None:
            print("Warning: overwriting already saved full-precision parameters"
                  " in CoolChicEncoder _store_full_precision_param().")
        for what, d in (("nn_q_step", self.nn_q_step), ("nn_expgol_cnt", self.nn_expgol_cnt)):
            assert all(v is None for dd in d.values() for v in dd.values()), (
                f"Trying to store full precision parameters, while CoolChicEncoder {what} attributes is not "
                "full of None. This means that the parameters have already been quantized... aborting!"
            )
        self.full_precision_param = self.get_param()

    def _load_full_precision_param(self) -> None:
        assert self.full_precision_param is not None, (
            "Trying to load full precision parameters but self.full_precision_param is None"
        )
        self.set_param(self.full_precision_param)
        self._reset_nn_descriptors()

    # ------------------------------------------------------------------ complexity / rates
    def get_flops(self) -> None:
        """MAC count per module.  The reference traces an eval forward with fvcore
        (coolchic.py:611-650); the same numbers follow in closed form from the architecture, as the
        reference EXECUTES it in eval mode (separable upsampling filters)."""
        p = self.param
        H, W = p.img_size
        sizes = [(s[2], s[3]) for s in self.size_per_latent]
        n_lat = sum(h * w for h, w in sizes)
        C, nh = p.dim_arm, p.n_hidden_layers_arm
        arm = n_lat * (nh * C * C + 2 * C)
        syn, cin = 0, len(sizes)
        for spec in self._layers_synthesis:
            o, k, _, _ = spec.split("-")
            syn += H * W * int(o) * cin * int(k) ** 2
            cin = int(o)
        ups, L = 0, len(sizes)
        k_up, k_pre = p.ups_k_size, p.ups_preconcat_k_size
        for idx in range(L - 1):
            hs, ws = sizes[L - 1 - idx]
            ht, wt = sizes[L - 2 - idx]
            c = idx + 1
            # horizontal transposed pass on the padded input, then vertical pass (upsampling.py:337-353)
            ups += c * (hs * (ws + k_up) * k_up + (hs + k_up) * (2 * ws) * k_up)
            ups += 2 * ht * wt * k_pre
        self.flops_per_module = {"arm": arm, "upsampling": ups, "synthesis": syn}
        self.total_flops = arm + ups + syn
        n_pix = H * W
        self.flops_str = "\n".join(
            [f"| {'module':<12} | {'MAC':>14} | {'MAC / pixel':>12} |"]
            + [f"| {k:<12} | {v:>14d} | {v / n_pix:>12.1f} |" for k, v in self.flops_per_module.items()]
        )

    def get_network_rate(self) -> DescriptorCoolChic:
        rate: DescriptorCoolChic = {m: {"weight": 0.0, "bias": 0.0} for m in self.modules_to_send}
        for m in self.modules_to_send:
            rate[m] = measure_expgolomb_rate(getattr(self, m), self.nn_q_step.get(m), self.nn_expgol_cnt.get(m))
        return rate

    def get_network_quantization_step(self) -> DescriptorCoolChic:
        return self.nn_q_step

    def get_network_expgol_count(self) -> DescriptorCoolChic:
        return self.nn_expgol_cnt

    def get_total_mac_per_pixel(self) -> float:
        if not self.flops_str:
            self.get_flops()
        return self.total_flops / (self.param.img_size[-2] * self.param.img_size[-1])

    def str_complexity(self) -> str:
        if not self.flops_str:
            self.get_flops()
        bar = "----------------------------------"
        return f"{self.flops_str}\n\n{bar}\nTotal MAC / decoded pixel: {self.get_total_mac_per_pixel():.1f}\n{bar}"

    def pretty_string(self) -> str:
        n_pix = self.param.img_size[-2] * self.param.img_size[-1]
        tot = self.get_total_mac_per_pixel()
        s = f"\nCool-chic architecture {tot:.0f} MAC / pixel\n\n"
        for m in ("upsampling", "arm", "synthesis"):
            c = self.flops_per_module[m] / n_pix
            s += f"{m}: {c:.0f} MAC/pixel ; {100 * c / tot:.1f} % of the complexity\n"
            s += "\n".join(f"    {k:<70} {tuple(v.shape)}" for k, v in getattr(self, m).named_parameters()) + "\n"
        return s

    def to_device(self, device) -> None:
        """(coolchic.py:733-750).  Kept for API parity; forward itself requires "cuda:0"."""
        assert str(device) == "cpu" or str(device).startswith("cuda"), f"Unknown device {device}, should be 'cpu' or a CUDA device"
        self.to(device)


class CoolChicEncoder(_FusedCoolChic):
    """CoolChicEncoder for a single frame (coolchic.py:154-795)."""

    def __init__(self, param: CoolChicEncoderParameter):
        super().__init__()
        self.latent_grids = nn.ModuleList()
        assert param.img_size is not None, "Use param.set_image_size((H, W)) before instantiating the CoolChicEncoder."
        for i in range(param.latent_n_grids):
            h_grid, w_grid = [int(math.ceil(x / (2**i))) for x in param.img_size]
            self.latent_grids.append(CoolChicLatentGrid(torch.empty((1, param.n_ft_per_res[i], h_grid, w_grid))))
        self.initialize_latent_grids()
        self._build_networks(param)

    def forward(
        self,
        quantizer_noise_type: POSSIBLE_QUANTIZATION_NOISE_TYPE = "kumaraswamy",
        quantizer_type: POSSIBLE_QUANTIZER_TYPE = "softround",
        soft_round_temperature: Optional[Tensor] = torch.tensor(0.3),
        noise_parameter: Optional[Tensor] = torch.tensor(1.0),
        AC_MAX_VAL: int = -1,
        flag_additional_outputs: bool = False,
    ) -> Tuple[Tensor, Tensor, Dict[str, Any]]:
        """raw_out [B,3,H,W], rate [B,M] (bits), additional_data (coolchic.py:291-479)."""
        latents = [g.data for g in self.latent_grids]
        return self._fused_forward(latents, quantizer_noise_type, quantizer_type, soft_round_temperature,
                                   noise_parameter, AC_MAX_VAL, flag_additional_outputs)

    # ------- Getter / Setter and Initializer (coolchic.py:482-549)
    def get_param(self) -> OrderedDict[str, Tensor]:
        from collections import OrderedDict as OD

        param = OD()
        param.update({f"latent_grids.{k}": v.detach().clone() for k, v in self.latent_grids.named_parameters()})
        param.update({f"arm.{k}": v for k, v in self.arm.get_param().items()})
        param.update({f"upsampling.{k}": v for k, v in self.upsampling.get_param().items()})
        param.update({f"synthesis.{k}": v for k, v in self.synthesis.get_param().items()})
        return param

    def set_param(self, param: OrderedDict[str, Tensor]):
        self.load_state_dict(param)

    def initialize_latent_grids(self, zeros: bool = True, random_seed: Optional[int] = None) -> None:
        generator = None
        if not zeros:
            assert random_seed is not None, (
                "Trying to initialize the latent grids with random values but random_seed is None."
            )
            generator = torch.Generator().manual_seed(random_seed)
        for i, latent in enumerate(self.latent_grids):
            dev = latent.data.device
            if zeros:
                new = torch.zeros_like(latent.data)
            else:
                new = (1e-2 * torch.randn(latent.data.shape, generator=generator)).to(dev)
            self.latent_grids[i] = CoolChicLatentGrid(new)

    def reinitialize_parameters(self):
        self.arm.reinitialize_parameters()
        self.upsampling.reinitialize_parameters()
        self.synthesis.reinitialize_parameters()
        self.initialize_latent_grids()
        self._reset_nn_descriptors()
