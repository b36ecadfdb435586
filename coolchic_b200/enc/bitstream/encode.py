"""Bitstream hand-off: what the reference's `encode_frame` needs from the trained encoder to
entropy-code the latents -- the sent symbols and, for every one of them, the fixed-point ARM's expectation and
log-scale -- produced by the integer ARM kernel (`ccb_arm_int`, csrc/ccb_armint.cu) in ONE launch over all latent
levels instead of the reference's CPU forward through `ArmInt` (enc/bitstream/encode.py:256-265, 486-567).

Scope: everything `encode_frame` / `encode_video` write for ONE intra frame -- headers (header.py:113-160, 255-467),
the network sub-streams and the latent sub-streams -- except the CABAC coder itself, which stays the reference's C++
(`CCLIB.ccencapi`, sequential by nature): `cc_code_wb_bac` / `cc_code_latent_layer_bac` are arguments (default: import
them from `CCLIB`, as the reference does, and fail loudly when it is not installed).  Video (inter frames, several
frames per file) is out of scope.

The decoder recomputes (mu, log-scale) with its own sequential integer ARM (cpp/arm_cpu.cpp) from the symbols it has
already decoded, so a single differing bit desynchronises the arithmetic decoder: tests/test_bitstream.py
round-trips every layer through the reference's coder and decoder (compiled from its own sources for the tests), and tests/test_gpu_int.py reproduces
the sub-streams of shipped .cool files byte for byte from this kernel's outputs.
"""

from __future__ import annotations

import os
from typing import Any, Callable, Dict, List, Optional, OrderedDict, Tuple

import torch
from torch import Tensor

from ... import ops
from ..utils.decoder_int import integer_latents

FIXED_POINT_FRACTIONAL_BITS = 8  # armint.py / encode.py: the ARM's fixed point
FIXED_POINT_FRACTIONAL_MULT = 1 << FIXED_POINT_FRACTIONAL_BITS


def get_sub_bitstream_path(root_bitstream_path: str, counter_2d_latent: int) -> str:
    """Sub-stream file of the `counter_2d_latent`-th 2-D feature map (enc/bitstream/utils.py:10-29)."""
    return f"{root_bitstream_path}_{counter_2d_latent}"


def _arm_fixed(v: Tensor, bits: int) -> Tensor:
    """v x 2^bits rounded half away from zero, `trunc(x + sign(x) / 2)`, int32 (armint.py:197-205)."""
    x = v.detach().to(torch.float32) * float(1 << bits)
    x = x + torch.sign(x) * 0.5
    return torch.where(x < 0, -((-x).to(torch.int32)), x.to(torch.int32)).contiguous()


@torch.no_grad()
def integerise_arm_param(float_param: OrderedDict[str, Tensor]) -> Tuple[List[Tensor], List[Tensor]]:
    """`ArmInt.set_param_from_float` (armint.py:190-209): weights x 2^8, biases x 2^16, rounded half away from zero,
    int32.  -> (weights per layer [out, in], biases per layer), input order kept."""
    ws, bs = [], []
    for name, v in float_param.items():
        if "weight" in name:
            ws.append(_arm_fixed(v, FIXED_POINT_FRACTIONAL_BITS))
        else:
            bs.append(_arm_fixed(v, 2 * FIXED_POINT_FRACTIONAL_BITS))
    if len(ws) != len(bs):
        raise ValueError("integerise_arm_param(): every ARM layer needs a weight and a bias")
    return ws, bs


@torch.no_grad()
def get_ac_max_val_latent(frame_encoder: Any) -> int:
    """ceil(max |sent latent|) + 2 (encode.py:128-163): symbols live in [-AC_MAX_VAL, AC_MAX_VAL - 1]."""
    cc = frame_encoder.coolchic_encoder if hasattr(frame_encoder, "coolchic_encoder") else frame_encoder
    return int(integer_latents(cc).abs().max().item()) + 2


@torch.no_grad()
def latent_layers_for_bac(frame_encoder: Any) -> List[Dict[str, Any]]:
    """One record per 2-D latent grid, finest first: `y`, `mu`, `log_scale` (CPU int32 [h*w]; mu and log_scale at
    8 fractional bits), `h`, `w`, `all_zero`.  These are the positional arguments of `cc_code_latent_layer_bac`
    (ccencapi.cpp:179-410) as encode.py:546-567 builds them."""
    cc = frame_encoder.coolchic_encoder if hasattr(frame_encoder, "coolchic_encoder") else frame_encoder
    if int(cc.latent_grids[0].data.shape[0]) != 1:
        raise ValueError("latent_layers_for_bac(): one frame at a time")
    if any(int(g.data.shape[1]) != 1 for g in cc.latent_grids):
        raise ValueError("latent_layers_for_bac(): one feature per resolution (the fused kernels' layout)")
    w_q, b_q = integerise_arm_param(cc.arm.get_param())
    y = integer_latents(cc)[None].contiguous()
    mu, ls = ops.arm_int(cc._spec(), y, w_q, b_q, indices=False)
    y, mu, ls = y[0].cpu(), mu[0].cpu(), ls[0].cpu()  # one D2H copy each for the whole pyramid
    out, off = [], 0
    for g in cc.latent_grids:
        h, w = int(g.data.shape[-2]), int(g.data.shape[-1])
        n = h * w
        yl = y[off:off + n]
        out.append({"y": yl, "mu": mu[off:off + n], "log_scale": ls[off:off + n], "h": h, "w": w,
                    "all_zero": bool((yl == 0).all())})
        off += n
    return out


def _reference_coder() -> Callable:
    try:
        from CCLIB.ccencapi import cc_code_latent_layer_bac  # type: ignore  # the reference's compiled coder
    except ImportError as e:  # no silent fallback: there is no other entropy coder in this package
        raise ImportError("encode_latents(): the reference's CABAC extension `CCLIB.ccencapi` is not importable; build "
                          "it from the reference checkout (setup.py) or pass cc_code_latent_layer_bac=") from e
    return cc_code_latent_layer_bac


def encode_latents(frame_encoder: Any, bitstream_path: str, hls_sig_blksize: int,
                   cc_code_latent_layer_bac: Optional[Callable] = None) -> List[int]:
    """The latent loop of `encode_frame` (encode.py:503-570): writes `<bitstream_path>_<k>` for every 2-D latent
    (an empty file when the grid is all zero, as the reference does) and returns the sub-stream sizes in bytes, the
    `n_bytes_per_latent` that `write_frame_header` records."""
    coder = cc_code_latent_layer_bac or _reference_coder()
    n_bytes: List[int] = []
    for k, layer in enumerate(latent_layers_for_bac(frame_encoder)):
        path = get_sub_bitstream_path(bitstream_path, k)
        if layer["all_zero"]:
            open(path, "wb").close()
            n_bytes.append(0)
            continue
        coder(path, layer["y"].tolist(), layer["mu"].tolist(), layer["log_scale"].tolist(), layer["h"], layer["w"],
              hls_sig_blksize)
        n_bytes.append(os.path.getsize(path))
    return n_bytes


# ---------------------------------------------------------------------------------------------------------------
# Whole-frame assembly: network sub-streams + headers, so that the file can be handed to the reference's decoder.
# ---------------------------------------------------------------------------------------------------------------
_HAVE_BIAS = {"arm": True, "upsampling": False, "synthesis": True}  # encode.py:240-245: upsampling biases are not sent
_FRAME_DATA_TYPE = ["rgb", "yuv420", "yuv444"]  # header.py:99-101
_POSSIBLE_BITDEPTH = [8, 9, 10, 11, 12, 13, 14, 15, 16]
_SYNTHESIS_MODE = ["linear", "residual"]
_SYNTHESIS_NON_LINEARITY = ["none", "relu"]


def _cc_of(frame_encoder: Any) -> Any:
    return frame_encoder.coolchic_encoder if hasattr(frame_encoder, "coolchic_encoder") else frame_encoder


def _q_step_index(cc: Any, module_name: str, kind: str) -> int:
    """Index of the module's step in POSSIBLE_Q_STEP (nearest, encode.py:291-299)."""
    from ..utils.misc import POSSIBLE_Q_STEP

    step = cc.nn_q_step.get(module_name, {}).get(kind)
    if step is None:
        raise ValueError(f"encode_frame(): {module_name}.{kind} has no quantisation step -- run quantize_model() first")
    possible = POSSIBLE_Q_STEP[module_name][kind]
    return int(torch.argmin((possible - float(step)).abs()).item())


def _toward_zero_shift(v: Tensor, n: int) -> Tensor:
    return torch.where(v < 0, -((-v) >> n), v >> n) if n > 0 else v


@torch.no_grad()
def network_symbols(frame_encoder: Any) -> Dict[str, Dict[str, Any]]:
    """Per module sent to the decoder: the integer symbols of its weights and of its biases (flat int32 on the CPU, in
    `named_parameters()` order), the step indices, and the values the DECODER will rebuild (`decoded`: name -> float
    tensor = symbol x step).  ARM symbols come from the fixed-point weights shifted down to the step, toward zero
    (encode.py:300-316, 331-347); the other modules send round(v / step) (encode.py:318-321)."""
    from ..utils.misc import POSSIBLE_Q_STEP

    cc = _cc_of(frame_encoder)
    out: Dict[str, Dict[str, Any]] = {}
    for module_name in cc.modules_to_send:
        module = getattr(cc, module_name)
        rec: Dict[str, Any] = {"weight": [], "bias": [], "decoded": {}, "q_step_index": {"weight": -1, "bias": 0}}
        for name, v in module.named_parameters():
            kind = "weight" if "weight" in name else "bias"
            if kind == "bias" and not _HAVE_BIAS[module_name]:
                continue
            idx = _q_step_index(cc, module_name, kind)
            step = float(POSSIBLE_Q_STEP[module_name][kind][idx])
            rec["q_step_index"][kind] = idx
            v = v.detach().cpu()
            if module_name == "arm":
                # the step of index i is 2^(i - bits): the fixed-point value drops its i low bits, toward zero
                sym = _toward_zero_shift(_arm_fixed(v, FIXED_POINT_FRACTIONAL_BITS * (1 if kind == "weight" else 2)), idx)
            else:
                sym = torch.round(v / step).to(torch.int32)
            rec[kind].append(sym.reshape(-1))
            rec["decoded"][name] = sym.to(torch.float32).reshape(v.shape) * step
        rec["weight"] = torch.cat(rec["weight"]).to(torch.int32)
        rec["bias"] = torch.cat(rec["bias"]).to(torch.int32) if rec["bias"] else None
        out[module_name] = rec
    return out


@torch.no_grad()
def get_ac_max_val_nn(frame_encoder: Any) -> int:
    """ceil(max |round(parameter / step)|) + 2 over every network parameter, the unsent upsampling biases included
    (encode.py:50-125)."""
    from ..utils.misc import POSSIBLE_Q_STEP

    cc = _cc_of(frame_encoder)
    worst = 0.0
    for module_name in cc.modules_to_send:
        for name, v in getattr(cc, module_name).named_parameters():
            kind = "weight" if "weight" in name else "bias"
            step = float(POSSIBLE_Q_STEP[module_name][kind][_q_step_index(cc, module_name, kind)])
            v = v.detach().cpu()
            if module_name == "arm":  # the reference measures the integerised ARM, scaled back to float
                bits = FIXED_POINT_FRACTIONAL_BITS * (1 if kind == "weight" else 2)
                v = _arm_fixed(v, bits).to(torch.float32) / float(1 << bits)
            worst = max(worst, float(torch.round(v / step).abs().max()))
    return int(worst) + 2


def frame_header_bytes(frame_encoder: Any, n_bytes_per_latent: List[int], q_step_index_nn: Dict, scale_index_nn: Dict,
                       n_bytes_nn: Dict, ac_max_val_nn: int, ac_max_val_latent: int, hls_sig_blksize: int,
                       display_order: int = 0, flow_gain: int = 1) -> bytes:
    """Frame header, field by field as header.py:255-467 lays it out (sizes in bytes): header size 2, display index
    1, ARM (dim/8 << 4 | hidden layers) 1, upsampling (kernels << 4 | size) 1, pre-concatenation filters likewise 1,
    synthesis branches 1, synthesis layers 1 + 3 each, flow gain 1, AC_MAX_VAL networks 2 / latents 2, significance
    block size 1 (signed), 6 step indices, 6 exp-Golomb counts, 6 x 2 network sub-stream sizes, latent resolutions 1,
    2-D grids 1, features per resolution 1 each, latent sub-stream sizes 3 each."""
    from ..utils.misc import MAX_AC_MAX_VAL

    cc = _cc_of(frame_encoder)
    p = cc.param
    if p.dim_arm // 8 >= 16 or p.n_hidden_layers_arm >= 16:
        raise ValueError("frame header: dim_arm / 8 and the number of hidden ARM layers must fit 4 bits each")
    if ac_max_val_nn > MAX_AC_MAX_VAL or ac_max_val_latent > MAX_AC_MAX_VAL:
        raise ValueError(f"frame header: AC_MAX_VAL ({ac_max_val_nn}, {ac_max_val_latent}) exceeds {MAX_AC_MAX_VAL}")
    n_grids = len(cc.latent_grids)
    body = bytearray()
    body += int(display_order).to_bytes(1, "big")
    body += ((p.dim_arm // 8) * 16 + p.n_hidden_layers_arm).to_bytes(1, "big")
    body += (((n_grids - 1) << 4) | p.ups_k_size).to_bytes(1, "big")
    body += (((n_grids - 1) << 4) | p.ups_preconcat_k_size).to_bytes(1, "big")
    body += (1).to_bytes(1, "big")  # synthesis branches
    body += len(p.layers_synthesis).to_bytes(1, "big")
    for spec in p.layers_synthesis:
        out_ft, k_size, mode, non_linearity = spec.split("-")
        body += int(out_ft).to_bytes(1, "big") + int(k_size).to_bytes(1, "big")
        body += (_SYNTHESIS_MODE.index(mode) * 16 + _SYNTHESIS_NON_LINEARITY.index(non_linearity)).to_bytes(1, "big")
    body += int(flow_gain).to_bytes(1, "big")
    body += int(ac_max_val_nn).to_bytes(2, "big") + int(ac_max_val_latent).to_bytes(2, "big")
    body += int(hls_sig_blksize).to_bytes(1, "big", signed=True)
    for table, width in ((q_step_index_nn, 1), (scale_index_nn, 1), (n_bytes_nn, 2)):
        for nn_name in ("arm", "upsampling", "synthesis"):
            for kind in ("weight", "bias"):
                body += int(table[nn_name][kind]).to_bytes(width, "big")
    body += n_grids.to_bytes(1, "big") + len(n_bytes_per_latent).to_bytes(1, "big")
    for g in cc.latent_grids:
        body += int(g.data.shape[1]).to_bytes(1, "big")
    for v in n_bytes_per_latent:
        body += int(v).to_bytes(3, "big")
    return (len(body) + 2).to_bytes(2, "big") + bytes(body)


def gop_header_bytes(frame_encoder: Any, intra_period: int = 0, p_period: int = 0) -> bytes:
    """GOP header (header.py:113-160): size 2, height 2, width 2, (bit-depth index << 4 | frame type) 1, intra period
    1, P period 1."""
    cc = _cc_of(frame_encoder)
    h, w = (int(v) for v in cc.param.img_size)
    bitdepth = int(getattr(frame_encoder, "bitdepth", 8))
    data_type = getattr(frame_encoder, "frame_data_type", "rgb")
    body = h.to_bytes(2, "big") + w.to_bytes(2, "big")
    body += (_POSSIBLE_BITDEPTH.index(bitdepth) * 16 + _FRAME_DATA_TYPE.index(data_type)).to_bytes(1, "big")
    body += int(intra_period).to_bytes(1, "big") + int(p_period).to_bytes(1, "big")
    return (len(body) + 2).to_bytes(2, "big") + body


def _reference_wb_coder() -> Callable:
    try:
        from CCLIB.ccencapi import cc_code_wb_bac  # type: ignore
    except ImportError as e:
        raise ImportError("encode_frame(): the reference's CABAC extension `CCLIB.ccencapi` is not importable; build it "
                          "from the reference checkout (setup.py) or pass cc_code_wb_bac=") from e
    return cc_code_wb_bac


@torch.no_grad()
def encode_frame(frame_encoder: Any, bitstream_path: str, hls_sig_blksize: int = 16, display_order: int = 0,
                 flow_gain: int = 1, cc_code_wb_bac: Optional[Callable] = None,
                 cc_code_latent_layer_bac: Optional[Callable] = None) -> Dict[str, Any]:
    """One intra frame -> `<bitstream_path>` = frame header + network sub-streams (ARM, upsampling, synthesis; weights
    then biases) + latent sub-streams, the layout of the reference's `encode_frame` (encode.py:223-623).

    Like the reference, the networks are first replaced by what the decoder will rebuild from the symbols
    (symbol x step; encode.py:443-484) -- a no-op after quantize_model() -- so the latents are coded with the ARM
    the decoder will run.  Returns the header fields and sub-stream sizes."""
    wb_coder = cc_code_wb_bac or _reference_wb_coder()
    cc = _cc_of(frame_encoder)
    ac_max_val_nn = get_ac_max_val_nn(frame_encoder)
    symbols = network_symbols(frame_encoder)
    q_step_index_nn, scale_index_nn, n_bytes_nn = {}, {}, {}
    nn_files = []
    for module_name in cc.modules_to_send:
        rec = symbols[module_name]
        q_step_index_nn[module_name] = dict(rec["q_step_index"])
        scale_index_nn[module_name] = {"weight": 0, "bias": 0}
        n_bytes_nn[module_name] = {"weight": 0, "bias": 0}
        for kind in ("weight", "bias"):
            if rec[kind] is None:
                continue
            path = f"{bitstream_path}_{module_name}_{kind}"
            count = cc.nn_expgol_cnt.get(module_name, {}).get(kind)
            used = wb_coder(path, rec[kind].tolist(), -1 if count is None else int(count))  # -1: the coder searches
            scale_index_nn[module_name][kind] = int(used)
            n_bytes_nn[module_name][kind] = os.path.getsize(path)
            nn_files.append((module_name, kind, path))
        live = dict(getattr(cc, module_name).named_parameters())
        for name, value in rec["decoded"].items():
            live[name].copy_(value.to(live[name].device))

    ac_max_val_latent = get_ac_max_val_latent(frame_encoder)
    n_bytes_per_latent = encode_latents(frame_encoder, bitstream_path, hls_sig_blksize, cc_code_latent_layer_bac)
    header = frame_header_bytes(frame_encoder, n_bytes_per_latent, q_step_index_nn, scale_index_nn, n_bytes_nn,
                                ac_max_val_nn, ac_max_val_latent, hls_sig_blksize, display_order, flow_gain)
    order = {m: i for i, m in enumerate(("arm", "upsampling", "synthesis"))}
    nn_files.sort(key=lambda f: (order[f[0]], 0 if f[1] == "weight" else 1))
    with open(bitstream_path, "wb") as out:
        out.write(header)
        for _, _, path in nn_files:
            with open(path, "rb") as f:
                out.write(f.read())
            os.remove(path)
        for k in range(len(n_bytes_per_latent)):
            path = get_sub_bitstream_path(bitstream_path, k)
            with open(path, "rb") as f:
                out.write(f.read())
            os.remove(path)
    return {"n_bytes": os.path.getsize(bitstream_path), "n_bytes_header": len(header), "n_bytes_nn": n_bytes_nn,
            "n_bytes_per_latent": n_bytes_per_latent, "q_step_index_nn": q_step_index_nn,
            "scale_index_nn": scale_index_nn, "ac_max_val_nn": ac_max_val_nn, "ac_max_val_latent": ac_max_val_latent}


def encode_image(frame_encoder: Any, bitstream_path: str, hls_sig_blksize: int = 16, **coders: Any) -> Dict[str, Any]:
    """A one-frame `.cool` file: GOP header + the intra frame (encode_video, encode.py:166-218, for a single image).
    Returns encode_frame()'s record plus the total size and the rate in bits per pixel."""
    frame_path = f"{bitstream_path}_idx_coding_order"
    info = encode_frame(frame_encoder, frame_path, hls_sig_blksize, **coders)
    with open(bitstream_path, "wb") as out, open(frame_path, "rb") as f:
        out.write(gop_header_bytes(frame_encoder))
        out.write(f.read())
    os.remove(frame_path)
    h, w = (int(v) for v in _cc_of(frame_encoder).param.img_size)
    info["n_bytes"] = os.path.getsize(bitstream_path)
    info["rate_bpp"] = info["n_bytes"] * 8.0 / (h * w)
    return info
