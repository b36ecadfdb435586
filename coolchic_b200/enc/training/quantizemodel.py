"""quantize_model -- the NN weight quantisation search that follows the STE phase of a preset
(`quantize_model: True`, preset_cfg/c3x.yaml), mirror of the reference's enc/training/quantizemodel.py:34-616.

Contract (same as the reference): for each network sent to the decoder, in sorted name order, try every pair of
power-of-two steps (weights, biases) of POSSIBLE_Q_STEP, quantise the network (`round(p / step) * step`, a pair is
skipped when a sent value exceeds MAX_AC_MAX_VAL), pick per kind the exp-Golomb exponent with the fewest bits,
evaluate `loss_function(decoded, rate, target, lmbda, rate_mlp_bit)` with an eval-mode forward
(`quantizer_type="hardround"`, no noise), keep the best pair, and leave the network quantised with it; the chosen
steps / exponents end up in `nn_q_step` / `nn_expgol_cnt`.  About 490 full-size eval forwards per image: on the
fused kernels that is a second or two instead of minutes.

What differs from the reference's implementation (not from its results): the two entry points share one search
routine, the exp-Golomb exponents of a candidate are scored in one vectorised expression instead of 2 x 13 calls,
and an optional `trace` list receives one record per candidate (used by the parity tests)."""

from __future__ import annotations

import itertools
import time
from typing import Any, Callable, List, Optional, OrderedDict, Tuple

import torch
from torch import Tensor

from ..component.coolchic import CoolChicEncoder
from ..component.frame import FrameEncoder
from ..utils.misc import (MAX_AC_MAX_VAL, POSSIBLE_EXP_GOL_COUNT, POSSIBLE_Q_STEP, DescriptorNN,
                          get_q_step_from_parameter_name)
from .loss import loss_function


def _quantize_parameters(fp_param: OrderedDict[str, Tensor], q_step: DescriptorNN) -> Optional[OrderedDict[str, Tensor]]:
    """`round(v / step) * step` per tensor, None when a sent value does not fit the header field
    (quantizemodel.py:34-68).  An infinite step sends zeros."""
    out = type(fp_param)()
    for name, value in fp_param.items():
        step = get_q_step_from_parameter_name(name, q_step)
        sent = torch.round(value / step)
        if sent.abs().max() > MAX_AC_MAX_VAL:
            return None
        out[name] = torch.zeros_like(value) if step == float("inf") else sent * step
    return out


def _best_expgol_counts(param: OrderedDict[str, Tensor], q_step: DescriptorNN, counts: DescriptorNN,
                        kinds: Tuple[str, ...] = ("weight", "bias")):
    """Per kind, the exp-Golomb exponent with the fewest bits for the sent symbols (first minimum, like the
    reference's strict `<` scan: quantizemodel.py:237-263) and those bits; all exponents scored at once."""
    best, bits = {}, {}
    for kind in kinds:
        sent = [(v / q_step[kind]).reshape(-1) for k, v in param.items() if kind in k]
        v = torch.cat(sent)
        c = counts[kind].to(v.device)[:, None]  # [n_counts, 1]
        nb = (2 * torch.floor(torch.log2(v.abs()[None, :] / (2.0**c) + 1)) + c + 1 + (v != 0)[None, :]).sum(dim=1)
        i = int(torch.argmin(nb))  # argmin returns the first minimum
        best[kind] = int(counts[kind][i])
        bits[kind] = nb[i]
    return best, bits


class _CandidateScore:
    """loss_function(evaluate(), target, lmbda, rate_mlp_bit) as a float, with the eval forward, the frame rounding and
    the loss arithmetic captured ONCE in a CUDA graph: a candidate then costs one replay and one read-back instead of
    ~25 launches driven from Python (0.45 ms of host time around 0.1 ms of kernels).  The parameters are read at fixed
    addresses (the search updates them in place) and `rate_mlp_bit` is a device scalar, so every candidate runs exactly
    the kernels, in the order and with the operands, of the eager call."""

    def __init__(self, evaluate: Callable[[], Tuple[Tensor, Tensor]], target: Tensor, lmbda: float):
        self.evaluate, self.target, self.lmbda = evaluate, target, lmbda
        self.graph = None
        self.rate_mlp = torch.zeros((), device=target.device, dtype=torch.float32)
        self.loss = None

    def _eager(self) -> Tensor:
        decoded, rate = self.evaluate()
        return loss_function(decoded, rate, self.target, lmbda=self.lmbda, rate_mlp_bit=self.rate_mlp,
                             compute_logs=False).loss

    def __call__(self, rate_mlp: float) -> float:
        self.rate_mlp.fill_(rate_mlp)
        if self.graph is None and self.target.is_cuda:
            self._eager()  # lazy initialisation outside the capture
            cur = torch.cuda.current_stream(self.target.device)
            from ... import ops

            side = ops.new_stream(self.target.device)  # private: pooled streams are shared between host threads
            side.wait_stream(cur)
            graph = torch.cuda.CUDAGraph()
            with ops.CAPTURE_LOCK, torch.cuda.stream(side):
                graph.capture_begin(capture_error_mode="thread_local")
                try:
                    self.loss = self._eager()
                finally:
                    graph.capture_end()
            cur.wait_stream(side)
            self.graph = graph
        if self.graph is None:
            return float(self._eager())
        self.graph.replay()
        return float(self.loss)


def _kind(name: str) -> str:
    return "weight" if ".weight" in name else "bias"


@torch.no_grad()
def _search(cc: Any, evaluate: Callable[[], Tuple[Tensor, Tensor]], target: Tensor, lmbda: float,
            trace: Optional[List[dict]]) -> None:
    """The candidate bookkeeping (quantisation of <= 1 k values, exp-Golomb scoring) runs on HOST copies of the
    parameters -- on the device it would be ~100 tiny launches and a dozen synchronisations per candidate, ten times
    the cost of the full-size eval forward it surrounds.  A candidate is a PAIR of steps, but the quantised weights
    (and their exp-Golomb score) depend on the weight step alone and the biases on the bias step alone: each is
    computed and uploaded once per step (13-17 + 1-25 times) instead of once per pair (up to 325 times), and a
    candidate only copies the tensors of the kind whose step changed into the live parameters."""
    score = _CandidateScore(evaluate, target, lmbda)
    for module_name in sorted(cc.modules_to_send):
        module = getattr(cc, module_name)
        steps, counts = POSSIBLE_Q_STEP[module_name], POSSIBLE_EXP_GOL_COUNT[module_name]
        live = dict(module.named_parameters())
        fp_param = type(module.get_param())((k, v.cpu()) for k, v in module.get_param().items())
        assert set(live) == set(fp_param)
        by_kind = {kind: type(fp_param)((k, v) for k, v in fp_param.items() if _kind(k) == kind)
                   for kind in ("weight", "bias")}
        # bits of the OTHER networks (already quantised, or 0 while their steps are None): constant during this search
        cc.nn_q_step[module_name] = {"weight": None, "bias": None}
        rate_others = 0.0
        for per_module in cc.get_network_rate().values():
            for r in per_module.values():
                rate_others += float(r)

        cache: dict = {}

        def per_step(kind: str, step):
            """(quantised tensors on the device | None if a value does not fit, best exponent, its bits)"""
            key = (kind, float(step))
            if key not in cache:
                q = _quantize_parameters(by_kind[kind], {kind: step})
                if q is None:
                    cache[key] = (None, None, None)
                else:
                    expgol, bits = _best_expgol_counts(q, {kind: step}, counts, kinds=(kind,))
                    dev_q = {k: v.to(live[k].device, non_blocking=True) for k, v in q.items()}
                    cache[key] = (dev_q, expgol[kind], float(bits[kind]))
            return cache[key]

        best = (1e6, {}, {})  # loss, steps, exponents
        loaded = {"weight": None, "bias": None}
        for q_w, q_b in itertools.product(steps["weight"], steps["bias"]):
            cand: DescriptorNN = {"weight": q_w, "bias": q_b}
            per_kind = {kind: per_step(kind, cand[kind]) for kind in ("weight", "bias")}
            if per_kind["weight"][0] is None or per_kind["bias"][0] is None:
                continue
            for kind in ("weight", "bias"):
                if loaded[kind] != float(cand[kind]):
                    for k, v in per_kind[kind][0].items():
                        live[k].copy_(v)
                    loaded[kind] = float(cand[kind])
            cc.nn_q_step[module_name] = cand
            expgol = {kind: per_kind[kind][1] for kind in ("weight", "bias")}
            cc.nn_expgol_cnt[module_name] = expgol
            rate_mlp = rate_others + per_kind["weight"][2] + per_kind["bias"][2]  # integer-valued: order-free
            loss = score(rate_mlp)
            if trace is not None:
                trace.append({"module": module_name, "q_w": float(q_w), "q_b": float(q_b), "loss": loss,
                              "expgol_w": expgol["weight"], "expgol_b": expgol["bias"], "rate_mlp": float(rate_mlp)})
            if loss < best[0]:
                best = (loss, cand, expgol)
        cc.nn_q_step[module_name], cc.nn_expgol_cnt[module_name] = best[1], best[2]
        q_param = _quantize_parameters(fp_param, best[1])
        assert q_param is not None, f"_quantize_parameters() failed with q_step {best[1]}"
        module.set_param(q_param)


def quantize_model_frame_encoder(frame_encoder: FrameEncoder, frame: Any, frame_encoder_manager: Any,
                                 trace: Optional[List[dict]] = None, verbose: bool = True) -> FrameEncoder:
    """(quantizemodel.py:118-314)  `frame` is the reference's Frame (`.data.data`, `.refs_data`) or the target tensor;
    `frame_encoder_manager.total_training_time_sec` is advanced in place."""
    from .train import _target_of

    start = time.time()
    frame_encoder.set_to_eval()
    target = _target_of(frame)
    refs = [r.data for r in getattr(frame, "refs_data", [])]

    def evaluate():
        out = frame_encoder.forward(reference_frames=refs, quantizer_noise_type="none", quantizer_type="hardround",
                                    AC_MAX_VAL=-1, flag_additional_outputs=False)
        return out.decoded_image, out.rate

    _search(frame_encoder.coolchic_encoder, evaluate, target.to(next(frame_encoder.parameters()).device),
            frame_encoder_manager.lmbda, trace)
    dt = time.time() - start
    if verbose:  # the reference prints this line (quantizemodel.py:311)
        print(f"\nTime quantize_model(): {dt:4.1f} seconds\n")
    if hasattr(frame_encoder_manager, "total_training_time_sec"):
        frame_encoder_manager.total_training_time_sec += dt
    return frame_encoder


def quantize_model_coolchic_encoder(cc_enc: CoolChicEncoder, input_img: Tensor, lmbda: float,
                                    trace: Optional[List[dict]] = None) -> CoolChicEncoder:
    """(quantizemodel.py:317-447)  Scores the raw synthesis output (no frame clamp / 8-bit rounding)."""
    cc_enc.eval()

    def evaluate():
        raw, rate, _ = cc_enc.forward(quantizer_noise_type="none", quantizer_type="hardround", AC_MAX_VAL=-1,
                                      flag_additional_outputs=False)
        return raw, rate

    _search(cc_enc, evaluate, input_img, lmbda, trace)
    return cc_enc


def quantize_model(encoder: Any, frame: Any = None, frame_encoder_manager: Any = None, input_img: Optional[Tensor] = None,
                   lmbda: Optional[float] = None, trace: Optional[List[dict]] = None, verbose: bool = True):
    """Dispatch on the encoder type, same keyword arguments as the reference (quantizemodel.py:93-115)."""
    if isinstance(encoder, FrameEncoder):
        assert frame is not None, "frame must be provided if encoder is a FrameEncoder"
        assert frame_encoder_manager is not None, "frame_encoder_manager must be provided if encoder is a FrameEncoder"
        return quantize_model_frame_encoder(encoder, frame, frame_encoder_manager, trace=trace, verbose=verbose)
    if isinstance(encoder, CoolChicEncoder):
        assert input_img is not None, "input_img must be provided if encoder is a CoolChicEncoder"
        assert lmbda is not None, "lmbda must be provided if encoder is a CoolChicEncoder"
        return quantize_model_coolchic_encoder(encoder, input_img, lmbda, trace=trace)
    raise ValueError("encoder must be either a FrameEncoder or a CoolChicEncoder")
