"""train() / test() -- drop-in for the reference's enc/training/train.py:60-374 and the loss part of
enc/training/test.py:370-438, with the inner iteration (train.py:244-268: forward, loss, backward,
clip_grad_norm_(0.1), Adam) executed by ONE fused call (csrc/ccb_api.cu: ccb_train_step) captured in a CUDA graph.

Host-side control flow is the reference's: validation every `frequency_validation` iterations, the
record rule (loss lower AND (delta bpp < 0.001 OR delta PSNR > 0.001), train.py:284-293), the patience
mechanism that reloads the best model + optimiser state under cosine scheduling (train.py:228-242), the
cosine learning-rate step per validation (CosineAnnealingLR with T_max = max_iterations / frequency_validation,
train.py:190-198, 367-368) and the linear softround-temperature / noise-parameter schedules refreshed at
validations (train.py:348-365).  Schedules live in device scalars so the captured graph never changes.

There is no CPU fallback: the frame encoder has to be on a CUDA device.
"""

from __future__ import annotations

import math
import time
from dataclasses import dataclass
from typing import Any, List, Optional, Sequence, Tuple

import torch
from torch import Tensor

from ... import _native as N
from ... import ops
from ..component.frame import FrameEncoder
from .loss import LossFunctionOutput, loss_function

MODULE_MASK = {"latent": N.OPT_LATENT, "arm": N.OPT_ARM, "upsampling": N.OPT_UPS, "synthesis": N.OPT_SYN,
               "all": N.OPT_ALL}


def _linear_schedule(initial_value: float, final_value: float, cur_itr: float, max_itr: float) -> float:
    """train.py:35-57."""
    assert cur_itr >= 0 and cur_itr <= max_itr, (
        f"Linear scheduling from 0 to {max_itr} iterations except to have a current iterations between those two "
        f"values. Found cur_itr = {cur_itr}.")
    return cur_itr * (final_value - initial_value) / max_itr + initial_value


def _cosine_lr(start_lr: float, end_lr: float, n_steps_done: int, t_max: float) -> float:
    """Closed form of torch.optim.lr_scheduler.CosineAnnealingLR after n_steps_done .step() calls."""
    return end_lr + (start_lr - end_lr) * (1.0 + math.cos(math.pi * n_steps_done / t_max)) / 2.0


def _target_of(frame: Any) -> Tensor:
    """Accepts the reference's Frame (frame.data.data), a FrameData-like (.data) or a plain [B,3,H,W] tensor."""
    t = frame
    for _ in range(2):
        if isinstance(t, Tensor):
            break
        t = t.data
    if not isinstance(t, Tensor):
        raise TypeError("frame must be a Tensor or expose .data(.data) as a Tensor (rgb / yuv444)")
    return t


@dataclass
class FrameEncoderLogs(LossFunctionOutput):
    """The subset of test.py:30-120's FrameEncoderLogs that the training loop consumes."""

    lmbda: float = 0.0
    encoding_time_second: float = 0.0
    encoding_iterations_cnt: int = 0
    seq_name: Optional[str] = None

    def pretty_string(self, show_col_name: bool = False, mode: str = "short", additional_data: Optional[dict] = None):
        cols = {"loss": f"{float(self.loss) * 1e3:.6f}", "psnr_db": f"{self.psnr_db:.4f}",
                "rate_latent_bpp": f"{self.rate_latent_bpp:.5f}", "iter": str(self.encoding_iterations_cnt),
                "time_s": f"{self.encoding_time_second:.1f}"}
        cols.update(additional_data or {})
        head = " ".join(f"{k:>14s}" for k in cols) + "\n" if show_col_name else ""
        return head + " ".join(f"{str(v):>14s}" for v in cols.values())


def network_rate_bit(frame_encoder: FrameEncoder) -> float:
    """Bits of the quantised network parameters, summed over modules and weight / bias (test.py:394-399); 0 until
    quantize_model() has chosen the steps."""
    total = 0.0
    for module_rate in frame_encoder.coolchic_encoder.get_network_rate().values():
        for param_rate in module_rate.values():
            total += float(param_rate)
    return total


_call_counter = 0  # train() calls of this process (default noise seed, see train())


@torch.no_grad()
def test(frame_encoder: FrameEncoder, frame: Any, frame_encoder_manager: Any) -> FrameEncoderLogs:
    """Evaluate in eval mode (hard rounding, 8-bit output rounding; test.py:391-437)."""
    was_training = frame_encoder.training
    rate_mlp = network_rate_bit(frame_encoder)
    frame_encoder.set_to_eval()
    target = _target_of(frame)
    out = frame_encoder.forward(quantizer_noise_type="none", quantizer_type="hardround", AC_MAX_VAL=-1,
                                flag_additional_outputs=False)
    lo = loss_function(out.decoded_image, out.rate, target, lmbda=frame_encoder_manager.lmbda, rate_mlp_bit=rate_mlp,
                       compute_logs=True)
    if was_training:
        frame_encoder.set_to_train()
    return FrameEncoderLogs(loss=lo.loss, mse=lo.mse, rate_nn_bpp=lo.rate_nn_bpp, rate_latent_bpp=lo.rate_latent_bpp,
                            lmbda=frame_encoder_manager.lmbda,
                            encoding_time_second=getattr(frame_encoder_manager, "total_training_time_sec", 0.0),
                            encoding_iterations_cnt=getattr(frame_encoder_manager, "iterations_counter", 0),
                            seq_name=getattr(frame, "seq_name", None))


class _FusedTrainer:
    """FusedState + one captured iteration for a FrameEncoder."""

    def __init__(self, frame_encoder: FrameEncoder, target: Tensor, lmbda: float, q_type: str, noise_type: str,
                 mask: int, seed: int):
        cc = frame_encoder.coolchic_encoder
        self.cc = cc
        dev = next(cc.parameters()).device
        if dev.type != "cuda":
            raise N.CcbError("train(): the frame encoder must live on a CUDA device (coolchic_b200 has no CPU path)")
        self.spec = cc._spec()
        # constant during a train() call: weights that are optimised are not quantised yet (rate 0), and once
        # quantize_model() has run only the latents move (train.py:165-182)
        self.rate_mlp_bit = network_rate_bit(frame_encoder)
        # more trainers alive than weight slots (threaded warm-ups of several images in flight): wait for one to
        # finish rather than run the slot-less kernels, so that results do not depend on what else is running
        self.st = ops.FusedState(self.spec, int(target.shape[0]), dev, slot_wait_s=120.0)
        self.target = target.to(dev, torch.float32).contiguous()
        self.push()
        st = self.st
        self._step = lambda: st.train_step(self.target, lmbda, q_type, noise_type, optimize_mask=mask, seed=seed)
        self._step()  # warm-up outside capture (module load, lazy allocations) ...
        # both streams of the step: the caller's and the state's auxiliary one (other host threads may be training
        # other encoders on their own streams meanwhile: no device-wide synchronisation on this path)
        torch.cuda.current_stream(dev).synchronize()
        st.aux_stream.synchronize()
        self.push()  # ... then undo it
        st.arena[1:].zero_()
        st.step.zero_()
        st.philox_offset.zero_()
        # capture_begin / capture_end directly: the `torch.cuda.graph` context manager also runs gc.collect() and
        # torch.cuda.empty_cache() on entry, 0.03-0.5 s per train() call in a process that holds other cached blocks
        self.graph = torch.cuda.CUDAGraph()
        cur = torch.cuda.current_stream(dev)
        side = ops.new_stream(dev)  # private: a pooled stream may be another thread's current stream
        side.wait_stream(cur)
        with ops.CAPTURE_LOCK, torch.cuda.stream(side):
            self.graph.capture_begin(capture_error_mode="thread_local")
            try:
                self._step()
            finally:
                self.graph.capture_end()
        cur.wait_stream(side)
        self.push()
        st.arena[1:].zero_()
        st.step.zero_()
        st.philox_offset.zero_()

    def close(self) -> None:
        """Deterministic release at the end of train(): wait for this trainer's queued work, then give back the weight
        slot, the graph and the private streams HERE, on this thread and outside any capture (the object is part of a
        reference cycle through `_step`, and the cyclic collector may run much later, on any thread)."""
        dev = self.st.device
        torch.cuda.current_stream(dev).synchronize()
        self.st.aux_stream.synchronize()
        self._step = None
        self.graph = None
        self.st.close()

    def push(self) -> None:
        """module parameters -> arena"""
        sd = {k: v for k, v in self.cc.state_dict().items()}
        for name, view in self.st.params.items():
            view.copy_(sd[name].reshape(view.shape))

    def pull(self) -> None:
        """arena -> module parameters (in place, names are the reference's state_dict names)"""
        sd = self.cc.state_dict()
        with torch.no_grad():
            for name, view in self.st.params.items():
                sd[name].copy_(view.reshape(sd[name].shape))

    @torch.no_grad()
    def validate(self, frame_encoder: FrameEncoder, frame: Any, frame_encoder_manager: Any) -> FrameEncoderLogs:
        """test() on the arena: the same eval-mode kernels and the same frame rounding / loss as
        `test(frame_encoder, ...)` after a pull(), without the ~40 tensor copies into the module and back
        (the parameter views ARE the arena).  Numerically identical to test(): same inputs, same launches."""
        v = self.st.params
        raw, rate = ops.coolchic_forward(self.spec, [v[n] for n in self.spec.latent_names()],
                                         [v[n] for n in self.spec.weight_names()], training=False, q_type="hardround",
                                         noise_type="none", ac_max_val=-1)
        max_dynamic = 2 ** (frame_encoder.bitdepth) - 1
        decoded = torch.clamp(torch.round(raw * max_dynamic) / max_dynamic, 0.0, 1.0)  # frame.py:175-183 (eval mode)
        lo = loss_function(decoded, rate, self.target, lmbda=frame_encoder_manager.lmbda,
                           rate_mlp_bit=self.rate_mlp_bit, compute_logs=True)
        return FrameEncoderLogs(loss=lo.loss, mse=lo.mse, rate_nn_bpp=lo.rate_nn_bpp, rate_latent_bpp=lo.rate_latent_bpp,
                                lmbda=frame_encoder_manager.lmbda,
                                encoding_time_second=getattr(frame_encoder_manager, "total_training_time_sec", 0.0),
                                encoding_iterations_cnt=getattr(frame_encoder_manager, "iterations_counter", 0),
                                seq_name=getattr(frame, "seq_name", None))

    def snapshot(self):
        return self.st.arena.clone(), self.st.step.clone()

    def restore(self, snap) -> None:
        self.st.arena.copy_(snap[0])
        self.st.step.copy_(snap[1])


def train(
    frame_encoder: FrameEncoder,
    frame: Any,
    frame_encoder_manager: Any,
    start_lr: float = 1e-2,
    end_lr: float = 0.00001,
    cosine_scheduling_lr: bool = True,
    max_iterations: int = 10000,
    frequency_validation: int = 100,
    patience: int = 10,
    optimized_module: Sequence[str] = ("all",),
    quantizer_type: str = "softround",
    quantizer_noise_type: str = "kumaraswamy",
    softround_temperature: Tuple[float, float] = (0.3, 0.2),
    noise_parameter: Tuple[float, float] = (2.0, 1.0),
    val_logs: Optional[List[Any]] = None,
    seed: Optional[int] = None,
    verbose: bool = False,
) -> FrameEncoder:
    """Same contract as the reference's train() (train.py:60-374): optimises `frame_encoder` in place and returns
    it holding the best parameters seen at a validation.  `frame_encoder_manager` needs `.lmbda`; its
    `.iterations_counter` and `.total_training_time_sec` are updated in place when present."""
    start_time = time.time()
    target = _target_of(frame)
    global _call_counter
    _call_counter += 1
    if seed is None:
        # the reference draws from torch's global generator, so successive calls never replay a noise stream: the
        # default Philox key advances per call and follows torch.manual_seed()
        seed = (torch.initial_seed() + 0x9E3779B97F4A7C15 * _call_counter) & 0xFFFFFFFFFFFFFFFF
    mask = 0
    for m in optimized_module:
        if m not in MODULE_MASK:
            raise ValueError(f"unknown module to optimise: {m}")
        mask |= MODULE_MASK[m]

    initial_logs = test(frame_encoder, frame, frame_encoder_manager)
    best_logs = initial_logs
    frame_encoder.set_to_train()
    tr = _FusedTrainer(frame_encoder, target, frame_encoder_manager.lmbda, quantizer_type, quantizer_noise_type, mask,
                       seed)
    try:
        st = tr.st
        best = tr.snapshot()

        t_max = max_iterations / frequency_validation
        n_sched = 0
        cur_lr = start_lr
        st.lr.fill_(cur_lr)
        st.T.fill_(_linear_schedule(softround_temperature[0], softround_temperature[1], 0, max_iterations))
        st.sigma.fill_(_linear_schedule(noise_parameter[0], noise_parameter[1], 0, max_iterations))

        cnt_record = 0
        show_col_name = True
        for cnt in range(max_iterations):
            if cnt - cnt_record > patience:
                if cosine_scheduling_lr:
                    tr.restore(best)  # best model + its Adam moments / step; the learning rate stays the current one
                    cnt_record = cnt
                else:
                    break
            tr.graph.replay()
            if hasattr(frame_encoder_manager, "iterations_counter"):
                frame_encoder_manager.iterations_counter += 1

            if ((cnt + 1) % frequency_validation == 0) or (cnt + 1 == max_iterations):
                if hasattr(frame_encoder_manager, "total_training_time_sec"):
                    torch.cuda.current_stream().synchronize()
                    frame_encoder_manager.total_training_time_sec += time.time() - start_time
                    start_time = time.time()
                logs = tr.validate(frame_encoder, frame, frame_encoder_manager)
                new_record = False
                if logs.loss < best_logs.loss:
                    delta_psnr = logs.psnr_db - best_logs.psnr_db
                    delta_bpp = logs.rate_latent_bpp - best_logs.rate_latent_bpp
                    new_record = delta_bpp < 0.001 or delta_psnr > 0.001
                record = ""
                if new_record:
                    best = tr.snapshot()
                    record = (f"{logs.rate_latent_bpp - initial_logs.rate_latent_bpp:+6.3f} bpp "
                              f"{logs.psnr_db - initial_logs.psnr_db:+6.3f} db")
                    best_logs = logs
                    cnt_record = cnt
                if verbose:
                    print(logs.pretty_string(show_col_name=show_col_name, additional_data={
                        "lr": f"{cur_lr:.8f}", "optim": ",".join(optimized_module),
                        "patience": (patience - cnt + cnt_record) // frequency_validation, "q_type": quantizer_type,
                        "sr_temp": f"{float(st.T):.5f}", "n_type": quantizer_noise_type, "noise": f"{float(st.sigma):.2f}",
                        "record": record}))
                    show_col_name = False
                if val_logs is not None:
                    val_logs.append(logs)
                st.T.fill_(_linear_schedule(softround_temperature[0], softround_temperature[1], cnt, max_iterations))
                st.sigma.fill_(_linear_schedule(noise_parameter[0], noise_parameter[1], cnt, max_iterations))
                if cosine_scheduling_lr:
                    n_sched += 1
                    cur_lr = _cosine_lr(start_lr, end_lr, n_sched, t_max)
                    st.lr.fill_(cur_lr)
                frame_encoder.set_to_train()

        tr.restore(best)
        tr.pull()
    finally:
        tr.close()
    return frame_encoder
