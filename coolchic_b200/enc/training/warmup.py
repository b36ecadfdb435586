"""warmup() -- the candidate search that opens a preset (reference: enc/training/warmup.py:21-158; described by
`preset.warmup.phases`, e.g. preset_cfg/c3x.yaml: 5 candidates x 400 iterations, then the best 2 x 400 more).

Same contract as the reference: every phase keeps the `candidates` best encoders of the previous one (all of them in
the first phase), trains each with the phase's `training_phase` through train(), scores it with test(), and sorts by
loss; the winner is returned as a deep copy.  `frame_encoder_manager` (its `.preset.warmup`, `.lmbda`, and the
iteration / time counters train() advances) is the reference's object or anything shaped like it; phases and training
phases may be attribute objects (the reference's pydantic models) or plain dicts (a parsed YAML).

The candidates are independent until they are ranked, and one iteration uses well under half of the
GPU's issue slots, so a phase trains ALL its candidates at the same time: one host thread and one CUDA stream per
candidate, each replaying its own iteration graph on its own weight slot (csrc: ccb_weight_slot_acquire).  Results are
those of the sequential order bit for bit (tests/test_gpu_boundary.py); `concurrent=False` runs them one after the
other.  Unlike the reference the candidates are not shuttled to the CPU between phases: a 768x512 encoder is 2 MB.
"""

from __future__ import annotations

import copy
import threading
import time
import types
from typing import Any, List

import torch

from ..component.frame import FrameEncoder
from .train import test, train


def _get(obj: Any, name: str, default: Any = None) -> Any:
    if isinstance(obj, dict):
        return obj.get(name, default)
    return getattr(obj, name, default)


def _train_kwargs(tp: Any) -> dict:
    """TrainerPhase -> train() keywords, with the reference's defaults (presets.py:25-120)."""
    return dict(
        start_lr=float(_get(tp, "lr", 1e-2)),
        cosine_scheduling_lr=bool(_get(tp, "schedule_lr", False)),
        max_iterations=int(_get(tp, "max_itr", 10000)),
        patience=int(_get(tp, "patience", 1000)),
        frequency_validation=int(_get(tp, "freq_valid", 100)),
        optimized_module=tuple(_get(tp, "optimized_module", ["all"])),
        quantizer_type=_get(tp, "quantizer_type", "softround"),
        quantizer_noise_type=_get(tp, "quantizer_noise_type", "kumaraswamy"),
        softround_temperature=tuple(_get(tp, "softround_temperature", (0.3, 0.3))),
        noise_parameter=tuple(_get(tp, "noise_parameter", (2.0, 1.0))),
    )


def _train_candidates(cands: List[dict], frame: Any, mgr: Any, device: Any, idx_phase: int, kwargs: dict,
                      concurrent: bool) -> None:
    """train() + test() of every candidate of one phase; concurrent: one thread + one stream per candidate."""

    def work(cand: dict, my_mgr: Any, stream) -> None:
        enc = cand["encoder"]
        enc.to_device(device)
        with torch.cuda.stream(stream) if stream is not None else _NullCtx():
            enc = train(frame_encoder=enc, frame=frame, frame_encoder_manager=my_mgr,
                        seed=1000 * idx_phase + cand["id"], **kwargs)
            cand["encoder"] = enc
            cand["metrics"] = test(enc, frame, my_mgr)
            if stream is not None:
                stream.synchronize()

    if not concurrent or len(cands) == 1 or not torch.cuda.is_available():
        for cand in cands:
            work(cand, mgr, None)
        return
    # a private counter object per thread (`x += 1` on a shared attribute is not atomic), folded back afterwards
    shims = [types.SimpleNamespace(lmbda=mgr.lmbda, iterations_counter=0, total_training_time_sec=0.0) for _ in cands]
    dev = torch.device(device)
    cur = torch.cuda.current_stream(dev)
    from ... import ops

    streams = [ops.new_stream(dev) for _ in cands]  # private streams (torch's pool of 32 is shared by every thread)
    errors: List[BaseException] = []

    def guarded(*a):
        try:
            work(*a)
        except BaseException as e:  # re-raised on the calling thread
            errors.append(e)

    threads = []
    for cand, shim, s in zip(cands, shims, streams):
        s.wait_stream(cur)
        t = threading.Thread(target=guarded, args=(cand, shim, s), daemon=True)
        t.start()
        threads.append(t)
    for t in threads:
        t.join()
    for s in streams:
        cur.wait_stream(s)
    if errors:
        raise errors[0]
    if hasattr(mgr, "iterations_counter"):
        mgr.iterations_counter += sum(sh.iterations_counter for sh in shims)
    if hasattr(mgr, "total_training_time_sec"):  # wall time of the phase, not the sum over candidates
        mgr.total_training_time_sec += max(sh.total_training_time_sec for sh in shims)


class _NullCtx:
    def __enter__(self):
        return None

    def __exit__(self, *a):
        return False


def warmup(frame_encoder_manager: Any, list_candidates: List[FrameEncoder], frame: Any, device: Any,
           verbose: bool = False, concurrent: bool = True) -> FrameEncoder:
    start_time = time.time()
    phases = _get(_get(_get(frame_encoder_manager, "preset"), "warmup"), "phases", [])
    all_candidates = [{"metrics": None, "id": i, "encoder": enc} for i, enc in enumerate(list_candidates)]
    if not all_candidates:
        raise ValueError("warmup(): no candidate")
    for idx_phase, phase in enumerate(phases):
        n_keep = int(_get(phase, "candidates"))
        if n_keep > len(all_candidates):
            raise ValueError(f"warmup(): phase {idx_phase} wants {n_keep} candidates, {len(all_candidates)} are left")
        if idx_phase != 0:
            del all_candidates[n_keep:]  # sorted by loss at the end of the previous phase: the worst go
        kwargs = _train_kwargs(_get(phase, "training_phase"))
        _train_candidates(all_candidates[:n_keep], frame, frame_encoder_manager, device, idx_phase, kwargs, concurrent)
        all_candidates.sort(key=lambda c: float(c["metrics"].loss))
        if verbose:
            print(f"\nPerformance at the end of warm-up phase {idx_phase}:\n")
            print(f'{"ID":^6}|{"loss x 1e3":^14}|{"rate_bpp":^14}|{"psnr_db":^14}|')
            for c in all_candidates:
                m = c["metrics"]
                print(f'{c["id"]:^6}|{float(m.loss) * 1e3:^14.4f}|{m.rate_latent_bpp:^14.4f}|{m.psnr_db:^14.4f}|')
    winner = all_candidates[0]
    frame_encoder = copy.deepcopy(winner["encoder"])
    if verbose:
        print(f"Intra warm-up time [s]: {time.time() - start_time:.2f}   winner ID: {winner['id']}")
    return frame_encoder
