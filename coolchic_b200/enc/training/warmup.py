"""warmup() -- the candidate search that opens a preset (reference: enc/training/warmup.py:21-158; described by
`preset.warmup.phases`, e.g. preset_cfg/c3x.yaml: 5 candidates x 400 iterations, then the best 2 x 400 more).

Same contract as the reference: every phase keeps the `candidates` best encoders of the previous one (all of them in
the first phase), trains each with the phase's `training_phase` through train(), scores it with test(), and sorts by
loss; the winner is returned as a deep copy.  `frame_encoder_manager` (its `.preset.warmup`, `.lmbda`, and the
iteration / time counters train() advances) is the reference's object or anything shaped like it; phases and training
phases may be attribute objects (the reference's pydantic models) or plain dicts (a parsed YAML).

The candidates are independent until they are ranked; here each runs on the fused iteration graph one after the
other (SURVEY.md 8f-2 lists running them as one batch with per-sample weights as the next step).  Unlike the
reference the candidates are not shuttled to the CPU between phases: a 768x512 encoder is 2 MB.
"""

from __future__ import annotations

import copy
import time
from typing import Any, List

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


def warmup(frame_encoder_manager: Any, list_candidates: List[FrameEncoder], frame: Any, device: Any,
           verbose: bool = False) -> FrameEncoder:
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
        for i in range(n_keep):
            cand = all_candidates[i]
            enc = cand["encoder"]
            enc.to_device(device)
            enc = train(frame_encoder=enc, frame=frame, frame_encoder_manager=frame_encoder_manager,
                        seed=1000 * idx_phase + cand["id"], **kwargs)
            cand["encoder"] = enc
            cand["metrics"] = test(enc, frame, frame_encoder_manager)
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
