"""Where does the REFERENCE's own encode loop end?  (needs /root/reference; run in the build container)

    python tests/golden/make_convergence_golden.py

Runs the reference's `warmup()` + `train()` per phase + `quantize_model()` + `test()` exactly as its encode loop chains
them (enc/component/video.py:281-312) on a 64x96 synthetic image, on the CPU, for

  * `debug`: preset_cfg/debug.yaml verbatim (110 iterations: recorded, not gated -- its end point is mostly the noise
    stream one happened to draw);
  * `mini`: the same shape of schedule, long enough to converge (warm-up 3 x 100 + 2 x 100, then 1500 softround /
    gaussian iterations with cosine learning rate, 200 ste iterations + quantize_model, 100 latent-only iterations),

each for several torch seeds: the file keeps every run's final loss / PSNR / rates, i.e. the reference's OWN run-to-run
spread.  tests/test_gpu_convergence.py runs the repo's warmup() / train() / quantize_model() over the same schedule
dicts (tests/common.py:CONVERGENCE_PRESETS) and must land inside that spread (the noise comes from another generator,
so only the distribution of end points can agree).  This pins the host logic end to end: schedules, validation and
record rule, patience reload, best-model bookkeeping, candidate selection, the quantisation search.
"""

import copy
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from refimport import import_reference  # noqa: E402

import_reference()

from coolchic.enc.component.coolchic import CoolChicEncoderParameter  # noqa: E402
from coolchic.enc.component.frame import FrameEncoder  # noqa: E402
from coolchic.enc.training.quantizemodel import quantize_model  # noqa: E402
from coolchic.enc.training.test import test  # noqa: E402
from coolchic.enc.training.train import train  # noqa: E402
from coolchic.enc.training.warmup import warmup  # noqa: E402
from coolchic.enc.utils.codingstructure import Frame, FrameData  # noqa: E402
from coolchic.enc.utils.manager import FrameEncoderManager  # noqa: E402
from coolchic.utils.types import PresetConfig  # noqa: E402

import common as Cm  # noqa: E402

# mini: 18 seeds.  Six looked like enough until the next twelve showed the reference's real spread (35.4 .. 38.0 dB at
# equal loss); the committed file holds seeds 0-5 from this script's first run and 6-17 from the same run() later
# (their loss rebuilt as mse + lambda * rate from the logged PSNR / rates).
SEEDS = {"debug": [0, 1, 2, 3, 4, 5], "mini": list(range(18))}


def run(preset_name: str, seed: int):
    H, W = Cm.CONVERGENCE_HW
    a = Cm.ARCHS["hop"]
    torch.manual_seed(seed)
    torch.set_num_threads(2)
    preset = PresetConfig(**copy.deepcopy(Cm.CONVERGENCE_PRESETS[preset_name]))
    mgr = FrameEncoderManager(preset_config=preset, lmbda=Cm.CONVERGENCE_LMBDA)
    target = Cm.O.synthetic_image(Cm.CONVERGENCE_IMAGE, H, W)
    frame = Frame(coding_order=0, display_order=0, data=FrameData(bitdepth=8, frame_data_type="rgb", data=target))
    prm = CoolChicEncoderParameter(layers_synthesis=list(a["syn"]), n_ft_per_res=[1] * 7, dim_arm=a["dim_arm"],
                                   n_hidden_layers_arm=a["nh"], encoder_gain=16)
    prm.set_image_size((H, W))
    cands = [FrameEncoder(prm) for _ in range(preset.warmup.phases[0].candidates)]
    fe = warmup(frame_encoder_manager=mgr, list_candidates=cands, frame=frame, device="cpu")
    fe.to_device("cpu")
    for ph in preset.all_phases:  # enc/component/video.py:281-312
        fe = train(frame_encoder=fe, frame=frame, frame_encoder_manager=mgr, start_lr=ph.lr,
                   cosine_scheduling_lr=ph.schedule_lr, max_iterations=ph.max_itr, frequency_validation=ph.freq_valid,
                   patience=ph.patience, optimized_module=ph.optimized_module, quantizer_type=ph.quantizer_type,
                   quantizer_noise_type=ph.quantizer_noise_type, softround_temperature=ph.softround_temperature,
                   noise_parameter=ph.noise_parameter)
        if ph.quantize_model:
            fe = quantize_model(encoder=fe, frame=frame, frame_encoder_manager=mgr)
    logs = test(fe, frame, mgr)
    lo = logs.loss_function_output
    return dict(loss=float(lo.loss), psnr_db=float(lo.psnr_db), rate_latent_bpp=float(lo.rate_latent_bpp),
                rate_nn_bpp=float(lo.rate_nn_bpp), iterations=int(mgr.iterations_counter))


if __name__ == "__main__":
    import contextlib
    import io

    import wandb

    wandb.init(mode="disabled")  # the reference's train() calls wandb.log() unconditionally (train.py:337)

    out = {}
    for name in ("debug", "mini"):
        rows = []
        for seed in SEEDS[name]:
            t0 = time.time()
            with contextlib.redirect_stdout(io.StringIO()):
                r = run(name, seed)
            rows.append(r)
            print(name, seed, {k: round(v, 5) for k, v in r.items()}, f"{time.time() - t0:.0f} s", flush=True)
        for k in rows[0]:
            out[f"{name}/{k}"] = np.array([r[k] for r in rows], dtype=np.float64)
        out[f"{name}/seeds"] = np.array(SEEDS[name])
    np.savez_compressed(os.path.join(HERE, "conv_reference_runs.npz"), **out)
    print("wrote conv_reference_runs.npz")
