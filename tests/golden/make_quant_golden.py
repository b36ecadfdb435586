"""Golden vectors for quantize_model (the NN weight quantisation search) from the REAL reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_quant_golden.py

Builds the reference's FrameEncoder with a seeded parameter set, runs the reference's own
`quantize_model(frame_encoder, frame, frame_encoder_manager)` (coolchic/enc/training/quantizemodel.py) and records,
for EVERY candidate pair of quantisation steps it scores, the loss it computed (tapped from its `loss_function`
call), plus the chosen steps / exp-Golomb exponents and the final quantised parameters.  Stored in
tests/golden/quant_<case>.npz next to the initial parameters and the target.
"""

import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from refimport import import_reference  # noqa: E402

import_reference()

import coolchic.enc.training.quantizemodel as RQ  # noqa: E402

from make_golden import build_reference  # noqa: E402
from oracle import float_ref as O  # noqa: E402

# (name, arch, H, W, lmbda, seed)
CASES = [("quant_hop_40x56", "hop", 40, 56, 1e-3, 11), ("quant_lop_24x40", "lop", 24, 40, 2e-2, 5)]


def run_case(name, arch_name, H, W, lmbda, seed):
    torch.manual_seed(99)
    fe, arch, p0 = build_reference(arch_name, H, W, 1, 16, seed)
    target = O.synthetic_image(3, H, W, 1)
    frame = types.SimpleNamespace(data=types.SimpleNamespace(data=target), refs_data=[])
    mgr = types.SimpleNamespace(lmbda=lmbda, total_training_time_sec=0.0)
    trace = []
    real_loss = RQ.loss_function

    def tapped(*a, **k):
        out = real_loss(*a, **k)
        cc = fe.coolchic_encoder
        trace.append((float(out.loss), float(k.get("rate_mlp_bit", 0.0))))
        return out

    RQ.loss_function = tapped
    try:
        RQ.quantize_model(fe, frame, mgr)
    finally:
        RQ.loss_function = real_loss
    cc = fe.coolchic_encoder
    out = {"target": target.numpy(), "H": H, "W": W, "arch": arch_name, "lmbda": lmbda,
           "trace_loss": np.array([t[0] for t in trace], dtype=np.float64),
           "trace_rate_mlp": np.array([t[1] for t in trace], dtype=np.float64)}
    for k, v in p0.items():
        out["p0/" + k] = v.numpy()
    for m in cc.modules_to_send:
        out[f"q_step/{m}"] = np.array([float(cc.nn_q_step[m]["weight"]), float(cc.nn_q_step[m]["bias"])])
        out[f"expgol/{m}"] = np.array([int(cc.nn_expgol_cnt[m]["weight"]), int(cc.nn_expgol_cnt[m]["bias"])])
    for k, v in fe.named_parameters():
        out["p_final/" + k.removeprefix("coolchic_encoder.")] = v.detach().numpy()
    fe.set_to_eval()
    with torch.no_grad():
        o = fe.forward(quantizer_noise_type="none", quantizer_type="hardround", AC_MAX_VAL=-1)
        rate_mlp = sum(float(r) for mod in cc.get_network_rate().values() for r in mod.values())
        l = RQ.loss_function(o.decoded_image, o.rate, target, lmbda=lmbda, rate_mlp_bit=rate_mlp, compute_logs=True)
    out["final_loss"] = np.float64(float(l.loss))
    out["final_rate_mlp"] = np.float64(rate_mlp)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    return len(trace)


if __name__ == "__main__":
    for c in CASES:
        n = run_case(*c)
        print("wrote", c[0], n, "candidates,", os.path.getsize(os.path.join(HERE, c[0] + ".npz")) // 1024, "kB")
