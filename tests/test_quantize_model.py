"""quantize_model (enc/training/quantizemodel.py) against the REAL reference's own search, recorded candidate by
candidate in tests/golden/quant_*.npz (tests/golden/make_quant_golden.py).

CPU tests: the host logic (step quantisation, exp-Golomb exponent choice, NN rate) reproduces the reference's final
parameters and choices exactly.  GPU test: the whole search through the fused eval forward scores every candidate
within 1e-4 (relative) of the reference's loss, picks the same steps unless the reference's own losses tie within
that tolerance, and ends at the same loss."""

import collections
import os
import types

import numpy as np
import pytest
import torch

from common import ARCHS, GOLDEN_DIR

from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
from coolchic_b200.enc.component.frame import FrameEncoder
from coolchic_b200.enc.training import quantizemodel as Q
from coolchic_b200.enc.utils.misc import POSSIBLE_EXP_GOL_COUNT, measure_expgolomb_rate

CASES = ["quant_hop_40x56", "quant_lop_24x40"]
MODULES = ["arm", "synthesis", "upsampling"]  # sorted: the order of the search


def _load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False)
    a = ARCHS[str(z["arch"])]
    prm = CoolChicEncoderParameter(layers_synthesis=list(a["syn"]), n_ft_per_res=[1] * 7, dim_arm=a["dim_arm"],
                                   n_hidden_layers_arm=a["nh"], encoder_gain=16)
    prm.set_image_size((int(z["H"]), int(z["W"])))
    fe = FrameEncoder(prm)
    fe.load_state_dict({"coolchic_encoder." + k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("p0/")})
    return z, fe


@pytest.mark.parametrize("name", CASES)
def test_host_logic_reproduces_the_reference_choice(name):
    z, fe = _load(name)
    cc = fe.coolchic_encoder
    for m in MODULES:
        module = getattr(cc, m)
        step = {"weight": torch.tensor(z[f"q_step/{m}"][0], dtype=torch.float32),
                "bias": torch.tensor(z[f"q_step/{m}"][1], dtype=torch.float32)}
        q = Q._quantize_parameters(module.get_param(), step)
        assert q is not None
        for k, v in q.items():
            ref = torch.from_numpy(z[f"p_final/{m}.{k}"])
            assert torch.equal(v.reshape(ref.shape), ref), (m, k)
        module.set_param(q)
        expgol, bits = Q._best_expgol_counts(module.get_param(), step, POSSIBLE_EXP_GOL_COUNT[m])
        assert [expgol["weight"], expgol["bias"]] == z[f"expgol/{m}"].tolist(), m
        cc.nn_q_step[m], cc.nn_expgol_cnt[m] = step, expgol
        rate = measure_expgolomb_rate(module, step, expgol)
        assert float(rate["weight"]) == float(bits["weight"]) and float(rate["bias"]) == float(bits["bias"])
    total = sum(float(r) for mod in cc.get_network_rate().values() for r in mod.values())
    assert total == float(z["final_rate_mlp"])


def test_step_too_small_is_rejected():
    p = collections.OrderedDict({"mlp.0.weight": torch.tensor([70000.0 * 2.0**-4]), "mlp.0.bias": torch.tensor([0.5])})
    assert Q._quantize_parameters(p, {"weight": 2.0**-4, "bias": 1.0}) is None
    q = Q._quantize_parameters(p, {"weight": 1.0, "bias": float("inf")})
    assert q["mlp.0.bias"].item() == 0.0 and q["mlp.0.weight"].item() == float(round(70000.0 * 2.0**-4))


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_search_matches_the_reference_candidate_by_candidate(name):
    z, fe = _load(name)
    fe.to_device("cuda:0")
    target = torch.from_numpy(z["target"]).to("cuda:0")
    mgr = types.SimpleNamespace(lmbda=float(z["lmbda"]), total_training_time_sec=0.0)
    trace = []
    Q.quantize_model(fe, target, mgr, trace=trace)
    ref_loss, ref_rate = z["trace_loss"], z["trace_rate_mlp"]
    assert len(trace) == len(ref_loss)
    got = np.array([t["loss"] for t in trace])
    assert np.array_equal(np.array([t["rate_mlp"] for t in trace]), ref_rate)  # integer bit counts: exact
    rel = np.abs(got - ref_loss) / np.abs(ref_loss)
    assert rel.max() <= 1e-4, (rel.max(), int(rel.argmax()))
    cc = fe.coolchic_encoder
    # the chosen pair is the reference's, unless the reference's own candidate losses tie within the tolerance
    pos = 0
    for m in MODULES:
        n = sum(1 for t in trace if t["module"] == m)
        seg = ref_loss[pos:pos + n]
        mine = [t for t in trace if t["module"] == m]
        chosen = [i for i, t in enumerate(mine) if t["q_w"] == float(cc.nn_q_step[m]["weight"])
                  and t["q_b"] == float(cc.nn_q_step[m]["bias"])][0]
        assert seg[chosen] <= seg.min() * (1 + 2e-4), (m, chosen, int(seg.argmin()))
        pos += n
        if seg[chosen] == seg.min():
            assert [float(cc.nn_q_step[m]["weight"]), float(cc.nn_q_step[m]["bias"])] == z[f"q_step/{m}"].tolist()
    fe.set_to_eval()
    with torch.no_grad():
        o = fe.forward(quantizer_noise_type="none", quantizer_type="hardround", AC_MAX_VAL=-1)
        rate_mlp = sum(float(r) for mod in cc.get_network_rate().values() for r in mod.values())
        from coolchic_b200.enc.training.loss import loss_function

        l = loss_function(o.decoded_image, o.rate, target, lmbda=mgr.lmbda, rate_mlp_bit=rate_mlp)
    assert abs(float(l.loss) - float(z["final_loss"])) <= 2e-4 * float(z["final_loss"])
    assert mgr.total_training_time_sec > 0
    # test() itself reports the reference's R-D cost: the network rate is part of it (test.py:394-420)
    from coolchic_b200.enc.training.train import test as cc_test

    logs = cc_test(fe, target, mgr)
    n_pix = target.shape[-2] * target.shape[-1]
    assert logs.rate_nn_bpp > 0 and abs(logs.rate_nn_bpp - float(z["final_rate_mlp"]) / n_pix) <= 1e-9
    assert abs(float(logs.loss) - float(z["final_loss"])) <= 2e-4 * float(z["final_loss"])
    assert abs(logs.total_rate_bpp - (logs.rate_nn_bpp + logs.rate_latent_bpp)) <= 1e-12
