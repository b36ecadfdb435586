"""CPU: pin the oracle (oracle/float_ref.py) against vectors produced by the real reference
(tests/golden/make_golden.py).  Tolerances are the float-path tolerances of tests/common.py."""

import pytest
import torch

from common import Golden, assert_close, golden_cases, O


@pytest.mark.parametrize("name", golden_cases())
def test_oracle_matches_reference_golden(name):
    g = Golden(name)
    p = g.tensors("p0/")
    if g.n_steps == 0:
        with torch.no_grad():
            raw, rate, ex = O.coolchic_forward(p, g.arch, False, detailed=True)
            dec = O.frame_output(raw, False)
            loss, mse, bpp = O.loss_ref(dec, rate, g.target, g.lmbda)
        assert_close(dec, g.get("decoded"), "decoded")
        assert_close(rate, g.get("rate"), "rate")
        assert_close(ex["mu"].reshape(-1), g.get("detailed_mu"), "mu")
        assert_close(ex["scale"].reshape(-1), g.get("detailed_scale"), "scale")
        assert_close(ex["log_scale"].reshape(-1), g.get("detailed_log_scale"), "log_scale")
        assert torch.equal(ex["yhat"].reshape(-1), g.get("detailed_sent_latent"))
        assert abs(float(loss) - float(g.z["loss"])) <= 1e-5 * abs(float(g.z["loss"]))
        assert abs(float(mse) - float(g.z["mse"])) <= 1e-5 * abs(float(g.z["mse"]))
        assert abs(float(bpp) - float(g.z["bpp"])) <= 1e-5 * abs(float(g.z["bpp"]))
        return

    if g.latent_free:
        leaves = {k: v.clone().requires_grad_(True) for k, v in p.items()}
        lats = [leaves[f"latent_grids.{i}.data"] for i in range(7)]
        raw, rate, _ = O.coolchic_forward(leaves, g.arch, True, g.noise_type, g.q_type, g.T, g.sigma,
                                          g.get("s0/noise"), latents=lats)
        loss, _, _ = O.loss_ref(raw, rate, g.target, g.lmbda)
        loss.backward()
        assert_close(raw, g.get("decoded"), "raw_out")
        assert_close(rate, g.get("rate"), "rate")
        for k, gref in g.tensors("s0/grad/").items():
            assert_close(leaves[k].grad, gref, "grad " + k)
        return

    state = {}
    for step in range(g.n_steps):
        noise = g.get(f"s{step}/noise")
        if step == 0:
            with torch.no_grad():
                raw, rate, _ = O.coolchic_forward(p, g.arch, True, g.noise_type, g.q_type, g.T, g.sigma, noise)
            assert_close(O.frame_output(raw, True), g.get("decoded"), "decoded")
            assert_close(rate, g.get("rate"), "rate")
        loss, mse, bpp, grads, norm = O.train_step_ref(
            p, g.arch, g.target, state, g.lmbda, g.lr, g.noise_type, g.q_type, g.T, g.sigma, noise)
        gref = g.tensors(f"s{step}/grad/")
        for k, v in grads.items():
            if v is None:
                assert k not in gref, f"oracle has no grad for {k} but the reference does"
            else:
                assert k in gref, f"oracle has a grad for {k} but the reference does not"
                assert_close(v, gref[k], f"step {step} grad {k}", rtol=2e-4)
        assert abs(loss - float(g.z[f"s{step}/loss"])) <= 1e-5 * abs(loss)
        assert abs(norm - float(g.z[f"s{step}/norm"])) <= 1e-4 * abs(norm)
    for k, v in g.tensors("p_final/").items():
        assert_close(p[k], v, "final " + k, rtol=2e-4)
