"""GPU parity: the CUDA path (through the C ABI) against the golden vectors produced by the REAL
reference (tests/golden/make_golden.py) -- forward outputs, every gradient tensor, and the
parameters after clip_grad_norm_ + Adam steps.  Tolerance: tests/common.py (1e-4 norm-wise)."""

import pytest
import torch

from common import Golden, assert_close, golden_cases, rel_err

pytestmark = pytest.mark.gpu


def _param(g: Golden):
    from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter

    a = g.arch
    prm = CoolChicEncoderParameter(layers_synthesis=list(a.layers_synthesis), n_ft_per_res=[1] * 7,
                                   dim_arm=a.dim_arm, n_hidden_layers_arm=a.n_hidden_layers_arm,
                                   encoder_gain=a.encoder_gain)
    prm.set_image_size((g.H, g.W))
    return prm


def _frame_encoder(g: Golden):
    from coolchic_b200.enc.component.frame import FrameEncoder

    fe = FrameEncoder(_param(g))
    p0 = g.tensors("p0/")
    fe.load_state_dict({"coolchic_encoder." + k: v for k, v in p0.items()})
    fe.to_device("cuda:0")
    return fe


class Report:
    def __init__(self):
        self.bad = []

    def check(self, a, b, what, rtol=1e-4):
        e = rel_err(a, b)
        flag = "" if e <= rtol else "   <-- FAIL"
        print(f"  {what:<75} rel {e:.2e}{flag}")
        if e > rtol:
            self.bad.append((what, e))

    def done(self):
        assert not self.bad, f"{len(self.bad)} mismatches: {self.bad[:6]}"


@pytest.mark.parametrize("name", golden_cases())
def test_module_api_matches_reference(name):
    """CoolChicEncoder / FrameEncoder modules + torch loss / clip / Adam, as train.py drives them."""
    from coolchic_b200.enc.training.loss import loss_function

    g = Golden(name)
    dev = torch.device("cuda:0")
    target = g.target.to(dev)
    rep = Report()
    if g.n_steps == 0:
        fe = _frame_encoder(g)
        fe.set_to_eval()
        with torch.no_grad():
            o = fe.forward(quantizer_noise_type="none", quantizer_type="hardround", AC_MAX_VAL=-1,
                           flag_additional_outputs=True)
            l = loss_function(o.decoded_image, o.rate, target, lmbda=g.lmbda, compute_logs=True)
        rep.check(o.decoded_image, g.get("decoded"), "decoded")
        rep.check(o.rate, g.get("rate"), "rate")
        for key in ("detailed_mu", "detailed_scale", "detailed_log_scale"):
            got = torch.cat([t.reshape(-1) for t in o.additional_data[key]])
            rep.check(got, g.get(key), key)
        sent = torch.cat([t.reshape(-1) for t in o.additional_data["detailed_sent_latent"]]).cpu()
        assert torch.equal(sent, g.get("detailed_sent_latent")), "quantised latents must be bit-exact"
        assert abs(l.loss.item() - float(g.z["loss"])) <= 1e-4 * abs(float(g.z["loss"]))
        rep.done()
        return

    if g.latent_free:
        from coolchic_b200.enc.component.nlcoolchic import LatentFreeCoolChicEncoder

        lf = LatentFreeCoolChicEncoder(_param(g))
        p0 = g.tensors("p0/")
        lf.load_state_dict({k: v for k, v in p0.items() if not k.startswith("latent_grids")})
        lf.to(dev)
        lf.train()
        lats = [p0[f"latent_grids.{i}.data"].to(dev).requires_grad_(True) for i in range(7)]
        noise = g.get("s0/noise")
        lf._injected_noise = None if noise is None else noise.to(dev)
        raw, rate, _ = lf.forward(lats, quantizer_noise_type=g.noise_type, quantizer_type=g.q_type,
                                  soft_round_temperature=g.T, noise_parameter=g.sigma)
        l = loss_function(raw, rate, target, lmbda=g.lmbda)
        l.loss.backward()
        rep.check(raw, g.get("decoded"), "raw_out")
        rep.check(rate, g.get("rate"), "rate")
        named = dict(lf.named_parameters())
        for k, gref in g.tensors("s0/grad/").items():
            if k.startswith("latent_grids."):
                got = lats[int(k.split(".")[1])].grad
            else:
                got = named[k].grad
            assert got is not None, f"no gradient for {k}"
            rep.check(got, gref, "grad " + k)
        rep.done()
        return

    fe = _frame_encoder(g)
    fe.set_to_train()
    params = list(fe.parameters())
    opt = torch.optim.Adam(params, lr=g.lr)
    T, S = g.T.to(dev), g.sigma.to(dev)
    for step in range(g.n_steps):
        for q in params:
            q.grad = None
        noise = g.get(f"s{step}/noise")
        fe.coolchic_encoder._injected_noise = None if noise is None else noise.to(dev)
        o = fe.forward(quantizer_noise_type=g.noise_type, quantizer_type=g.q_type, soft_round_temperature=T,
                       noise_parameter=S)
        l = loss_function(o.decoded_image, o.rate, target, lmbda=g.lmbda)
        l.loss.backward()
        if step == 0:
            rep.check(o.decoded_image, g.get("decoded"), "decoded")
            rep.check(o.rate, g.get("rate"), "rate")
        gref = g.tensors(f"s{step}/grad/")
        for k, v in fe.named_parameters():
            k = k.removeprefix("coolchic_encoder.")
            if k in gref:
                assert v.grad is not None, f"step {step}: no gradient for {k}"
                rep.check(v.grad, gref[k], f"step {step} grad {k}", rtol=2e-4)
            else:
                assert v.grad is None, f"step {step}: unexpected gradient for {k} (reference has None)"
        lref = float(g.z[f"s{step}/loss"])
        assert abs(l.loss.item() - lref) <= 1e-4 * abs(lref), (l.loss.item(), lref)
        norm = torch.nn.utils.clip_grad_norm_(params, 1e-1, norm_type=2.0, error_if_nonfinite=False)
        nref = float(g.z[f"s{step}/norm"])
        assert abs(norm.item() - nref) <= 2e-4 * abs(nref), (norm.item(), nref)
        opt.step()
    final = g.tensors("p_final/")
    for k, v in fe.named_parameters():
        rep.check(v, final[k.removeprefix("coolchic_encoder.")], "final " + k, rtol=2e-4)
    rep.done()


@pytest.mark.parametrize("name", [n for n in golden_cases() if "eval" not in n and "b2" not in n])
def test_fused_train_step_matches_reference(name):
    """ccb_train_step (forward + loss + backward + clip + Adam in the library) with the reference's noise."""
    from coolchic_b200 import ops

    g = Golden(name)
    dev = torch.device("cuda:0")
    a = g.arch
    spec = ops.ArchSpec(g.H, g.W, a.dim_arm, a.n_hidden_layers_arm, list(a.layers_synthesis), 7, a.ups_k_size,
                        a.ups_preconcat_k_size, a.encoder_gain)
    st = ops.FusedState(spec, g.B, dev)
    st.load(g.tensors("p0/"))
    st.lr.fill_(g.lr)
    st.T.fill_(float(g.T))
    st.sigma.fill_(float(g.sigma))
    target = g.target.to(dev).contiguous()
    rep = Report()
    for step in range(g.n_steps):
        noise = g.get(f"s{step}/noise")
        noise = None if noise is None else noise.to(dev).contiguous()
        st.train_step(target, g.lmbda, g.q_type, g.noise_type, noise=noise)
        out4 = st.out4.cpu()
        lref, nref = float(g.z[f"s{step}/loss"]), float(g.z[f"s{step}/norm"])
        print(f"  step {step}: loss {out4[0]:.6f} (ref {lref:.6f})  norm {out4[3]:.6f} (ref {nref:.6f})")
        assert abs(out4[0].item() - lref) <= 1e-4 * abs(lref)
        assert abs(out4[3].item() - nref) <= 2e-4 * abs(nref)
        gref = g.tensors(f"s{step}/grad/")
        for k, v in gref.items():
            rep.check(st.grads[k], v, f"step {step} grad {k}", rtol=2e-4)
    for k, v in g.tensors("p_final/").items():
        if k in st.params:
            rep.check(st.params[k], v, "final " + k, rtol=2e-4)
    rep.done()
