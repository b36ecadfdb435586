"""Golden vectors of the REAL reference at BASELINE.json's full shapes and with per-sample weights.

Run in the build container only (needs /root/reference):

    python tests/golden/make_fullsize_golden.py [case ...]

Cases are listed in tests/common.py (FULLSIZE_CASES).  The inputs are NOT stored: tests regenerate them from
seeds with tests/common.py:fullsize_inputs (CPU torch generators are deterministic).  Stored per case: loss, mse,
bpp, gradient norm, per-sample rate sums, and -- on the fixed sample of tests/common.py:sample_indices (4096 random
positions + the two outermost rows / columns of every plane) -- the decoded image, the rate and every latent
gradient, plus max|g| and ||g||_2 of each; weight gradients are stored in full.  Every stored value exists twice:
from the reference as it runs (float32: val/, full/) and from the SAME reference modules cast to float64 (val64/,
full64/).  The float64 copy tells the tests where the reference's own fp32 result is off (a ReLU / clamp branch taken
differently from exact arithmetic moves a patch of latent gradients by 1e-3 of the maximum).

  * B = 1 cases run the reference's FrameEncoder exactly as train.py:244-266 does (forward -> loss_function ->
    backward).
  * shared-weight batches run the reference's LatentFreeCoolChicEncoder (hypernet path, NOWholeNet shape).
  * per-sample cases run it under torch.vmap(functional_call(...), randomness="different") with one weight set
    per batch element, the way DeltaWholeNet.forward does (hypernet/hypernet.py:1266-1288,
    hypernet/common.py:227-253); the noise the reference would draw is replaced by the seeded tensor of
    fullsize_inputs through a patched torch.randn_like / torch.rand_like.
"""

import os
import sys

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
from coolchic.enc.component.nlcoolchic import LatentFreeCoolChicEncoder  # noqa: E402
from coolchic.enc.training.loss import loss_function  # noqa: E402
from torch.func import functional_call  # noqa: E402

import common as Cm  # noqa: E402


class FixedNoise:
    """torch.randn_like / rand_like return the prepared tensor (under vmap: the current batched slice).  Module
    constructors also call them for their initialisation: untouched while no noise is armed."""

    def __init__(self):
        self.value = None

    def __enter__(self):
        self._randn, self._rand = torch.randn_like, torch.rand_like
        pick = lambda x: self.value.reshape(x.shape).to(x.dtype)  # noqa: E731
        torch.randn_like = lambda x, **kw: self._randn(x, **kw) if self.value is None else pick(x)
        torch.rand_like = lambda x, **kw: self._rand(x, **kw) if self.value is None else pick(x)
        return self

    def __exit__(self, *a):
        torch.randn_like, torch.rand_like = self._randn, self._rand


def store_sampled(out, key, t, seed):
    t = t.detach()
    # the rate is [B, M], a concatenation of flattened planes, not a plane: random positions only
    idx = Cm.sample_indices(tuple(t.shape) if t.dim() == 4 else (t.numel(),), seed)
    out["idx/" + key] = idx.astype(np.int32)
    out["val/" + key] = t.reshape(-1)[torch.from_numpy(idx)].numpy()
    out["max/" + key] = np.float64(t.abs().max().item())
    out["l2/" + key] = np.float64(t.double().norm().item())


def evaluate(name, dtype):
    """The reference's forward -> loss_function -> backward in `dtype`.  -> dec, rate, grads, loss output."""
    inp = Cm.fullsize_inputs(name)
    p, B, H, W = inp["p"], inp["B"], inp["H"], inp["W"]
    a = Cm.ARCHS[Cm.FULLSIZE_CASES[name][0]]
    prm = CoolChicEncoderParameter(layers_synthesis=list(a["syn"]), n_ft_per_res=[1] * 7, dim_arm=a["dim_arm"],
                                   n_hidden_layers_arm=a["nh"], encoder_gain=inp["gain"])
    prm.set_image_size((H, W))
    Tt, St = torch.tensor(inp["T"], dtype=dtype), torch.tensor(inp["sigma"], dtype=dtype)
    target = inp["target"].to(dtype)
    grads = {}
    with FixedNoise() as fx:
        if B == 1 and inp["per_sample"] is None:
            fe = FrameEncoder(prm)
            fe.load_state_dict({("coolchic_encoder." + k): v.clone() for k, v in p.items()})
            fe.to(dtype)
            fe.set_to_train()
            fx.value = inp["noise"]
            o = fe.forward(quantizer_noise_type=inp["noise_type"], quantizer_type=inp["q_type"],
                           soft_round_temperature=Tt, noise_parameter=St)
            dec, rate = o.decoded_image, o.rate
            l = loss_function(dec, rate, target, lmbda=inp["lmbda"], compute_logs=True)
            l.loss.backward()
            for k, v in fe.named_parameters():
                if v.grad is not None:
                    grads[k.removeprefix("coolchic_encoder.")] = v.grad
        else:
            lf = LatentFreeCoolChicEncoder(prm)
            lf.load_state_dict({k: v.clone() for k, v in p.items() if not k.startswith("latent_grids")})
            lf.to(dtype)
            lf.train()
            lats = [p[f"latent_grids.{i}.data"].to(dtype).clone().requires_grad_(True) for i in range(7)]
            kw = dict(quantizer_noise_type=inp["noise_type"], quantizer_type=inp["q_type"],
                      soft_round_temperature=Tt, noise_parameter=St, AC_MAX_VAL=-1, flag_additional_outputs=False)
            if inp["per_sample"] is None:
                fx.value = inp["noise"]
                dec, rate, _ = lf.forward(lats, **kw)
                leaves = dict(lf.named_parameters())
            else:
                # every named parameter gets a batch dimension, as add_deltas does (hypernet/common.py:227-253)
                leaves = {}
                for k, v in lf.named_parameters():
                    if k in inp["per_sample"]:
                        leaves[k] = inp["per_sample"][k].to(dtype).clone().requires_grad_(True)
                    else:  # the never-used upsampling biases
                        leaves[k] = v.detach().unsqueeze(0).expand(B, *v.shape)

                def fwd(lat_list, params, nz):
                    fx.value = nz
                    return functional_call(lf, params, (lat_list,), kwargs=kw)

                dec, rate, _ = torch.vmap(fwd, randomness="different")(
                    [t.unsqueeze(1) for t in lats], leaves, inp["noise"].unsqueeze(1))
                dec, rate = dec.squeeze(1), rate.squeeze(1)
            l = loss_function(dec, rate, target, lmbda=inp["lmbda"], compute_logs=True)
            l.loss.backward()
            for i, t in enumerate(lats):
                grads[f"latent_grids.{i}.data"] = t.grad
            for k, v in leaves.items():
                if getattr(v, "grad", None) is not None:
                    grads[k] = v.grad
    return dec.detach(), rate.detach(), grads, l


def run_case(name):
    dec, rate, grads, l = evaluate(name, torch.float32)
    dec64, rate64, grads64, l64 = evaluate(name, torch.float64)
    out = {}
    out["loss"] = np.float64(l.loss.item())
    out["mse"] = np.float64(l.mse)
    out["bpp"] = np.float64(l.rate_latent_bpp)
    out["loss64"] = np.float64(l64.loss.item())
    out["rate_sum"] = rate.double().sum(dim=1).numpy()
    # share of positions on the 2^-16 rate floor (coolchic.py:419-423): should be small for a well-conditioned case
    out["frac_rate_floor"] = np.float64((rate >= 15.999).double().mean().item())
    out["norm"] = np.float64(sum(float(g.double().pow(2).sum()) for g in grads.values()) ** 0.5)

    def both(key, t32, t64, seed):
        store_sampled(out, key, t32, seed)
        out["val64/" + key] = t64.detach().reshape(-1)[torch.from_numpy(out["idx/" + key])].numpy()

    both("decoded", dec, dec64, 1)
    both("rate", rate, rate64, 2)
    worst = 0.0
    for n, k in enumerate(sorted(grads)):
        g, g64 = grads[k], grads64[k]
        if k.startswith("latent_grids"):
            both("grad/" + k, g, g64, 100 + n)
            d = np.abs(out["val/grad/" + k].astype(np.float64) - out["val64/grad/" + k]).max() / float(out["max/grad/" + k])
        else:
            out["full/grad/" + k] = g.detach().numpy()
            out["full64/grad/" + k] = g64.detach().numpy()
            d = float((g.double() - g64).abs().max() / g64.abs().max())
        worst = max(worst, float(d))
    out["ref32_vs_ref64_worst_grad"] = np.float64(worst)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    return out


if __name__ == "__main__":
    names = sys.argv[1:] or list(Cm.FULLSIZE_CASES)
    torch.set_num_threads(os.cpu_count())
    for n in names:
        o = run_case(n)
        print("wrote", n, os.path.getsize(os.path.join(HERE, n + ".npz")) // 1024, "kB",
              f"bpp {float(o['bpp']):.3f} rate-floor share {float(o['frac_rate_floor']):.4f} "
              f"reference fp32 vs fp64, worst gradient tensor: {float(o['ref32_vs_ref64_worst_grad']):.2e}", flush=True)
