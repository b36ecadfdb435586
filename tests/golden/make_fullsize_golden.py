"""Golden vectors of the REAL reference at BASELINE.json's full shapes and with per-sample weights.

Run in the build container only (needs /root/reference):

    python tests/golden/make_fullsize_golden.py [case ...]

Cases are listed in tests/common.py (FULLSIZE_CASES).  The inputs are NOT stored: tests regenerate them from
seeds with tests/common.py:fullsize_inputs (CPU torch generators are deterministic).  Stored per case: loss, mse,
bpp, gradient norm, per-sample rate sums, and -- on the fixed sample of tests/common.py:sample_indices (4096 random
positions + the two outermost rows / columns of every plane) -- the decoded image, the rate and every latent
gradient, plus max|g| and ||g||_2 of each; weight gradients are stored in full.

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
        torch.randn_like = lambda x, **kw: self._randn(x, **kw) if self.value is None else self.value.reshape(x.shape)
        torch.rand_like = lambda x, **kw: self._rand(x, **kw) if self.value is None else self.value.reshape(x.shape)
        return self

    def __exit__(self, *a):
        torch.randn_like, torch.rand_like = self._randn, self._rand


def store_sampled(out, key, t, seed):
    t = t.detach()
    idx = Cm.sample_indices(tuple(t.shape), seed)
    out["idx/" + key] = idx.astype(np.int64)
    out["val/" + key] = t.reshape(-1)[torch.from_numpy(idx)].numpy()
    out["max/" + key] = np.float64(t.abs().max().item())
    out["l2/" + key] = np.float64(t.double().norm().item())


def run_case(name):
    inp = Cm.fullsize_inputs(name)
    p, B, H, W = inp["p"], inp["B"], inp["H"], inp["W"]
    a = Cm.ARCHS[Cm.FULLSIZE_CASES[name][0]]
    prm = CoolChicEncoderParameter(layers_synthesis=list(a["syn"]), n_ft_per_res=[1] * 7, dim_arm=a["dim_arm"],
                                   n_hidden_layers_arm=a["nh"], encoder_gain=inp["gain"])
    prm.set_image_size((H, W))
    Tt, St = torch.tensor(inp["T"]), torch.tensor(inp["sigma"])
    out = {}
    grads = {}
    with FixedNoise() as fx:
        if B == 1 and inp["per_sample"] is None:
            fe = FrameEncoder(prm)
            fe.load_state_dict({("coolchic_encoder." + k): v.clone() for k, v in p.items()})
            fe.set_to_train()
            fx.value = inp["noise"]
            o = fe.forward(quantizer_noise_type=inp["noise_type"], quantizer_type=inp["q_type"],
                           soft_round_temperature=Tt, noise_parameter=St)
            dec, rate = o.decoded_image, o.rate
            l = loss_function(dec, rate, inp["target"], lmbda=inp["lmbda"], compute_logs=True)
            l.loss.backward()
            for k, v in fe.named_parameters():
                if v.grad is not None:
                    grads[k.removeprefix("coolchic_encoder.")] = v.grad
        else:
            lf = LatentFreeCoolChicEncoder(prm)
            lf.load_state_dict({k: v.clone() for k, v in p.items() if not k.startswith("latent_grids")})
            lf.train()
            lats = [p[f"latent_grids.{i}.data"].clone().requires_grad_(True) for i in range(7)]
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
                        leaves[k] = inp["per_sample"][k].clone().requires_grad_(True)
                    else:  # the never-used upsampling biases
                        leaves[k] = v.detach().unsqueeze(0).expand(B, *v.shape)

                def fwd(lat_list, params, nz):
                    fx.value = nz
                    return functional_call(lf, params, (lat_list,), kwargs=kw)

                dec, rate, _ = torch.vmap(fwd, randomness="different")(
                    [t.unsqueeze(1) for t in lats], leaves, inp["noise"].unsqueeze(1))
                dec, rate = dec.squeeze(1), rate.squeeze(1)
            l = loss_function(dec, rate, inp["target"], lmbda=inp["lmbda"], compute_logs=True)
            l.loss.backward()
            for i, t in enumerate(lats):
                grads[f"latent_grids.{i}.data"] = t.grad
            for k, v in leaves.items():
                if getattr(v, "grad", None) is not None:
                    grads[k] = v.grad
    out["loss"] = np.float64(l.loss.item())
    out["mse"] = np.float64(l.mse)
    out["bpp"] = np.float64(l.rate_latent_bpp)
    out["rate_sum"] = rate.detach().double().sum(dim=1).numpy()
    # share of positions on the 2^-16 rate floor (coolchic.py:419-423): should be small for a well-conditioned case
    out["frac_rate_floor"] = np.float64((rate.detach() >= 15.999).double().mean().item())
    sq = 0.0
    for g in grads.values():
        sq += float(g.double().pow(2).sum())
    out["norm"] = np.float64(sq ** 0.5)
    store_sampled(out, "decoded", dec, 1)
    store_sampled(out, "rate", rate, 2)
    for n, (k, g) in enumerate(sorted(grads.items())):
        if k.startswith("latent_grids"):
            store_sampled(out, "grad/" + k, g, 100 + n)
        else:
            out["full/grad/" + k] = g.detach().numpy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    return out


if __name__ == "__main__":
    names = sys.argv[1:] or list(Cm.FULLSIZE_CASES)
    torch.set_num_threads(os.cpu_count())
    for n in names:
        o = run_case(n)
        print("wrote", n, os.path.getsize(os.path.join(HERE, n + ".npz")) // 1024, "kB",
              f"bpp {float(o['bpp']):.3f} rate-floor share {float(o['frac_rate_floor']):.4f}", flush=True)
