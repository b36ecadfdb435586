"""Generate golden input/output vectors from the REAL reference (pepborrell/cool-chic).

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

For every case it instantiates the reference's own `FrameEncoder`
(coolchic/enc/component/frame.py), loads a seeded parameter set, runs exactly the
loop body of coolchic/enc/training/train.py:244-266 (forward -> loss_function ->
backward -> clip_grad_norm_ -> Adam.step) and stores inputs, the noise tensor the
reference drew, outputs, raw gradients and the parameters after the optimiser steps
in tests/golden/<case>.npz.  The vectors are small (a few 10 kB each) and are
committed; they pin both the oracle (tests/test_oracle_golden.py, CPU) and the CUDA
path (tests/test_gpu_golden.py).
"""

import os
import sys

import zlib

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from refimport import import_reference  # noqa: E402

import_reference()

from coolchic.enc.component.coolchic import CoolChicEncoderParameter  # noqa: E402
from coolchic.enc.component.frame import FrameEncoder  # noqa: E402
from coolchic.enc.training.loss import loss_function  # noqa: E402
from torch.nn.utils import clip_grad_norm_  # noqa: E402

from oracle import float_ref as O  # noqa: E402

ARCHS = {
    "hop": dict(dim_arm=16, nh=2, syn=["48-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"]),
    "default": dict(dim_arm=24, nh=2, syn=["40-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"]),
    "lop": dict(dim_arm=8, nh=2, syn=["16-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"]),
    "vlop": dict(dim_arm=8, nh=1, syn=["8-1-linear-relu", "3-1-linear-none", "3-3-residual-none"]),
    "arm32": dict(dim_arm=32, nh=2, syn=["16-1-linear-relu", "3-1-linear-none", "3-3-residual-relu", "3-3-residual-none"]),
}

# (name, arch, H, W, B, q_type, noise_type, T, sigma, n_steps, gain)
CASES = [
    ("hop_sr_gauss_40x56", "hop", 40, 56, 1, "softround", "gaussian", 0.3, 0.25, 3, 16),
    ("hop_sr_kuma_37x53", "hop", 37, 53, 1, "softround", "kumaraswamy", 0.3, 2.0, 3, 16),
    ("hop_ste_40x56", "hop", 40, 56, 1, "ste", "none", 1e-4, 1.0, 2, 16),
    ("hop_trueste_b2_32x32", "hop", 32, 32, 2, "true_ste", "none", 0.3, 1.0, 2, 1),
    ("default_sr_gauss_33x47", "default", 33, 47, 1, "softround", "gaussian", 0.2, 0.15, 2, 16),
    ("lop_sralone_24x40", "lop", 24, 40, 1, "softround_alone", "none", 0.3, 1.0, 2, 16),
    ("vlop_sr_kuma_19x35", "vlop", 19, 35, 1, "softround", "kumaraswamy", 0.25, 1.5, 2, 16),
    ("arm32_none_gauss_24x24", "arm32", 24, 24, 1, "none", "gaussian", 0.3, 0.3, 2, 16),
    ("hop_eval_40x56", "hop", 40, 56, 1, "hardround", "none", 0.3, 1.0, 0, 16),
]


class NoiseTap:
    """Records what torch.randn_like / torch.rand_like returned inside quantize()."""

    def __enter__(self):
        self.captured = []
        self._randn, self._rand = torch.randn_like, torch.rand_like

        def randn_like(x, **kw):
            n = self._randn(x, **kw)
            self.captured.append(n.detach().clone())
            return n

        def rand_like(x, **kw):
            n = self._rand(x, **kw)
            self.captured.append(n.detach().clone())
            return n

        torch.randn_like, torch.rand_like = randn_like, rand_like
        return self

    def __exit__(self, *a):
        torch.randn_like, torch.rand_like = self._randn, self._rand


def build_reference(arch_name, H, W, B, gain, seed):
    a = ARCHS[arch_name]
    prm = CoolChicEncoderParameter(
        layers_synthesis=list(a["syn"]), n_ft_per_res=[1] * 7, dim_arm=a["dim_arm"],
        n_hidden_layers_arm=a["nh"], encoder_gain=gain,
    )
    prm.set_image_size((H, W))
    fe = FrameEncoder(prm)
    arch = O.Arch(img_size=(H, W), dim_arm=a["dim_arm"], n_hidden_layers_arm=a["nh"],
                  layers_synthesis=list(a["syn"]), encoder_gain=float(gain))
    p = O.init_params(arch, seed=seed, B=B, random=True)
    if gain == 1:
        for i in range(7):
            p[f"latent_grids.{i}.data"] = p[f"latent_grids.{i}.data"] * 12.0
    # B>1: the fork batches by giving the latent parameters a leading batch dim
    for i in range(7):
        fe.coolchic_encoder.latent_grids[i].data = torch.nn.Parameter(p[f"latent_grids.{i}.data"].clone())
    sd = {("coolchic_encoder." + k): v.clone() for k, v in p.items()}
    fe.load_state_dict(sd)
    return fe, arch, p


def run_case(name, arch_name, H, W, B, q_type, noise_type, T, sigma, n_steps, gain):
    torch.manual_seed(1234)
    fe, arch, p0 = build_reference(arch_name, H, W, B, gain, seed=zlib.crc32(name.encode()) % 1000)
    target = O.synthetic_image(7, H, W, B)
    lmbda, lr = 1e-3, 1e-2
    out = {"target": target.numpy(), "H": H, "W": W, "B": B, "gain": gain, "arch": arch_name,
           "q_type": q_type, "noise_type": noise_type, "T": np.float32(T), "sigma": np.float32(sigma),
           "lmbda": lmbda, "lr": lr, "n_steps": n_steps}
    for k, v in p0.items():
        out["p0/" + k] = v.numpy()
    Tt, St = torch.tensor(T), torch.tensor(sigma)

    if n_steps == 0:  # eval-mode forward with detailed outputs (test.py:402-411)
        fe.set_to_eval()
        with torch.no_grad():
            o = fe.forward(quantizer_noise_type="none", quantizer_type="hardround", AC_MAX_VAL=-1,
                           flag_additional_outputs=True)
            l = loss_function(o.decoded_image, o.rate, target, lmbda=lmbda, compute_logs=True)
        out["decoded"] = o.decoded_image.numpy()
        out["rate"] = o.rate.numpy()
        out["loss"] = np.float64(l.loss.item())
        out["mse"] = np.float64(l.mse)
        out["bpp"] = np.float64(l.rate_latent_bpp)
        for key in ("detailed_mu", "detailed_scale", "detailed_log_scale", "detailed_sent_latent"):
            out[key] = np.concatenate([t.reshape(-1).numpy() for t in o.additional_data[key]])
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        return

    if B > 1:
        # The fork's batched path is LatentFreeCoolChicEncoder (nlcoolchic.py:151-336;
        # CoolChicEncoder.forward itself cannot view a batched flat latent back:
        # coolchic.py:388).  Latents are inputs, the loss is on the raw synthesis
        # output without clamp, as hypernet/training.py:296-313 does.
        from coolchic.enc.component.nlcoolchic import LatentFreeCoolChicEncoder
        lf = LatentFreeCoolChicEncoder(fe.coolchic_encoder.param)
        lf.load_state_dict({k: v.clone() for k, v in p0.items() if not k.startswith("latent_grids")})
        lf.train()
        lats = [p0[f"latent_grids.{i}.data"].clone().requires_grad_(True) for i in range(7)]
        with NoiseTap() as tap:
            raw, rate, _ = lf.forward(lats, quantizer_noise_type=noise_type, quantizer_type=q_type,
                                      soft_round_temperature=Tt, noise_parameter=St)
        l = loss_function(raw, rate, target, lmbda=lmbda)
        l.loss.backward()
        if tap.captured:
            out["s0/noise"] = tap.captured[0].numpy()
        out["decoded"] = raw.detach().numpy()
        out["rate"] = rate.detach().numpy()
        out["s0/loss"] = np.float64(l.loss.item())
        out["latent_free"] = 1
        out["n_steps"] = 1
        for i, t in enumerate(lats):
            out[f"s0/grad/latent_grids.{i}.data"] = t.grad.numpy()
        for k, v in lf.named_parameters():
            if v.grad is not None:
                out["s0/grad/" + k] = v.grad.detach().clone().numpy()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        return

    fe.set_to_train()
    all_params = [x for x in fe.parameters()]
    opt = torch.optim.Adam(all_params, lr=lr)
    for step in range(n_steps):
        for q in all_params:
            q.grad = None
        with NoiseTap() as tap:
            o = fe.forward(quantizer_noise_type=noise_type, quantizer_type=q_type,
                           soft_round_temperature=Tt, noise_parameter=St)
        l = loss_function(o.decoded_image, o.rate, target, lmbda=lmbda)
        l.loss.backward()
        if tap.captured:
            out[f"s{step}/noise"] = tap.captured[0].numpy()
        if step == 0:
            out["decoded"] = o.decoded_image.detach().numpy()
            out["rate"] = o.rate.detach().numpy()
        out[f"s{step}/loss"] = np.float64(l.loss.item())
        for k, v in fe.named_parameters():
            k = k.removeprefix("coolchic_encoder.")
            if v.grad is not None:
                out[f"s{step}/grad/" + k] = v.grad.detach().clone().numpy()
        norm = clip_grad_norm_(all_params, 1e-1, norm_type=2.0, error_if_nonfinite=False)
        out[f"s{step}/norm"] = np.float64(norm.item())
        opt.step()
    for k, v in fe.named_parameters():
        out["p_final/" + k.removeprefix("coolchic_encoder.")] = v.detach().numpy()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)


if __name__ == "__main__":
    for c in CASES:
        run_case(*c)
        print("wrote", c[0], os.path.getsize(os.path.join(HERE, c[0] + ".npz")) // 1024, "kB")
