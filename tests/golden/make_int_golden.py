"""Generate the integer-path golden vectors.  Run in the BUILD container only:

    python tests/golden/make_int_golden.py

Two families, both produced by the reference itself:

  int_armint_c{C}_h{nh}.npz   ArmInt(pure_int=True) (coolchic/enc/bitstream/armint.py:114-261) imported from
                              /root/reference, fed the 2-D contexts of `_get_neighbor`
                              (coolchic/enc/component/core/arm.py:308-352) of a random integer latent grid.
                              Stored: grid, float ARM weights, the integerised weights the module ended up with,
                              mu*256 and log_scale*256 as int32.
  int_cool_<name>.npz         a shipped bitstream (results/image/kodak/bitstreams/<name>.cool) decoded by the
                              reference's C++ (oracle/_ref/libccref.so, built in place by oracle/build.sh) up to
                              the ARM stage: the 7 decoded latent grids (int8), the decoded fixed-point ARM
                              weights, every latent sub-stream's bytes, and the sha256 of the (mu, log-scale,
                              mu_rounded, mu_index, sig_index) arrays that reproduce those bytes exactly through the
                              reference's own cc_code_latent_layer_bac.
"""

import hashlib
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

from refimport import REFERENCE_ROOT, import_reference  # noqa: E402

from oracle import int_ref as I  # noqa: E402


def make_armint_cases():
    import_reference()
    from coolchic.enc.bitstream.armint import ArmInt
    from coolchic.enc.component.core.arm import Arm, _get_neighbor, _get_non_zero_pixel_ctx_index
    from coolchic.enc.component.coolchic import MAX_ARM_MASK_SIZE  # noqa: F401

    try:
        from coolchic.enc.utils.misc import FIXED_POINT_FRACTIONAL_MULT as FPFM
    except Exception:
        FPFM = 256
    assert FPFM == 256
    cases = [(8, 0), (8, 2), (16, 2), (16, 1), (24, 2), (32, 3), (32, 2)]
    for C, nh in cases:
        g = torch.Generator().manual_seed(1000 + C * 10 + nh)
        H, W = 21 + C // 8, 37
        arm = Arm(C, nh)
        fp = arm.get_param()
        # trained-looking weights: perturb the identity-ish init so every rounding branch is hit
        for k in fp:
            fp[k] = fp[k] + 0.35 * torch.randn(fp[k].shape, generator=g) * (0.3 if "weight" in k else 1.5)
        ai = ArmInt(C, nh, FPFM, pure_int=True)
        ai.set_param_from_float(fp)
        y = torch.randint(-14, 16, (1, 1, H, W), generator=g).to(torch.float32)
        y[0, 0, : H // 3] = torch.randint(-2, 3, (H // 3, W), generator=g).to(torch.float32)
        ctx = _get_neighbor(y, MAX_ARM_MASK_SIZE, _get_non_zero_pixel_ctx_index(C))
        if ctx.dim() == 3:  # the fork returns [B, HW, C]; ArmInt indexes dim 1 as the output -> feed 2-D
            ctx = ctx[0]
        mu, scale, ls = ai(ctx)
        out = {
            "grid": y[0, 0].numpy().astype(np.int32),
            "mu_q8": torch.round(mu * 256).to(torch.int32).numpy(),
            "ls_q8": torch.round(ls * 256).to(torch.int32).numpy(),
            "C": np.int32(C), "nh": np.int32(nh),
        }
        assert np.array_equal(out["mu_q8"] / 256.0, mu.numpy()) and np.array_equal(out["ls_q8"] / 256.0, ls.numpy())
        ip = ai.get_param()
        for li in range(nh + 1):
            out[f"w{li}"] = fp[f"mlp.{2 * li}.weight"].numpy()
            out[f"b{li}"] = fp[f"mlp.{2 * li}.bias"].numpy()
            out[f"wq{li}"] = ip[f"mlp.{2 * li}.weight"].numpy().astype(np.int32)
            out[f"bq{li}"] = ip[f"mlp.{2 * li}.bias"].numpy().astype(np.int32)
        path = os.path.join(HERE, f"int_armint_c{C}_h{nh}.npz")
        np.savez_compressed(path, **out)
        print("wrote", path, os.path.getsize(path))


def sha(*arrs):
    h = hashlib.sha256()
    for a in arrs:
        h.update(np.ascontiguousarray(a, dtype=np.int32).tobytes())
    return h.hexdigest()


def make_cool_cases():
    names = ["kodim01-lmbda-0001", "kodim02-lmbda-002", "kodim13-lmbda-00001"]
    tmp = tempfile.mktemp()
    for name in names:
        path = os.path.join(REFERENCE_ROOT, "results", "image", "kodak", "bitstreams", name + ".cool")
        f = I.CoolFile(path)
        out = {
            "img_hw": np.array([f.img_h, f.img_w], np.int32), "C": np.int32(f.dim_arm), "nh": np.int32(f.n_hidden),
            "hls_sig_blksize": np.int32(f.hls_sig_blksize), "ac_max_val_latent": np.int32(f.ac_max_val_latent),
            "n_latents": np.int32(f.n_latents),
        }
        for li in range(f.n_hidden + 1):
            out[f"wq{li}"] = f.w_q[li]
            out[f"bq{li}"] = f.b_q[li]
        mus, lss = [], []
        for l, g in enumerate(f.grids):
            assert not f.zero_layer[l]
            assert g.min() >= -128 and g.max() <= 127
            out[f"grid{l}"] = g.astype(np.int8)
            out[f"bytes{l}"] = np.frombuffer(f.layer_bytes[l], dtype=np.uint8)
            mu, ls = I.arm_int_grid(g, f.w_q, f.b_q)
            # the pin: the reference encoder fed these (mu, ls) must give back the shipped sub-stream
            again = I.ref_encode_layer(g, mu, ls, f.hls_sig_blksize, tmp)
            assert again == f.layer_bytes[l], (name, l)
            mus.append(mu.reshape(-1))
            lss.append(ls.reshape(-1))
        mu, ls = np.concatenate(mus), np.concatenate(lss)
        mr, mi, si = I.mu_sig_indices(mu, ls)
        out["sha_mu_ls"] = np.array(sha(mu, ls))
        out["sha_indices"] = np.array(sha(mr, mi, si))
        out["head_mu"] = mu[:4096]
        out["head_ls"] = ls[:4096]
        p = os.path.join(HERE, f"int_cool_{name}.npz")
        np.savez_compressed(p, **out)
        print("wrote", p, os.path.getsize(p))
    if os.path.exists(tmp):
        os.remove(tmp)


if __name__ == "__main__":
    make_armint_cases()
    make_cool_cases()
