"""Golden vectors for the fixed-point decoder's upsampling + synthesis from the REAL reference.

Run in the build container only (needs /root/reference and oracle/_ref/libccref.so from oracle/build.sh):

    python tests/golden/make_dec_golden.py

For each shipped bitstream the reference's own C++ (compiled in place) parses the file, entropy-decodes the latents
and the integer filters / synthesis weights, and runs its run_ups + run_syn.  Stored in tests/golden/int_dec_<name>.npz:
the integer weights, the latent grids (int8; omitted when tests/golden/int_cool_<name>.npz already holds them), SHA-256
digests of the upsampled planes, of the synthesis output and of the 8-bit pixels, and a strided sample of the output.
"""

import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import dec_int_ref as D  # noqa: E402
from oracle import int_ref as I  # noqa: E402

REFERENCE_ROOT = os.environ.get("COOLCHIC_REFERENCE", "/root/reference")
CASES = [("kodak", "kodim01-lmbda-0001"), ("kodak", "kodim02-lmbda-002"), ("kodak", "kodim13-lmbda-00001"),
         ("clic20-pro-valid", "amy-zhang-15940-lmbda-0004"),      # 915 x 1632: odd height, ragged pyramid
         ("clic20-pro-valid", "davide-ragusa-716-lmbda-0004")]    # 1292 x 1945: odd width


def sha(a, dtype=np.int32):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=dtype).tobytes()).hexdigest()


def main():
    for ds, name in CASES:
        f = I.CoolFile(os.path.join(REFERENCE_ROOT, "results", "image", ds, "bitstreams", name + ".cool"))
        prm = f.decoder_params()
        assert prm["n_branches"] == 1
        ups, out = f.run_ups_syn()
        o = {"img_hw": np.array([f.img_h, f.img_w], np.int32), "n_latents": np.int32(f.n_latents),
             "ups_w": prm["ups_w"], "pre_w": prm["pre_w"],
             "syn_desc": np.array([[l["w"].shape[0], l["ks"], int(l["residual"]), int(l["relu"])] for l in prm["syn"]],
                                  np.int32),
             "sha_ups": np.array(sha(ups)), "sha_out": np.array(sha(out)),
             "sha_pix": np.array(sha(D.to_pixels(out), np.uint8)),
             "sample_out": out[:, ::37, ::41].copy(), "sample_ups": ups[:, ::37, ::41].copy()}
        for i, l in enumerate(prm["syn"]):
            o[f"syn{i}_w"], o[f"syn{i}_b"] = l["w"], l["b"]
        if not os.path.exists(os.path.join(HERE, f"int_cool_{name}.npz")):
            for k, g in enumerate(f.grids):
                assert g.min() >= -128 and g.max() <= 127
                o[f"grid{k}"] = g.astype(np.int8)
        # the restatement must agree before the vectors are trusted as ITS pin
        r = D.decode(f.grids, prm)
        assert np.array_equal(r["ups"], ups) and np.array_equal(r["out"], out), name
        p = os.path.join(HERE, f"int_dec_{name}.npz")
        np.savez_compressed(p, **o)
        print("wrote", p, os.path.getsize(p) // 1024, "kB", (f.img_h, f.img_w))
        f.close()


if __name__ == "__main__":
    main()
