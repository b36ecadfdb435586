import sys, numpy as np
sys.path.insert(0, "tests")
import test_gpu_convergence as T
z = np.load(T.GOLDEN)
ours = T._runs("mini", [10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21])
np.set_printoptions(precision=4, suppress=True, linewidth=200)
for k in ("loss", "psnr_db", "rate_latent_bpp", "rate_nn_bpp", "iterations"):
    print(k, "ref ", z["mini/" + k] * (1e3 if k == "loss" else 1))
    print(k, "ours", ours[k] * (1e3 if k == "loss" else 1))
print("mean loss ref/ours", z["mini/loss"].mean(), ours["loss"].mean())
print("mean psnr ref/ours", z["mini/psnr_db"].mean(), ours["psnr_db"].mean())
print("mean total rate ref/ours", (z["mini/rate_latent_bpp"] + z["mini/rate_nn_bpp"]).mean(), (ours["rate_latent_bpp"] + ours["rate_nn_bpp"]).mean())
