"""The noise the TIMED path draws: in-kernel Philox4x32-10 + Box-Muller / inverse Kumaraswamy CDF (csrc/ccb_quant.cu).

Every parity test injects the reference's own noise tensor; bench.py and train() do not (the reference draws from
torch's generator, enc/component/core/quantizer.py:186-199, which no other implementation can replay).  What can be
pinned is the DISTRIBUTION: moments and a Kolmogorov-Smirnov distance against the closed-form CDFs for 524 256 draws,
independence of consecutive iterations (the Philox offset advances by one per step), and reproducibility for equal
(seed, offset).  With quantizer "none" and zero latents, y_hat IS the shaped noise (quantizer.py:200-203), and the fused
step leaves y_hat at the start of its workspace.
"""

import math

import numpy as np
import pytest
import torch

import common as Cm

pytestmark = pytest.mark.gpu

HOP = Cm.ARCHS["hop"]
H, W = 512, 768


def _draw(noise_type, sigma, seed, n_steps=1, offset0=0):
    """-> list of n_steps arrays [M] of the noise the kernel adds at consecutive iterations."""
    from coolchic_b200 import ops

    dev = torch.device("cuda:0")
    spec = ops.ArchSpec(H, W, HOP["dim_arm"], HOP["nh"], list(HOP["syn"]), 7, 8, 7, 16.0)
    st = ops.FusedState(spec, 1, dev)
    p = Cm.O.init_params(Cm.make_arch("hop", H, W), seed=1, B=1, random=True)
    for i in range(7):
        p[f"latent_grids.{i}.data"].zero_()
    st.load(p)
    st.lr.fill_(0.0)  # the latents stay zero
    st.T.fill_(0.3)
    st.sigma.fill_(sigma)
    st.philox_offset.fill_(offset0)
    target = torch.zeros(1, 3, H, W, device=dev)
    M = st.plan.M
    out = []
    for _ in range(n_steps):
        st.train_step(target, 1e-3, "none", noise_type, seed=seed)
        torch.cuda.synchronize()
        out.append(st.ws[:4 * M].view(torch.float32).double().cpu().numpy().copy())
    assert int(st.philox_offset) == offset0 + n_steps
    st.close()
    return out


def _ks(sample, cdf):
    x = np.sort(sample)
    n = x.size
    f = cdf(x)
    return max(np.max(np.arange(1, n + 1) / n - f), np.max(f - np.arange(0, n) / n))


def _phi(x):
    return 0.5 * (1.0 + np.vectorize(math.erf)(x / math.sqrt(2.0)))


def test_gaussian_noise_moments_and_ks():
    sigma = 0.25
    (n,) = _draw("gaussian", sigma, seed=11)
    z = n / sigma
    N = z.size
    assert N == 524256
    assert abs(z.mean()) < 5 / math.sqrt(N)                      # 5 standard errors
    assert abs(z.var() - 1.0) < 5 * math.sqrt(2.0 / N)
    assert abs((z ** 3).mean()) < 5 * math.sqrt(15.0 / N)         # skewness
    assert abs((z ** 4).mean() - 3.0) < 5 * math.sqrt(96.0 / N)   # kurtosis
    assert np.abs(z).max() > 4.0 and np.abs(z).max() < 6.5        # tails are there, nothing absurd
    sub = z[:: 8]                                                 # KS on 65 532 draws: critical value at p = 1e-3
    assert _ks(sub, _phi) < 1.95 / math.sqrt(sub.size)
    # the same test on torch's own generator, for scale
    ref = torch.randn(sub.size, generator=torch.Generator().manual_seed(0)).double().numpy()
    assert _ks(ref, _phi) < 1.95 / math.sqrt(sub.size)


def test_kumaraswamy_noise_matches_its_cdf():
    a = 2.0                                                       # noise_parameter (quantizer.py:62-107)
    b = (2 ** a * (a - 1) + 1) / a
    (n,) = _draw("kumaraswamy", a, seed=12)
    u = n + 0.5                                                   # the kernel returns the draw shifted to [-0.5, 0.5]
    assert u.min() >= 0.0 and u.max() <= 1.0
    cdf = lambda x: 1.0 - (1.0 - np.clip(x, 0, 1) ** a) ** b      # noqa: E731
    sub = u[:: 8]
    assert _ks(sub, cdf) < 1.95 / math.sqrt(sub.size)
    # mean and variance of Kumaraswamy(a, b): m_k = b * B(1 + k / a, b)
    mom = lambda k: b * math.gamma(1 + k / a) * math.gamma(b) / math.gamma(1 + k / a + b)  # noqa: E731
    N = u.size
    var = mom(2) - mom(1) ** 2
    assert abs(u.mean() - mom(1)) < 5 * math.sqrt(var / N)
    assert abs(u.var() - var) < 0.01 * var
    # the reference's mode is at 0.5: the shifted noise is centred (quantizer.py:95-99)
    hist, edges = np.histogram(u, bins=50, range=(0, 1))
    assert abs(edges[hist.argmax()] + 0.01 - 0.5) <= 0.03


def test_consecutive_iterations_are_independent_and_reproducible():
    a, b2, c = _draw("gaussian", 1.0, seed=21, n_steps=3)
    N = a.size
    for x, y in ((a, b2), (b2, c), (a, c)):
        r = np.corrcoef(x, y)[0, 1]
        assert abs(r) < 5 / math.sqrt(N), r                       # same positions, different offsets: uncorrelated
    # neighbouring positions within one draw (the four outputs of one Philox block, Box-Muller pairs)
    for lag in (1, 2, 3, 4):
        r = np.corrcoef(a[:-lag], a[lag:])[0, 1]
        assert abs(r) < 5 / math.sqrt(N), (lag, r)
    assert not np.array_equal(a, b2)
    # equal (seed, offset) -> equal noise; another seed -> another stream
    (again,) = _draw("gaussian", 1.0, seed=21, offset0=1)
    assert np.array_equal(again, b2)
    (other,) = _draw("gaussian", 1.0, seed=22)
    assert not np.array_equal(other, a)
    assert abs(np.corrcoef(other, a)[0, 1]) < 5 / math.sqrt(N)
