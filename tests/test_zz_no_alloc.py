"""ccb_loss_noalloc (include/ccb.h): the allocation-free variant of ccb_loss, against ccb_loss itself, through the C ABI.

The file name sorts last on purpose: a failure here must not hide other tests.
Two calls in a row on different inputs (the second finds the first call's contents in the shared static scratch),
the size limit, and that no device memory is taken by the calls.  Passes alone in a fresh process and after the others.
"""

import ctypes as C

import pytest
import torch

import common as Cm

pytestmark = pytest.mark.gpu

HOP = Cm.ARCHS["hop"]


def _inputs(B, H, W, M, seed, dev):
    g = torch.Generator(device="cuda").manual_seed(seed)
    raw = torch.rand(B, 3, H, W, device=dev, generator=g) * 1.4 - 0.2  # some values outside [0, 1]: the clamp mask
    target = torch.rand(B, 3, H, W, device=dev, generator=g)
    rate = torch.rand(B, M, device=dev, generator=g) * 8.0
    return raw, rate, target


def _reference_loss(spec, B, raw, rate, target, lmbda):
    """The existing entry point: ccb_loss on a plan workspace."""
    from coolchic_b200 import ops

    plan = spec.plan(B, raw.device)
    ws = torch.empty(plan.ws_bytes, dtype=torch.uint8, device=raw.device)
    out4, d_raw = ops.fused_loss(spec, plan, ws, raw, rate, target, lmbda)
    torch.cuda.synchronize()
    return out4, d_raw


def _noalloc(B, H, W, M, raw, rate, target, lmbda, d_raw, out4):
    from coolchic_b200 import _native as N

    return N.lib().ccb_loss_noalloc(0, B, H, W, M, raw.data_ptr(), rate.data_ptr(), target.data_ptr(), float(lmbda),
                                    None if d_raw is None else d_raw.data_ptr(), out4.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("B,H,W", [(1, 135, 241), (2, 256, 256)])
def test_noalloc_loss_matches_ccb_loss_twice_in_a_row(B, H, W):
    from coolchic_b200 import _native as N
    from coolchic_b200 import ops

    dev = torch.device("cuda:0")
    spec = ops.ArchSpec(H, W, HOP["dim_arm"], HOP["nh"], list(HOP["syn"]), 7, 8, 7, 16.0)
    M = spec.plan(B, dev).M
    lmbda = 1e-3
    cases = []
    for seed in (1, 2):  # two different inputs: the second call must not see anything of the first
        raw, rate, target = _inputs(B, H, W, M, seed, dev)
        ref_out4, ref_draw = _reference_loss(spec, B, raw, rate, target, lmbda)
        cases.append((raw, rate, target, ref_out4, ref_draw))
    outs = [(torch.full((4,), 7.0, device=dev), torch.full((B, 3, H, W), 7.0, device=dev)) for _ in cases]
    torch.cuda.synchronize()
    N.check(_noalloc(B, H, W, M, *cases[0][:3], lmbda, outs[0][1], outs[0][0]), "ccb_loss_noalloc (set-up call)")
    torch.cuda.synchronize()
    free0, alloc0 = torch.cuda.mem_get_info(dev)[0], torch.cuda.memory_allocated(dev)
    for (raw, rate, target, _, _), (out4, d_raw) in zip(cases, outs):  # back to back, no sync in between
        N.check(_noalloc(B, H, W, M, raw, rate, target, lmbda, d_raw, out4), "ccb_loss_noalloc")
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] == free0 and torch.cuda.memory_allocated(dev) == alloc0
    for (_, _, _, ref_out4, ref_draw), (out4, d_raw) in zip(cases, outs):
        assert torch.equal(d_raw, ref_draw)  # element-wise, same expression
        Cm.assert_close(out4[:3], ref_out4[:3], "loss / mse / bpp", rtol=1e-5)  # sums in a different order
        assert float(out4[3]) == 0.0
    assert not torch.equal(outs[0][0], outs[1][0])


def test_noalloc_loss_without_gradient_output_and_size_limit():
    from coolchic_b200 import _native as N

    dev = torch.device("cuda:0")
    B, H, W, M = 1, 64, 96, 8191
    raw, rate, target = _inputs(B, H, W, M, 5, dev)
    out4 = torch.zeros(4, device=dev)
    N.check(_noalloc(B, H, W, M, raw, rate, target, 1e-3, None, out4), "ccb_loss_noalloc")
    torch.cuda.synchronize()
    mse = ((raw.clamp(0, 1) - target) ** 2).mean()
    bpp = rate.sum() / (B * H * W)
    Cm.assert_close(out4[:3], torch.stack([mse + 1e-3 * bpp, mse, bpp]), "loss / mse / bpp vs torch", rtol=1e-5)
    # one element beyond the static scratch: refused with the dedicated status, nothing launched, out4 untouched
    before = out4.clone()
    too_big_h = N.NOALLOC_MAX_ELEMENTS // (3 * 4096) + 1
    rc = _noalloc(1, too_big_h, 4096, M, raw, rate, target, 1e-3, None, out4)
    assert rc == -5 and b"static scratch" in N.lib().ccb_strerror(rc)
    torch.cuda.synchronize()
    assert torch.equal(out4, before)
    with pytest.raises(N.CcbError):
        N.check(rc, "ccb_loss_noalloc")
