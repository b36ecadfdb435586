"""The drop-in boundary beyond plain eager calls:

  * torch.compile(frame_encoder, dynamic=False, mode="reduce-overhead", fullgraph=True), exactly what the reference
    does for every non-debug preset (enc/component/video.py:271-279): the forward is a custom operator, so Dynamo
    does not break the graph; results equal the eager ones bit for bit;
  * no process-global weight state: encoders trained at the same time on different streams (each FusedState owns a
    constant-bank weight slot) end bitwise where they end when trained one after the other;
  * run-to-run reproducibility of the fused iteration (the ARM halo accumulation is order-independent).
"""

import pytest
import torch

import common as Cm

pytestmark = pytest.mark.gpu

HOP = Cm.ARCHS["hop"]


def _frame_encoder(H, W, seed, arch=HOP):
    from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
    from coolchic_b200.enc.component.frame import FrameEncoder

    prm = CoolChicEncoderParameter(layers_synthesis=list(arch["syn"]), n_ft_per_res=[1] * 7, dim_arm=arch["dim_arm"],
                                   n_hidden_layers_arm=arch["nh"], encoder_gain=16)
    prm.set_image_size((H, W))
    fe = FrameEncoder(prm)
    p = Cm.O.init_params(Cm.make_arch("hop", H, W), seed=seed, B=1, random=True)
    fe.load_state_dict({"coolchic_encoder." + k: v for k, v in p.items()})
    fe.to_device("cuda:0")
    return fe


def test_torch_compile_fullgraph_matches_eager():
    from coolchic_b200.enc.training.loss import loss_function

    H, W = 40, 56
    fe = _frame_encoder(H, W, 5)
    target = Cm.O.synthetic_image(3, H, W).cuda()
    kw_eval = dict(quantizer_noise_type="none", quantizer_type="hardround", AC_MAX_VAL=-1, flag_additional_outputs=False)
    # ---- eval forward (what test() runs, test.py:402-411)
    fe.set_to_eval()
    with torch.no_grad():
        ref = fe(**kw_eval)
    cfe = torch.compile(fe, dynamic=False, mode="reduce-overhead", fullgraph=True)
    with torch.no_grad():
        for _ in range(3):  # warm-up, CUDA-graph capture of the surrounding partitions, replay
            out = cfe(**kw_eval)
            dec, rate = out.decoded_image.clone(), out.rate.clone()
    assert torch.equal(dec, ref.decoded_image) and torch.equal(rate, ref.rate)
    # ---- one training iteration (train.py:244-266) with the same noise in both
    fe.set_to_train()
    M = ref.rate.shape[1]
    noise = torch.rand(1, M, device="cuda", generator=torch.Generator("cuda").manual_seed(1))
    fe.coolchic_encoder._injected_noise = noise
    kw = dict(quantizer_noise_type="kumaraswamy", quantizer_type="softround",
              soft_round_temperature=torch.tensor(0.3, device="cuda"), noise_parameter=torch.tensor(2.0, device="cuda"))

    def grads_of(module):
        for q in fe.parameters():
            q.grad = None
        o = module(**kw)
        lo = loss_function(o.decoded_image, o.rate, target, lmbda=1e-3)
        lo.loss.backward()
        return lo.loss.detach().clone(), {k: v.grad.clone() for k, v in fe.named_parameters() if v.grad is not None}

    loss_e, g_e = grads_of(fe)
    ctrain = torch.compile(fe, dynamic=False, fullgraph=True)
    loss_c, g_c = grads_of(ctrain)
    assert torch.equal(loss_e, loss_c)
    assert g_e.keys() == g_c.keys() and len(g_e) > 20
    for k in g_e:
        assert torch.equal(g_e[k], g_c[k]), k


def _train_states(n, H, W, iters, concurrent):
    """n independent encoders x `iters` fused iterations; concurrent: each on its own stream, interleaved."""
    from coolchic_b200 import ops

    dev = torch.device("cuda:0")
    spec = ops.ArchSpec(H, W, HOP["dim_arm"], HOP["nh"], list(HOP["syn"]), 7, 8, 7, 16.0)
    states, targets, streams = [], [], []
    for i in range(n):
        st = ops.FusedState(spec, 1, dev)
        st.load(Cm.O.init_params(Cm.make_arch("hop", H, W), seed=20 + i, B=1, random=True))
        st.lr.fill_(1e-2)
        st.T.fill_(0.3)
        st.sigma.fill_(2.0)
        states.append(st)
        targets.append(Cm.O.synthetic_image(40 + i, H, W).to(dev))
        streams.append(torch.cuda.Stream(dev))
    assert len({st.slot.index for st in states}) == n and all(st.slot.index >= 0 for st in states)
    torch.cuda.synchronize()
    if concurrent:
        for _ in range(iters):
            for st, tg, s in zip(states, targets, streams):
                with torch.cuda.stream(s):
                    st.train_step(tg, 1e-3, "softround", "kumaraswamy", seed=7)
    else:
        for st, tg in zip(states, targets):
            for _ in range(iters):
                st.train_step(tg, 1e-3, "softround", "kumaraswamy", seed=7)
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    return [st.arena[0].clone() for st in states], [st.out4.clone() for st in states]


def test_two_streams_two_images_equal_sequential_runs():
    """Three encoders at once on different streams == the same encoders one after the other, bit for bit."""
    H, W = 96, 128
    seq_p, seq_o = _train_states(3, H, W, 40, concurrent=False)
    con_p, con_o = _train_states(3, H, W, 40, concurrent=True)
    for a, b in zip(seq_p, con_p):
        assert torch.equal(a, b)
    for a, b in zip(seq_o, con_o):
        assert torch.equal(a, b)
    assert not torch.equal(seq_p[0], seq_p[1])


def test_fused_iteration_is_bitwise_reproducible():
    """Equal seeds, 300 iterations, twice: identical parameters (no order-dependent float accumulation is left)."""
    a_p, a_o = _train_states(1, 135, 241, 300, concurrent=False)
    b_p, b_o = _train_states(1, 135, 241, 300, concurrent=False)
    assert torch.equal(a_p[0], b_p[0]) and torch.equal(a_o[0], b_o[0])


def test_weight_slots_are_exclusive_and_returned():
    from coolchic_b200 import _native as N

    held = [N.WeightSlot(0) for _ in range(N.N_WEIGHT_SLOTS + 2)]
    idx = [h.index for h in held]
    taken = [i for i in idx if i >= 0]
    assert len(taken) == len(set(taken)) and idx[-1] == -1  # exhausted: -1, never a shared slot
    for h in held:
        h.release()
    again = N.WeightSlot(0)
    assert again.index >= 0
    again.release()


def test_train_step_without_slot_gives_the_same_iteration():
    """weight_slot = -1 (all slots taken) runs the tensor-reading kernels: same loss, same gradients to fp32 noise."""
    from coolchic_b200 import ops

    dev = torch.device("cuda:0")
    H, W = 64, 96
    spec = ops.ArchSpec(H, W, HOP["dim_arm"], HOP["nh"], list(HOP["syn"]), 7, 8, 7, 16.0)
    tg = Cm.O.synthetic_image(1, H, W).to(dev)
    res = []
    for use_slot in (True, False):
        st = ops.FusedState(spec, 1, dev)
        if not use_slot:
            st.slot.release()
        st.load(Cm.O.init_params(Cm.make_arch("hop", H, W), seed=2, B=1, random=True))
        st.lr.fill_(1e-2)
        st.T.fill_(0.3)
        st.sigma.fill_(2.0)
        st.train_step(tg, 1e-3, "softround", "kumaraswamy", seed=3)
        torch.cuda.synchronize()
        res.append((st.out4.clone(), st.arena[1].clone()))
    assert torch.allclose(res[0][0], res[1][0], rtol=1e-5, atol=0)
    Cm.assert_close(res[0][1], res[1][1], "gradient arena, slot vs no slot", rtol=1e-4)


def test_images_in_flight_with_threaded_warmups_do_not_share_streams():
    """Three images in flight on one GPU, each running a threaded warm-up (4 candidates) and a train() phase: 15+ host
    threads that capture and replay CUDA graphs at the same time.  torch.cuda.Stream() draws from a round-robin pool
    of 32 streams per device, so pooled streams end up shared between threads and one thread's capture breaks another's
    replay ("Cannot prepare for replay during capturing stage"); every stream of this path is private
    (ops.new_stream).  Results equal the one-image-at-a-time run bit for bit."""
    import types

    from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
    from coolchic_b200.enc.component.frame import FrameEncoder
    from coolchic_b200.enc.training import sharding
    from coolchic_b200.enc.training.train import test as evaluate
    from coolchic_b200.enc.training.train import train
    from coolchic_b200.enc.training.warmup import warmup

    dev = torch.device("cuda:0")
    H, W = 48, 64
    _pool = [torch.cuda.Stream(dev) for _ in range(40)]  # walk through torch's pool once: later pooled streams repeat
    tp = dict(lr=1e-2, max_itr=30, freq_valid=30, patience=100000, schedule_lr=False, softround_temperature=[0.3, 0.3],
              noise_parameter=[2.0, 2.0], quantizer_noise_type="kumaraswamy", quantizer_type="softround",
              optimized_module=["all"])
    preset = {"warmup": {"phases": [{"candidates": 4, "training_phase": tp}, {"candidates": 2, "training_phase": tp}]}}

    def encode_one(i):
        torch.manual_seed(500 + i)
        prm = CoolChicEncoderParameter(layers_synthesis=list(HOP["syn"]), n_ft_per_res=[1] * 7, dim_arm=16,
                                       n_hidden_layers_arm=2, encoder_gain=16)
        prm.set_image_size((H, W))
        cands = []
        for c in range(4):
            fe = FrameEncoder(prm)
            p = Cm.O.init_params(Cm.make_arch("hop", H, W), seed=10 * i + c, B=1, random=True)
            fe.load_state_dict({"coolchic_encoder." + k: v for k, v in p.items()})
            cands.append(fe)
        target = Cm.O.synthetic_image(300 + i, H, W).to(dev)
        mgr = types.SimpleNamespace(lmbda=1e-3, iterations_counter=0, total_training_time_sec=0.0, preset=preset)
        fe = warmup(mgr, cands, target, dev)
        train(fe, target, mgr, start_lr=1e-2, cosine_scheduling_lr=True, max_iterations=40, frequency_validation=10,
              patience=1000, quantizer_type="softround", quantizer_noise_type="gaussian",
              softround_temperature=(0.3, 0.2), noise_parameter=(0.25, 0.1), seed=9000 + i)
        logs = evaluate(fe, target, mgr)
        return dict(loss=float(logs.loss), mse=float(logs.mse), psnr_db=float(logs.psnr_db),
                    rate_latent_bpp=float(logs.rate_latent_bpp), iterations=mgr.iterations_counter, seconds=0.0)

    seq = sharding.encode_images(5, encode_one, rank=0, world_size=1, device=dev, concurrency=1)
    con = sharding.encode_images(5, encode_one, rank=0, world_size=1, device=dev, concurrency=3)
    assert [r["iterations"] for r in con] == [4 * 30 + 2 * 30 + 40] * 5
    for a, b in zip(seq, con):
        assert a["index"] == b["index"] and a["loss"] == b["loss"] and a["mse"] == b["mse"], (a, b)
    del _pool
