"""GPU tests of the train() / test() mirror (enc/training/train.py) and of an end-to-end two-image encode through
the sharding driver (single process)."""

import math
import types

import pytest
import torch

from common import O
from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
from coolchic_b200.enc.component.frame import FrameEncoder
from coolchic_b200.enc.training import sharding as S
from coolchic_b200.enc.training.train import test as evaluate
from coolchic_b200.enc.training.train import train

pytestmark = pytest.mark.gpu
SYN = ["16-1-linear-relu", "X-1-linear-none", "X-3-residual-relu", "X-3-residual-none"]


def make_encoder(H, W, dev, seed=0):
    # the reference's random initialisation draws from torch's global generator: pin it, an unlucky draw
    # (dead ReLU in the last synthesis layer) leaves any implementation stuck at the initial loss
    torch.manual_seed(1234 + seed)
    p = CoolChicEncoderParameter(layers_synthesis=list(SYN), n_ft_per_res=[1] * 7, dim_arm=8, n_hidden_layers_arm=2,
                                 ups_k_size=8, ups_preconcat_k_size=7, encoder_gain=16)
    p.set_image_size((H, W))
    fe = FrameEncoder(p)
    fe.to_device(dev)
    return fe


def manager(lmbda=1e-3):
    return types.SimpleNamespace(lmbda=lmbda, iterations_counter=0, total_training_time_sec=0.0)


def test_train_improves_and_returns_best():
    dev = torch.device("cuda:0")
    H, W = 64, 96
    target = O.synthetic_image(5, H, W).to(dev)
    fe = make_encoder(H, W, dev)
    mgr = manager()
    before = evaluate(fe, target, mgr)
    logs = []
    train(fe, target, mgr, start_lr=1e-2, max_iterations=400, frequency_validation=50, patience=100000,
          val_logs=logs, seed=3)
    after = evaluate(fe, target, mgr)
    assert mgr.iterations_counter == 400 and len(logs) == 8
    assert float(after.loss) < 0.5 * float(before.loss), (float(before.loss), float(after.loss))
    assert after.psnr_db > before.psnr_db + 3.0
    # the encoder holds the best validated model: its loss is the minimum over the records
    best = min(float(l.loss) for l in logs)
    assert float(after.loss) <= best * (1 + 1e-5) + 1e-9
    assert math.isfinite(after.rate_latent_bpp) and after.rate_latent_bpp > 0


def test_patience_without_cosine_stops_early_and_module_mask():
    dev = torch.device("cuda:0")
    H, W = 32, 48
    target = O.synthetic_image(6, H, W).to(dev)
    fe = make_encoder(H, W, dev)
    mgr = manager()
    # zero learning rate: never a record -> the loop must stop right after `patience` iterations
    train(fe, target, mgr, start_lr=0.0, cosine_scheduling_lr=False, max_iterations=300, frequency_validation=10,
          patience=20)
    assert mgr.iterations_counter == 21
    # only the synthesis is optimised: everything else must stay bit-identical
    p0 = fe.get_param()
    train(fe, target, mgr, start_lr=1e-2, max_iterations=60, frequency_validation=20, patience=1000,
          optimized_module=["synthesis"])
    p1 = fe.get_param()
    changed = {k for k in p0 if not torch.equal(p0[k], p1[k])}
    assert changed and all(k.startswith("coolchic_encoder.synthesis.") or k.startswith("synthesis.") for k in changed), changed
    with pytest.raises(ValueError):
        train(fe, target, mgr, max_iterations=1, optimized_module=["nope"])


def test_encode_two_images_through_the_sharding_driver():
    dev = torch.device("cuda:0")
    H, W = 32, 48

    def encode_one(i):
        target = O.synthetic_image(10 + i, H, W).to(dev)
        fe = make_encoder(H, W, dev, seed=i)
        mgr = manager()
        train(fe, target, mgr, max_iterations=100, frequency_validation=50, patience=1000, seed=i)
        l = evaluate(fe, target, mgr)
        return dict(loss=float(l.loss), mse=l.mse, psnr_db=l.psnr_db, rate_latent_bpp=l.rate_latent_bpp,
                    iterations=mgr.iterations_counter, seconds=mgr.total_training_time_sec)

    rows = S.encode_images(2, encode_one)
    assert [r["index"] for r in rows] == [0, 1] and all(r["iterations"] == 100 for r in rows)
    assert all(math.isfinite(r["loss"]) and r["psnr_db"] > 10 for r in rows)


def test_validation_on_the_arena_equals_test_on_the_module():
    """train() validates straight from the arena (no copy into the module); the record it keeps must be exactly what
    test() reports once the parameters are back in the module."""
    dev = torch.device("cuda:0")
    H, W = 48, 80
    target = O.synthetic_image(9, H, W).to(dev)
    fe = make_encoder(H, W, dev, seed=2)
    mgr = manager()
    logs = []
    train(fe, target, mgr, start_lr=1e-2, max_iterations=120, frequency_validation=120, patience=100000, val_logs=logs,
          seed=11)
    assert len(logs) == 1
    after = evaluate(fe, target, mgr)
    assert float(after.loss) == float(logs[0].loss)
    assert after.mse == logs[0].mse and after.rate_latent_bpp == logs[0].rate_latent_bpp


def test_warmup_keeps_the_best_candidates():
    """warmup() (reference: enc/training/warmup.py:21-158): 4 random candidates x 60 iterations, the best 2 x 60 more;
    the winner is the candidate with the lowest test() loss, the manager's iteration counter saw every iteration, and
    training the winner's starting point directly through train() with the same seeds lands on the same loss, bit for
    bit; the candidates of a phase train concurrently (one thread + stream each) with the sequential search's result."""
    import copy
    import types

    from coolchic_b200.enc.component.coolchic import CoolChicEncoderParameter
    from coolchic_b200.enc.component.frame import FrameEncoder
    from coolchic_b200.enc.training.train import test as evaluate
    from coolchic_b200.enc.training.train import train
    from coolchic_b200.enc.training.warmup import warmup
    from oracle import float_ref as O

    dev = torch.device("cuda:0")
    H, W = 48, 64
    target = O.synthetic_image(5, H, W).to(dev)
    cands = []
    for i in range(4):
        torch.manual_seed(100 + i)
        prm = CoolChicEncoderParameter(layers_synthesis=["16-1-linear-relu", "3-1-linear-none"], n_ft_per_res=[1] * 7,
                                       dim_arm=8, n_hidden_layers_arm=1, ups_k_size=8, ups_preconcat_k_size=7,
                                       encoder_gain=16)
        prm.set_image_size((H, W))
        cands.append(FrameEncoder(prm))
    start = [copy.deepcopy(c) for c in cands]
    seq = [copy.deepcopy(c) for c in cands]  # untouched starting points for the sequential search below
    tp = dict(lr=1e-2, max_itr=60, freq_valid=60, patience=100000, schedule_lr=False, softround_temperature=[0.3, 0.3],
              noise_parameter=[2.0, 2.0], quantizer_noise_type="kumaraswamy", quantizer_type="softround",
              optimized_module=["all"])
    preset = {"warmup": {"phases": [{"candidates": 4, "training_phase": tp}, {"candidates": 2, "training_phase": tp}]}}
    mgr = types.SimpleNamespace(lmbda=1e-3, iterations_counter=0, total_training_time_sec=0.0, preset=preset)
    winner = warmup(mgr, cands, target, dev)
    assert mgr.iterations_counter == 4 * 60 + 2 * 60
    losses = [float(evaluate(c, target, mgr).loss) for c in cands]
    w_loss = float(evaluate(winner, target, mgr).loss)
    # the two finalists trained 120 iterations, the others 60; the winner is one of the candidate objects' copies
    assert w_loss == min(losses)
    wid = losses.index(w_loss)
    # replay the winner's history by hand: same seeds (1000 * phase + id), same phases
    again = start[wid]
    again.to_device(dev)
    m2 = types.SimpleNamespace(lmbda=1e-3, iterations_counter=0, total_training_time_sec=0.0)
    for phase in range(2):
        train(again, target, m2, start_lr=1e-2, cosine_scheduling_lr=False, max_iterations=60, patience=100000,
              frequency_validation=60, optimized_module=("all",), quantizer_type="softround",
              quantizer_noise_type="kumaraswamy", softround_temperature=(0.3, 0.3), noise_parameter=(2.0, 2.0),
              seed=1000 * phase + wid)
    # bitwise: the fused iteration has no order-dependent accumulation left (the ARM halo gradients go through two
    # accumulators that each receive at most two adds), and the candidates trained side by side on their own streams
    # and weight slots do not see each other
    assert float(evaluate(again, target, m2).loss) == w_loss
    # the same search with the candidates one after the other: same winner, same parameters
    m3 = types.SimpleNamespace(lmbda=1e-3, iterations_counter=0, total_training_time_sec=0.0, preset=preset)
    winner_seq = warmup(m3, seq, target, dev, concurrent=False)
    assert m3.iterations_counter == mgr.iterations_counter
    pa, pb = winner.get_param(), winner_seq.get_param()
    assert pa.keys() == pb.keys()
    for k in pa:
        assert torch.equal(pa[k], pb[k]), k
    with pytest.raises(ValueError):
        warmup(types.SimpleNamespace(lmbda=1e-3, preset={"warmup": {"phases": [{"candidates": 9, "training_phase": tp}]}}),
               cands, target, dev)
