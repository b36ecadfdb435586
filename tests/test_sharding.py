"""Multi-process (gloo, world_size 2, CPU) test of the image-sharded driver: round-robin partition, no collective
in the loop, one gather of the rate-distortion rows at the end."""

import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from coolchic_b200.enc.training import sharding as S


def test_shard_indices_partition():
    for n in (0, 1, 5, 8, 17):
        for world in (1, 2, 3, 8):
            parts = [S.shard_indices(n, r, world) for r in range(world)]
            assert sorted(i for p in parts for i in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        S.shard_indices(4, 2, 2)


def test_single_process_gather_is_identity():
    rows = S.encode_images(3, lambda i: dict(loss=i * 0.5, mse=0.1, psnr_db=10.0, rate_latent_bpp=1.0, iterations=7,
                                             seconds=0.0))
    assert [r["index"] for r in rows] == [0, 1, 2] and rows[2]["loss"] == 1.0


def _worker(rank, world, port, n_images, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    seen = []

    def encode_one(i):
        seen.append(i)
        return dict(loss=1.0 / (1 + i), mse=0.01 * i, psnr_db=20.0 + i, rate_latent_bpp=0.5 * i, iterations=100 + i,
                    seconds=float(rank))

    rows = S.encode_images(n_images, encode_one)
    q.put((rank, seen, rows))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_images", [5, 2, 1])
def test_two_ranks_gloo(n_images):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_images, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    got.sort()
    assert got[0][1] == list(range(0, n_images, 2)) and got[1][1] == list(range(1, n_images, 2))
    for _, _, rows in got:
        assert [r["index"] for r in rows] == list(range(n_images))
        for r in rows:
            i = r["index"]
            assert r["iterations"] == 100 + i and abs(r["psnr_db"] - (20.0 + i)) < 1e-12 and r["seconds"] == float(i % 2)
    assert got[0][2] == got[1][2]


def test_concurrent_images_per_rank_return_every_row_once():
    """concurrency > 1 (several images in flight per rank): every owned image is encoded exactly once, rows come back
    sorted, and an exception inside one encode reaches the caller."""
    import threading
    import time

    seen, lock = [], threading.Lock()
    in_flight = {"now": 0, "max": 0}

    def encode_one(i):
        with lock:
            seen.append(i)
            in_flight["now"] += 1
            in_flight["max"] = max(in_flight["max"], in_flight["now"])
        time.sleep(0.02)
        with lock:
            in_flight["now"] -= 1
        return dict(loss=float(i), mse=0.0, psnr_db=0.0, rate_latent_bpp=0.0, iterations=10 * i, seconds=0.0)

    rows = S.encode_images(9, encode_one, rank=1, world_size=2, concurrency=3)
    assert [r["index"] for r in rows] == [1, 3, 5, 7] and sorted(seen) == [1, 3, 5, 7]
    assert [r["iterations"] for r in rows] == [10, 30, 50, 70]
    assert 2 <= in_flight["max"] <= 3

    def failing(i):
        if i == 4:
            raise RuntimeError("boom")
        return encode_one(i)

    with pytest.raises(RuntimeError, match="boom"):
        S.encode_images(8, failing, rank=0, world_size=1, concurrency=2)
