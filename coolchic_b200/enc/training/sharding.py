"""Multi-GPU encoding of independent images.

Per-image overfitting shares nothing between images, so the path shards by image and nothing else: one process
per GPU (torchrun), image i goes to rank i mod world_size, NO collective inside the loop, and one gather of the
per-image rate-distortion rows at the end (the reference's equivalent is one SLURM / shell job per image:
`scripts/`, `coolchic/encode.py:29-188` writing one results row per sequence, `eval/results.py`).

`gather_rows` works on any initialised torch.distributed backend (NCCL over NVLink on the GPU box, gloo in the
CPU tests); rows are fixed-width float64 so the exchange is a single all_gather of a padded tensor.
"""

from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence

import torch

ROW_FIELDS = ("index", "loss", "mse", "psnr_db", "rate_latent_bpp", "iterations", "seconds")


def shard_indices(n_items: int, rank: int, world_size: int) -> List[int]:
    """Round-robin partition: item i is owned by rank i mod world_size."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return list(range(rank, n_items, world_size))


def rows_to_tensor(rows: Sequence[Dict[str, float]], capacity: int, device) -> torch.Tensor:
    """[capacity, len(ROW_FIELDS)] float64, unused rows flagged by index = -1."""
    t = torch.full((capacity, len(ROW_FIELDS)), -1.0, dtype=torch.float64)
    for i, r in enumerate(rows):
        t[i] = torch.tensor([float(r[f]) for f in ROW_FIELDS], dtype=torch.float64)
    return t.to(device)


def gather_rows(rows: Sequence[Dict[str, float]], n_items: int, device=None, group=None) -> List[Dict[str, float]]:
    """All ranks receive every rank's rows, sorted by image index.  Without an initialised process group (single
    process) the local rows are returned sorted."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return sorted((dict(r) for r in rows), key=lambda r: r["index"])
    world = dist.get_world_size(group)
    capacity = (n_items + world - 1) // world
    if len(rows) > capacity:
        raise ValueError("a rank holds more rows than the round-robin partition allows")
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else "cpu"
    mine = rows_to_tensor(rows, capacity, device)
    out = torch.empty((world * capacity, len(ROW_FIELDS)), dtype=torch.float64, device=device)
    dist.all_gather_into_tensor(out, mine, group=group)
    res = []
    for row in out.cpu().tolist():
        if row[0] >= 0:
            res.append({f: (int(v) if f in ("index", "iterations") else v) for f, v in zip(ROW_FIELDS, row)})
    res.sort(key=lambda r: r["index"])
    if [r["index"] for r in res] != sorted(set(r["index"] for r in res)):
        raise RuntimeError("duplicate image index in gathered rows")
    return res


def encode_images(n_images: int, encode_one: Callable[[int], Dict[str, float]], rank: Optional[int] = None,
                  world_size: Optional[int] = None, device=None, concurrency: int = 1) -> List[Dict[str, float]]:
    """Run `encode_one(i)` for the images this rank owns, then gather every row on every rank.
    `encode_one` returns a dict with the ROW_FIELDS keys (index is filled in here).

    concurrency > 1 keeps that many of this rank's images in flight on ITS GPU: one host thread and one CUDA stream
    per image in flight (SURVEY.md 8e "in-GPU concurrency (still replicas)").  An iteration uses well under half of
    a B200's issue slots, each encoder owns its weight slot and workspace, and per-image results do not depend on
    what else runs (tests/test_gpu_boundary.py), so this raises images/hour without touching the path."""
    import torch.distributed as dist

    if rank is None or world_size is None:
        if dist.is_available() and dist.is_initialized():
            rank, world_size = dist.get_rank(), dist.get_world_size()
        else:
            rank, world_size = 0, 1
    mine = shard_indices(n_images, rank, world_size)
    rows: List[Dict[str, float]] = []
    if concurrency <= 1 or len(mine) <= 1:
        for i in mine:
            r = dict(encode_one(i))
            r["index"] = i
            rows.append(r)
        return gather_rows(rows, n_images, device=device)

    import threading

    use_cuda = torch.cuda.is_available()
    lock = threading.Lock()
    todo = list(mine)
    errors: List[BaseException] = []

    def worker():
        stream = None
        if use_cuda:
            from ... import ops

            stream = ops.new_stream(device if device is not None else torch.device("cuda", torch.cuda.current_device()))
        while True:
            with lock:
                if not todo or errors:
                    return
                i = todo.pop(0)
            try:
                if stream is not None:
                    with torch.cuda.stream(stream):
                        r = dict(encode_one(i))
                        stream.synchronize()
                else:
                    r = dict(encode_one(i))
            except BaseException as e:  # re-raised on the calling thread
                with lock:
                    errors.append(e)
                return
            r["index"] = i
            with lock:
                rows.append(r)

    threads = [threading.Thread(target=worker, daemon=True) for _ in range(min(concurrency, len(mine)))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
    rows.sort(key=lambda r: r["index"])
    return gather_rows(rows, n_images, device=device)
