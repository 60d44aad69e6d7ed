"""N > 1 host logic on CPU: two gloo ranks shard one batch by image (no collective on the data path; the only
collectives are the bench's barrier and max-over-ranks timing reduce)."""
import hashlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import ssad_b200  # noqa: F401  (registers the package)
from ssad_b200 import GeneratorConfig, sample_params
from ssad_b200.params import shard_bounds


def test_shard_bounds_partition_the_batch():
    for batch in (0, 1, 7, 96, 512, 1025):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_bounds(batch, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == batch
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(8, 2, 2)


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, batch: int, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = GeneratorConfig(subject="bottle", imsize=(64, 64))
        full = sample_params(cfg, batch, np.random.default_rng(99))          # same seed on every rank
        mine = full.shard(rank, world)
        lo, hi = shard_bounds(batch, rank, world)
        # what travels between ranks: nothing on the data path.  For the check, gather digests and sizes.
        digest = hashlib.sha256(mine.records.tobytes()).digest()
        digests = [None] * world
        dist.all_gather_object(digests, (lo, hi, digest))
        # the bench's timing reduce: max over ranks
        t = torch.tensor([10.0 + rank])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.barrier()
        if rank == 0:
            ok = all(hashlib.sha256(full.records[a:b].tobytes()).digest() == dg for a, b, dg in digests)
            covered = sorted((a, b) for a, b, _ in digests)
            out.put((ok, covered, float(t.item())))
    finally:
        dist.destroy_process_group()


def test_two_gloo_ranks_shard_one_batch():
    world, batch = 2, 37
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, batch, out)) for r in range(world)]
    for p in procs:
        p.start()
    ok, covered, tmax = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok
    assert covered == [(0, 19), (19, 37)]
    assert tmax == 11.0
