"""N > 1 host logic on CPU: two gloo ranks deal the shards, "map" their pairs, and rank 0 gets the per-pair
results back in input order; timings reduce as max over ranks (the bench.py contract)."""
import os
import socket

import numpy as np
import pytest

from circminer_b200 import sharding


def test_deal_shards_covers_every_pair_once():
    for n, w, sp in ((0, 2, 4), (10, 1, 3), (10, 2, 3), (1000003, 8, 4096), (5, 8, 1)):
        plan = sharding.deal_shards(n, w, sp)
        idx = np.concatenate([sharding.local_indices(p) for p in plan]) if n else np.zeros(0, np.int64)
        assert sorted(idx.tolist()) == list(range(n))
        for p in plan:
            assert all(b - a <= sp and a < b for a, b in p)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, shard_pairs, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        plan = sharding.deal_shards(n, world, shard_pairs)
        mine = sharding.local_indices(plan[rank])
        # stand-in for cm_gpu_process_batch_pe: a structured per-pair record that depends on the global index only
        rec = np.zeros(len(mine), dtype=[("type", "<i4"), ("spos", "<u4"), ("tlen", "<i4")])
        rec["type"] = mine % 14
        rec["spos"] = (mine * 2654435761) % (1 << 32)
        rec["tlen"] = -mine.astype(np.int64) % 1000
        full = sharding.gather_in_input_order(rec, n, rank, world, shard_pairs, dist=dist)
        ms = sharding.max_over_ranks([10.0 + rank, 5.0 - rank], world, dist=dist)
        if rank == 0:
            g = np.arange(n, dtype=np.int64)
            ok = (full["type"] == g % 14).all() and (full["spos"] == (g * 2654435761) % (1 << 32)).all() \
                and (full["tlen"] == -g % 1000).all()
            q.put((bool(ok), ms))
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,shard_pairs", [(1000, 64), (7, 100), (4096, 1024)])
def test_two_ranks_gather_in_input_order(n, shard_pairs):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, shard_pairs, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, ms = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok
    assert ms == [11.0, 5.0]
