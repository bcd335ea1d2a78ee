"""Read sharding across GPUs: every pair is mapped independently against a replicated index,
so the only multi-GPU logic is (a) dealing contiguous shards of the input to the ranks, (b) bringing the
per-pair results back to rank 0 in input order for the unchanged host stages (output writers, stage 2), and
(c) the max-over-ranks of the device timings.  No collective sits on the data path of the kernels.

One process per GPU; `torch.distributed` (nccl on the GPU box, gloo in the CPU tests) is the plumbing."""
from __future__ import annotations

import numpy as np

DEFAULT_SHARD_PAIRS = 262144


def deal_shards(n_pairs: int, world_size: int, shard_pairs: int = DEFAULT_SHARD_PAIRS) -> list[list[tuple[int, int]]]:
    """Contiguous shards [start, end) of the input dealt round-robin to the ranks."""
    if world_size < 1 or shard_pairs < 1:
        raise ValueError("world_size and shard_pairs must be >= 1")
    out: list[list[tuple[int, int]]] = [[] for _ in range(world_size)]
    for k, start in enumerate(range(0, n_pairs, shard_pairs)):
        out[k % world_size].append((start, min(n_pairs, start + shard_pairs)))
    return out


def local_indices(shards: list[tuple[int, int]]) -> np.ndarray:
    """Global pair indices a rank owns, in the order it processes them."""
    if not shards:
        return np.zeros(0, np.int64)
    return np.concatenate([np.arange(a, b, dtype=np.int64) for a, b in shards])


def gather_in_input_order(local: np.ndarray, n_pairs: int, rank: int, world_size: int,
                          shard_pairs: int = DEFAULT_SHARD_PAIRS, dist=None):
    """Collect the per-pair result rows of every rank on rank 0, ordered like the input.

    `local` holds this rank's rows in `local_indices(deal_shards(...)[rank])` order (any numpy dtype,
    structured ones included: rows travel as bytes).  Returns the full array on rank 0, None elsewhere."""
    plan = deal_shards(n_pairs, world_size, shard_pairs)
    local = np.ascontiguousarray(local)
    if len(local) != sum(b - a for a, b in plan[rank]):
        raise ValueError("local result count does not match this rank's shards")
    if world_size == 1:
        parts = [local]
    else:
        import torch
        raw = torch.from_numpy(local.view(np.uint8).reshape(-1).copy())
        if dist.get_backend() == "nccl":
            raw = raw.cuda()
        sizes = [sum(b - a for a, b in plan[r]) * local.dtype.itemsize for r in range(world_size)]
        bufs = [torch.empty(s, dtype=torch.uint8, device=raw.device) for s in sizes] if rank == 0 else None
        # ragged gather as point-to-point copies (results are ~100 B per pair, once per batch)
        if rank == 0:
            bufs[0].copy_(raw)
            for r in range(1, world_size):
                if sizes[r]:
                    dist.recv(bufs[r], src=r)
        elif sizes[rank]:
            dist.send(raw, dst=0)
        if rank != 0:
            return None
        parts = [b.cpu().numpy().view(local.dtype) for b in bufs]
    full = np.empty(n_pairs, dtype=local.dtype)
    for r, part in enumerate(parts):
        full[local_indices(plan[r])] = part
    return full


def max_over_ranks(values, world_size: int, dist=None, device=None) -> list[float]:
    """Element-wise max of a small list of floats over all ranks (device timings of a step)."""
    if world_size == 1:
        return [float(v) for v in values]
    import torch
    t = torch.tensor([float(v) for v in values], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]
