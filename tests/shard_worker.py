"""Worker of tests/test_gpu_sharding.py (one process per rank, launched by torch.distributed.run): deal the pairs of T1 to
the ranks (sharding.deal_shards), map every rank's pairs on the GPU through the rounds API, gather the per-pair rows on rank 0
in input order (sharding.gather_in_input_order) and require them to equal the single-context run of the whole input.
The ranks share cuda:0 (the test box has one GPU), so the gather runs over gloo; on a multi-GPU box pass --nccl."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from circminer_b200 import sharding, workload
    from circminer_b200.api import Mapper, default_params, fresh_mreads
    data, shard_pairs = sys.argv[1], int(sys.argv[2])
    nccl = "--nccl" in sys.argv
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = int(os.environ.get("LOCAL_RANK", "0")) if nccl else 0
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl" if nccl else "gloo")
    w = workload.prepare("T1", os.path.join(data, f"t1_r{rank}"), contig_size=3_500_000, n_pairs=12_000)
    n = len(w["s1"])
    nc = len(w["contigs"])
    m = Mapper(default_params(), device=dev)
    workload.load_into(m, w)
    plan = sharding.deal_shards(n, world, shard_pairs)
    mine = sharding.local_indices(plan[rank])
    res = m.map_rounds(w["s1"][mine], w["s2"][mine], fresh_mreads(len(mine)), 0, nc)
    rows = np.zeros(len(mine), np.dtype([("mr", res["out"].dtype), ("state", "<i4"), ("round", "<i4")]))
    rows["mr"], rows["state"], rows["round"] = res["out"], res["state"], res["final_round"]
    full = sharding.gather_in_input_order(rows, n, rank, world, shard_pairs, dist=dist)
    ok = True
    if rank == 0:
        one = m.map_rounds(w["s1"], w["s2"], fresh_mreads(n), 0, nc)
        ok = bool((full["mr"] == one["out"]).all() and (full["state"] == one["state"]).all() and (full["round"] == one["final_round"]).all())
        open(os.path.join(data, "result.txt"), "w").write(f"{'OK' if ok else 'MISMATCH'} {n} pairs {world} ranks {len(plan[0])} shards on rank 0\n")
    m.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
