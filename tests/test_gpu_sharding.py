"""Read sharding on hardware: two ranks deal the input in contiguous shards, map their pairs on the GPU and
rank 0 gathers the rows in input order; the result must be the single-context run's, pair for pair."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu


def test_two_ranks_reproduce_the_single_gpu_rows(tmp_path):
    import torch
    nccl = torch.cuda.device_count() >= 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29731", os.path.join(ROOT, "tests", "shard_worker.py"), str(tmp_path), "1000"] + (["--nccl"] if nccl else [])
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    assert open(tmp_path / "result.txt").read().startswith("OK 12000 pairs 2 ranks 6 shards")
