#!/usr/bin/env python
"""bench.py -- paired-end reads/s of the mapping hot path on B200 (BASELINE.json metric).

A step = one pass of the hot path (seed lookup -> chaining -> extension -> classification, every
packed contig in order) over one batch of synthetic pairs.  `value` times the batch resident in HBM
(CUDA events on the library's stream, max over ranks); `e2e` times cm_gpu_process_batch_pe with host
buffers (H2D + kernels + D2H).  `--impl reference` times the unmodified reference (oracle/_ref) on
the host cores.  One process per GPU; reads are sharded, the index is replicated, no collective on
the data path."""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--pairs", type=int, default=0, help="pairs per GPU per step (0 = the config's)")
    ap.add_argument("--data", default="/tmp/cm_bench")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--upto", type=int, default=3, help="1 seed, 2 +chain, 3 whole path")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def mark(self):
        """Start of the timed region: only samples taken after this count (the sampler itself starts earlier,
        nvidia-smi takes a while to come up)."""
        self.t_mark = time.time()

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                pass
        sm, mx, reasons = [], 0, set()
        t_mark = getattr(self, "t_mark", 0.0)
        rows = [r for t, r in self.rows if t >= t_mark] or [r for _, r in self.rows[-3:]]
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def run_b200(a):
    import torch
    import torch.distributed as dist
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, default_params, fresh_mreads

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # keep stdout to the one JSON line: NCCL prints its version banner on stdout from NCCL_DEBUG=VERSION upwards
        if "CM_NCCL_DEBUG" in os.environ:
            os.environ["NCCL_DEBUG"] = os.environ["CM_NCCL_DEBUG"]
        else:
            os.environ.pop("NCCL_DEBUG", None)
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    t_prep = time.time()
    w = workload.prepare(a.workload, os.path.join(a.data, f"{a.workload}_r{rank}"), n_pairs=a.pairs or None,
                         seed_shift=rank)
    n = len(w["s1"])
    L = w["s1"].shape[1]
    m = Mapper(default_params(), device=local)
    workload.load_into(m, w)
    n_contigs = len(w["contigs"])
    t_prep = time.time() - t_prep

    # pinned host buffers for the e2e leg
    def pinned(arr):
        t = torch.empty(arr.nbytes, dtype=torch.uint8, pin_memory=True)
        v = t.numpy().view(arr.dtype).reshape(arr.shape)
        v[...] = arr
        return t, v
    keep = []
    hb = {}
    for k, arr in (("s1", w["s1"]), ("s2", w["s2"]), ("mr", fresh_mreads(n))):
        t, v = pinned(arr)
        keep.append(t)
        hb[k] = v

    align_acc = np.zeros(2)

    def step_resident():
        ms = 0.0
        for c in range(n_contigs):
            m.run_staged(c, a.upto)
            ms += m.stage_ms()[4]
            if a.upto >= 2:
                align_acc[:] += m.align_ms()
        return ms

    # e2e: two contexts sharing the resident index (cm_gpu_create_shared), one host thread each, alternating
    # batches -- one batch's H2D / D2H overlaps the other's kernels.  Inputs and results are pinned host buffers.
    m2 = Mapper(parent=m)
    outs = []
    for _ in range(2):
        to, vo = pinned(fresh_mreads(n))
        ts, vs = pinned(np.zeros(n, np.int32))
        keep += [to, ts]
        outs.append((vo, vs))

    def step_e2e(mp, ob):
        mr = hb["mr"]
        for c in range(n_contigs):
            if a.upto >= 3:
                mr, _ = mp.process_read(c, hb["s1"], hb["s2"], mr, out=ob[0], state=ob[1])
            else:
                mp.stage(hb["s1"], hb["s2"], mr)
                mp.run_staged(c, a.upto)
        return mr

    def run_e2e(steps):
        def worker(k):
            for s in range(k, steps, 2):
                step_e2e((m, m2)[k], outs[k])
        th = [threading.Thread(target=worker, args=(k,)) for k in range(2)]
        for t in th:
            t.start()
        for t in th:
            t.join()

    m.stage(hb["s1"], hb["s2"], hb["mr"])
    clk = ClockSampler(local)
    clk.start()
    for _ in range(a.warmup):
        step_resident()
    m.counters(reset=True)
    align_acc[:] = 0
    barrier()
    clk.mark()
    t0 = time.time()
    dev_ms = 0.0
    stage_acc = np.zeros(5)
    for _ in range(a.steps):
        dev_ms += step_resident()
        stage_acc += np.array(m.stage_ms())
    barrier()
    wall_ms = (time.time() - t0) * 1e3
    clocks = clk.stop()
    ctr = m.counters()

    # e2e through the C-ABI with host buffers
    run_e2e(2)
    barrier()
    t0 = time.time()
    run_e2e(a.steps)
    barrier()
    e2e_ms = (time.time() - t0) * 1e3

    from circminer_b200 import sharding
    dev_ms, wall_ms, e2e_ms = sharding.max_over_ranks([dev_ms, wall_ms, e2e_ms], world, dist=dist, device="cuda")
    reads_total = 2.0 * n * world * a.steps

    if rank == 0:
        peak, how = measured_peak()
        seed_ms = stage_acc[1] / a.steps
        seed_bytes = 32.0 * ctr["seed_sectors"] / a.steps
        ach = seed_bytes / (seed_ms * 1e-3) / 1e9 if seed_ms > 0 else 0.0
        traffic = None   # dram read + write bytes of k_seed per launch, from the committed ncu capture of this workload
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r1_kernel_traffic.json")))
            if tr["workload"] == a.workload and tr["pairs_per_launch"] == n and a.upto >= 1:
                traffic = tr["kernels"]["k_seed"]["dram_bytes_read"] + tr["kernels"]["k_seed"]["dram_bytes_write"]
        except Exception:
            pass
        line = {
            "metric": "paired-end reads/sec", "value": reads_total / (dev_ms * 1e-3), "unit": "reads/s",
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": dev_ms / a.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": f"{a.workload}: {sum(len(c) for c in w['contigs'])} bp packed reference, "
                                   f"{len(w['genes'])} genes, {n} pairs 2x{L} per GPU, k=20",
                       "stages": {1: "seed", 2: "seed+chain", 3: "seed+chain+extend+classify"}[a.upto],
                       "contigs": n_contigs, "l2": "inputs larger than L2 (index+reads > 1.5 GB per pass)",
                       "sharding": "reads sharded per rank, index replicated, no collective"},
            "wall_ms_per_step": wall_ms / a.steps,
            "stage_ms": {"seed": stage_acc[1] / a.steps, "chain": stage_acc[2] / a.steps,
                         "extend_classify": stage_acc[3] / a.steps},
            "e2e": {"value": reads_total / (e2e_ms * 1e-3), "unit": "reads/s",
                    "h2d_bytes_per_step": int(n_contigs * (hb["s1"].nbytes + hb["s2"].nbytes + hb["mr"].nbytes + 4 * n)),
                    "d2h_bytes_per_step": int(n_contigs * (hb["mr"].nbytes + 4 * n)),
                    "how": "cm_gpu_process_batch_pe with pinned host buffers; two contexts sharing the index, "
                           "one host thread each, alternating steps"},
            "gpu_launches": int(ctr["kernel_launches"]),
            "clocks": clocks,
            "roofline": {"kernel": "k_seed", "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak, "traffic": traffic, "peak_source": how,
                         "algorithmic_bytes_per_launch": seed_bytes},
            "counters": {k: v / a.steps for k, v in ctr.items()},
            "alignment": alignment_line(ctr, align_acc, stage_acc, a.steps),
            "chaining": chaining_line(ctr, align_acc, stage_acc, a.steps),
            "prep_s": t_prep,
        }
        if not a.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(a, os.path.join(a.data, f"{a.workload}_r0"), n, threads=1)
        print(json.dumps(line), flush=True)
    m2.close()
    m.close()
    if world > 1:
        dist.destroy_process_group()


def alignment_line(ctr, align_acc, stage_acc, steps):
    """GCUPS of the banded DPs (X-drop + banded edit distance, src/align.cpp:166-723).  Cells are counted on the
    device by every DP call (cm_counters); nearly all of them run in k_chain_spec / k_chain_walk, whose time is taken
    with CUDA events on the launching stream.  INT32 roofline: one cell costs ~10 integer instructions (three adds,
    compare/select chain, the character compare, band bookkeeping) -> peak cells/s = SMs x 128 lanes x clock / 10."""
    cells = (ctr["xdrop_cells"] + ctr["edit_cells"]) / steps
    ms = float(align_acc.sum()) / steps
    if ms <= 0:
        return None
    gcups = cells / (ms * 1e-3) / 1e9
    int32_peak_gcups = 148 * 128 * 1.965 / 10.0
    return {"gcups": gcups, "cells_per_step": cells, "dp_calls_per_step": (ctr["xdrop_calls"] + ctr["edit_calls"]) / steps,
            "k_chain_spec_ms": float(align_acc[0]) / steps, "k_chain_walk_ms": float(align_acc[1]) / steps,
            "int32_peak_gcups": int32_peak_gcups, "frac_int32": gcups / int32_peak_gcups,
            "note": "band 3 (7 cells per row), thread per DP, band in registers; peak assumes 10 INT32 instr per cell"}


def chaining_line(ctr, align_acc, stage_acc, steps):
    """Links = candidate evaluations reaching src/chain.cpp:173, counted on the device; time = the chain stage minus the
    dense alignment kernels that run in the same stage (sorts and k_chain / k_chain_warp / k_chain_big remain)."""
    ms = (float(stage_acc[2]) - float(align_acc.sum())) / steps
    if ms <= 0:
        return None
    links = ctr["chain_links"] / steps
    return {"mlinks_per_s": links / (ms * 1e-3) / 1e6, "links_per_step": links, "ms": ms, "seed_hits_per_step": ctr["seed_hits"] / steps,
            "note": "FP64 scores (IEEE, no contraction), thread per strand + warp per repeat-rich strand; latency bound, no HBM claim"}


def cpu_baseline(a, datadir, n_pairs, threads, target_s=12.0):
    """The unmodified reference (oracle/_ref) on this box's host cores, bounded sample, in a
    subprocess: checker code never runs inside the product process."""
    cmd = [sys.executable, "-m", "oracle.cpu_baseline", "--workload", a.workload, "--data", datadir,
           "--pairs", str(n_pairs), "--threads", str(threads), "--target-s", str(target_s)]
    try:
        out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=1500)
        for ln in reversed(out.stdout.strip().splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)
        return {"value": None, "error": (out.stderr or out.stdout)[-400:]}
    except Exception as e:  # noqa: BLE001
        return {"value": None, "error": repr(e)[:400]}


def run_reference(a):
    rank, world, _ = dist_env()
    if rank != 0:
        return
    ncores = os.cpu_count() or 1
    n = a.pairs or 0
    res = []
    datadir = os.path.join(a.data, f"{a.workload}_r0")
    for _ in range(a.warmup + a.steps):
        res.append(cpu_baseline(a, datadir, n, threads=ncores, target_s=10.0))
    res = res[a.warmup:]
    ok = [r for r in res if r.get("value")]
    if not ok:
        print(json.dumps({"impl": "reference", "unavailable": (res[-1].get("error") or "no result")[:200]}))
        return
    v = float(np.mean([r["value"] for r in ok]))
    ms = float(np.mean([r["seconds"] for r in ok])) * 1e3
    cb = dict(ok[-1])
    cb["value"] = v
    print(json.dumps({
        "impl": "reference", "metric": "paired-end reads/sec", "value": v, "unit": "reads/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic",
        "config": {"workload": f"{a.workload} (bounded sample per step: {cb.get('sample')})"},
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
