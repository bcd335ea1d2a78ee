#!/usr/bin/env python
"""bench.py -- paired-end reads/s of the mapping hot path on B200 (BASELINE.json metric, 2x150 bp, GRCh38-scale).

A step = the reference's round loop for one batch of synthetic pairs with every packed contig resident: contig 1 -> 2 -> 3,
each seed lookup -> chaining -> extension -> classification, with the skip rule and the carried MatchedRead between contigs
(cm_gpu_map_rounds_pe).  Default workload C3: 3.09 Gbp in three packed contigs, 2.5 M pairs per GPU (the per-GPU shard of
the config's 20 M pairs on 8 GPUs).  `value` times the batch resident in HBM (CUDA events on the library's stream, max over
ranks); `e2e` times cm_gpu_map_rounds_pe with pinned host buffers (H2D + kernels + D2H).  `--impl reference` times the
unmodified reference binary (oracle/_ref) on the host cores.  One process per GPU; reads are sharded, the index is
replicated, no collective on the data path."""
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
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--pairs", type=int, default=0, help="pairs per GPU per step (0 = the config's)")
    ap.add_argument("--data", default="/tmp/cm_bench")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--same-sample", action="store_true", help="every rank maps the same read sample (separates data skew from GPU skew)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="strong: --pairs is the whole job, split over the ranks, results gathered on rank 0 inside the e2e clock")
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
        """Start of the timed region: only samples taken after this count."""
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


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            j = json.load(open(p))
            return float(j["hbm_gbs"]), float(j.get("sm_max_mhz", 1965.0)), "measured"
        except Exception:
            pass
    return 6650.0, 1965.0, "fallback"


def cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def workload_text(a, w, n, L, n_contigs):
    from circminer_b200 import workload
    p = workload.params_for(a.workload)
    bp = sum(len(c) for c in w["contigs"])
    what = f"{a.workload}: {bp} bp packed reference in {n_contigs} contig(s), {len(w['genes'])} genes, {n} pairs 2x{L} per GPU, k={p.kmer}"
    if workload.is_scale(a.workload):
        what += " (GRCh38 chromosome lengths; rounds with skip rule and carried MatchedRead on the device)"
    return what


def ensure_ref_index_file(m, w, workload_name):
    """<ref>.packed.fa.index for the CPU baseline / reference arm of a GRCh38-scale config: written from the resident
    device index in the reference's on-disk format (13 s instead of the 8 minutes of `circminer --index`; the file
    is checked against the reference's own builder in tests/test_gpu_index_io.py)."""
    import fcntl
    from circminer_b200 import workload
    if not workload.is_scale(workload_name):
        return None
    path = w["fa"] + ".packed.fa.index"
    with open(path + ".lock", "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        if not os.path.exists(path + ".done"):
            t = time.time()
            m.write_index(path, len(w["contigs"]), compact=bool(workload.config(workload_name).compact_index))
            open(path + ".done", "w").write(f"{time.time() - t:.1f}\n")
        fcntl.flock(lk, fcntl.LOCK_UN)
    return path


def run_b200(a):
    import torch
    import torch.distributed as dist
    from circminer_b200 import sharding, workload
    from circminer_b200.api import Mapper, fresh_mreads

    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU path")
    torch.cuda.set_device(local)
    # stdout carries the one JSON line and nothing else: NCCL prints its version banner (and, under NCCL_DEBUG=INFO, its log)
    # on file descriptor 1 whatever NCCL_DEBUG_FILE says, so fd 1 points at stderr until the line is printed.  NCCL_DEBUG
    # is left as the caller set it.
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    cfg = workload.config(a.workload)
    if getattr(cfg, "single_end", False):
        raise SystemExit(f"bench.py: {a.workload} is single-end (function-level parity only, SURVEY 0.5b); see scripts/config_lines.py")
    per_gpu = a.pairs or cfg.n_pairs
    if a.scaling == "strong":
        per_gpu = (per_gpu + world - 1) // world
    t_prep = time.time()
    w = workload.prepare(a.workload, os.path.join(a.data, f"{a.workload}_r{rank}"), n_pairs=per_gpu,
                         seed_shift=0 if a.same_sample else rank)
    n, L = w["s1"].shape
    m = Mapper(workload.params_for(a.workload), device=local)
    workload.load_into(m, w)
    n_contigs = len(w["contigs"])
    if rank == 0:
        ensure_ref_index_file(m, w, a.workload)
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

    rounds_acc = np.zeros(8)
    pairs_round = np.zeros(n_contigs)

    def step_resident():
        m.run_staged_rounds(0, n_contigs)
        ms, per_round = m.rounds_ms(cap=n_contigs)
        rounds_acc[:] += ms
        pairs_round[:len(per_round)] += per_round
        return ms[5]

    # e2e: two contexts sharing the resident index (cm_gpu_create_shared), one host thread each, alternating
    # batches -- one batch's H2D / D2H overlaps the other's kernels.  Inputs and results are pinned host buffers.
    m2 = Mapper(parent=m)
    outs = []
    for _ in range(2):
        to, vo = pinned(fresh_mreads(n))
        ts, vs = pinned(np.zeros(n, np.int32))
        keep += [to, ts]
        outs.append((vo, vs))

    def run_e2e(steps):
        errs = []

        def worker(k):
            try:
                for _ in range(k, steps, 2):
                    (m, m2)[k].map_rounds(hb["s1"], hb["s2"], hb["mr"], 0, n_contigs, out=outs[k][0], state=outs[k][1])
            except Exception as e:  # noqa: BLE001
                errs.append(e)
        th = [threading.Thread(target=worker, args=(k,)) for k in range(2)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if errs:
            raise errs[0]   # a step that did not run is not a measurement

    m.stage(hb["s1"], hb["s2"], hb["mr"])
    clk = ClockSampler(local) if rank == 0 else None   # one poller per box, not one per rank
    if clk:
        clk.start()
    for _ in range(a.warmup):
        step_resident()
    m.counters(reset=True)
    rounds_acc[:] = 0
    pairs_round[:] = 0
    barrier()
    if clk:
        clk.mark()
    t0 = time.time()
    dev_ms = 0.0
    for _ in range(a.steps):
        dev_ms += step_resident()
    barrier()
    wall_ms = (time.time() - t0) * 1e3
    clocks = clk.stop() if clk else None
    ctr = m.counters()
    launches_step = ctr["kernel_launches"] / a.steps

    # e2e through the C-ABI with host buffers
    run_e2e(2)
    barrier()
    t0 = time.time()
    run_e2e(a.steps)
    if a.scaling == "strong" and world > 1:   # the job's results on one host, input order: inside the e2e clock
        local_out = outs[(a.steps - 1) % 2][0]
        sharding.gather_in_input_order(local_out, n * world, rank, world, shard_pairs=n, dist=dist)
    barrier()
    e2e_ms = (time.time() - t0) * 1e3

    my = [dev_ms, wall_ms, e2e_ms]
    dev_ms, wall_ms, e2e_ms = sharding.max_over_ranks(my, world, dist=dist, device="cuda")
    per_rank = None
    if world > 1:
        t = torch.tensor(my + [rounds_acc[1], rounds_acc[2], rounds_acc[3]], dtype=torch.float64, device="cuda")
        allr = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allr, t)
        per_rank = [{"rank": r, "ms_per_step": float(x[0]) / a.steps, "e2e_ms_per_step": float(x[2]) / a.steps,
                     "seed_ms": float(x[3]) / a.steps, "chain_ms": float(x[4]) / a.steps, "extend_ms": float(x[5]) / a.steps}
                    for r, x in enumerate(allr)]
    reads_total = 2.0 * n * world * a.steps

    if rank == 0:
        peak, sm_max, how = measured_peaks()
        seed_ms = rounds_acc[1] / a.steps
        seed_bytes = 32.0 * ctr["seed_sectors"] / a.steps
        ach = seed_bytes / (seed_ms * 1e-3) / 1e9 if seed_ms > 0 else 0.0
        traffic = None   # dram read + write bytes of k_seed per step, from the committed ncu capture of this workload
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "r2_kernel_traffic.json")))
            if tr["workload"] == a.workload and tr["pairs_per_step"] == n:
                traffic = tr["kernels"]["k_seed"]["dram_bytes_read"] + tr["kernels"]["k_seed"]["dram_bytes_write"]
        except Exception:
            pass
        h2d = hb["s1"].nbytes + hb["s2"].nbytes + hb["mr"].nbytes + 4 * n
        d2h = hb["mr"].nbytes + 8 * n
        line = {
            "metric": "paired-end reads/sec", "value": reads_total / (dev_ms * 1e-3), "unit": "reads/s",
            "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": dev_ms / a.steps,
            "higher_is_better": True, "scaling": a.scaling, "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": workload_text(a, w, n, L, n_contigs),
                       "stages": "seed+chain+extend+classify, every contig in order",
                       "contigs": n_contigs, "pairs_per_round": (pairs_round / a.steps).tolist(),
                       "l2": "inputs larger than L2 (three indexes of 12 GB each; 0.5 GB of reads per pass)",
                       "sharding": "reads sharded per rank, index replicated, no collective on the data path"},
            "wall_ms_per_step": wall_ms / a.steps,
            "stage_ms": {"seed": rounds_acc[1] / a.steps, "chain": rounds_acc[2] / a.steps,
                         "extend_classify": rounds_acc[3] / a.steps,
                         "commit_compact": (rounds_acc[5] - rounds_acc[4]) / a.steps},
            "e2e": {"value": reads_total / (e2e_ms * 1e-3), "unit": "reads/s",
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "how": "cm_gpu_map_rounds_pe with pinned host buffers (reads up once, survivors compacted on the device, "
                           "final MatchedReads down once); two contexts sharing the index, one host thread each, alternating steps"
                           + ("; results of all ranks gathered on rank 0 in input order inside the clock" if a.scaling == "strong" and world > 1 else "")},
            "gpu_launches": int(ctr["kernel_launches"]),
            "gpu_launches_per_step": launches_step,
            "clocks": clocks,
            "roofline": {"kernel": "k_seed", "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak, "traffic": traffic, "peak_source": how,
                         "algorithmic_bytes_per_step": seed_bytes, "launches_per_step": n_contigs,
                         "note": "one 128-byte line per seed is this memory system's floor for a random read (profiles/r1_gather_probe.txt); "
                                 "SURVEY's sector model credits a seed with 64-96 bytes"},
            "counters": {k: v / a.steps for k, v in ctr.items()},
            "chaining": chaining_line(ctr, rounds_acc, a.steps),
            "prep_s": t_prep, "host": {"cpu": cpu_model(), "cores": os.cpu_count()},
        }
        if per_rank:
            line["per_rank"] = per_rank
        try:
            line["alignment"] = alignment_line(m, ctr, rounds_acc, a.steps, hb, n, n_contigs, sm_max)
        except Exception as e:  # noqa: BLE001
            line["alignment"] = {"error": repr(e)[:300]}
        if not a.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(a, w, n, threads=1)
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    m2.close()
    m.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def alignment_line(m, ctr, rounds_acc, steps, hb, n, n_contigs, sm_max_mhz):
    """GCUPS of the banded DPs.  Two numerators:
      cells_executed  every DP cell the device ran in the timed steps (the dense pre-alignment is speculative);
      cells_reference the cells of exactly the DP calls the reference makes: the same pairs run once, outside the timed
                      region, on a context created with CM_GPU_NO_SPEC=1, where every DP is issued inline by the per-pair
                      code at the point the reference issues it (sample of <= 250 k pairs, scaled).
    INT32 fraction = GCUPS x ops_per_cell (6 X-drop, 5 edit) / min(SMs x 128 x f_max, measured add+max lane-ops/s)."""
    from circminer_b200.api import Mapper, fresh_mreads
    align_ms = float(rounds_acc[6] + rounds_acc[7]) / steps
    if align_ms <= 0:
        return None
    xc, ec = ctr["xdrop_cells"] / steps, ctr["edit_cells"] / steps
    k = min(n, 250_000)
    os.environ["CM_GPU_NO_SPEC"] = "1"
    try:
        mx = Mapper(parent=m)
    finally:
        os.environ.pop("CM_GPU_NO_SPEC", None)
    mx.counters(reset=True)
    mx.map_rounds(hb["s1"][:k], hb["s2"][:k], fresh_mreads(k), 0, n_contigs)
    cx = mx.counters()
    mx.close()
    scale = n / k
    rx, re_ = cx["xdrop_cells"] * scale, cx["edit_cells"] * scale
    nominal = 148 * 128 * sm_max_mhz * 1e6
    measured = m.int32_peak()
    peak_ops = min(nominal, measured)
    ops_exec = 6.0 * xc + 5.0 * ec
    ops_ref = 6.0 * rx + 5.0 * re_
    return {"gcups_executed": (xc + ec) / (align_ms * 1e-3) / 1e9, "gcups_useful": (rx + re_) / (align_ms * 1e-3) / 1e9,
            "cells_executed_per_step": xc + ec, "cells_reference_per_step": rx + re_,
            "dp_calls_executed_per_step": (ctr["xdrop_calls"] + ctr["edit_calls"]) / steps,
            "dp_calls_reference_per_step": (cx["xdrop_calls"] + cx["edit_calls"]) * scale,
            "dense_alignment_ms": align_ms, "k_chain_spec_ms": float(rounds_acc[6]) / steps, "k_chain_walk_ms": float(rounds_acc[7]) / steps,
            "int32_peak_lane_ops": peak_ops, "int32_peak_nominal": nominal, "int32_peak_measured": measured,
            "frac_int32_executed": ops_exec / (align_ms * 1e-3) / peak_ops, "frac_int32_useful": ops_ref / (align_ms * 1e-3) / peak_ops,
            "reference_sample_pairs": k,
            "note": "band 3, thread per DP, band in registers; time = the dense alignment kernels (k_chain_spec + k_chain_walk)"}


def chaining_line(ctr, rounds_acc, steps):
    """Links = candidate evaluations reaching src/chain.cpp:173, counted on the device; time = the chain stage minus the
    dense alignment kernels that run in the same stage (sorts and k_chain / k_chain_warp / k_chain_big remain)."""
    ms = (float(rounds_acc[2]) - float(rounds_acc[6] + rounds_acc[7])) / steps
    if ms <= 0:
        return None
    links = ctr["chain_links"] / steps
    return {"mlinks_per_s": links / (ms * 1e-3) / 1e6, "links_per_step": links, "ms": ms, "seed_hits_per_step": ctr["seed_hits"] / steps,
            "note": "FP64 scores (IEEE, no contraction), thread per strand + warp per repeat-rich strand; latency bound, no HBM claim"}


def cpu_baseline(a, w, n_pairs, threads, target_s=12.0, seed_shift=0):
    """The unmodified reference (oracle/_ref) on this box's host cores, bounded sample, in a
    subprocess: checker code never runs inside the product process."""
    cmd = [sys.executable, "-m", "oracle.cpu_baseline", "--workload", a.workload, "--data", os.path.join(a.data, f"{a.workload}_r0"),
           "--pairs", str(n_pairs), "--threads", str(threads), "--target-s", str(target_s)]
    try:
        out = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=2400)
        for ln in reversed(out.stdout.strip().splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)
        return {"value": None, "error": (out.stderr or out.stdout)[-400:]}
    except Exception as e:  # noqa: BLE001
        return {"value": None, "error": repr(e)[:400]}


def run_reference(a):
    """The reference binary, `-t <all cores>`, mapping stage, on a bounded FASTQ sample of the same workload.  A GRCh38-scale
    run costs about a minute (three index loads from the page cache + the rounds), so at most 1 warm-up and 3 timed runs
    are made whatever --steps says; `steps_run` states it.  The rounds' own timer (src/circminer.cpp:302) is the clock:
    index and GTF load are outside it, FASTQ parse and PAM / remain-file writing inside, as in the reference."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    from circminer_b200 import workload
    ncores = os.cpu_count() or 1
    scale = workload.is_scale(a.workload)
    n_warm, n_run = (min(a.warmup, 1), min(a.steps, 3)) if scale else (min(a.warmup, 2), min(a.steps, 5))
    res = []
    w = None
    for _ in range(n_warm + n_run):
        res.append(cpu_baseline(a, w, a.pairs or 0, threads=ncores, target_s=10.0))
    res = res[n_warm:]
    ok = [r for r in res if r.get("value")]
    if not ok:
        print(json.dumps({"impl": "reference", "unavailable": (res[-1].get("error") or "no result")[:200]}))
        return
    v = float(np.mean([r["value"] for r in ok]))
    ms = float(np.mean([r["seconds"] for r in ok])) * 1e3
    cb = dict(ok[-1])
    cb["value"] = v
    cb["cpu"] = cpu_model()
    print(json.dumps({
        "impl": "reference", "metric": "paired-end reads/sec", "value": v, "unit": "reads/s", "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "steps_run": len(ok), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
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
