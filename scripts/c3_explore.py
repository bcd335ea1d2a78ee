"""First look at a GRCh38-scale workload on the GPU: prepare, build the three contigs, run the rounds, print stage
times / pairs per round / counters; optionally write the reference-format index and time it."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--pairs", type=int, nargs="+", default=[1_000_000, 2_500_000])
    ap.add_argument("--data", default="/tmp/cm_bench")
    ap.add_argument("--write-index", action="store_true")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--variants", nargs="*", default=[], help="ENV=VAL[,ENV=VAL] settings, each run on a shared context created under them")
    ap.add_argument("--no-per-contig", action="store_true")
    a = ap.parse_args()
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, fresh_mreads
    log = {}
    t = time.time()
    w = workload.prepare(a.workload, os.path.join(a.data, f"{a.workload}_r0"), n_pairs=max(a.pairs))
    log["prepare_s"] = time.time() - t
    m = Mapper(workload.params_for(a.workload))
    t = time.time()
    build = []
    for c, seq in enumerate(w["contigs"]):
        t1 = time.time()
        m.build_contig(c, np.ascontiguousarray(seq))
        t2 = time.time()
        m.load_annotation(c, w["flat"][c])
        build.append({"contig": c, "len": len(seq), "build_s": t2 - t1, "annot_s": time.time() - t2, "entries": m.contig_stats(c)[0]})
    log["build"] = build
    log["build_s"] = time.time() - t
    print(json.dumps(log), flush=True)
    nc = len(w["contigs"])
    if a.write_index:
        t = time.time()
        path = os.path.join(a.data, "written.index")
        m.write_index(path, nc)
        print(json.dumps({"write_index_s": time.time() - t, "bytes": os.path.getsize(path)}), flush=True)
    for n in a.pairs:
        s1, s2 = w["s1"][:n], w["s2"][:n]
        m.stage(s1, s2, fresh_mreads(n))
        for rep in range(a.reps):
            m.counters(reset=True)
            t = time.time()
            m.run_staged_rounds(0, nc)
            wall = time.time() - t
            ms, per_round = m.rounds_ms()
            ctr = m.counters()
            out, st, fr = m.fetch_rounds()
            print(json.dumps({"pairs": n, "rep": rep, "wall_s": wall, "ms": ms, "pairs_per_round": per_round,
                              "reads_per_s": 2 * n / (ms[5] * 1e-3), "counters": ctr,
                              "types": np.bincount(out["type"], minlength=14).tolist(),
                              "final_round": np.bincount(fr, minlength=nc).tolist()}), flush=True)
        for var in a.variants:
            kv = dict(x.split("=", 1) for x in var.split(",") if x)
            os.environ.update(kv)
            mv = Mapper(parent=m)
            for k in kv:
                os.environ.pop(k, None)
            mv.stage(s1, s2, fresh_mreads(n))
            for rep in range(a.reps):
                mv.run_staged_rounds(0, nc)
                ms, per_round = mv.rounds_ms()
            o2, st2, fr2 = mv.fetch_rounds()
            print(json.dumps({"pairs": n, "variant": var, "ms": ms, "reads_per_s": 2 * n / (ms[5] * 1e-3),
                              "same_results": bool((o2 == out).all() and (st2 == st).all())}), flush=True)
            mv.close()
        # per contig, every pair (what the old bench line timed)
        for c in range(nc if not a.no_per_contig else 0):
            m.run_staged(c, 3)
            print(json.dumps({"pairs": n, "contig": c, "all_pairs_stage_ms": m.stage_ms(), "align_ms": m.align_ms()}), flush=True)
    m.close()


if __name__ == "__main__":
    main()
