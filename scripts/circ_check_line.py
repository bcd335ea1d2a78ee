"""Timed line for cm_gpu_circ_check_batch (ProcessCirc::check_split_map and everything under it): every check_split_map call the
reference makes on the back-splice pairs of T1 (20 000 pairs, oracle/refdump.py --circ-single), replicated to about 10^5 rows for
the device, with the reference's own check_split_map timed on the recorded calls (one core) and equality of all results.

usage (GPU box): python scripts/circ_check_line.py [--out FILE.json]      TEST / BENCH INFRASTRUCTURE (runs oracle/_ref)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=20000)
    ap.add_argument("--rows", type=int, default=100_000)
    ap.add_argument("--data", default="/tmp/cm_circ_check")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    from parity import live_dump
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, default_params
    os.makedirs(a.data, exist_ok=True)
    w = workload.prepare("T1", os.path.join(a.data, "t1"), contig_size=3_500_000, n_pairs=a.pairs)
    dump = live_dump(os.path.join(a.data, "cs.npz"), workload="T1", contig_size=3_500_000, pairs=a.pairs, circ_single=True, no_stage_dumps=True)
    m = Mapper(default_params())
    workload.load_into(m, w)
    calls, ref_s, dev_rows, dev_s, equal = 0, 0.0, 0, 0.0, True
    types = {}
    for c in range(int(dump["n_contigs"])):
        rec = {k[len(f"c{c}_circs_"):]: dump[k] for k in dump.files if k.startswith(f"c{c}_circs_")}
        n = len(rec["ret"])
        if not n:
            continue
        calls += n
        ref_s += float(rec["check_seconds"])
        for t, k in zip(*np.unique(rec["ret"], return_counts=True)):
            types[int(t)] = types.get(int(t), 0) + int(k)
        rep = max(1, a.rows // (2 * n))
        big = {k: np.concatenate([v] * rep) for k, v in rec.items() if k not in ("best_type", "check_seconds")}
        m.circ_check(c, rec)                                   # warm-up
        t0 = time.time()
        got = m.circ_check(c, big)
        dev_s += time.time() - t0
        dev_rows += n * rep
        for nm, k in (("ret", "ret"), ("cr", "cr"), ("r1", "out_r1"), ("r2", "out_r2"), ("pm", "out_pm")):
            equal = equal and bool((got[nm] == big[k]).all())
    m.close()
    out = {"workload": f"T1, {a.pairs} pairs", "check_split_map_calls": calls, "types": types,
           "device": {"rows": dev_rows, "wall_s_with_copies": dev_s, "calls_per_s": dev_rows / dev_s},
           "reference": {"cores": 1, "calls": calls, "seconds": ref_s, "calls_per_s": calls / ref_s}, "equal_to_reference": equal}
    line = json.dumps(out)
    print(line)
    if a.out:
        open(a.out, "w").write(line + "\n")
    if not equal:
        raise SystemExit("device and reference differ")


if __name__ == "__main__":
    main()
