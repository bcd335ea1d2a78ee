"""Timed line for cm_gpu_circ_single_batch (ProcessCirc::call_circ_single_split for a batch) on the back-splice pairs of T1, with
the unmodified ProcessCirc::call_circ timed on the same pairs (one core, per-gene tables rebuilt as get_hash_table_smart does)
and equality of the attempts and final results.

usage (GPU box): python scripts/circ_single_line.py [--pairs N] [--out FILE.json]     TEST / BENCH INFRASTRUCTURE (runs oracle/_ref)."""
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
    ap.add_argument("--data", default="/tmp/cm_circ_single")
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
    n, L = w["s1"].shape
    lens = np.full(n, L, np.uint16)
    cand, att, dev_s, ref_s, equal = 0, 0, 0.0, 0.0, True
    for c in range(int(dump["n_contigs"])):
        rec = {k[len(f"c{c}_circs_"):]: dump[k] for k in dump.files if k.startswith(f"c{c}_circs_")}
        is_c = rec["mr_con"]["type"] == 3
        cand += int(is_c.sum())
        ref_s += float(dump[f"c{c}_circreal_seconds"])
        args = (c, w["s1"][is_c], w["s2"][is_c], lens[is_c], lens[is_c], rec["mr_con"][is_c], rec["gene_cnt"][is_c], rec["gene_se"][is_c])
        m.circ_single(*args)                      # warm-up (builds the 8-mer index of the contig)
        t0 = time.time()
        got = m.circ_single(*args)
        dev_s += time.time() - t0
        att += len(got["ret"])
        equal = equal and len(got["ret"]) == len(rec["ret"]) and bool((got["ret"] == rec["ret"]).all() and (got["cr"] == rec["cr"]).all()
                                                                      and (got["best_type"] == rec["best_type"][is_c]).all())
    m.close()
    out = {"workload": f"T1, {a.pairs} pairs", "candidate_pairs": cand, "check_split_map_attempts": att,
           "device": {"wall_s_with_copies": dev_s, "pairs_per_s": cand / dev_s},
           "reference": {"cores": 1, "seconds": ref_s, "pairs_per_s": cand / ref_s, "what": "unmodified ProcessCirc::call_circ per pair"},
           "equal_to_reference": equal}
    line = json.dumps(out)
    print(line)
    if a.out:
        open(a.out, "w").write(line + "\n")
    if not equal:
        raise SystemExit("device and reference differ")


if __name__ == "__main__":
    main()
