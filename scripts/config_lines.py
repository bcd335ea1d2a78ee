"""One timed line per BASELINE config that is not the bench's: C4 (single-end 1x100, k = 22, scan-lev 2) and C5 (2x250, k = 14
compact index, seed-lim 500), on the GRCh38-scale reference with every packed contig resident.  Device time from the
library's own events; prints one JSON object per config (kept under profiles/).
  C4: FilterRead::process_read (SE) has no file output in the reference: every read is mapped against every
      contig, the line is reads/s over the three contigs.
  C5: the rounds call, as bench.py does for C3."""
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
    ap.add_argument("--configs", nargs="+", default=["C4", "C5"])
    ap.add_argument("--data", default="/tmp/cm_bench")
    ap.add_argument("--pairs", type=int, default=0)
    ap.add_argument("--steps", type=int, default=3)
    a = ap.parse_args()
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, fresh_mreads
    for name in a.configs:
        cfg = workload.config(name)
        n = a.pairs or cfg.n_pairs
        w = workload.prepare(name, os.path.join(a.data, f"{name}_r0"), n_pairs=n)
        p = workload.params_for(name)
        m = Mapper(p)
        t = time.time()
        workload.load_into(m, w)
        build_s = time.time() - t
        nc = len(w["contigs"])
        L = w["s1"].shape[1]
        line = {"config": name, "kmer": p.kmer, "scan_level": p.scan_level, "seed_lim": p.seed_lim, "contigs": nc,
                "reference_bp": int(sum(len(c) for c in w["contigs"])), "index_build_s": build_s,
                "index_entries": [m.contig_stats(c)[0] for c in range(nc)]}
        if cfg.single_end:
            CH = 1 << 20   # reads per call (arenas sized by the call)
            ms = np.zeros(5)
            states = []
            for step in range(a.steps + 1):
                ms[:] = 0
                wall = time.time()
                for c in range(nc):
                    for lo in range(0, n, CH):
                        st = m.process_read_se(c, w["s1"][lo:lo + CH])
                        ms += np.array(m.stage_ms())
                        if step == a.steps and c == 0:
                            states.append(st)
                wall = time.time() - wall
            st = np.concatenate(states)
            line.update(reads=n, read_len=L, device_ms={"seed": ms[1], "chain": ms[2], "extend_classify": ms[3], "total": ms[4]},
                        reads_per_s_device=n / (ms[4] * 1e-3), reads_per_s_wall_with_copies=n / wall,
                        states_contig0=np.bincount(st, minlength=14).tolist(),
                        note="every read against all three contigs; device time = events around seed..classify, wall = C-ABI call with host buffers")
        else:
            m.stage(w["s1"], w["s2"], fresh_mreads(n))
            for _ in range(a.steps + 1):
                m.counters(reset=True)
                m.run_staged_rounds(0, nc)
                ms, per_round = m.rounds_ms(cap=nc)
            ctr = m.counters()
            out, st, fr = m.fetch_rounds()
            line.update(pairs=n, read_len=L, pairs_per_round=per_round,
                        device_ms={"seed": ms[1], "chain": ms[2], "extend_classify": ms[3], "whole_call": ms[5], "k_chain_spec": ms[6], "k_chain_walk": ms[7]},
                        reads_per_s_device=2 * n / (ms[5] * 1e-3), hits_per_seed=ctr["seed_hits"] / max(1, ctr["seeds"]),
                        links_per_pair=ctr["chain_links"] / n, types=np.bincount(out["type"], minlength=14).tolist(),
                        seed_gbs_algorithmic=32.0 * ctr["seed_sectors"] / (ms[1] * 1e-3) / 1e9)
        print(json.dumps(line), flush=True)
        m.close()


if __name__ == "__main__":
    main()
