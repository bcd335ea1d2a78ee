"""TEST INFRASTRUCTURE ONLY.  FilterRead::process_read of the unmodified reference (oracle/_ref/libcm_ref.so) on reads held
in .npy files, contig by contig, for parity tests at sizes where the index comes from a file that is already on disk.
  python -m oracle.ref_states --fa F --gtf G --params JSON --s1 a.npy [--s2 b.npy] --contigs N --out res.npz
PE: fresh MatchedReads into contig 0, then the reference's skip rule + header round trip between contigs (oracle/rounds.py);
writes out/state/final_round and per-round snapshots.  SE: the return value per read per contig."""
from __future__ import annotations

import argparse
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--fa", required=True)
    ap.add_argument("--gtf", required=True)
    ap.add_argument("--params", default="{}")
    ap.add_argument("--s1", required=True)
    ap.add_argument("--s2")
    ap.add_argument("--contigs", type=int, default=1)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--out", required=True)
    a = ap.parse_args()
    from oracle import rounds
    from oracle.refharness import MREAD_DTYPE, RefHarness, default_params, fresh_mreads
    p = default_params(**json.loads(a.params))
    h = RefHarness(p, threads=a.threads)   # threads: only the compact-index rebuild at load time uses them
    h.open(a.fa, a.gtf, a.out + ".tmp")
    s1 = np.load(a.s1)
    n, L = s1.shape
    lens = np.full(n, L, np.uint16)
    res = {}
    if a.s2 is None:
        st = np.zeros((a.contigs, n), np.int32)
        for c in range(a.contigs):
            h.next_contig()
            st[c] = h.process_se(s1, lens)
        res["se_state"] = st
    else:
        s2 = np.load(a.s2)
        out = np.zeros(n, MREAD_DTYPE)
        state = np.zeros(n, np.int32)
        fr = np.zeros(n, np.int32)
        r_out = np.zeros((a.contigs, n), MREAD_DTYPE)
        r_state = np.full((a.contigs, n), -1, np.int32)
        ids = np.arange(n)
        mr = fresh_mreads(n, p.max_ed)
        for c in range(a.contigs):
            h.next_contig()
            o, st = h.process_pe(s1[ids], s2[ids], lens[ids], lens[ids], mr)
            r_out[c, ids] = o
            r_state[c, ids] = st
            fin = rounds.skip_rule(p.scan_level, st, o, 2 * L) | (c == a.contigs - 1)
            out[ids[fin]] = o[fin]
            state[ids[fin]] = st[fin]
            fr[ids[fin]] = c
            ids, mr = ids[~fin], rounds.header_round_trip(o[~fin], p.max_ed)
            if not len(ids):
                break
        res.update(out=out, state=state, final_round=fr, round_out=r_out, round_state=r_state)
    np.savez(a.out, **res)
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
