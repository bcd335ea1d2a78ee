"""Debug helper (test infra): hostsim vs the reference on a synthetic config, contig by contig."""
import os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from circminer_b200 import workload
from circminer_b200.api import MREAD_COMPARE_FIELDS, fresh_mreads, default_params
from oracle.refharness import RefHarness, default_params as ref_params
from hostsim_api import HostSim

name = sys.argv[1] if len(sys.argv) > 1 else "T0"
contig_size = int(sys.argv[2]) if len(sys.argv) > 2 else 250_000
use_ref_annot = os.environ.get("REF_ANNOT", "1") == "1"
scan = int(os.environ.get("SCAN", "0"))
tmp = tempfile.mkdtemp(prefix="cm_dbg_")
w = workload.prepare(name, tmp, contig_size=contig_size)
h = RefHarness(ref_params(scan_level=scan))
h.build_index(w["fa"], contig_size=contig_size)
h.open(w["fa"], w["gtf"], os.path.join(tmp, "ref"))
hs = HostSim(default_params(scan_level=scan))
s1, s2 = w["s1"], w["s2"]
n, L = s1.shape
lens = np.full(n, L, np.uint16)
mr_h = fresh_mreads(n); mr_r = fresh_mreads(n)
more = 1
while more:
    more = h.next_contig(); c = h.contig
    tab, mem, memsz, cref, rlen = h.index_pointers()
    hs.load_contig_from_reference(c, tab, mem, w["contigs"][c])
    hs.load_annotation(c, h.annotation() if use_ref_annot else w["flat"][c])
    for nm, s in (("r1", s1), ("r2", s2)):
        want = h.chain(s, lens); got = hs.get_best_chains(c, s, lens)
        for k in ("chain_count", "high_hits", "score", "chain_len", "frag_rpos", "frag_qpos"):
            bad = np.nonzero((got[k] != want[k]).reshape(n, -1).any(1))[0]
            if len(bad):
                print(f"contig {c} {nm} chain {k}: {len(bad)} rows differ, first {bad[:5]}")
    mr_r2, st_r = h.process_pe(s1, s2, lens, lens, mr_r)
    mr_h2, st_h = hs.process_read(c, s1, s2, mr_h)
    bad = np.nonzero(st_r != st_h)[0]
    print(f"contig {c}: state mismatches {len(bad)} / {n}", bad[:10], st_r[bad[:10]], st_h[bad[:10]])
    for f in MREAD_COMPARE_FIELDS:
        b = np.nonzero(mr_r2[f] != mr_h2[f])[0]
        if len(b):
            print(f"   field {f}: {len(b)} differ, first {b[:6]} ref {mr_r2[f][b[:6]]} got {mr_h2[f][b[:6]]}")
    print("   type histogram ref:", np.bincount(mr_r2["type"], minlength=14))
    print("   counters:", hs.ctr.as_dict())
    # carry the REFERENCE state forward on both sides so one mismatch does not cascade
    mr_r = mr_r2; mr_h = mr_r2.copy()
