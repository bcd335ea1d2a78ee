"""smoke(): one small batch of the hot path on cuda:0, checked against the unmodified reference
(oracle/_ref) contig by contig.  The oracle is only the checker here."""
from __future__ import annotations

import os
import tempfile

import numpy as np


def run() -> None:
    from . import synth, workload
    from .api import MREAD_COMPARE_FIELDS, Mapper, default_params, fresh_mreads
    from oracle.refharness import RefHarness
    from oracle.refharness import default_params as ref_params

    contig_size = 250_000
    tmp = tempfile.mkdtemp(prefix="cm_smoke_")
    w = workload.prepare("T0", tmp, contig_size=contig_size)
    m = Mapper(default_params(), device=0)
    workload.load_into(m, w)
    h = RefHarness(ref_params())
    h.build_index(w["fa"], contig_size=contig_size)
    h.open(w["fa"], w["gtf"], os.path.join(tmp, "ref"))
    s1, s2 = w["s1"], w["s2"]
    n, L = s1.shape
    lens = np.full(n, L, np.uint16)
    mr_gpu = fresh_mreads(n)
    mr_ref = fresh_mreads(n)
    more = 1
    while more:
        more = h.next_contig()
        c = h.contig
        want = h.seed(s1, lens)
        got = m.split_match_hash(c, s1, lens)
        for k in ("frag_count", "has_frags", "first_info", "last_info", "sum_info"):
            assert (got[k] == want[k]).all(), f"seed parity: {k} differs on contig {c}"
        wantc = h.chain(s1, lens)
        gotc = m.get_best_chains(c, s1, lens)
        for k in ("chain_count", "high_hits", "score", "chain_len", "frag_rpos", "frag_qpos"):
            assert (gotc[k] == wantc[k]).all(), f"chain parity: {k} differs on contig {c}"
        mr_ref, st_ref = h.process_pe(s1, s2, lens, lens, mr_ref)
        mr_gpu, st_gpu = m.process_read(c, s1, s2, mr_gpu)
        assert (st_ref == st_gpu).all(), f"process_read state differs on contig {c}"
        for f in MREAD_COMPARE_FIELDS:
            assert (mr_ref[f] == mr_gpu[f]).all(), f"process_read {f} differs on contig {c}"
    cnt = m.counters()
    assert cnt["kernel_launches"] > 0
    print(f"smoke ok: {n} pairs x {c + 1} contigs, {cnt['kernel_launches']} kernel launches, "
          f"{cnt['seeds']} seeds, {cnt['chain_links']} chain links")
    m.close()
    os._exit(0) if os.environ.get("CM_SMOKE_HARD_EXIT") else None
