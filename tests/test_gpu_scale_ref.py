"""Parity against the LIVE reference at the size the metric is quoted on.

The reference's own index builder needs about 8 minutes for the 3.09 Gbp genome, so the index file it reads here is
written from the device index by cm_gpu_write_index (tests/test_gpu_index_io.py shows that file to be the reference
builder's); everything that is compared comes out of the unmodified reference reading that file:
  C3  the whole BINARY, three packed contigs of real size, 100 k pairs 2x150: mapping.pam and the per-round remain
      FASTQ headers against cm_gpu_map_rounds_pe (final rows, per-round snapshots, who visits which contig);
  C4  single-end 1x100, k = 22 (int16 checksum truncation), scan-lev 2: process_read return values on the first
      packed contig (1.06 Gbp) through libcm_ref.so;
  C5  2x250, k = 14 compact index, seed-lim 500: process_read (PE) on the first packed contig.
Data lives under CM_SCALE_DATA (default /tmp/cm_bench, shared with bench.py on the same box)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import has_ref
import refbinary
from test_rounds import check_rounds_against_binary

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.environ.get("CM_SCALE_DATA", "/tmp/cm_bench")

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(os.environ.get("CM_SKIP_SCALE") == "1", reason="CM_SKIP_SCALE=1"),
              pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")]


def _first_contig_files(w, tag, mapper, compact):
    """A one-contig reference for the function-level tests: index of packed contig 0 only, its rows of the index.info
    file, the GTF rows of its chromosomes."""
    d = os.path.join(w["shared"], tag)
    os.makedirs(d, exist_ok=True)
    fa = os.path.join(d, "ref.fa")
    rows = [r for r in w["info"] if r[1] == 1]
    with open(fa + ".packed.fa.index.info", "w") as f:
        for nm, cid, s, e in rows:
            f.write(f"{cid}\t{s}\t{e}\t{nm}\n")
    names = {r[0] for r in rows}
    gtf = os.path.join(d, "ann.gtf")
    with open(w["gtf"]) as fi, open(gtf, "w") as fo:
        for ln in fi:
            if ln.startswith("#") or ln.split("\t", 1)[0] in names:
                fo.write(ln)
    mapper.write_index(fa + ".packed.fa.index", 1, compact=compact)
    return fa, gtf


def _ref_states(fa, gtf, params, s1, s2, contigs, out):
    np.save(out + "_s1.npy", s1)
    cmd = [sys.executable, "-m", "oracle.ref_states", "--fa", fa, "--gtf", gtf, "--params", json.dumps(params), "--s1", out + "_s1.npy",
           "--contigs", str(contigs), "--out", out + ".npz"]
    if s2 is not None:
        np.save(out + "_s2.npy", s2)
        cmd += ["--s2", out + "_s2.npy"]
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=3000)
    assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
    return np.load(out + ".npz")


def test_c3_rounds_against_the_reference_binary():
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, fresh_mreads
    import bench
    n = 100_000
    w = workload.prepare("C3", os.path.join(DATA, "C3_t"), n_pairs=n, write_reads=True)
    m = Mapper(workload.params_for("C3"))
    workload.load_into(m, w)
    bench.ensure_ref_index_file(m, w, "C3")
    nc = len(w["contigs"])
    assert nc == 3 and sum(len(c) for c in w["contigs"]) > 3_000_000_000
    res = m.map_rounds(w["s1"], w["s2"], fresh_mreads(n), 0, nc, snapshots=True)
    out = os.path.join(DATA, "C3_t", "refbin")
    refbinary.run_reference(w["fa"], w["gtf"], w["fq1"], w["fq2"], out, threads=os.cpu_count() or 1, timeout=3000)
    survivors = check_rounds_against_binary(res, w, out, nc, 0)
    per_round = [(res["round_state"][r] >= 0).sum() for r in range(nc)]
    assert per_round[0] == n and per_round[1] < n and per_round[2] < per_round[1] and survivors > 0
    # a sub-batch gives the same rows (batch independence of the rounds)
    sub = np.random.default_rng(3).permutation(n)[:5000]
    r2 = m.map_rounds(w["s1"][sub], w["s2"][sub], fresh_mreads(len(sub)), 0, nc)
    assert (r2["out"] == res["out"][sub]).all() and (r2["state"] == res["state"][sub]).all()
    m.close()


def test_c4_single_end_k22_first_contig():
    from circminer_b200 import workload
    from circminer_b200.api import Mapper
    n = 100_000
    w = workload.prepare("C4", os.path.join(DATA, "C4_t"), n_pairs=n)
    p = workload.params_for("C4")
    m = Mapper(p)
    m.build_contig(0, np.ascontiguousarray(w["contigs"][0]))
    m.load_annotation(0, w["flat"][0])
    fa, gtf = _first_contig_files(w, "c4_k22", m, compact=False)
    got = m.process_read_se(0, w["s1"])
    want = _ref_states(fa, gtf, workload.config("C4").params, w["s1"], None, 1, os.path.join(DATA, "C4_t", "ref"))["se_state"][0]
    bad = np.nonzero(got != want)[0]
    assert len(bad) == 0, f"{len(bad)} SE states differ, first {bad[:5]}: {got[bad[:5]]} vs {want[bad[:5]]}"
    assert (got == 0).sum() > n // 10 and len(np.unique(got)) >= 3
    m.close()


def test_c5_k14_compact_first_contig():
    from circminer_b200 import workload
    from circminer_b200.api import MREAD_COMPARE_FIELDS, Mapper, fresh_mreads
    n = 4_000
    w = workload.prepare("C5", os.path.join(DATA, "C5_t"), n_pairs=n)
    p = workload.params_for("C5")
    m = Mapper(p)
    m.build_contig(0, np.ascontiguousarray(w["contigs"][0]))
    m.load_annotation(0, w["flat"][0])
    fa, gtf = _first_contig_files(w, "c5_k14", m, compact=True)
    out, st = m.process_read(0, w["s1"], w["s2"], fresh_mreads(n))
    ref = _ref_states(fa, gtf, workload.config("C5").params, w["s1"], w["s2"], 1, os.path.join(DATA, "C5_t", "ref"))
    bad = np.nonzero(st != ref["state"])[0]
    assert len(bad) == 0, f"{len(bad)} states differ, first {bad[:5]}: {st[bad[:5]]} vs {ref['state'][bad[:5]]}"
    for f in MREAD_COMPARE_FIELDS:
        bad = np.nonzero(out[f] != ref["out"][f])[0]
        assert len(bad) == 0, f"field {f}: {len(bad)} differ, first {bad[:5]}"
    ctr = m.counters()
    assert ctr["seed_hits"] / ctr["seeds"] > 8, "C5 is the long-hit-list stress (k = 14: every entry of a bucket is a hit)"
    m.close()
