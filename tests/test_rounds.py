"""Round semantics (skip rule + MatchedRead carried through the remain-FASTQ header) against the whole reference
BINARY: its mapping.pam and the per-round remain FASTQs of a `-t 1 -P -z -q 0` run are parsed back and compared
with cm_gpu_map_rounds_pe (GPU) and with the same CM_HD round functions compiled for the host (hostsim, CPU)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import has_ref
import refbinary

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_ref_index(fa, contig_size, kmer=20, compact=False):
    """The reference's own index builder with a chosen CONTIG_SIZE (its CLI always packs at 1.1 Gbp); subprocess:
    the reference keeps process-global state."""
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from oracle.refharness import RefHarness, default_params\n"
            "RefHarness(default_params(kmer=%d)).build_index(%r, contig_size=%d, compact=%r)" % (ROOT, kmer, fa, contig_size, compact))
    subprocess.run([sys.executable, "-c", code], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=1800)


def check_rounds_against_binary(res, w, out_prefix, n_contigs, scan):
    """res: map_rounds(..., snapshots=True) of the engine under test."""
    from circminer_b200.api import fresh_mreads
    n = len(w["s1"])
    names = w["chr_names"]
    seen, typ, pam = refbinary.parse_pam(out_prefix + ".mapping.pam", n, names)
    assert seen.all(), "every pair is printed exactly once"
    refbinary.assert_rows_equal(res["out"], pam, np.arange(n), f"scan {scan}: mapping.pam")
    alive = np.ones(n, bool)
    for r in range(n_contigs):
        visited = res["round_state"][r] >= 0
        assert (visited == alive).all(), f"round {r + 1}: set of pairs that visit the contig differs"
        path = f"{out_prefix}_{r + 1}_remain_R1.fastq"
        present, hdr = refbinary.parse_remain(path, n, names)
        last = r == n_contigs - 1
        if not last:
            # survivors of this round = pairs in the remain file; their header = this round's MatchedRead
            gone = alive & ~present
            assert (res["final_round"][gone] == r).all(), f"round {r + 1}: pairs printed in another round"
            assert (res["final_round"][present] > r).all()
            refbinary.assert_rows_equal(res["round_out"][r], hdr, np.nonzero(present)[0], f"scan {scan}: remain headers of round {r + 1}",
                                        with_contig=True)
            alive = present
        else:
            # last round: only back-splice pairs are written (src/circminer.cpp:395-397)
            bsj = alive & np.isin(res["out"]["type"], (3, 4))
            assert (present == bsj).all(), "last-round remain file holds exactly the CHIBSJ / CHI2BSJ pairs"
            refbinary.assert_rows_equal(res["out"], hdr, np.nonzero(present)[0], f"scan {scan}: last-round remain headers", with_contig=True)
    return int(alive.sum())


@pytest.fixture(scope="module")
def t1_rounds(tmp_path_factory):
    if not has_ref():
        pytest.skip("oracle/_ref not built")
    from circminer_b200 import workload
    d = str(tmp_path_factory.mktemp("t1r") / "t1")
    w = workload.prepare("T1", d, contig_size=3_500_000, write_reads=True, n_pairs=6000)
    build_ref_index(w["fa"], 3_500_000)
    outs = {}
    for scan in (0, 1, 2):
        out = os.path.join(d, f"ref_a{scan}")
        refbinary.run_reference(w["fa"], w["gtf"], w["fq1"], w["fq2"], out, extra=("-a", str(scan)))
        outs[scan] = out
    return w, outs


@pytest.mark.parametrize("scan", [0, 1, 2])
def test_rounds_hostsim_vs_reference_binary(t1_rounds, scan):
    """CPU: the shared round functions (cm_rounds.h) + the per-pair host build against the reference binary's files."""
    from hostsim_api import HostSim
    from circminer_b200.api import default_params, fresh_mreads
    w, outs = t1_rounds
    hs = HostSim(default_params(scan_level=scan))
    for c, seq in enumerate(w["contigs"]):
        hs.build_contig(c, seq)
        hs.load_annotation(c, w["flat"][c])
    n = len(w["s1"])
    res = hs.map_rounds(w["s1"], w["s2"], fresh_mreads(n), 0, len(w["contigs"]), snapshots=True)
    survivors = check_rounds_against_binary(res, w, outs[scan], len(w["contigs"]), scan)
    if scan == 0:
        assert (res["final_round"] == 0).sum() > n // 4 and survivors > 0   # both paths are exercised
    if scan == 2:
        assert (res["final_round"] == len(w["contigs"]) - 1).all()


@pytest.mark.gpu
@pytest.mark.parametrize("scan,chunk", [(0, None), (1, None), (2, None), (0, "2048")])
def test_rounds_gpu_vs_reference_binary(t1_rounds, scan, chunk, monkeypatch):
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, default_params, fresh_mreads
    w, outs = t1_rounds
    if chunk:   # several pipeline passes per round (the arenas are sized by the chunk)
        monkeypatch.setenv("CM_GPU_CHUNK", chunk)
    m = Mapper(default_params(scan_level=scan))
    workload.load_into(m, w)
    n = len(w["s1"])
    res = m.map_rounds(w["s1"], w["s2"], fresh_mreads(n), 0, len(w["contigs"]), snapshots=True)
    check_rounds_against_binary(res, w, outs[scan], len(w["contigs"]), scan)
    # the staged variant gives the same per-pair results and leaves the staged batch intact (run twice)
    m.stage(w["s1"], w["s2"], fresh_mreads(n))
    for _ in range(2):
        m.run_staged_rounds(0, len(w["contigs"]))
        out, st, fr = m.fetch_rounds()
        assert (out == res["out"]).all() and (st == res["state"]).all() and (fr == res["final_round"]).all()
    ms, per_round = m.rounds_ms()
    assert per_round[0] == n and len(per_round) >= 1 and ms[5] > 0
    m.close()


# ---- C1 at its stated shape: one 2 Mb chromosome, 40 genes, 10 000 pairs 2x100, -k 20 -t 1 -P -z --------------
@pytest.fixture(scope="module")
def c1_binary(tmp_path_factory):
    if not has_ref():
        pytest.skip("oracle/_ref not built")
    from circminer_b200 import workload
    d = str(tmp_path_factory.mktemp("c1r") / "c1")
    w = workload.prepare("C1", d, write_reads=True)
    assert len(w["contigs"]) == 1 and len(w["contigs"][0]) == 2_000_000 and w["s1"].shape == (10_000, 100)
    build_ref_index(w["fa"], 1_100_000_000)
    out = os.path.join(d, "ref")
    refbinary.run_reference(w["fa"], w["gtf"], w["fq1"], w["fq2"], out)
    return w, out


def test_c1_hostsim_vs_reference_binary(c1_binary):
    from hostsim_api import HostSim
    from circminer_b200.api import default_params, fresh_mreads
    w, out = c1_binary
    hs = HostSim(default_params())
    hs.build_contig(0, w["contigs"][0])
    hs.load_annotation(0, w["flat"][0])
    n = len(w["s1"])
    res = hs.map_rounds(w["s1"], w["s2"], fresh_mreads(n), 0, 1, snapshots=True)
    check_rounds_against_binary(res, w, out, 1, 0)
    types = set(np.unique(res["out"]["type"]).tolist())
    assert {0, 3} <= types, types   # concordant and back-splice pairs


@pytest.mark.gpu
def test_c1_gpu_vs_reference_binary(c1_binary):
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, default_params, fresh_mreads
    w, out = c1_binary
    m = Mapper(default_params())
    workload.load_into(m, w)
    n = len(w["s1"])
    res = m.map_rounds(w["s1"], w["s2"], fresh_mreads(n), 0, 1, snapshots=True)
    check_rounds_against_binary(res, w, out, 1, 0)
    m.close()
