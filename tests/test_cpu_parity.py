"""CPU suite (no GPU): the golden fixtures are pinned to the reference itself, the device headers
(g++ build, tests/hostsim) reproduce them bit-exactly, the host-side annotation flattening equals the
reference's tables, and the C-ABI library exports what include/cm_gpu.h declares."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, has_ref
from parity import GOLDEN, check_reads, compare_align, compare_contig, live_dump

T0_CONTIG = 250_000


@pytest.fixture(scope="module")
def w_t0(tmp_path_factory):
    from circminer_b200 import workload
    return workload.prepare("T0", str(tmp_path_factory.mktemp("t0w")), contig_size=T0_CONTIG)


def _hostsim(w, **params):
    from circminer_b200.api import default_params
    from hostsim_api import HostSim
    hs = HostSim(default_params(**params))
    for c, seq in enumerate(w["contigs"]):
        hs.build_contig(c, seq)
        hs.load_annotation(c, w["flat"][c])
    return hs


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("name,kw", [("T0_k20.npz", dict(align=400, se=True)),
                                     ("T0_k22_scan2.npz", dict(kmer=22, scan_level=2))])
def test_golden_is_what_the_reference_produces(tmp_path, name, kw):
    gold = np.load(os.path.join(GOLDEN, name))
    live = live_dump(str(tmp_path / "live.npz"), workload="T0", contig_size=T0_CONTIG, **kw)
    assert sorted(gold.files) == sorted(live.files)
    for k in gold.files:
        assert np.array_equal(gold[k], live[k]), k


def test_annotation_flatten_equals_reference_tables(w_t0):
    gold = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    for c in range(int(gold["n_contigs"])):
        flat = w_t0["flat"][c]
        for k in ("n_iv", "n_seg", "n_tid", "n_trans", "n_genes", "n_chr", "iv_spos", "iv_epos", "iv_max_end", "iv_min_end",
                  "iv_max_next_exon", "iv_seg_off", "seg_start", "seg_end", "seg_next_exon_beg", "seg_gene_id", "seg_tid_off",
                  "tid", "trans_start_ind", "t2s_off", "t2s", "gene_start", "gene_end", "chr_shift", "chr_id"):
            assert np.array_equal(np.asarray(flat[k]), gold[f"c{c}_annot_{k}"]), f"contig {c} {k}"
        nb = int(gold[f"c{c}_annot_n_bits"])
        assert nb == flat["n_bits"]
        for k in ("near_border", "intronic"):
            assert np.array_equal(flat[k], gold[f"c{c}_annot_{k}"]), f"contig {c} {k}"


def test_pack_genome_matches_reference_index_info(w_t0):
    gold = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    assert [str(x) for x in gold["chr_names"]] == [r[0] for r in w_t0["info"]]
    assert int(gold["n_contigs"]) == len(w_t0["contigs"])


def test_hostsim_matches_golden_k20(w_t0):
    from circminer_b200.api import fresh_mreads
    gold = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    check_reads(w_t0, gold)
    hs = _hostsim(w_t0)
    mr = fresh_mreads(len(w_t0["s1"]))
    for c in range(int(gold["n_contigs"])):
        mr = compare_contig(hs, c, w_t0, gold, mr, se=True)
    compare_align(hs, gold, 400)


def test_hostsim_matches_golden_k22_scan2(w_t0):
    """k=22 exercises the int16 checksum truncation (src/match_read.cpp:77); scan-lev 2 never exits early."""
    from circminer_b200.api import fresh_mreads
    gold = np.load(os.path.join(GOLDEN, "T0_k22_scan2.npz"))
    hs = _hostsim(w_t0, kmer=22, scan_level=2)
    mr = fresh_mreads(len(w_t0["s1"]))
    for c in range(int(gold["n_contigs"])):
        mr = compare_contig(hs, c, w_t0, gold, mr)
    # the bug is observable: k=22 loses seeds that k=20 finds
    g20 = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    assert (gold["c0_pe_mread"]["type"] != g20["c0_pe_mread"]["type"]).any()


def test_cabi_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "cm_gpu.h")).read()
    names = set(re.findall(r"\b(cm_(?:gpu|params|mread)_\w+)\s*\(", hdr))
    assert len(names) >= 20
    so = os.path.join(ROOT, "circminer_b200", "csrc", "libcm_gpu.so")
    assert os.path.exists(so), "build first: python __graft_entry__.py"
    lib = C.CDLL(so)
    for n in sorted(names):
        assert hasattr(lib, n), f"{n} declared in include/cm_gpu.h but not exported"


def test_no_cpu_fallback():
    """Without a device the library must refuse to create a context (no host path behind the ABI)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from circminer_b200.api import CmGpuError, Mapper
    with pytest.raises(CmGpuError, match="no CUDA device|no CPU path|cannot initialise"):
        Mapper()


def test_product_package_never_imports_oracle():
    """Only tests/, smoke() and bench.py's baseline legs may touch oracle/."""
    pkg = os.path.join(ROOT, "circminer_b200")
    pat = re.compile(r"^\s*(from\s+oracle|import\s+oracle|#\s*include\s+.*oracle)", re.M)
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")) and f != "smoke.py":
                assert not pat.search(open(os.path.join(dp, f)).read()), (dp, f)


@pytest.mark.parametrize("kmer,big_th", [(20, 5), (22, 5), (14, 5)])
def test_fused_directory_lookup_equals_plain_lookup(t0_packed, kmer, big_th):
    """cm_seed_lookup_x (32-byte directory entry with 12 inline checksums + coarse checksum tables for large buckets) is the
    same map as the plain directory -> bucket bisection of GenomeSeeder::get_exact_locs_hash (src/match_read.cpp:54-110)."""
    from hostsim_api import HostSim
    from circminer_b200.api import default_params
    hs = HostSim(default_params(kmer=kmer))
    for c, seq in enumerate(t0_packed["contigs"][:1]):
        hs.build_contig(c, seq)
        bad = hs.seed_x_check(c, big_th=big_th, n_probe=100000, seed=c + 1)
        assert bad == 0 or (kmer == 14 and bad == 1000000), f"contig {c}: {bad} look-ups differ"


def test_annotation_file_round_trip(t0_packed, tmp_path):
    """cm_annot_file_write / _read (host only): the build-once annotation artefact gives back every array bit for bit;
    a truncated file is refused."""
    from circminer_b200 import api
    for c, a in enumerate(t0_packed["flat"]):
        path = str(tmp_path / f"annot{c}.bin")
        api.write_annotation_file(path, a, max_read_len=300, tag=1234 + c)
        b = api.read_annotation_file(path)
        assert b["max_read_len"] == 300 and b["tag"] == 1234 + c
        for k in ("n_iv", "n_seg", "n_tid", "n_trans", "n_genes", "n_chr", "n_bits"):
            assert int(a[k]) == b[k], k
        for k in api._ANNOT_U4 + tuple(n for n, _ in api._ANNOT_OTHER):
            want = np.asarray(a[k])
            assert want.shape == b[k].shape and (want == b[k]).all(), k
        raw = open(path, "rb").read()
        open(path, "wb").write(raw[:len(raw) // 2])
        with pytest.raises(api.CmGpuError):
            api.read_annotation_file(path)
