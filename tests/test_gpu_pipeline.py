"""GPU parity through the C-ABI (libcm_gpu.so): chaining, alignment kernels, process_read PE/SE against
the committed golden fixtures (T0) and against the reference itself run live on the box (T1)."""
import os

import numpy as np
import pytest

from conftest import has_ref
from parity import GOLDEN, check_reads, compare_align, compare_contig, live_dump

pytestmark = pytest.mark.gpu
T0_CONTIG = 250_000


@pytest.fixture(scope="module")
def w_t0(tmp_path_factory):
    from circminer_b200 import workload
    return workload.prepare("T0", str(tmp_path_factory.mktemp("t0w")), contig_size=T0_CONTIG)


def _mapper(w, **params):
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, default_params
    m = Mapper(default_params(**params))
    workload.load_into(m, w)
    return m


def test_t0_golden_k20(w_t0):
    from circminer_b200.api import fresh_mreads
    gold = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    check_reads(w_t0, gold)
    m = _mapper(w_t0)
    mr = fresh_mreads(len(w_t0["s1"]))
    for c in range(int(gold["n_contigs"])):
        mr = compare_contig(m, c, w_t0, gold, mr, se=True)
    compare_align(m, gold, 400)
    cnt = m.counters()
    assert cnt["kernel_launches"] > 0 and cnt["xdrop_cells"] > 0 and cnt["chain_links"] > 0
    m.close()


def test_t0_golden_k22_scan2(w_t0):
    from circminer_b200.api import fresh_mreads
    gold = np.load(os.path.join(GOLDEN, "T0_k22_scan2.npz"))
    m = _mapper(w_t0, kmer=22, scan_level=2)
    mr = fresh_mreads(len(w_t0["s1"]))
    for c in range(int(gold["n_contigs"])):
        mr = compare_contig(m, c, w_t0, gold, mr)
    m.close()


def test_chain_overflow_path_is_exact(w_t0, monkeypatch):
    """Force every multi-hit strand through k_chain_big (record arena slice of 8) and compare again."""
    from circminer_b200.api import fresh_mreads
    monkeypatch.setenv("CM_GPU_REC_MULT", "0")
    gold = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    m = _mapper(w_t0)
    mr = fresh_mreads(len(w_t0["s1"]))
    for c in range(int(gold["n_contigs"])):
        mr = compare_contig(m, c, w_t0, gold, mr, stages=("chain", "pe"))
    m.close()


@pytest.mark.parametrize("heavy_th", ["0", "3", "254"])
def test_warp_chaining_is_exact(w_t0, monkeypatch, heavy_th):
    """CM_GPU_HEAVY_TH: 0 sends every strand with a hit through the warp-per-strand kernel (k_chain_warp), 254
    nearly none; chains and the per-pair results must not depend on which kernel chained a strand."""
    from circminer_b200.api import fresh_mreads
    monkeypatch.setenv("CM_GPU_HEAVY_TH", heavy_th)
    gold = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    m = _mapper(w_t0)
    mr = fresh_mreads(len(w_t0["s1"]))
    for c in range(int(gold["n_contigs"])):
        mr = compare_contig(m, c, w_t0, gold, mr, stages=("chain", "pe"))
    m.close()


@pytest.mark.parametrize("th", ["0", "2", "4294967295"])
def test_warp_pair_kernel_is_exact(w_t0, monkeypatch, th):
    """CM_GPU_HEAVY_PAIR_TH=0 sends every pair with chains through the warp-per-pair kernel (speculative evaluation of a
    row's MatePairs with in-order commit, unpaired chains one per lane), 4294967295 sends every pair -- the heavy ones
    too -- through the thread-per-pair kernel; results must equal the goldens, scan-lev 0 and 2."""
    from circminer_b200.api import fresh_mreads
    monkeypatch.setenv("CM_GPU_HEAVY_PAIR_TH", th)
    for name, kw in (("T0_k20.npz", {}), ("T0_k22_scan2.npz", dict(kmer=22, scan_level=2))):
        gold = np.load(os.path.join(GOLDEN, name))
        m = _mapper(w_t0, **kw)
        mr = fresh_mreads(len(w_t0["s1"]))
        for c in range(int(gold["n_contigs"])):
            mr = compare_contig(m, c, w_t0, gold, mr, stages=("pe",))
        m.close()


def test_warp_chaining_overflow_path(w_t0, monkeypatch):
    """k_chain_warp with a record slice of 8: overflowing strands are re-run by k_chain_big."""
    from circminer_b200.api import fresh_mreads
    monkeypatch.setenv("CM_GPU_HEAVY_TH", "0")
    monkeypatch.setenv("CM_GPU_REC_MULT", "0")
    gold = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    m = _mapper(w_t0)
    mr = fresh_mreads(len(w_t0["s1"]))
    for c in range(int(gold["n_contigs"])):
        mr = compare_contig(m, c, w_t0, gold, mr, stages=("chain", "pe"))
    m.close()


def test_inline_dp_path_is_exact(w_t0, monkeypatch):
    """CM_GPU_NO_SPEC=1: no pre-alignment memo, every DP runs inline in the per-pair kernel."""
    from circminer_b200.api import fresh_mreads
    monkeypatch.setenv("CM_GPU_NO_SPEC", "1")
    gold = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    m = _mapper(w_t0)
    mr = fresh_mreads(len(w_t0["s1"]))
    for c in range(int(gold["n_contigs"])):
        mr = compare_contig(m, c, w_t0, gold, mr, stages=("pe",), se=True)
    m.close()


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_t1_live_reference(tmp_path):
    """20 000 pairs 2x150 over two packed contigs, the reference run on this box as the checker."""
    from circminer_b200 import workload
    from circminer_b200.api import fresh_mreads
    dump = live_dump(str(tmp_path / "t1.npz"), workload="T1", contig_size=3_500_000, no_stage_dumps=True)
    w = workload.prepare("T1", str(tmp_path / "t1"), contig_size=3_500_000)
    check_reads(w, dump)
    m = _mapper(w)
    mr = fresh_mreads(len(w["s1"]))
    for c in range(int(dump["n_contigs"])):
        mr = compare_contig(m, c, w, dump, mr, stages=("pe",))
    assert m.counters()["pairs"] == len(w["s1"]) * int(dump["n_contigs"])
    types = np.bincount(mr["type"], minlength=14)
    assert types[0] > 10000 and types[3] > 100      # concordant and back-splice classes both populated
    m.close()


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_t2_k14_compact_2x250_live_reference(tmp_path):
    """C5's shape in small: 2x250 reads, k = 14 (checksum length 0: every bucket entry matches), the reference's
    compact (-m) index, 17 seeds per strand; seeds, chains and process_read over two packed contigs."""
    from circminer_b200 import workload
    from circminer_b200.api import fresh_mreads
    dump = live_dump(str(tmp_path / "t2.npz"), workload="T2", contig_size=1_800_000, kmer=14, compact=True)
    assert int(dump["kmer"]) == 14 and int(dump["n_contigs"]) == 2
    w = workload.prepare("T2", str(tmp_path / "t2"), contig_size=1_800_000)
    check_reads(w, dump)
    m = _mapper(w, kmer=14)
    mr = fresh_mreads(len(w["s1"]))
    for c in range(2):
        mr = compare_contig(m, c, w, dump, mr)
    m.close()


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_t1_single_end_k22_scan2_live_reference(tmp_path):
    """C4's shape in small: single-end, k = 22 (the int16_t checksum truncation of match_read.cpp:77 is part of the
    contract), scan-lev 2, max-ed 4, band 3.  The reference's SE mode writes no files, so the pinned result is the
    return value of FilterRead::process_read (SE overload) per read per contig."""
    from circminer_b200 import workload
    dump = live_dump(str(tmp_path / "t1se.npz"), workload="T1", contig_size=3_500_000, kmer=22, scan_level=2, se=True,
                     pairs=6000, no_stage_dumps=True)
    w = workload.prepare("T1", str(tmp_path / "t1"), contig_size=3_500_000, n_pairs=6000)
    check_reads(w, dump)
    m = _mapper(w, kmer=22, scan_level=2)
    for c in range(int(dump["n_contigs"])):
        st = m.process_read_se(c, w["s1"])
        want = dump[f"c{c}_se_state"]
        bad = np.nonzero(st != want)[0]
        assert len(bad) == 0, f"contig {c}: {len(bad)} SE states differ, first {bad[:5]}"
    assert len(np.unique(dump["c0_se_state"])) >= 3
    m.close()


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_c2_sample_live_reference(tmp_path):
    """The bench workload itself (C2: 100 Mb, 15 000 + 500 repeat families, N gaps), first 100 000 pairs, against the
    reference run on this box: the repeat-rich strands go through k_chain_warp, the many-candidate pairs through
    k_process_pe_warp (batch <= 300 k), everything else through the thread-per-unit kernels."""
    from circminer_b200 import workload
    from circminer_b200.api import fresh_mreads
    n = 100_000
    dump = live_dump(str(tmp_path / "c2.npz"), workload="C2", contig_size=1_100_000_000, pairs=n, no_stage_dumps=True)
    w = workload.prepare("C2", str(tmp_path / "c2"), n_pairs=n)
    check_reads(w, dump)
    m = _mapper(w)
    mr = fresh_mreads(n)
    for c in range(int(dump["n_contigs"])):
        mr = compare_contig(m, c, w, dump, mr, stages=("pe",))
    cnt = m.counters()
    assert cnt["pairs"] == n * int(dump["n_contigs"])
    types = np.bincount(mr["type"], minlength=14)
    assert types[0] > 0.5 * n and types[3] > 1000
    m.close()


def test_shared_context_gives_the_same_results(w_t0):
    """cm_gpu_create_shared: a second context over the parent's resident contigs, used from a second host thread."""
    import threading
    from circminer_b200.api import CmGpuError, Mapper, fresh_mreads
    with np.load(os.path.join(GOLDEN, "T0_k20.npz")) as z:
        gold = {k: z[k] for k in z.files}   # materialised: a lazily read NpzFile (zipfile) is not safe from two threads
    m = _mapper(w_t0)
    m2 = Mapper(parent=m)
    with pytest.raises(CmGpuError):
        m2.build_contig(0, w_t0["contigs"][0])
    res = {}

    def work(name, mp):
        mr = fresh_mreads(len(w_t0["s1"]))
        for c in range(int(gold["n_contigs"])):
            mr = compare_contig(mp, c, w_t0, gold, mr, stages=("pe",))
        res[name] = mr

    th = [threading.Thread(target=work, args=(k, mp)) for k, mp in (("a", m), ("b", m2))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert "a" in res and "b" in res and (res["a"] == res["b"]).all()
    m2.close()
    m.close()


def test_ragged_and_degenerate_reads(w_t0):
    """Ragged lengths, reads shorter than k, N / lower-case bases: states must equal the all-contig golden
    run restricted to untouched rows, and nothing may fault."""
    from circminer_b200.api import fresh_mreads
    m = _mapper(w_t0)
    s1, s2 = w_t0["s1"][:64].copy(), w_t0["s2"][:64].copy()
    l1 = np.full(64, s1.shape[1], np.uint16)
    l2 = l1.copy()
    l1[0] = 19
    l2[1] = 1
    l1[2] = 57
    s1[3, :] = ord("N")
    s2[4, 10:30] = np.frombuffer(bytes(s2[4, 10:30]).lower(), np.uint8)
    out, st = m.process_read(0, s1, s2, fresh_mreads(64), l1, l2)
    gold = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    keep = np.ones(64, bool)
    keep[:5] = False
    assert (st[keep] == gold["c0_pe_state"][:64][keep]).all()
    assert st[3] in (10, 13)          # one mate all-N: OEANCH or NOPROC_NOMATCH
    # empty batch
    out, st = m.process_read(0, s1[:0], s2[:0], fresh_mreads(0))
    assert len(st) == 0
    m.close()


def test_host_buffers_warmup_and_alignment_timers(w_t0):
    """cm_gpu_host_alloc / cm_gpu_host_free (pinned batch buffers of the drop-in's host feed), cm_gpu_warmup and
    cm_gpu_last_align_ms: results from pinned caller buffers equal the goldens; the alignment kernels report a time."""
    import ctypes as C
    from circminer_b200.api import fresh_mreads, load_library, MREAD_DTYPE
    lib = load_library()
    assert lib.cm_gpu_warmup(0) == 0
    gold = np.load(os.path.join(GOLDEN, "T0_k20.npz"))
    m = _mapper(w_t0)
    n = len(w_t0["s1"])
    bufs = []

    def pinned_like(arr):
        p = C.c_void_p()
        assert lib.cm_gpu_host_alloc(C.c_size_t(arr.nbytes), C.byref(p)) == 0 and p.value
        bufs.append(p)
        v = np.frombuffer((C.c_char * arr.nbytes).from_address(p.value), dtype=arr.dtype).reshape(arr.shape)
        v[...] = arr
        return v

    s1, s2 = pinned_like(w_t0["s1"]), pinned_like(w_t0["s2"])
    mr = pinned_like(fresh_mreads(n))
    out = pinned_like(np.zeros(n, MREAD_DTYPE))
    state = pinned_like(np.zeros(n, np.int32))
    want = fresh_mreads(n)
    for c in range(int(gold["n_contigs"])):
        want = compare_contig(m, c, w_t0, gold, want, stages=("pe",))
        got, st = m.process_read(c, s1, s2, mr, out=out, state=state)
        mr[...] = got
        a = m.align_ms()
        assert len(a) == 2 and a[0] >= 0 and a[1] >= 0
    assert (mr == want).all()
    m.close()
    for p in bufs:
        assert lib.cm_gpu_host_free(p) == 0
