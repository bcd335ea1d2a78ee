"""Stage 2, first piece: ProcessCirc::chaining -- RegionalHashTable look-ups inside one gene, chain_seeds_sorted_kbest2
and the missing-k-mer cut -- against the reference run live (oracle/refdump.py --circ: the reference's own ProcessCirc::chaining on
its own RegionalHashTable).  CPU: the device functions of csrc/cm_circ.h in the host build; GPU: cm_gpu_circ_chain_batch."""
import numpy as np
import pytest

from conftest import has_ref
from parity import live_dump

KEYS = ("chain_count", "score", "chain_len", "frag_rpos", "frag_qpos")
REF_F = (300 - 8) // 3 + 1   # fragment slots of the reference's chain lists: (maxReadLength - window_size) / step + 1
XKEYS = ("seqs", "r1_spos", "r1_epos", "r2_spos", "r2_epos", "dir", "qspos", "rlen", "whole_len", "chain_len", "frag_rpos", "frag_qpos")
MATE_FIELDS = ("spos", "epos", "qspos", "qepos", "matched_len", "left_ed", "middle_ed", "right_ed", "sclen_left", "sclen_right", "dir", "type",
               "is_concord", "left_ok", "right_ok")


_T1_DUMPS = {}


def _t1_single_split_dump(pairs):
    """T1 and the reference's record of its stage 2 (oracle/refdump.py --circ-single), made once per test session and pair count:
    several tests read the same dump."""
    if pairs not in _T1_DUMPS:
        import atexit
        import os
        import shutil
        import tempfile
        from circminer_b200 import workload
        d = tempfile.mkdtemp(prefix="cm_t1_circs_")
        atexit.register(shutil.rmtree, d, ignore_errors=True)
        w = workload.prepare("T1", os.path.join(d, "t1"), contig_size=3_500_000, n_pairs=pairs)
        dump = live_dump(os.path.join(d, "cs.npz"), workload="T1", contig_size=3_500_000, pairs=pairs, circ_single=True, no_stage_dumps=True)
        _T1_DUMPS[pairs] = (w, dump)
    return _T1_DUMPS[pairs]


def _compare_exact(engine, dump, c, min_concord):
    """find_exact_coord on every chain ProcessCirc would try: the return value and every MatchedMate field of partial_mm."""
    prob = {k: dump[f"c{c}_circx_p_{k}"] for k in XKEYS}
    got = engine.circ_exact(c, prob)
    want_ret, want = dump[f"c{c}_circx_r_ret"], dump[f"c{c}_circx_r_pmm"]
    bad = np.nonzero(got["ret"] != want_ret)[0]
    assert len(bad) == 0, f"contig {c}: find_exact_coord return differs on {len(bad)} of {len(want_ret)} problems, first {bad[:5]}"
    for f in MATE_FIELDS:
        bad = np.nonzero(got["pmm"][f] != want[f])[0]
        assert len(bad) == 0, f"contig {c} partial_mm.{f}: {len(bad)} of {len(want)} differ, first {bad[:5]}: {got['pmm'][f][bad[:5]]} vs {want[f][bad[:5]]}"
    assert (want_ret == 1).sum() >= min_concord and (want_ret == 0).sum() > 10, (int((want_ret == 1).sum()), len(want_ret))
    return want


def _compare(got, dump, c):
    want = {k: dump[f"c{c}_circ_r_{k}"] for k in KEYS}
    for k in KEYS:
        g, w = got[k], want[k]
        bad = np.nonzero((g != w).reshape(len(w), -1).any(axis=1))[0]
        assert len(bad) == 0, f"contig {c} {k}: {len(bad)} of {len(w)} problems differ, first {bad[:5]}: {g[bad[:2]]} vs {w[bad[:2]]}"
    return want


def _shape_checks(want, prob):
    cnt = want["chain_count"]
    assert (cnt > 0).sum() > len(cnt) // 3 and (cnt == 0).sum() > 10      # chains found, and problems without any
    assert (cnt > 1).sum() > 5                                            # several chains survive the cut
    assert want["chain_len"].max() >= 10                                  # long chains of overlapping 8-mers
    short = prob["qepos"].astype(np.int64) - prob["qspos"].astype(np.int64) + 1 < 8
    assert short.sum() > 3 and (cnt[short] == 0).all()


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_host_build_equals_process_circ_chaining(tmp_path):
    from circminer_b200 import workload
    from circminer_b200.api import default_params
    from hostsim_api import HostSim
    w = workload.prepare("T0", str(tmp_path / "t0"), contig_size=250_000)
    dump = live_dump(str(tmp_path / "circ.npz"), workload="T0", contig_size=250_000, circ=700, no_stage_dumps=True)
    hs = HostSim(default_params())
    for c, seq in enumerate(w["contigs"]):
        hs.build_contig(c, seq)
        hs.load_annotation(c, w["flat"][c])
    for c in range(int(dump["n_contigs"])):
        prob = {k: dump[f"c{c}_circ_p_{k}"] for k in ("seqs", "qspos", "qepos", "gene_start", "gene_end")}
        got = hs.circ_chaining(c, w["contigs"][c], **prob)
        want = _compare(got, dump, c)
        _shape_checks(want, prob)
        _compare_exact(hs, dump, c, 5)


@pytest.mark.gpu
@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_gpu_equals_process_circ_chaining(tmp_path):
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, default_params
    w = workload.prepare("T1", str(tmp_path / "t1"), contig_size=3_500_000, n_pairs=2000)
    dump = live_dump(str(tmp_path / "circ.npz"), workload="T1", contig_size=3_500_000, pairs=2000, circ=3000, no_stage_dumps=True)
    m = Mapper(default_params())
    workload.load_into(m, w)
    for c in range(int(dump["n_contigs"])):
        prob = {k: dump[f"c{c}_circ_p_{k}"] for k in ("seqs", "qspos", "qepos", "gene_start", "gene_end")}
        got = m.circ_chaining(c, F=REF_F, **prob)
        want = _compare(got, dump, c)
        _shape_checks(want, prob)
        pm = _compare_exact(m, dump, c, 20)
        assert {0, 9} <= set(np.unique(pm["type"]).tolist())   # CONCRD and CANDID partial mates
        again = m.circ_chaining(c, F=REF_F, **prob)       # the index is kept; a second call gives the same
        for k in KEYS:
            assert np.array_equal(got[k], again[k])
    # a shared context reads the root's ws-mer index
    child = Mapper(parent=m)
    prob = {k: dump[f"c0_circ_p_{k}"][:200] for k in ("seqs", "qspos", "qepos", "gene_start", "gene_end")}
    got = child.circ_chaining(0, F=REF_F, **prob)
    for k in KEYS:
        assert np.array_equal(got[k], dump[f"c0_circ_r_{k}"][:200]), k
    child.close()
    m.close()


def _repeat_rich(tmp_path, run):
    """C3's shape at 1/50 (Alu / L1-like families and tandem repeats inside genes): hundreds of equal-score chains per problem,
    where chain_seeds_sorted_kbest2's `repeats` filter matters -- it looks up the gene-relative position in a set of shifted
    ones (src/chain.cpp:494-495 vs :509-510), which is kept."""
    from circminer_b200 import synth, workload
    from oracle.cpu_baseline import ensure_ref_index
    w = workload.prepare("C3s", str(tmp_path / "c3s"), n_pairs=1000)
    p = workload.params_for("C3s")
    flat = w["flat"][0]
    prob = synth.split_read_problems(w["contigs"][0], flat["gene_start"], flat["gene_end"], w["s1"].shape[1], 1500, 77)
    order = np.argsort(prob["gene_start"], kind="stable")
    prob = {k: v[order] for k, v in prob.items()}
    ensure_ref_index(w, "C3s", p.kmer)
    np.savez(str(tmp_path / "prob.npz"), **prob)
    code = ("import sys, json, numpy as np; sys.path.insert(0, %r)\n"
            "from circminer_b200 import workload\n"
            "from oracle.refharness import RefHarness\n"
            "w = json.load(open(%r)); prob = dict(np.load(%r))\n"
            "h = RefHarness(workload.params_for('C3s')); h.open(w['fa'], w['gtf'], %r); h.next_contig()\n"
            "np.savez(%r, **h.circ_chain(prob))\n")
    import json
    import subprocess
    import sys
    from conftest import ROOT
    json.dump(dict(fa=w["fa"], gtf=w["gtf"]), open(str(tmp_path / "w.json"), "w"))
    subprocess.run([sys.executable, "-c", code % (ROOT, str(tmp_path / "w.json"), str(tmp_path / "prob.npz"), str(tmp_path / "ref"),
                                                  str(tmp_path / "want.npz"))], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    want = np.load(str(tmp_path / "want.npz"))
    got = run(w, p, prob)
    for k in KEYS:
        bad = np.nonzero((got[k] != want[k]).reshape(len(want[k]), -1).any(axis=1))[0]
        assert len(bad) == 0, f"{k}: {len(bad)} of {len(want[k])} problems differ, first {bad[:5]}"
    assert (want["chain_count"] == p.max_chain_len).sum() > 20   # problems with a full list of tied chains


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_host_build_chaining_in_repeat_rich_genes(tmp_path):
    def run(w, p, prob):
        from hostsim_api import HostSim
        hs = HostSim(p)
        hs.build_contig(0, w["contigs"][0])
        hs.load_annotation(0, w["flat"][0])
        return hs.circ_chaining(0, w["contigs"][0], **prob)
    _repeat_rich(tmp_path, run)


@pytest.mark.gpu
@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_gpu_chaining_in_repeat_rich_genes(tmp_path):
    def run(w, p, prob):
        from circminer_b200.api import Mapper
        m = Mapper(p)
        m.build_contig(0, np.ascontiguousarray(w["contigs"][0]))
        m.load_annotation(0, w["flat"][0])
        got = m.circ_chaining(0, F=REF_F, **prob)
        m.close()
        return got
    _repeat_rich(tmp_path, run)


# ---- third piece: check_split_map and what is under it ------------------------------------------------------------------------
def _check_split_map(tmp_path, make_engine, pairs):
    """Every check_split_map call the reference makes on the single-split path of the back-splice pairs of T1 (oracle/refdump.py
    --circ-single: its own call_circ_single_split control flow over its own member functions, every call recorded): the return
    value, the CircRes it fills (breakpoints, signals, reference bases) and the three mates as it leaves them."""
    from circminer_b200 import workload
    w, dump = _t1_single_split_dump(pairs)
    eng = make_engine(w)
    seen = set()
    total = total4 = 0
    for c in range(int(dump["n_contigs"])):
        rec = {k[len(f"c{c}_circs_"):]: dump[k] for k in dump.files if k.startswith(f"c{c}_circs_")}
        n = len(rec["ret"])
        total += n
        if not n:
            continue
        got = eng.circ_check(c, rec)
        bad = np.nonzero(got["ret"] != rec["ret"])[0]
        assert len(bad) == 0, f"contig {c}: check_split_map differs on {len(bad)} of {n} calls, first {bad[:5]}: {got['ret'][bad[:5]]} vs {rec['ret'][bad[:5]]}"
        for nm, k in (("r1", "out_r1"), ("r2", "out_r2"), ("pm", "out_pm"), ("cr", "cr")):
            bad = np.nonzero(got[nm] != rec[k])[0]
            assert len(bad) == 0, f"contig {c} {nm}: {len(bad)} of {n} differ, first {bad[:3]}: {got[nm][bad[:1]]} vs {rec[k][bad[:1]]}"
        seen |= set(np.unique(rec["ret"]).tolist())
        # the four-mate function on rows made from these calls (oracle/refdump.py: circ_check4_problems): one read cut in two regular
        # pieces, the other back-spliced -- its two single-BSJ arms, which merge the pieces and end in final_check
        r4 = {k[len(f"c{c}_circ4_"):]: dump[k] for k in dump.files if k.startswith(f"c{c}_circ4_")}
        if r4:
            got4 = eng.circ_check4(c, r4)
            bad = np.nonzero((got4["ret"] != r4["ret"]) | (got4["cr"] != r4["cr"]))[0]
            assert len(bad) == 0, f"contig {c}: four-mate check_split_map differs on {len(bad)} of {len(r4['ret'])} rows, first {bad[:5]}"
            total4 += len(r4["ret"])
    assert total > 100 and {0, 20, 30} <= seen, (total, seen)   # FR, CR and UD calls
    assert total4 > 50, total4
    return eng


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_host_build_equals_check_split_map(tmp_path):
    def make(w):
        from circminer_b200.api import default_params
        from hostsim_api import HostSim
        hs = HostSim(default_params())
        for c, seq in enumerate(w["contigs"]):
            hs.build_contig(c, seq)
            hs.load_annotation(c, w["flat"][c])
        return hs
    hs = _check_split_map(tmp_path, make, 8000)
    import ctypes as C
    tr = (C.c_long * 16)()
    hs.lib.hs_circ_trace(tr)
    # both arms of final_check ran, and through the made-up four-mate rows its two single-BSJ arms (8, 9) and rescue_overlapping_bsj (3, 4)
    assert tr[0] > 10 and tr[1] > 50 and tr[8] > 5 and tr[9] > 5 and tr[4] > 5, list(tr)


@pytest.mark.gpu
@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_gpu_equals_check_split_map(tmp_path):
    def make(w):
        from circminer_b200 import workload
        from circminer_b200.api import Mapper, default_params
        m = Mapper(default_params())
        workload.load_into(m, w)
        return m
    m = _check_split_map(tmp_path, make, 20000)
    m.close()


# ---- the single-split path end to end -------------------------------------------------------------------------------------------
def _single_split(tmp_path, make_engine, run, pairs):
    """call_circ_single_split for a batch (csrc/cm_circ_orch.h over the three pieces) against the reference in two steps:
    (1) the UNMODIFIED ProcessCirc::call_circ on every back-splice pair of T1 -- the type column it writes to candidates.pam per
    attempt and the CircRes it pushes to circ_res; (2) its control flow restated over the reference's own member functions with
    every check_split_map call recorded (oracle/ref_harness.cpp: cmref_circ_single), which must agree with (1) and gives the
    mates of every attempt.  The batch entry must reproduce both: attempts in order with their mates before and after, types,
    CircRes, and per pair what the function ends with."""
    from circminer_b200 import workload
    w, dump = _t1_single_split_dump(pairs)
    eng = make_engine(w)
    n, L = w["s1"].shape
    lens = np.full(n, L, np.uint16)
    total = 0
    for c in range(int(dump["n_contigs"])):
        rec = {k[len(f"c{c}_circs_"):]: dump[k] for k in dump.files if k.startswith(f"c{c}_circs_")}
        real = {k[len(f"c{c}_circreal_"):]: dump[k] for k in dump.files if k.startswith(f"c{c}_circreal_")}
        cnt = np.bincount(rec["call_pair"], minlength=n)
        off = np.concatenate([[0], np.cumsum(cnt)])
        # (2) agrees with (1)
        assert (real["n_types"] == cnt).all()
        for i in np.nonzero(cnt)[0]:
            assert (real["types"][i, :cnt[i]] == rec["ret"][off[i]:off[i + 1]]).all(), i
        rec["mr_con"] = rec["mr_con"].copy()
        rec["mr_con"]["type"][rec["mr_con"]["type"] == 4] = -1   # the entry takes CHI2BSJ pairs as well; this test is about the CHIBSJ ones
        got = run(eng, c, w, lens, rec)
        assert len(got["ret"]) == len(rec["ret"]), (len(got["ret"]), len(rec["ret"]))
        for k in ("call_pair", "r1_partial", "ret", "cr", "in_r1", "in_r2", "in_pm", "out_r1", "out_r2", "out_pm"):
            bad = np.nonzero(got[k] != rec[k])[0]
            assert len(bad) == 0, f"contig {c} {k}: {len(bad)} of {len(rec[k])} attempts differ, first {bad[:5]}"
        assert (got["best_type"] == rec["best_type"]).all()
        bt = got["best_type"]
        pushed = (bt >= 20) & (bt <= 22)
        assert (pushed == (real["pushed"] == 1)).all() and (real["push_type"][pushed] == bt[pushed]).all()
        shift = rec["mr_con"]["spos_r1"].astype(np.int64) - rec["mr_chr"]["spos_r1"].astype(np.int64)   # contig -> chromosome coordinates
        assert (got["best_cr"]["spos"][pushed].astype(np.int64) - shift[pushed] == real["push_spos"][pushed]).all()
        assert (got["best_cr"]["epos"][pushed].astype(np.int64) - shift[pushed] == real["push_epos"][pushed]).all()
        total += int(pushed.sum())
    assert total > 50
    return eng


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_host_build_equals_call_circ_single_split(tmp_path):
    def make(w):
        from circminer_b200.api import default_params
        from hostsim_api import HostSim
        hs = HostSim(default_params())
        for c, seq in enumerate(w["contigs"]):
            hs.build_contig(c, seq)
            hs.load_annotation(c, w["flat"][c])
        return hs
    _single_split(tmp_path, make, lambda hs, c, w, lens, rec: hs.circ_single(c, w["contigs"][c], w["s1"], w["s2"], lens, lens, rec["mr_con"],
                                                                             rec["gene_cnt"], rec["gene_se"]), 8000)


@pytest.mark.gpu
@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_gpu_equals_call_circ_single_split(tmp_path):
    def make(w):
        from circminer_b200 import workload
        from circminer_b200.api import Mapper, default_params
        m = Mapper(default_params())
        workload.load_into(m, w)
        return m
    m = _single_split(tmp_path, make, lambda m, c, w, lens, rec: m.circ_single(c, w["s1"], w["s2"], lens, lens, rec["mr_con"], rec["gene_cnt"],
                                                                               rec["gene_se"]), 20000)
    m.close()


# ---- the double-split path ------------------------------------------------------------------------------------------------------
def _mate5(m, sh):
    return "%u\t%u\t%d\t%d\t%d\t" % (int(m["spos"]) - sh, int(m["epos"]) - sh, int(m["qspos"]), int(m["matched_len"]), int(m["dir"]))


def _double_split(tmp_path, make_engine, single, double, pairs):
    """call_circ_double_split for a batch against the UNMODIFIED ProcessCirc::call_circ on every CHI2BSJ pair of T1: the text it writes
    to candidates.pam (print_split_mapping lines of its own four-mate attempts and of the call_circ_single_split calls it makes on
    the same pair, in order, byte for byte) and everything it pushes to circ_res."""
    from circminer_b200 import workload
    w, dump = _t1_single_split_dump(pairs)
    eng = make_engine(w)
    n, L = w["s1"].shape
    lens = np.full(n, L, np.uint16)
    names = [str(x) for x in dump["chr_names"]]
    n_pairs = n_kind1 = n_single_calls = 0
    for c in range(int(dump["n_contigs"])):
        rec = {k[len(f"c{c}_circs_"):]: dump[k] for k in dump.files if k.startswith(f"c{c}_circs_")}
        real = {k[len(f"c{c}_circreal2_"):]: dump[k] for k in dump.files if k.startswith(f"c{c}_circreal2_")}
        idx = np.nonzero(rec["mr_con"]["type"] == 4)[0]
        args = (w["s1"][idx], w["s2"][idx], lens[idx], lens[idx], rec["mr_con"][idx], rec["gene_cnt"][idx], rec["gene_se"][idx])
        ev = double(eng, c, w, *args)
        sg = single(eng, c, w, *args)
        shift = (rec["mr_con"]["spos_r1"].astype(np.int64) - rec["mr_chr"]["spos_r1"].astype(np.int64))[idx]
        s_off = np.concatenate([[0], np.cumsum(np.bincount(sg["call_pair"], minlength=len(idx)))])
        e_off = np.concatenate([[0], np.cumsum(np.bincount(ev["pair"], minlength=len(idx)))])
        for k, i in enumerate(idx):
            sh = int(shift[k])
            s_text = "".join("r\t%s\t" % names[c] + _mate5(sg["out_pm"][a]["m"], sh) + _mate5(sg["out_r1"][a]["m"], sh) + _mate5(sg["out_r2"][a]["m"], sh)
                             + "%d\n" % sg["ret"][a] for a in range(s_off[k], s_off[k + 1]))
            bt = int(sg["best_type"][k])
            s_push = [(bt, int(sg["best_cr"]["spos"][k]) - sh, int(sg["best_cr"]["epos"][k]) - sh)] if 20 <= bt <= 22 else []
            text, pushes = "", []
            for a in range(e_off[k], e_off[k + 1]):
                if ev["kind"][a] == 0:
                    text += s_text
                    pushes += s_push
                    n_single_calls += 1
                else:
                    text += ("r\t%s\t" % names[c] + _mate5(ev["p1"][a]["m"], sh) + _mate5(ev["p2"][a]["m"], sh) + _mate5(ev["r1"][a]["m"], sh)
                             + _mate5(ev["r2"][a]["m"], sh) + "%d\n" % ev["ret"][a])
                    n_kind1 += 1
            if ev["push_type"][k] >= 20:
                pushes.append((int(ev["push_type"][k]), int(ev["push_cr"]["spos"][k]) - sh, int(ev["push_cr"]["epos"][k]) - sh))
            want = bytes(real["text"][int(real["text_off"][i]):int(real["text_off"][i + 1])]).decode()
            assert text == want, f"contig {c} pair {i}: candidates.pam text differs\n got {text!r}\nwant {want!r}"
            wp = [(int(real["push_type"][i][q]), int(real["push_spos"][i][q]), int(real["push_epos"][i][q])) for q in range(real["n_push"][i])]
            assert pushes == wp, f"contig {c} pair {i}: circ_res pushes differ: {pushes} vs {wp}"
            n_pairs += 1
    assert n_pairs > 50 and n_kind1 > 20 and n_single_calls > 20, (n_pairs, n_kind1, n_single_calls)
    return eng


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_host_build_equals_call_circ_double_split(tmp_path):
    def make(w):
        from circminer_b200.api import default_params
        from hostsim_api import HostSim
        hs = HostSim(default_params())
        for c, seq in enumerate(w["contigs"]):
            hs.build_contig(c, seq)
            hs.load_annotation(c, w["flat"][c])
        return hs
    _double_split(tmp_path, make, lambda hs, c, w, *a: hs.circ_single(c, w["contigs"][c], *a), lambda hs, c, w, *a: hs.circ_double(c, w["contigs"][c], *a),
                  20000)


@pytest.mark.gpu
@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_gpu_equals_call_circ_double_split(tmp_path):
    def make(w):
        from circminer_b200 import workload
        from circminer_b200.api import Mapper, default_params
        m = Mapper(default_params())
        workload.load_into(m, w)
        return m
    m = _double_split(tmp_path, make, lambda m, c, w, *a: m.circ_single(c, *a), lambda m, c, w, *a: m.circ_double(c, *a), 20000)
    m.close()


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_list_overflow_hands_the_pair_back(tmp_path):
    """A check_split_map that needs more than a fixed-size list of the device code holds (CM_CIRC_TID_CAP transcripts at a breakpoint,
    CM_CIRC_JUNC_CAP junctions) must not fail the batch: the pair comes back as type -2 (the drop-in redoes it with the reference's
    call_circ) and every other pair is what the normal build gives.  Host build with the caps at 2 / 1 against the normal one, on the
    single- and double-split pairs of T1."""
    from circminer_b200 import workload
    from circminer_b200.api import default_params
    from hostsim_api import HostSim
    pairs = 8000
    w, dump = _t1_single_split_dump(pairs)
    sims = []
    for variant in ("", "smallcap"):
        hs = HostSim(default_params(), variant=variant)
        for c, seq in enumerate(w["contigs"]):
            hs.build_contig(c, seq)
            hs.load_annotation(c, w["flat"][c])
        sims.append(hs)
    n, L = w["s1"].shape
    lens = np.full(n, L, np.uint16)
    handed_back = handed_back2 = kept = 0
    for c in range(int(dump["n_contigs"])):
        rec = {k[len(f"c{c}_circs_"):]: dump[k] for k in dump.files if k.startswith(f"c{c}_circs_")}
        args = (w["s1"], w["s2"], lens, lens, rec["mr_con"], rec["gene_cnt"], rec["gene_se"])
        a, b = (hs.circ_single(c, w["contigs"][c], *args) for hs in sims)
        over = b["best_type"] == -2
        handed_back += int(over.sum())
        assert (a["best_type"] != -2).all()
        assert (a["best_type"][~over] == b["best_type"][~over]).all() and (a["best_cr"][~over] == b["best_cr"][~over]).all()
        # attempts of the pairs that were not handed back are the same, in the same order
        ka, kb = ~over[a["call_pair"]], ~over[b["call_pair"]]
        assert ka.sum() == kb.sum()
        for f in ("call_pair", "ret", "cr", "in_r1", "in_r2", "in_pm", "out_r1", "out_r2", "out_pm"):
            assert (a[f][ka] == b[f][kb]).all(), f
        kept += int((~over & (a["best_type"] >= 0)).sum())
        idx = np.nonzero(rec["mr_con"]["type"] == 4)[0]
        args2 = tuple(x[idx] for x in args)
        ea, eb = (hs.circ_double(c, w["contigs"][c], *args2) for hs in sims)
        over2 = eb["push_type"] == -2
        handed_back2 += int(over2.sum())
        assert (ea["push_type"] != -2).all()
        assert (ea["push_type"][~over2] == eb["push_type"][~over2]).all() and (ea["push_cr"][~over2] == eb["push_cr"][~over2]).all()
        ka, kb = ~over2[ea["pair"]], ~over2[eb["pair"]]
        assert ka.sum() == kb.sum()
        for f in ("pair", "kind", "ret", "cr", "r1", "r2", "p1", "p2"):
            assert (ea[f][ka] == eb[f][kb]).all(), f
    assert handed_back > 5 and handed_back2 > 0 and kept > 100, (handed_back, handed_back2, kept)
