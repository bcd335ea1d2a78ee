"""The extension stage on its own (cm_gpu_extend_batch) against the reference's TransExtension:
extend_both_mates on chain pairs of real read pairs (with and without a common transcript set) and extend_chain_both_sides on
single chains, every MatchedMate field the reference writes.  A wrong intermediate mate that a later tie-break of
process_read would hide shows here."""
import numpy as np
import pytest

from conftest import has_ref
from parity import live_dump

pytestmark = pytest.mark.gpu

FIELDS = ("spos", "epos", "qspos", "qepos", "matched_len", "left_ed", "middle_ed", "right_ed", "sclen_left", "sclen_right", "dir",
          "type", "is_concord", "left_ok", "right_ok")


def _cmp(got, want, what):
    for f in FIELDS:
        bad = np.nonzero(got[f] != want[f])[0]
        assert len(bad) == 0, f"{what} {f}: {len(bad)} of {len(want)} differ, first {bad[:5]}: {got[f][bad[:5]]} vs {want[f][bad[:5]]}"


@pytest.mark.skipif(not has_ref(), reason="oracle/_ref not built")
def test_extension_stage_equals_transextension(tmp_path):
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, default_params
    w = workload.prepare("T1", str(tmp_path / "t1"), contig_size=3_500_000, n_pairs=8000)
    dump = live_dump(str(tmp_path / "ext.npz"), workload="T1", contig_size=3_500_000, pairs=8000, extend=True, no_stage_dumps=True)
    m = Mapper(default_params())
    workload.load_into(m, w)
    seen_types = set()
    for c in range(len(w["contigs"])):
        names = ("lseq", "rseq", "llen", "rlen", "l_chain_len", "l_rpos", "l_qpos", "r_chain_len", "r_rpos", "r_qpos", "iv_a", "iv_b")
        p0 = {k: dump[f"c{c}_ext_p0_{k}"] for k in names}
        p0["dir"] = None
        got = m.extend(c, 0, p0)
        bad = np.nonzero(got["ret"] != dump[f"c{c}_ext_r0_ret"])[0]
        assert len(bad) == 0, f"contig {c}: extend_both_mates return differs on {len(bad)} problems, first {bad[:5]}"
        _cmp(got["lmm"], dump[f"c{c}_ext_r0_lmm"], f"contig {c} lmm")
        _cmp(got["rmm"], dump[f"c{c}_ext_r0_rmm"], f"contig {c} rmm")
        seen_types |= set(np.unique(got["lmm"]["type"]).tolist())
        assert (p0["iv_a"] >= 0).sum() > 100 and (p0["iv_a"] < 0).sum() > 10
        p1 = {k: None for k in names}
        for k in ("lseq", "llen", "l_chain_len", "l_rpos", "l_qpos", "dir"):
            p1[k] = dump[f"c{c}_ext_p1_{k}"]
        got1 = m.extend(c, 1, p1)
        assert (got1["ret"] == dump[f"c{c}_ext_r1_ret"]).all()
        _cmp(got1["lmm"], dump[f"c{c}_ext_r1_lmm"], f"contig {c} single chain")
    assert {0, 9} <= seen_types   # CONCRD and CANDID mates both occur
    m.close()
