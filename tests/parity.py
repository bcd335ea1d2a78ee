"""Shared comparison code: an engine (api.Mapper on the GPU, hostsim on the CPU) against a reference dump
(oracle/refdump.py .npz: golden fixture or generated live from oracle/_ref)."""
from __future__ import annotations

import os
import subprocess
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

SEED_KEYS = ("frag_count", "has_frags", "first_info", "last_info", "sum_info")
CHAIN_KEYS = ("chain_count", "high_hits", "score", "chain_len", "frag_rpos", "frag_qpos")


def live_dump(path, **kw):
    """Run oracle/refdump.py in a subprocess (the reference keeps process-global state)."""
    cmd = [sys.executable, "-m", "oracle.refdump", "--out", path]
    for k, v in kw.items():
        flag = "--" + k.replace("_", "-")
        if v is True:
            cmd.append(flag)
        elif v not in (False, None):
            cmd += [flag, str(v)]
    subprocess.run(cmd, cwd=ROOT, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return np.load(path)


def check_reads(w, dump):
    crc = zlib.crc32(w["s1"].tobytes() + w["s2"].tobytes())
    assert crc == int(dump["reads_crc"]), "synthetic reads differ from the ones the dump was made on"


def compare_contig(engine, c, w, dump, mr_in, stages=("seed", "chain", "pe"), se=False):
    """Compare one contig round; returns the engine's updated MatchedReads."""
    from circminer_b200.api import MREAD_COMPARE_FIELDS
    s1, s2 = w["s1"], w["s2"]
    for nm, s in (("r1", s1), ("r2", s2)):
        if "seed" in stages and f"c{c}_seed_{nm}_frag_count" in dump and hasattr(engine, "split_match_hash"):
            got = engine.split_match_hash(c, s)
            for k in SEED_KEYS:
                assert (got[k] == dump[f"c{c}_seed_{nm}_{k}"]).all(), f"contig {c} {nm} seed {k}"
        if "chain" in stages and f"c{c}_chain_{nm}_score" in dump:
            got = engine.get_best_chains(c, s)
            for k in CHAIN_KEYS:
                want = dump[f"c{c}_chain_{nm}_{k}"]
                bad = np.nonzero((got[k] != want).reshape(len(want), -1).any(1))[0]
                assert len(bad) == 0, f"contig {c} {nm} chain {k}: {len(bad)} rows differ, first {bad[:5]}"
    if se and f"c{c}_se_state" in dump:
        st = engine.process_read_se(c, s1)
        assert (st == dump[f"c{c}_se_state"]).all(), f"contig {c} SE states differ"
    if "pe" in stages:
        out, st = engine.process_read(c, s1, s2, mr_in)
        want, wst = dump[f"c{c}_pe_mread"], dump[f"c{c}_pe_state"]
        bad = np.nonzero(st != wst)[0]
        assert len(bad) == 0, f"contig {c}: {len(bad)} states differ, first {bad[:5]}: {st[bad[:5]]} vs {wst[bad[:5]]}"
        for f in MREAD_COMPARE_FIELDS:
            bad = np.nonzero(out[f] != want[f])[0]
            assert len(bad) == 0, f"contig {c} field {f}: {len(bad)} differ, first {bad[:5]}"
        return out
    return mr_in


def compare_align(engine, dump, n):
    from oracle.refdump import align_problems
    probs = align_problems(77, n)
    for kind, (s, ns, t, nt, aux) in probs.items():
        got = engine.align(kind, s, ns, t, nt, aux)
        for k in ("ed", "sclen", "indel", "score"):
            want = dump[f"align{kind}_{k}"]
            bad = np.nonzero(got[k] != want)[0]
            assert len(bad) == 0, f"align kind {kind} {k}: {len(bad)} differ, first {bad[:5]}: {got[k][bad[:5]]} vs {want[bad[:5]]}"
