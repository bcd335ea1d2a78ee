"""TEST INFRASTRUCTURE ONLY: the unmodified reference BINARY (oracle/_ref/circminer) run on FASTQ files, and its text
outputs parsed back into cm_mread rows: mapping.pam (src/output.cpp:280-299) and the per-round remain-FASTQ headers
(src/filter.cpp:413-455) of a `-z` run.  Read names must be synth's `r<index>`."""
from __future__ import annotations

import os
import subprocess

import numpy as np

from circminer_b200.api import MREAD_DTYPE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "circminer")
MAPPED = (0, 1, 2, 3, 4, 5, 7)


def run_reference(fa, gtf, fq1, fq2, out, kmer=20, threads=1, extra=(), stage0=True, timeout=3000):
    cmd = [REF_BIN, "-r", fa, "-g", gtf, "-1", fq1, "-2", fq2, "-k", str(kmer), "-t", str(threads), "-P", "-z", "-o", out, *extra]
    if stage0:
        cmd += ["-q", "0"]   # mapping stage only (src/commandline_parser.cpp:189-195)
    p = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, LC_ALL="C"), timeout=timeout)
    assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
    return p.stdout + p.stderr


def _fill(m, i, chr_index, f):
    """f: chr1 s1 e1 mlen1 qs1 qe1 dir1 ed1 chr2 s2 e2 mlen2 qs2 qe2 dir2 ed2 tlen junc gm"""
    m["chr_r1"][i] = chr_index[f[0]]
    m["spos_r1"][i], m["epos_r1"][i], m["mlen_r1"][i], m["qspos_r1"][i], m["qepos_r1"][i] = (int(x) for x in f[1:6])
    m["r1_forward"][i] = f[6] == "+"
    m["ed_r1"][i] = int(f[7])
    m["chr_r2"][i] = chr_index[f[8]]
    m["spos_r2"][i], m["epos_r2"][i], m["mlen_r2"][i], m["qspos_r2"][i], m["qepos_r2"][i] = (int(x) for x in f[9:14])
    m["r2_forward"][i] = f[14] == "+"
    m["ed_r2"][i] = int(f[15])
    m["tlen"][i], m["junc_num"][i], m["gm_compatible"][i] = int(f[16]), int(f[17]), int(f[18])


def parse_pam(path, n, chr_names):
    """-> (seen[n] bool, type[n], mread[n] with the coordinate fields of mapped rows)."""
    chr_index = {c: i for i, c in enumerate(chr_names)}
    m = np.zeros(n, MREAD_DTYPE)
    seen = np.zeros(n, bool)
    typ = np.full(n, -1, np.int32)
    with open(path) as fh:
        for ln in fh:
            f = ln.rstrip("\n").split("\t")
            i = int(f[0][1:])
            assert not seen[i], f"pair {i} printed twice"
            seen[i] = True
            typ[i] = int(f[-1])
            if typ[i] in MAPPED:
                assert len(f) == 21
                _fill(m, i, chr_index, f[1:20])
            else:
                assert len(f) == 23 and all(x == "*" for x in f[1:22])
    m["type"] = typ
    return seen, typ, m


def parse_remain(path, n, chr_names):
    """Headers of one remain FASTQ -> (present[n] bool, mread[n] incl. contig_num; unmapped rows carry only type)."""
    chr_index = {c: i for i, c in enumerate(chr_names)}
    m = np.zeros(n, MREAD_DTYPE)
    present = np.zeros(n, bool)
    with open(path) as fh:
        for k, ln in enumerate(fh):
            if k % 4:
                continue
            f = ln.rstrip("\n").split(" ")
            assert f[0][0] == "@" and len(f) == 23, ln
            i = int(f[0][2:])
            present[i] = True
            t = int(f[2])
            m["type"][i] = t
            if t in MAPPED:
                _fill(m, i, chr_index, f[3:22])
                m["contig_num"][i] = int(f[22])
    return present, m


COORD_FIELDS = ["spos_r1", "spos_r2", "epos_r1", "epos_r2", "qspos_r1", "qspos_r2", "qepos_r1", "qepos_r2", "mlen_r1", "mlen_r2",
                "ed_r1", "ed_r2", "tlen", "chr_r1", "chr_r2", "junc_num", "r1_forward", "r2_forward", "gm_compatible"]


def assert_rows_equal(got, want, rows, what, with_contig=False):
    """type everywhere; coordinates where the type carries them."""
    rows = np.asarray(rows)
    bad = rows[got["type"][rows] != want["type"][rows]]
    assert len(bad) == 0, f"{what}: {len(bad)} types differ, first {bad[:5]}: {got['type'][bad[:5]]} vs {want['type'][bad[:5]]}"
    mp = rows[np.isin(want["type"][rows], MAPPED)]
    for f in COORD_FIELDS + (["contig_num"] if with_contig else []):
        bad = mp[got[f][mp] != want[f][mp]]
        assert len(bad) == 0, f"{what}: field {f}: {len(bad)} differ, first {bad[:5]}: {got[f][bad[:5]]} vs {want[f][bad[:5]]}"
