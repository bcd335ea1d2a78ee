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


def parse_index(path):
    """The reference's on-disk index (src/mrsfast/HashTable.c:106-254) -> (header dict, [contig dict]).  The six-byte table of a
    full index is returned canonical: the bytes the reference never initialises and never reads (the checksum field of every
    bucket's header slot and the slots behind a bucket's last entry; its table is malloc'ed, HashTable.c:346) are zeroed."""
    import struct
    b = np.fromfile(path, np.uint8)
    raw = b.tobytes() if len(b) < (1 << 28) else None
    mv = memoryview(b)

    def unpack(fmt, o):
        return struct.unpack_from(fmt, mv, o)
    magic, win, csl = int(b[0]), int(b[1]), int(b[2])
    maxmem, iobuf, cms = unpack("<III", 3)
    o = 15
    nchr, = unpack("<i", o)
    o += 4
    chrs = []
    for _ in range(nchr):
        nl, = unpack("<i", o)
        o += 4
        name = bytes(mv[o:o + nl]).decode()
        o += nl
        gl, = unpack("<i", o)
        o += 4
        chrs.append((name, gl))
    hdr = dict(magic=magic, window=win, checksum_len=csl, max_mem=maxmem, io_buffer=iobuf, contig_max_size=cms, chrs=chrs)
    contigs = []
    while o < len(b):
        more = int(b[o])
        o += 1
        nl, = unpack("<h", o)
        o += 2
        name = bytes(mv[o:o + nl]).decode()
        o += nl
        off, ln = unpack("<iI", o)
        o += 8
        nw = ln // 21 + (1 if ln % 21 else 0)
        cref = b[o:o + 8 * nw].view(np.uint64).copy()
        o += 8 * nw
        nne, = unpack("<I", o)
        o += 4
        blocks = []
        seen = 0
        while seen < nne:
            nb, = unpack("<i", o)
            o += 4
            blk = b[o:o + nb]
            o += nb
            blocks.append(blk.copy())
            seen += int((blk >> 7).sum()) // 2
        c = dict(more=more, name=name, offset=off, len=ln, cref=cref, n_nonempty=nne, blocks=blocks)
        if magic == 3:
            mem, = unpack("<I", o)
            o += 4
            tab = b[o:o + 6 * mem].reshape(mem, 6)
            o += 6 * mem
            cks = tab[:, 0].astype(np.uint16) | (tab[:, 1].astype(np.uint16) << 8)
            info = np.ascontiguousarray(tab[:, 2:]).view(np.uint32)[:, 0].copy()
            # decode the counts -> slot offsets of every non-empty bucket
            allb = np.concatenate(blocks)
            ends = np.nonzero(allb >> 7)[0]
            starts = np.concatenate([[0], ends[:-1] + 1])
            vals = np.zeros(len(ends), np.uint64)
            for k in range(5):   # a 32-bit value takes at most 5 bytes
                idx = starts + k
                ok = idx <= ends
                vals[ok] |= (allb[idx[ok]].astype(np.uint64) & 127) << np.uint64(7 * k)
            cnt = vals[1::2].astype(np.int64)
            slot0 = np.concatenate([[0], np.cumsum(cnt + 1)[:-1]])
            assert int((cnt + 1).sum()) == mem
            n_used = info[slot0].astype(np.int64)
            assert (n_used <= cnt).all()
            keep = np.zeros(mem + 1, np.int64)
            np.add.at(keep, slot0 + 1, 1)
            np.add.at(keep, slot0 + 1 + n_used, -1)
            used = np.cumsum(keep[:mem]) > 0      # entry slots
            head = np.zeros(mem, bool)
            head[slot0] = True
            cks = np.where(used, cks, 0).astype(np.uint16)
            info = np.where(used | head, info, 0).astype(np.uint32)
            c.update(mem=mem, cks=cks, info=info, bucket_windows=cnt)
        contigs.append(c)
    return hdr, contigs


def assert_same_index(mine, ref):
    h1, c1 = parse_index(mine)
    h2, c2 = parse_index(ref)
    assert h1 == h2, (h1, h2)
    assert len(c1) == len(c2)
    for a, b in zip(c1, c2):
        for k in ("more", "name", "offset", "len", "n_nonempty"):
            assert a[k] == b[k], (k, a[k], b[k])
        assert (a["cref"] == b["cref"]).all(), "packed text differs"
        assert len(a["blocks"]) == len(b["blocks"]) and all((x.shape == y.shape) and (x == y).all() for x, y in zip(a["blocks"], b["blocks"])), \
            "bucket-count blocks differ"
        if "mem" in b:
            assert a["mem"] == b["mem"]
            assert (a["info"] == b["info"]).all(), "table positions / bucket sizes differ"
            assert (a["cks"] == b["cks"]).all(), "table checksums differ"
