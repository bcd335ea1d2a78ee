"""GTF -> flat per-contig annotation (the `cm_annot_flat` arrays of include/cm_gpu.h).

Host-side restatement of GTFParser::load_gtf (src/gene_annotation.cpp:191-399),
FlatIntervalTree::build / build_trans2seg_table (src/interval_tree_impl.h:98-241) and
GenomePacker::pack_genome / load_index_info (src/genome.cpp:96-167).  A maintainer who keeps
the reference's GTFParser would instead fill cm_annot_flat from its members;
this module exists so that the CUDA path can be driven without the reference, and is checked
against the reference's own tables in tests/test_annotation.py.
"""
from __future__ import annotations

import re

import numpy as np

DEF_CONTIG_SIZE = 1_100_000_000   # src/common.h:81
MIDNCNT = 50                       # src/genome.cpp:13


def pack_genome(chroms, contig_size: int = DEF_CONTIG_SIZE):
    """chroms: list of (name, uint8 ASCII array).  Returns (contigs, info) where contigs is a list of
    uint8 arrays (packed contig text) and info rows are (name, contig_id(1-based), start, end) in FASTA
    order -- the rows of <ref>.packed.fa.index.info."""
    contigs, info = [], []
    cur = None
    size = 0
    mid = np.full(MIDNCNT, ord("N"), np.uint8)
    for name, seq in chroms:
        L = len(seq)
        if size == 0 or L + MIDNCNT + size > contig_size:
            if cur is not None:
                contigs.append(np.concatenate(cur))
            cur = [seq]
            info.append((name, len(contigs) + 1, 0, L))
            size = L
        else:
            cur += [mid, seq]
            info.append((name, len(contigs) + 1, size + MIDNCNT, size + MIDNCNT + L))
            size += MIDNCNT + L
    if cur is not None:
        contigs.append(np.concatenate(cur))
    return contigs, info


_TOK = {}


def _tokenize(s: str, delims: str):
    """GTFParser::tokenize (gene_annotation.cpp:84-104): split on any delimiter, skip empties."""
    rx = _TOK.get(delims)
    if rx is None:
        rx = _TOK[delims] = re.compile("[^" + re.escape(delims) + "]+")
    return rx.findall(s)


def parse_gtf(path: str):
    """Yield (chr, type, start, end, forward, gene_id, trans_id) for gene/transcript/exon rows."""
    with open(path) as f:
        for line in f:
            if line.startswith("#"):
                continue
            fields = _tokenize(line.rstrip("\r\n"), "\t")
            if len(fields) < 9 or fields[2] not in ("gene", "transcript", "exon"):
                continue
            attr = _tokenize(fields[8], ' ;"')
            gid = tid = "-"
            for i in range(0, len(attr) - 1, 2):
                if attr[i] == "gene_id":
                    gid = attr[i + 1]
                elif attr[i] == "transcript_id":
                    tid = attr[i + 1]
            yield fields[0], fields[2], int(fields[3]), int(fields[4]), fields[6] == "+", gid, tid


def flatten(gtf_path: str, info, contig_lens, max_read_len: int):
    """Returns a list (one per packed contig) of dicts with the cm_annot_flat arrays.
    info: rows (name, contig_id, start, end) as from pack_genome; contig_lens[c] = refGenLen of contig c."""
    n_contigs = max(r[1] for r in info)
    chr2con = {r[0]: (r[1] - 1, r[2]) for r in info}
    chr_index = {r[0]: i for i, r in enumerate(info)}

    n_bits = [int(contig_lens[c]) + 64 for c in range(n_contigs)]
    near = [np.zeros(n_bits[c], np.bool_) for c in range(n_contigs)]
    intr = [np.zeros(n_bits[c], np.bool_) for c in range(n_contigs)]
    genes = [[] for _ in range(n_contigs)]          # gid2ginfo: (start, end)
    n_trans = [0] * n_contigs
    merged = [dict() for _ in range(n_contigs)]     # (start, end, gene, next) -> [tid, ...]

    def add(con, seg, tid):
        merged[con].setdefault(seg, []).append(tid)

    prev = None   # dict of the pending exon

    def flush_prev(final_next_zero):
        nonlocal prev
        if prev is None:
            return
        if final_next_zero and prev["fw"]:
            prev["next"] = 0
        add(prev["con"], (prev["s"], prev["e"], prev["gene"], prev["next"]), prev["tid"])
        prev = None

    for chrom, typ, s, e, fw, gid, tid in parse_gtf(gtf_path):
        if chrom not in chr2con:
            continue
        con, shift = chr2con[chrom]
        s += shift
        e += shift
        if typ == "gene":
            genes[con].append((s, e))
            intr[con][s:e + 1] = True
        if typ == "transcript":
            n_trans[con] += 1
        if typ == "exon":
            intr[con][s:e + 1] = False
            R = max_read_len
            if s >= R:                       # uint32 wrap of `start - maxReadLength` (gene_annotation.cpp:273)
                near[con][s - R:s] = True
            if e + 1 >= R:
                near[con][e - R + 1:e + 1] = True
            cur = dict(con=con, s=s, e=e, fw=fw, gene=len(genes[con]) - 1, tid=n_trans[con] - 1, next=0)
            if prev is None:
                prev = cur
            else:
                if prev["fw"]:
                    prev["next"] = s
                else:
                    cur["next"] = prev["s"]
                add(prev["con"], (prev["s"], prev["e"], prev["gene"], prev["next"]), prev["tid"])
                prev = cur
        else:
            flush_prev(True)
    flush_prev(True)

    out = []
    for c in range(n_contigs):
        out.append(_flatten_contig(c, merged[c], genes[c], n_trans[c], near[c], intr[c], n_bits[c], info, chr_index))
    return out


def _flatten_contig(c, merged, genes, n_trans, near, intr, n_bits, info, chr_index):
    # std::map order of UniqSeg::operator< (common.cpp:110-118): start, end, gene_id, next_exon_beg DESC
    segs = sorted(merged.keys(), key=lambda k: (k[0], k[1], k[2], -k[3]))
    iv_spos, iv_epos, iv_seg_off = [], [], [0]
    iv_max_end, iv_min_end, iv_max_next = [], [], []
    seg_start, seg_end, seg_next, seg_gene, seg_tid_off, tids = [], [], [], [], [0], []

    if segs:
        # elementary intervals over covered regions; each lists its covering segments in map order
        bps = sorted(set([s[0] for s in segs] + [s[1] + 1 for s in segs]))
        order = {s: i for i, s in enumerate(segs)}
        import heapq
        active = []           # heap of (end, order) ; lazily removed
        live = set()
        k = 0                 # next seg to activate
        for bi in range(len(bps) - 1):
            a, b = bps[bi], bps[bi + 1] - 1
            while k < len(segs) and segs[k][0] <= a:
                live.add(k)
                heapq.heappush(active, (segs[k][1], k))
                k += 1
            while active and active[0][0] < a:
                live.discard(heapq.heappop(active)[1])
            if not live:
                continue
            cover = sorted(live)
            iv_spos.append(a)
            iv_epos.append(b)
            mx, mn, mnx = 0, 1_000_000_000, 0
            for si in cover:
                s = segs[si]
                seg_start.append(s[0]); seg_end.append(s[1]); seg_next.append(s[3]); seg_gene.append(s[2])
                tids.extend(merged[s])
                seg_tid_off.append(len(tids))
                mx = max(mx, s[1]); mn = min(mn, s[1]); mnx = max(mnx, s[3])
            iv_max_end.append(mx); iv_min_end.append(mn); iv_max_next.append(mnx)
            iv_seg_off.append(len(seg_start))
    n_real_iv = len(iv_spos)

    # trans2seg (interval_tree_impl.h:184-241)
    starts = [1_000_000_000] * n_trans
    ends = [0] * n_trans
    for i in range(n_real_iv):
        for sj in range(iv_seg_off[i], iv_seg_off[i + 1]):
            for t in tids[seg_tid_off[sj]:seg_tid_off[sj + 1]]:
                if i < starts[t]:
                    starts[t] = i
                if i > ends[t]:
                    ends[t] = i
    t2s_off = [0]
    for t in range(n_trans):
        t2s_off.append(t2s_off[-1] + max(0, ends[t] - starts[t] + 1))
    t2s = np.zeros(max(1, t2s_off[-1]), np.uint8)
    for i in range(n_real_iv):
        for sj in range(iv_seg_off[i], iv_seg_off[i + 1]):
            state = 1 if iv_spos[i] == seg_start[sj] else (3 if iv_epos[i] == seg_end[sj] else 2)
            for t in tids[seg_tid_off[sj]:seg_tid_off[sj + 1]]:
                t2s[t2s_off[t] + i - starts[t]] = state

    if n_real_iv == 0:
        # dummy interval (gene_annotation.cpp:368-382); max_end etc. stay unset in the reference
        iv_spos, iv_epos = [0xFFFFFFFF], [0xFFFFFFFF]
        iv_max_end, iv_min_end, iv_max_next = [0], [0], [0]
        seg_start, seg_end, seg_next, seg_gene = [0xFFFFFFFF], [0xFFFFFFFF], [0], [0]
        seg_tid_off = [0, 0]
        iv_seg_off = [0, 1]

    rows = [r for r in info if r[1] - 1 == c]
    u4 = lambda x: np.asarray(x, np.uint32)
    return dict(
        n_iv=len(iv_spos), iv_spos=u4(iv_spos), iv_epos=u4(iv_epos), iv_max_end=u4(iv_max_end),
        iv_min_end=u4(iv_min_end), iv_max_next_exon=u4(iv_max_next), iv_seg_off=u4(iv_seg_off),
        n_seg=len(seg_start), seg_start=u4(seg_start), seg_end=u4(seg_end), seg_next_exon_beg=u4(seg_next),
        seg_gene_id=u4(seg_gene), seg_tid_off=u4(seg_tid_off), n_tid=len(tids), tid=u4(tids),
        n_trans=n_trans, trans_start_ind=np.asarray(starts, np.int32), t2s_off=u4(t2s_off),
        t2s=t2s[:t2s_off[-1]], n_t2s=t2s_off[-1],
        n_genes=len(genes), gene_start=u4([g[0] for g in genes]), gene_end=u4([g[1] for g in genes]),
        n_chr=len(rows), chr_shift=u4([r[2] for r in rows]),
        chr_id=np.asarray([chr_index[r[0]] for r in rows], np.int32),
        n_bits=n_bits, near_border=np.packbits(np.pad(near, (0, (-n_bits) % 64)), bitorder="little").view(np.uint64),
        intronic=np.packbits(np.pad(intr, (0, (-n_bits) % 64)), bitorder="little").view(np.uint64))
