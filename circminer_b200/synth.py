"""Seeded synthetic inputs for the mapping hot path.

Writes a reference FASTA, an Ensembl-style GTF (gene / transcript / exon rows in
that order -- the reference's parser relies on the order,
/root/reference/src/gene_annotation.cpp:221-326) and paired FASTQ files whose
pairs are drawn from linear transcripts, back-spliced (circular) exon runs and
plain genomic fragments.  Everything is numpy so that the 100 Mb / 1 M-pair
configuration is generated in well under a minute on the GPU box.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field

import numpy as np

_ACGT = np.frombuffer(b"ACGTN", dtype=np.uint8)
_COMP = np.array([3, 2, 1, 0, 4], dtype=np.uint8)


@dataclass
class SynthConfig:
    name: str
    chrom_lens: list
    n_genes: int
    n_pairs: int
    read_len: int
    frag_lo: int
    frag_hi: int
    seed_ref: int
    seed_reads: int
    n_repeats: int = 200          # diverged copies of a 300-bp repeat family
    n_long_repeats: int = 0       # copies of a 6-kb repeat
    n_gaps: int = 0               # 10-kb N blocks
    frac_linear: float = 0.85
    frac_circ: float = 0.08
    sub_rate: float = 0.01
    indel_rate: float = 0.0002
    single_end: bool = False
    extra: dict = field(default_factory=dict)


CONFIGS = {
    # C1: README test-run shape, the reference's own CPU-runnable case
    "C1": SynthConfig("C1", [2_000_000], 40, 10_000, 100, 220, 320, 101, 201),
    # C2: the single-GPU bench workload
    "C2": SynthConfig("C2", [30_000_000, 25_000_000, 20_000_000, 15_000_000, 10_000_000], 2000,
                      1_000_000, 150, 250, 450, 102, 202, n_repeats=15_000, n_long_repeats=500, n_gaps=20),
    # tiny smoke / unit-test shape
    "T0": SynthConfig("T0", [200_000, 120_000], 12, 400, 100, 220, 320, 11, 12, n_repeats=40),
    # mid-size parity shape (seconds on the CPU oracle)
    "T1": SynthConfig("T1", [3_000_000, 2_000_000], 120, 20_000, 150, 250, 450, 21, 22, n_repeats=1500,
                      n_long_repeats=20, n_gaps=2),
    # C5's shape in small: 2x250 reads for the k=14 compact index (17 seeds per strand, long hit lists)
    "T2": SynthConfig("T2", [1_500_000, 1_000_000], 60, 4_000, 250, 400, 700, 31, 32, n_repeats=800,
                      n_long_repeats=10, n_gaps=1),
}


def revcomp_codes(a: np.ndarray) -> np.ndarray:
    return _COMP[a[..., ::-1]]


def make_genome(cfg: SynthConfig):
    """Return list of uint8 code arrays (0..3 = ACGT, 4 = N), one per chromosome."""
    rng = np.random.default_rng(cfg.seed_ref)
    gc = 0.41
    p = np.array([(1 - gc) / 2, gc / 2, gc / 2, (1 - gc) / 2])
    chroms = []
    for L in cfg.chrom_lens:
        chroms.append(rng.choice(4, size=L, p=p).astype(np.uint8))
    total = sum(cfg.chrom_lens)

    def plant(n, rlen, div):
        if n <= 0:
            return
        fam = rng.integers(0, 4, size=rlen).astype(np.uint8)
        cidx = rng.choice(len(chroms), size=n, p=np.array(cfg.chrom_lens) / total)
        for c in cidx:
            L = len(chroms[c])
            if L < 4 * rlen:
                continue
            pos = int(rng.integers(1000, L - rlen - 1000))
            cp = fam.copy()
            m = rng.random(rlen) < div
            cp[m] = (cp[m] + rng.integers(1, 4, size=int(m.sum()))) % 4
            if rng.random() < 0.5:
                cp = revcomp_codes(cp)
            chroms[c][pos:pos + rlen] = cp

    plant(cfg.n_repeats, 300, 0.08)
    plant(cfg.n_long_repeats, 6000, 0.03)
    for _ in range(cfg.n_gaps):
        c = int(rng.integers(0, len(chroms)))
        L = len(chroms[c])
        if L < 200_000:
            continue
        pos = int(rng.integers(50_000, L - 60_000))
        chroms[c][pos:pos + 10_000] = 4
    return chroms


def make_annotation(cfg: SynthConfig, chroms):
    """Genes laid out left to right on each chromosome.  Returns a list of gene dicts:
    {chr, start, end, strand, transcripts: [[(s, e), ...] ascending]} with 1-based inclusive coords."""
    rng = np.random.default_rng(cfg.seed_ref + 7)
    total = sum(len(c) for c in chroms)
    genes = []
    for ci, ch in enumerate(chroms):
        n_here = max(1, int(round(cfg.n_genes * len(ch) / total)))
        span = len(ch) // (n_here + 1)
        for g in range(n_here):
            base = 2000 + g * span + int(rng.integers(0, max(1, span // 4)))
            n_ex = int(rng.integers(3, 11))
            exons = []
            pos = base
            for _ in range(n_ex):
                elen = int(rng.integers(80, 401))
                exons.append((pos, pos + elen - 1))
                pos += elen + int(rng.integers(200, 5001))
            if pos >= min(base + span - 500, len(ch) - 3000):
                # shrink: keep only exons that fit
                lim = min(base + span - 500, len(ch) - 3000)
                exons = [e for e in exons if e[1] < lim]
            if len(exons) < 2:
                continue
            # avoid N gaps inside exons
            if any((ch[s - 1:e] == 4).any() for s, e in exons):
                continue
            n_tx = int(rng.integers(1, 4))
            txs = []
            for t in range(n_tx):
                if t == 0:
                    tx = list(exons)
                else:
                    keep = rng.random(len(exons)) < 0.7
                    keep[0] = True
                    keep[-1] = True
                    tx = [e for e, k in zip(exons, keep) if k]
                    # alternative 5'/3' exon boundary on one exon -> partial-overlap segments
                    if len(tx) > 2 and rng.random() < 0.5:
                        j = int(rng.integers(1, len(tx) - 1))
                        s, e = tx[j]
                        d = int(rng.integers(5, 30))
                        if rng.random() < 0.5:
                            tx[j] = (s + d, e)
                        else:
                            tx[j] = (s, e - d)
                txs.append(tx)
            gs = min(t[0][0] for t in txs)
            ge = max(t[-1][1] for t in txs)
            genes.append(dict(chr=ci, start=gs, end=ge, strand="+" if rng.random() < 0.5 else "-",
                              transcripts=txs))
    return genes


def chrom_name(i: int) -> str:
    return str(i + 1)


def write_fasta(path, chroms):
    with open(path, "wb") as f:
        for i, c in enumerate(chroms):
            f.write(b">" + chrom_name(i).encode() + b"\n")
            f.write(_ACGT[c].tobytes())
            f.write(b"\n")


def write_gtf(path, genes, names=None):
    with open(path, "w") as f:
        f.write("#!synthetic annotation\n")
        tcount = 0
        for gi, g in enumerate(genes):
            ch = names[g["chr"]] if names else chrom_name(g["chr"])
            gid = f"SG{gi:06d}"
            attr_g = f'gene_id "{gid}"; gene_name "N{gi}";'
            f.write(f"{ch}\tsynth\tgene\t{g['start']}\t{g['end']}\t.\t{g['strand']}\t.\t{attr_g}\n")
            for tx in g["transcripts"]:
                tid = f"ST{tcount:07d}"
                tcount += 1
                attr_t = f'gene_id "{gid}"; transcript_id "{tid}"; gene_name "N{gi}";'
                f.write(f"{ch}\tsynth\ttranscript\t{tx[0][0]}\t{tx[-1][1]}\t.\t{g['strand']}\t.\t{attr_t}\n")
                order = tx if g["strand"] == "+" else tx[::-1]
                for en, (s, e) in enumerate(order):
                    attr_e = (f'gene_id "{gid}"; transcript_id "{tid}"; exon_number "{en + 1}"; '
                              f'gene_name "N{gi}";')
                    f.write(f"{ch}\tsynth\texon\t{s}\t{e}\t.\t{g['strand']}\t.\t{attr_e}\n")


def _build_pools(cfg, chroms, genes, rng):
    """Concatenate linear transcript sequences and tripled circular exon runs into one
    code array each, with (offset, length) tables."""
    lin_parts, lin_off, lin_len = [], [], []
    circ_parts, circ_off, circ_len = [], [], []
    lo = co = 0
    need = cfg.frag_hi + 8
    for g in genes:
        ch = chroms[g["chr"]]
        for tx in g["transcripts"]:
            seq = np.concatenate([ch[s - 1:e] for s, e in tx])
            if len(seq) >= need:
                lin_parts.append(seq)
                lin_off.append(lo)
                lin_len.append(len(seq))
                lo += len(seq)
            if len(tx) >= 3:
                a = int(rng.integers(0, len(tx) - 1))
                b = int(rng.integers(a + 1, len(tx)))
                cs = np.concatenate([ch[s - 1:e] for s, e in tx[a:b + 1]])
                reps = 2 + (need + len(cs) - 1) // len(cs)
                cs3 = np.tile(cs, reps)
                circ_parts.append(cs3)
                circ_off.append(co)
                circ_len.append(len(cs))
                co += len(cs3)
    lin = np.concatenate(lin_parts) if lin_parts else np.zeros(0, np.uint8)
    circ = np.concatenate(circ_parts) if circ_parts else np.zeros(0, np.uint8)
    return (lin, np.array(lin_off, np.int64), np.array(lin_len, np.int64),
            circ, np.array(circ_off, np.int64), np.array(circ_len, np.int64))


def make_reads(cfg: SynthConfig, chroms, genes, genome=None, avoid_n=False):
    """Returns (r1, r2) uint8 code arrays of shape (n_pairs, read_len) and a kind array
    (0 linear, 1 circ, 2 genomic).  genome: the chromosomes back to back when the caller already holds them as
    one array; avoid_n: redraw genomic fragments that start or end in an N block (GRCh38-scale references)."""
    rng = np.random.default_rng(cfg.seed_reads)
    n, L = cfg.n_pairs, cfg.read_len
    lin, lin_off, lin_len, circ, circ_off, circ_len = _build_pools(cfg, chroms, genes, rng)
    u = rng.random(n)
    kind = np.where(u < cfg.frac_linear, 0, np.where(u < cfg.frac_linear + cfg.frac_circ, 1, 2)).astype(np.int8)
    if len(lin_off) == 0:
        kind[kind == 0] = 2
    if len(circ_off) == 0:
        kind[kind == 1] = 2
    frag = rng.integers(cfg.frag_lo, cfg.frag_hi + 1, size=n)
    r1 = np.empty((n, L), np.uint8)
    r2 = np.empty((n, L), np.uint8)
    ar = np.arange(L)
    if genome is None:
        genome = np.concatenate(chroms)
    goff = np.concatenate([[0], np.cumsum([len(c) for c in chroms])])

    CH = 200_000
    for lo in range(0, n, CH):
        hi = min(n, lo + CH)
        k = kind[lo:hi]
        F = frag[lo:hi]
        m = hi - lo
        left = np.empty((m, L), np.uint8)
        right = np.empty((m, L), np.uint8)
        # linear
        sel = np.nonzero(k == 0)[0]
        if len(sel):
            t = rng.integers(0, len(lin_off), size=len(sel))
            Fs = np.minimum(F[sel], lin_len[t])
            st = (rng.random(len(sel)) * (lin_len[t] - Fs + 1)).astype(np.int64)
            base = lin_off[t] + st
            left[sel] = lin[base[:, None] + ar]
            right[sel] = lin[(base + Fs - L)[:, None] + ar]
        sel = np.nonzero(k == 1)[0]
        if len(sel):
            t = rng.integers(0, len(circ_off), size=len(sel))
            st = (rng.random(len(sel)) * circ_len[t]).astype(np.int64)
            base = circ_off[t] + st
            left[sel] = circ[base[:, None] + ar]
            right[sel] = circ[(base + F[sel] - L)[:, None] + ar]
        sel = np.nonzero(k == 2)[0]
        if len(sel):
            c = rng.integers(0, len(chroms), size=len(sel))
            clen = np.array([len(x) for x in chroms])[c]
            st = (rng.random(len(sel)) * (clen - F[sel] - 2000)).astype(np.int64) + 1000
            base = goff[c] + st
            if avoid_n:
                for _ in range(8):
                    bad = np.nonzero((genome[base] == 4) | (genome[base + F[sel] - 1] == 4))[0]
                    if not len(bad):
                        break
                    base[bad] = goff[c[bad]] + 1000 + (rng.random(len(bad)) * (clen[bad] - F[sel][bad] - 2000)).astype(np.int64)
            left[sel] = genome[base[:, None] + ar]
            right[sel] = genome[(base + F[sel] - L)[:, None] + ar]
        flip = rng.random(m) < 0.5
        a = np.where(flip[:, None], revcomp_codes(right), left)
        b = np.where(flip[:, None], left, revcomp_codes(right))
        r1[lo:hi] = a
        r2[lo:hi] = b

    # substitutions
    for r in (r1, r2):
        msk = rng.random(r.shape) < cfg.sub_rate
        cnt = int(msk.sum())
        v = r[msk]
        ok = v < 4
        v[ok] = (v[ok] + rng.integers(1, 4, size=cnt)[ok]) % 4
        r[msk] = v
    # indels (rare; row-wise)
    for r in (r1, r2):
        rows = np.nonzero(rng.random(n) < cfg.indel_rate * L)[0]
        for i in rows:
            p = int(rng.integers(5, L - 5))
            ln = int(rng.integers(1, 3))
            row = r[i].copy()
            if rng.random() < 0.5:   # deletion in read: shift left, pad with random bases
                row[p:L - ln] = row[p + ln:]
                row[L - ln:] = rng.integers(0, 4, size=ln)
            else:                    # insertion
                row[p + ln:] = row[p:L - ln].copy()
                row[p:p + ln] = rng.integers(0, 4, size=ln)
            r[i] = row
    return r1, r2, kind



def split_read_problems(text, gene_start, gene_end, L, n, seed):
    """Stage-2 (ProcessCirc::chaining) problems on one packed contig: rows shaped like the reads stage 2 re-maps (two pieces of
    one gene joined out of order, a few substitutions), one piece twice, random rows, rows with N, short and empty ranges; genes
    of every size, sometimes only a part of one.  text[p - 1] = base at 1-based contig position p (ASCII); genes in contig
    coordinates.  -> dict(seqs, qspos, qepos, gene_start, gene_end)."""
    rng = np.random.default_rng(seed)
    gs, ge = np.asarray(gene_start).astype(np.int64), np.asarray(gene_end).astype(np.int64)
    ref_len = len(text)
    ok = np.nonzero((ge - gs + 1 >= 60) & (ge <= ref_len))[0]
    acgt = np.frombuffer(b"ACGT", np.uint8)
    seqs = np.zeros((n, L), np.uint8)
    qs, qe, a_gs, a_ge = (np.zeros(n, np.uint32) for _ in range(4))
    for i in range(n):
        g = int(ok[rng.integers(len(ok))])
        s, e = int(gs[g]), int(ge[g])
        glen = e - s + 1
        u = rng.random()
        if u < 0.65:      # two pieces of the gene, second piece upstream of the first
            l1 = int(rng.integers(20, L - 20))
            l1 = min(l1, glen - 1)
            l2 = min(L - l1, glen - 1)
            p1 = s + int(rng.integers(0, glen - l1 + 1))
            p2 = s + int(rng.integers(0, glen - l2 + 1))
            row = np.concatenate([text[p1 - 1:p1 - 1 + l1], text[p2 - 1:p2 - 1 + l2]])
            row = np.concatenate([row, acgt[rng.integers(0, 4, size=L - len(row))]])
            mut = rng.random(L) < 0.01
            row = np.where(mut, acgt[rng.integers(0, 4, size=L)], row)
        elif u < 0.8:     # one piece, repeated twice (several hits per k-mer)
            l1 = min(L // 2, glen)
            p1 = s + int(rng.integers(0, glen - l1 + 1))
            piece = text[p1 - 1:p1 - 1 + l1]
            row = np.concatenate([piece, piece, acgt[rng.integers(0, 4, size=L)]])[:L]
        else:
            row = acgt[rng.integers(0, 4, size=L)]
        if rng.random() < 0.15:
            row = row.copy()
            row[rng.integers(0, L, size=int(rng.integers(1, 4)))] = ord("N")
        seqs[i] = row
        v = rng.random()
        if v < 0.3:
            qs[i], qe[i] = 1, L
        elif v < 0.6:
            qs[i], qe[i] = 1, int(rng.integers(8, L))
        elif v < 0.9:
            qs[i], qe[i] = int(rng.integers(2, L - 8)), L
        else:            # shorter than a window, or empty
            qs[i] = int(rng.integers(1, L))
            qe[i] = min(L, int(qs[i]) + int(rng.integers(-1, 8)))
        if rng.random() < 0.1:   # a window that is only part of the gene, down to a few bases
            e = min(e, s + int(rng.integers(3, 400)))
        a_gs[i], a_ge[i] = s, e
    return dict(seqs=seqs, qspos=qs, qepos=qe, gene_start=a_gs, gene_end=a_ge)


def read_names(n: int, kind=None):
    w = max(7, len(str(n)))
    return [f"r{str(i).zfill(w)}" for i in range(n)]


def write_fastq(path, codes: np.ndarray, mate: int):
    """Fixed-width records so the file is one 2-D uint8 array."""
    n, L = codes.shape
    w = max(7, len(str(n)))
    rec = 1 + 1 + w + 2 + 1 + L + 1 + 2 + L + 1   # @r<w>/m\n SEQ\n +\n QUAL\n
    out = np.empty((n, rec), np.uint8)
    out[:, 0] = ord("@")
    out[:, 1] = ord("r")
    idx = np.arange(n)
    for d in range(w):
        out[:, 2 + w - 1 - d] = (idx // (10 ** d)) % 10 + ord("0")
    p = 2 + w
    out[:, p] = ord("/")
    out[:, p + 1] = ord("0") + mate
    out[:, p + 2] = ord("\n")
    p += 3
    out[:, p:p + L] = _ACGT[codes]
    p += L
    out[:, p] = ord("\n")
    out[:, p + 1] = ord("+")
    out[:, p + 2] = ord("\n")
    p += 3
    out[:, p:p + L] = ord("I")
    p += L
    out[:, p] = ord("\n")
    out.tofile(path)


def generate(cfg: SynthConfig, outdir: str, write_reads: bool = True):
    """Generate (or reuse) the dataset under outdir.  Returns a dict of paths + arrays."""
    os.makedirs(outdir, exist_ok=True)
    chroms = make_genome(cfg)
    genes = make_annotation(cfg, chroms)
    fa = os.path.join(outdir, "ref.fa")
    gtf = os.path.join(outdir, "ann.gtf")
    if not os.path.exists(fa):
        write_fasta(fa, chroms)
    if not os.path.exists(gtf):
        write_gtf(gtf, genes)
    r1, r2, kind = make_reads(cfg, chroms, genes)
    fq1 = os.path.join(outdir, "reads_1.fastq")
    fq2 = os.path.join(outdir, "reads_2.fastq")
    if write_reads:
        write_fastq(fq1, r1, 1)
        write_fastq(fq2, r2, 2)
    return dict(cfg=cfg, chroms=chroms, genes=genes, fa=fa, gtf=gtf, fq1=fq1, fq2=fq2,
                r1=r1, r2=r2, kind=kind, names=read_names(cfg.n_pairs))


def codes_to_ascii(codes: np.ndarray) -> np.ndarray:
    return _ACGT[codes]


if __name__ == "__main__":
    import sys
    name = sys.argv[1] if len(sys.argv) > 1 else "T0"
    out = sys.argv[2] if len(sys.argv) > 2 else f"/tmp/cm_{name}"
    d = generate(CONFIGS[name], out)
    print(out, len(d["genes"]), "genes", d["r1"].shape)
