"""GRCh38-scale synthetic inputs.

One reference for the three configs: 24 chromosomes with the GRCh38 primary-assembly lengths
(3.09 Gbp -> three packed contigs: 1-5 | 6-12 | 13-22,X,Y, src/genome.cpp:96-145), an Alu-like
300-bp family (1 M copies, 10-15 % diverged), an L1-like 6-kb family (100 k copies), 2 % low-complexity
sequence (short tandem repeats), telomeric / centromeric N blocks, and 20 000 genes / ~60 000
transcripts with log-normal exon and intron lengths (median 130 bp / 1.5 kb, introns <= 1 Mb, so the
stateless chaining cursor of SURVEY Appendix A applies: no annotated intron beyond maxIntronLen).

Everything is vectorised numpy over ONE code array (0..3 = ACGT, 4 = N) so that the whole reference is
made in about a minute; `scale` shrinks every length and count by the same factor (tests use 1/50).
The read sampler is synth.make_reads (same three read classes as C1 / C2).
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field, replace

import numpy as np

from . import synth

GRCH38_NAMES = [str(i) for i in range(1, 23)] + ["X", "Y"]
GRCH38_LENS = [248956422, 242193529, 198295559, 190214555, 181538259, 170805979, 159345973, 145138636, 138394717,
               133797422, 135086622, 133275309, 114364328, 107043718, 101991189, 90338345, 83257441, 80373285,
               58617616, 64444167, 46709983, 50818468, 156040895, 57227415]


@dataclass
class ScaleConfig:
    name: str
    n_pairs: int
    read_len: int
    frag_lo: int
    frag_hi: int
    seed_reads: int
    kmer: int = 20
    compact_index: bool = False
    single_end: bool = False
    params: dict = field(default_factory=dict)   # cm_params overrides (scan_level, max_ed, ...)
    scale: float = 1.0
    seed_ref: int = 103
    n_genes: int = 20_000
    n_alu: int = 1_000_000
    n_l1: int = 100_000
    frac_lowcx: float = 0.02
    frac_linear: float = 0.85
    frac_circ: float = 0.08
    sub_rate: float = 0.01
    indel_rate: float = 0.0002
    contig_size: int = 1_100_000_000   # DEF_CONTIG_SIZE, src/common.h:81

    @property
    def chrom_lens(self):
        return [max(200_000, int(L * self.scale)) for L in GRCH38_LENS]

    @property
    def chrom_names(self):
        return GRCH38_NAMES


CONFIGS = {
    # the metric's configuration: 20 M pairs over 8 GPUs = 2.5 M pairs per GPU
    "C3": ScaleConfig("C3", 2_500_000, 150, 250, 450, 203),
    # single-end 1x100, k = 22 (int16 checksum truncation, src/match_read.cpp:77), best mapping
    "C4": ScaleConfig("C4", 2_500_000, 100, 220, 320, 204, kmer=22, single_end=True,
                      params=dict(kmer=22, scan_level=2, max_ed=4, band_width=3)),
    # 2x250, k = 14 compact index (every bucket entry matches), hit lists near seed_lim 500
    "C5": ScaleConfig("C5", 625_000, 250, 400, 700, 205, kmer=14, compact_index=True,
                      params=dict(kmer=14, max_read_len=300, seed_lim=500)),
}
# the same shapes at 1/50 of the size (63 Mbp, contig size 22 Mbp -> still three contigs): parity tests
for _k in ("C3", "C4", "C5"):
    _c = CONFIGS[_k]
    CONFIGS[_k + "s"] = replace(_c, name=_k + "s", scale=0.02, n_pairs=max(20_000, _c.n_pairs // 50),
                                n_genes=400, n_alu=20_000, n_l1=2_000, contig_size=22_000_000)


def _byte_lut(gc=0.41):
    """256-entry byte -> base code table with P(G) + P(C) = gc."""
    na = int(round((1 - gc) / 2 * 256))
    nc = 128 - na
    return np.concatenate([np.full(na, 0), np.full(nc, 1), np.full(nc, 2), np.full(na, 3)]).astype(np.uint8)


def make_genome(cfg: ScaleConfig):
    """Returns (g, off): g = all chromosomes back to back as codes, off[i]..off[i+1] = chromosome i."""
    rng = np.random.default_rng(cfg.seed_ref)
    lens = np.asarray(cfg.chrom_lens, np.int64)
    off = np.concatenate([[0], np.cumsum(lens)])
    total = int(off[-1])
    g = np.empty(total, np.uint8)
    lut = _byte_lut()
    CH = 1 << 27
    for lo in range(0, total, CH):
        hi = min(total, lo + CH)
        g[lo:hi] = lut[rng.integers(0, 256, size=hi - lo, dtype=np.uint8)]

    def positions(n, span):
        """n start positions of `span` bases, chromosome drawn by length, 20 kb clear of the ends."""
        c = rng.choice(len(lens), size=n, p=lens / total)
        room = np.maximum(lens[c] - span - 40_000, 1)
        return off[c] + 20_000 + (rng.random(n) * room).astype(np.int64)

    def plant_family(n, rlen, div_lo, div_hi):
        n = int(n)
        if n <= 0:
            return
        fam = rng.integers(0, 4, size=rlen, dtype=np.uint8)
        pos = positions(n, rlen)
        ar = np.arange(rlen, dtype=np.int64)
        step = max(1, (48 << 20) // rlen)
        for lo in range(0, n, step):
            hi = min(n, lo + step)
            m = hi - lo
            div = rng.uniform(div_lo, div_hi, size=m).astype(np.float32)
            mut = rng.random((m, rlen), dtype=np.float32) < div[:, None]
            add = rng.integers(1, 4, size=(m, rlen), dtype=np.uint8)
            cp = np.where(mut, (fam[None, :] + add) & 3, fam[None, :]).astype(np.uint8)
            rc = rng.random(m) < 0.5
            cp[rc] = 3 - cp[rc][:, ::-1]
            g[(pos[lo:hi, None] + ar).ravel()] = cp.ravel()

    plant_family(cfg.n_alu, 300, 0.10, 0.15)
    plant_family(cfg.n_l1, 6000, 0.03, 0.10)

    # low-complexity: tandem repeats of a 1..6-base unit, 100..700 bp per block
    n_blk = int(cfg.frac_lowcx * total / 400)
    if n_blk:
        blen = rng.integers(100, 701, size=n_blk)
        per = rng.integers(1, 7, size=n_blk)
        pos = positions(n_blk, 700)
        ar = np.arange(700, dtype=np.int64)
        for p in range(1, 7):
            sel = np.nonzero(per == p)[0]
            if not len(sel):
                continue
            unit = rng.integers(0, 4, size=(len(sel), p), dtype=np.uint8)
            seq = unit[:, ar % p]
            keep = ar[None, :] < blen[sel, None]
            idx = (pos[sel, None] + ar)[keep]
            g[idx] = seq[keep]

    # N blocks: 10 kb at both ends of every chromosome, a centromere of 2 % of its length at 40 %
    for i, L in enumerate(lens):
        o = int(off[i])
        L = int(L)
        g[o:o + 10_000] = 4
        g[o + L - 10_000:o + L] = 4
        c0 = o + int(0.4 * L)
        g[c0:c0 + int(0.02 * L)] = 4
    return g, off


def make_annotation(cfg: ScaleConfig, g, off):
    """Genes left to right on every chromosome; same dict layout as synth.make_annotation."""
    rng = np.random.default_rng(cfg.seed_ref + 7)
    lens = np.diff(off)
    total = int(off[-1])
    genes = []
    for ci in range(len(lens)):
        L = int(lens[ci])
        ch = g[off[ci]:off[ci + 1]]
        n_here = max(1, int(round(cfg.n_genes * L / total)))
        span = (L - 60_000) // (n_here + 1)
        for k in range(n_here):
            base = 30_000 + k * span + int(rng.integers(0, max(1, span // 4)))
            lim = min(base + span - 1000, L - 30_000)
            n_ex = int(np.clip(rng.geometric(0.12) + 1, 2, 40))
            elen = np.clip(np.exp(rng.normal(np.log(130), 0.6, size=n_ex)), 30, 3000).astype(np.int64)
            ilen = np.clip(np.exp(rng.normal(np.log(1500), 1.3, size=n_ex)), 70, 1_000_000).astype(np.int64)
            exons = []
            pos = base
            for j in range(n_ex):
                e = pos + int(elen[j]) - 1
                if e >= lim:
                    break
                exons.append((pos, e))
                pos = e + 1 + int(ilen[j])
            if len(exons) < 2:
                continue
            if (ch[exons[0][0] - 1:exons[-1][1]:997] == 4).any() or any((ch[s - 1:e] == 4).any() for s, e in exons):
                continue
            n_tx = int(rng.integers(1, 6))   # 3 on average
            txs = []
            for t in range(n_tx):
                if t == 0:
                    tx = list(exons)
                else:
                    keep = rng.random(len(exons)) < 0.7
                    keep[0] = True
                    keep[-1] = True
                    tx = [e for e, kp in zip(exons, keep) if kp]
                    if len(tx) > 2 and rng.random() < 0.5:
                        j = int(rng.integers(1, len(tx) - 1))
                        s, e = tx[j]
                        d = int(rng.integers(5, 30))
                        if e - s > 2 * d + 20:
                            tx[j] = (s + d, e) if rng.random() < 0.5 else (s, e - d)
                txs.append(tx)
            genes.append(dict(chr=ci, start=min(t[0][0] for t in txs), end=max(t[-1][1] for t in txs),
                              strand="+" if rng.random() < 0.5 else "-", transcripts=txs))
    return genes


def write_fasta(path, g, off, names):
    with open(path, "wb") as f:
        for i, nm in enumerate(names):
            f.write(b">" + nm.encode() + b"\n")
            CH = 1 << 26
            for lo in range(int(off[i]), int(off[i + 1]), CH):
                hi = min(int(off[i + 1]), lo + CH)
                f.write(synth._ACGT[g[lo:hi]].tobytes())
            f.write(b"\n")


def reference(cfg: ScaleConfig, outdir: str, need_fasta: bool = False) -> dict:
    """The reference of the config family (cached under outdir as .npy; genes as the GTF + a pickle)."""
    import pickle
    os.makedirs(outdir, exist_ok=True)
    gp, op, ap = (os.path.join(outdir, x) for x in ("genome.npy", "off.npy", "genes.pkl"))
    gtf = os.path.join(outdir, "ann.gtf")
    fa = os.path.join(outdir, "ref.fa")
    done = os.path.join(outdir, ".ref_done")
    if os.path.exists(done):
        g = np.load(gp, mmap_mode="r")
        off = np.load(op)
        genes = pickle.load(open(ap, "rb"))
    else:
        g, off = make_genome(cfg)
        genes = make_annotation(cfg, g, off)
        np.save(gp, g)
        np.save(op, off)
        pickle.dump(genes, open(ap, "wb"))
        synth.write_gtf(gtf, genes, names=cfg.chrom_names)
        open(done, "w").close()
    if need_fasta and not os.path.exists(fa + ".done"):
        write_fasta(fa, g, off, cfg.chrom_names)
        open(fa + ".done", "w").close()
    chroms = [g[off[i]:off[i + 1]] for i in range(len(off) - 1)]
    return dict(cfg=cfg, g=g, off=off, chroms=chroms, genes=genes, gtf=gtf, fa=fa, names=cfg.chrom_names)


def reads(cfg: ScaleConfig, ref: dict, n_pairs: int | None = None, seed_shift: int = 0):
    """(r1, r2, kind) code arrays; seed_shift picks another shard of the same sample."""
    sc = synth.SynthConfig(cfg.name, cfg.chrom_lens, cfg.n_genes, n_pairs or cfg.n_pairs, cfg.read_len, cfg.frag_lo,
                           cfg.frag_hi, cfg.seed_ref, cfg.seed_reads + seed_shift, frac_linear=cfg.frac_linear,
                           frac_circ=cfg.frac_circ, sub_rate=cfg.sub_rate, indel_rate=cfg.indel_rate)
    return synth.make_reads(sc, ref["chroms"], ref["genes"], genome=ref["g"], avoid_n=True)
