"""One packed contig at GRCh38 scale (1.09 Gbp, the reference's contig limit is 1.1 Gbp, positions cross 2^30):
device-side index build (1.09 G k-mers sorted by (bucket, checksum, position)), fused directory, 136 MB annotation
bitsets, then seeds / the whole path on reads cut from known places.  No reference run at this size (minutes of CPU);
the checks are the size-independent properties: every seed of an error-free read finds its own position, an error-free
genomic pair maps concordantly at the place it was cut from, and results do not depend on the batch they are in."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

L, K = 150, 20


def _contig(n, seed=7):
    rng = np.random.default_rng(seed)
    codes = rng.integers(0, 4, size=n, dtype=np.uint8)
    return np.frombuffer(b"ACGT", np.uint8)[codes]


def _revcomp(a):
    lut = np.zeros(256, np.uint8)
    for x, y in zip(b"ACGT", b"TGCA"):
        lut[x] = y
    return lut[a[..., ::-1]]


@pytest.mark.skipif(os.environ.get("CM_SKIP_SCALE") == "1", reason="CM_SKIP_SCALE=1")
def test_grch38_scale_contig(tmp_path):
    from circminer_b200 import annotation
    from circminer_b200.api import Mapper, default_params, fresh_mreads
    n = 1_090_000_000
    seq = _contig(n)
    gtf = tmp_path / "big.gtf"
    rows = []
    for g, (s, strand) in enumerate(((5_000_000, "+"), (600_000_000, "-"), (1_080_000_000, "+"))):   # the last one beyond 2^30
        e3 = s + 2500
        rows.append(f'1\tsyn\tgene\t{s}\t{e3}\t.\t{strand}\t.\tgene_id "G{g}"; gene_name "G{g}";')
        rows.append(f'1\tsyn\ttranscript\t{s}\t{e3}\t.\t{strand}\t.\tgene_id "G{g}"; transcript_id "T{g}"; gene_name "G{g}";')
        for x, (a, b) in enumerate(((s, s + 300), (s + 1000, s + 1400), (s + 2200, e3))):
            rows.append(f'1\tsyn\texon\t{a}\t{b}\t.\t{strand}\t.\tgene_id "G{g}"; transcript_id "T{g}"; exon_number "{x + 1}"; gene_name "G{g}";')
    gtf.write_text("\n".join(rows) + "\n")
    info = [("1", 1, 0, n)]
    flat = annotation.flatten(str(gtf), info, [n], 300)
    m = Mapper(default_params())
    m.build_contig(0, seq)
    m.load_annotation(0, flat[0])
    n_entries, ref_len = m.contig_stats(0)
    assert ref_len == n and n_entries > 0.99 * n
    rng = np.random.default_rng(11)
    npairs = 4000
    starts = np.concatenate([rng.integers(1000, n - 2000, size=npairs - 200), rng.integers((1 << 30) - 500, (1 << 30) + 500, size=100),
                             rng.integers(n - 1500, n - 700, size=100)])
    frag = rng.integers(250, 450, size=npairs)
    # 60 pairs inside the 400-bp middle exon of the gene beyond 2^30: transcript-concordant
    starts[:60] = 1_080_000_000 + 1000 - 1 + rng.integers(0, 10, size=60)
    frag[:60] = rng.integers(250, 380, size=60)
    idx = np.arange(L)
    r1 = seq[starts[:, None] + idx]                                   # forward, 0-based start
    r2 = _revcomp(seq[(starts + frag - L)[:, None] + idx])            # mate from the other end, reverse strand
    # seeds: slot t of R1 finds position start + t*K (1-based: + 1), uniquely in a random genome
    sd = m.split_match_hash(0, r1)
    nslots = (L - K) // K + 1
    fc = sd["frag_count"][:, 0:2 * nslots:2]
    fi = sd["first_info"][:, 0:2 * nslots:2]
    assert (fc >= 1).all()
    want = starts[:, None] + 1 + K * np.arange(nslots)
    assert ((fi == want) | (fc > 1)).all() and (fc == 1).mean() > 0.99
    out, st = m.process_read(0, r1, r2, fresh_mreads(npairs))
    # CONCRD (0) inside the exon, CONGNM (7, concordant on the genome) elsewhere
    assert (st[:60] == 0).all(), np.bincount(st[:60], minlength=14)
    assert (st[60:] == 7).mean() > 0.995, np.bincount(st, minlength=14)
    assert (out["type"] == st).all()
    ok = (st == 0) | (st == 7)
    assert (out["spos_r1"][ok] == starts[ok] + 1).all()
    assert (out["epos_r2"][ok] == (starts + frag)[ok]).all()
    assert (out["tlen"][ok] == frag[ok]).all()
    # batch independence: a sub-batch in another order gives the same rows
    sub = rng.permutation(npairs)[:500]
    out2, st2 = m.process_read(0, r1[sub], r2[sub], fresh_mreads(len(sub)))
    assert (st2 == st[sub]).all() and (out2 == out[sub]).all()
    m.close()
