"""TEST INFRASTRUCTURE ONLY.  Runs the UNMODIFIED reference (oracle/_ref/libcm_ref.so) over a synthetic
workload, contig by contig as mapping() does, and dumps stage-level results to an .npz:
seed lists, chains, process_read (PE, carried MatchedRead) and optionally SE states and alignment
kernels on random problems.  One process per dump (the reference lives in process-global singletons).

    python -m oracle.refdump --workload T0 --contig-size 250000 --out tests/golden/T0_k20.npz
"""
from __future__ import annotations

import argparse
import os
import sys
import tempfile
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)


def align_problems(seed: int, n: int, stride: int = 200):
    """Random (s, t) problems for the five alignment kinds: t derived from s by a few edits, plus junk tails."""
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    out = {}
    for kind in range(5):
        s = np.zeros((n, stride), np.uint8)
        t = np.zeros((n, stride), np.uint8)
        ns = np.zeros(n, np.uint16)
        nt = np.zeros(n, np.uint16)
        aux = np.zeros(n, np.int32)
        for i in range(n):
            L = int(rng.integers(1, 120))
            base = acgt[rng.integers(0, 4, size=L + 40)]
            q = list(base[:L])
            for _ in range(int(rng.integers(0, 5))):
                p = int(rng.integers(0, max(1, len(q))))
                r = rng.random()
                if r < 0.5 and q:
                    q[p] = int(acgt[rng.integers(0, 4)])
                elif r < 0.75 and len(q) > 1:
                    del q[p]
                else:
                    q.insert(p, int(acgt[rng.integers(0, 4)]))
            if rng.random() < 0.3:          # diverging tail -> soft clip / x-drop
                k = int(rng.integers(1, 12))
                q = q[:max(1, len(q) - k)] + [int(x) for x in acgt[rng.integers(0, 4, size=k)]]
            if rng.random() < 0.05:
                q[int(rng.integers(0, len(q)))] = ord("N")
            q = np.array(q[:stride - 8], np.uint8)
            if kind in (0, 1):       # s = reference text (len + band), t = read piece
                a, b = base[:min(L + 3, stride - 8)], q
                if kind == 1:
                    a, b = a[::-1], b[::-1]
            elif kind in (2, 3):     # s = read piece (<= m + band), t = exon text
                m = min(L, stride - 8)
                b = base[:m]
                a = np.concatenate([q, acgt[rng.integers(0, 4, size=8)]])[:m + 3]
                if kind == 3:
                    a, b = a[::-1], b[::-1]
            else:                    # one-side banded: m = n + w
                w = int(rng.integers(0, 4))
                a = q
                tail = acgt[rng.integers(0, 4, size=w + len(q))]
                b = np.concatenate([base, tail])[:len(q) + w]
                aux[i] = w
            s[i, :len(a)] = a
            t[i, :len(b)] = b
            ns[i] = len(a)
            nt[i] = len(b)
        out[kind] = (s, ns, t, nt, aux)
    return out


def extend_problems(h, s1, s2, lens):
    """Problems for the extension stage and the reference's answers: for every pair whose R1 (forward) and R2 (reverse
    complement) both chain on the current contig, up to two chain pairs (best x best, second x second), left = the chain
    that starts first; common_tid from the exon intervals under the two chain starts.  Mode 1: the best chain of every R1."""
    comp = np.zeros(256, np.uint8)
    for x, y in zip(b"ACGTNacgtn", b"TGCANTGCAN"):
        comp[x] = y
    n, L = s1.shape
    rc2 = np.ascontiguousarray(comp[s2[:, ::-1]])
    c1, c2 = h.chain(s1, lens), h.chain(rc2, lens)
    F = c1["frag_rpos"].shape[2]
    rows = []
    for k in range(2):
        ok = np.nonzero((c1["chain_count"] > k) & (c2["chain_count"] > k))[0]
        rows += [(i, k) for i in ok]
    idx = np.array([r[0] for r in rows], np.int64)
    kk = np.array([r[1] for r in rows], np.int64)
    a_first = c1["frag_rpos"][idx, kk, 0] <= c2["frag_rpos"][idx, kk, 0]

    def pick(x1, x2):
        return np.where(a_first.reshape((-1,) + (1,) * (x1.ndim - 1)), x1, x2)
    prob = dict(lseq=pick(s1[idx], rc2[idx]), rseq=pick(rc2[idx], s1[idx]), llen=lens[idx], rlen=lens[idx],
                l_chain_len=pick(c1["chain_len"][idx, kk], c2["chain_len"][idx, kk]), r_chain_len=pick(c2["chain_len"][idx, kk], c1["chain_len"][idx, kk]),
                l_rpos=pick(c1["frag_rpos"][idx, kk], c2["frag_rpos"][idx, kk]), r_rpos=pick(c2["frag_rpos"][idx, kk], c1["frag_rpos"][idx, kk]),
                l_qpos=pick(c1["frag_qpos"][idx, kk], c2["frag_qpos"][idx, kk]), r_qpos=pick(c2["frag_qpos"][idx, kk], c1["frag_qpos"][idx, kk]))
    iv_l = h.annot_lookup(prob["l_rpos"][:, 0], h.kmer, 100)["loc_iv"]
    iv_r = h.annot_lookup(prob["r_rpos"][:, 0], h.kmer, 100)["loc_iv"]
    both = (iv_l >= 0) & (iv_r >= 0)
    prob["iv_a"] = np.where(both, iv_l, -1).astype(np.int32)
    prob["iv_b"] = np.where(both, iv_r, -1).astype(np.int32)
    prob["dir"] = None
    out = {("p0_" + k): v for k, v in prob.items() if v is not None}
    out.update({("r0_" + k): v for k, v in h.extend(0, prob).items()})
    one = np.nonzero(c1["chain_count"] > 0)[0]
    p1 = dict(lseq=s1[one], rseq=None, llen=lens[one], rlen=None, l_chain_len=c1["chain_len"][one, 0], l_rpos=c1["frag_rpos"][one, 0],
              l_qpos=c1["frag_qpos"][one, 0], r_chain_len=None, r_rpos=None, r_qpos=None, iv_a=None, iv_b=None,
              dir=np.where(np.arange(len(one)) % 2 == 0, 1, -1).astype(np.int32))
    out.update({("p1_" + k): v for k, v in p1.items() if v is not None})
    r1 = h.extend(1, p1)
    out["r1_lmm"], out["r1_ret"] = r1["lmm"], r1["ret"]
    return out


def circ_problems(h, L, n, seed):
    """Stage-2 chaining problems on the current contig (synth.split_read_problems over the reference's own gene table and text)."""
    from circminer_b200 import synth
    ann = h.annotation()
    ref_len = h.index_pointers()[4]
    text = np.frombuffer(h.genome_chars(1, ref_len), np.uint8)   # text[p - 1] = base at 1-based p
    return synth.split_read_problems(text, ann["gene_start"], ann["gene_end"], L, n, seed)


def circ_exact_problems(h, prob, res, seed, top=10):
    """find_exact_coord problems from chaining results: every chain ProcessCirc would try (the first min(count, TOPCHAIN)),
    with the pair's two mates placed in exon intervals of the chain's neighbourhood (common transcripts exist) or anywhere
    in the gene (mostly none)."""
    rng = np.random.default_rng(seed)
    ann = h.annotation()
    ivs, ive = ann["iv_spos"].astype(np.int64), ann["iv_epos"].astype(np.int64)
    has_seg = np.diff(ann["iv_seg_off"].astype(np.int64)) > 0
    L = prob["seqs"].shape[1]
    rows = []
    for i in np.nonzero(res["chain_count"] > 0)[0]:
        for j in range(min(int(res["chain_count"][i]), top)):
            rows.append((int(i), j))
    n = len(rows)
    F = res["frag_rpos"].shape[2]
    out = dict(seqs=np.zeros((n, L), np.uint8), r1_spos=np.zeros(n, np.uint32), r1_epos=np.zeros(n, np.uint32), r2_spos=np.zeros(n, np.uint32),
               r2_epos=np.zeros(n, np.uint32), dir=np.zeros(n, np.int32), qspos=np.zeros(n, np.uint32), rlen=np.zeros(n, np.int32),
               whole_len=np.full(n, L, np.int32), chain_len=np.zeros(n, np.uint32), frag_rpos=np.zeros((n, F), np.uint32),
               frag_qpos=np.zeros((n, F), np.int32))

    def place(anchor, gs, ge):
        if rng.random() < 0.75:
            k = int(np.searchsorted(ivs, anchor, side="right")) - 1 + int(rng.integers(-2, 3))
            k = min(max(k, 0), len(ivs) - 1)
            if has_seg[k]:
                s = int(rng.integers(ivs[k], ive[k] + 1))
                return s, s + int(rng.integers(20, L))
        s = int(rng.integers(gs, max(gs + 1, ge - L)))
        return s, s + int(rng.integers(20, L))

    for r, (i, j) in enumerate(rows):
        cl = int(res["chain_len"][i, j])
        out["seqs"][r] = prob["seqs"][i]
        out["chain_len"][r] = cl
        out["frag_rpos"][r] = res["frag_rpos"][i, j]
        out["frag_qpos"][r] = res["frag_qpos"][i, j]
        out["qspos"][r] = prob["qspos"][i]
        out["rlen"][r] = int(prob["qepos"][i]) - int(prob["qspos"][i]) + 1
        out["dir"][r] = 1 if rng.random() < 0.5 else -1
        gs, ge = int(prob["gene_start"][i]), int(prob["gene_end"][i])
        anchor = int(res["frag_rpos"][i, j, 0])
        out["r1_spos"][r], out["r1_epos"][r] = place(anchor, gs, ge)
        out["r2_spos"][r], out["r2_epos"][r] = place(anchor, gs, ge)
    return out


def circ_check4_problems(rec, seed):
    """Four-mate check_split_map rows made from recorded three-mate calls: the fully mapped read (when it has no indel, intron or
    soft clip inside) cut in two adjacent pieces at a random base, with the split read's two mates -- R1 regular and R2 back-spliced,
    or the mirror image -- which is what the two single-BSJ arms of the four-mate function (process_circ.cpp:968-985) take; they merge
    the pieces back and end in final_check / rescue_overlapping_bsj.  fullmap_sel names the cut read."""
    from circminer_b200.api import CIRC_MATE_DTYPE
    rng = np.random.default_rng(seed)
    r1p = np.asarray(rec["r1_partial"]).astype(bool)
    full = np.where(r1p, rec["in_r2"], rec["in_r1"]).astype(CIRC_MATE_DTYPE)
    part = np.where(r1p, rec["in_r1"], rec["in_r2"]).astype(CIRC_MATE_DTYPE)
    pm = np.array(rec["in_pm"], CIRC_MATE_DTYPE)
    fm = full["m"]
    glen = fm["epos"].astype(np.int64) - fm["spos"] + 1
    qlen = fm["qepos"].astype(np.int64) - fm["qspos"] + 1
    idx = np.nonzero((glen == qlen) & (glen == fm["matched_len"]) & (glen >= 20))[0]
    h = np.array([rng.integers(8, glen[i] - 8) for i in idx], np.uint32)
    lo, hi = full[idx].copy(), full[idx].copy()
    lo["m"]["epos"] = lo["m"]["spos"] + h - 1
    lo["m"]["qepos"] = lo["m"]["qspos"] + h - 1
    lo["m"]["matched_len"] = h
    lo["m"]["right_ed"] = 0
    lo["m"]["sclen_right"] = 0
    hi["m"]["spos"] = hi["m"]["spos"] + h
    hi["m"]["qspos"] = hi["m"]["qspos"] + h
    hi["m"]["matched_len"] = hi["m"]["matched_len"] - h
    hi["m"]["left_ed"] = 0
    hi["m"]["middle_ed"] = 0
    hi["m"]["sclen_left"] = 0
    p = r1p[idx]
    return dict(src=idx.astype(np.int32),
                r1_rows=np.where(p[:, None], rec["remain_rows"][idx], rec["fullmap_rows"][idx]),
                r2_rows=np.where(p[:, None], rec["fullmap_rows"][idx], rec["remain_rows"][idx]),
                l1=np.where(p, rec["remain_len"][idx], rec["fullmap_len"][idx]).astype(np.int32),
                l2=np.where(p, rec["fullmap_len"][idx], rec["remain_len"][idx]).astype(np.int32),
                fullmap_sel=np.where(p, 2, 1).astype(np.uint8),
                r1_1=np.where(p, part[idx], lo), r1_2=np.where(p, pm[idx], hi), r2_1=np.where(p, lo, part[idx]), r2_2=np.where(p, hi, pm[idx]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="T0")
    ap.add_argument("--contig-size", type=int, default=250_000)
    ap.add_argument("--pairs", type=int, default=0)
    ap.add_argument("--kmer", type=int, default=20)
    ap.add_argument("--scan-level", type=int, default=0)
    ap.add_argument("--max-ed", type=int, default=4)
    ap.add_argument("--band", type=int, default=3)
    ap.add_argument("--se", action="store_true")
    ap.add_argument("--compact", action="store_true", help="build the reference index with -m (counts only)")
    ap.add_argument("--lookups", type=int, default=0, help="random positions per contig for the GTFParser look-up dump")
    ap.add_argument("--extend", action="store_true", help="extension-stage dump: extend_both_mates / extend_chain_both_sides on chain pairs")
    ap.add_argument("--circ", type=int, default=0, help="stage-2 chaining problems per contig (ProcessCirc::chaining)")
    ap.add_argument("--circ-single", action="store_true", help="stage 2: record every check_split_map call of the single-split path")
    ap.add_argument("--align", type=int, default=0, help="random alignment problems per kind")
    ap.add_argument("--no-stage-dumps", action="store_true", help="only process_read results (large workloads)")
    ap.add_argument("--out", required=True)
    a = ap.parse_args()

    from circminer_b200 import workload
    from oracle.refharness import REF_BIN, RefHarness, default_params, fresh_mreads
    import subprocess

    tmp = tempfile.mkdtemp(prefix="cm_refdump_")
    w = workload.prepare(a.workload, tmp, contig_size=a.contig_size, n_pairs=a.pairs or None)
    p = default_params(kmer=a.kmer, scan_level=a.scan_level, max_ed=a.max_ed, band_width=a.band)
    h = RefHarness(p)
    if a.kmer != 20:
        # the harness builds with checkSumLength from set_params (commandline_parser.cpp:247)
        pass
    h.build_index(w["fa"], contig_size=a.contig_size, compact=a.compact)
    h.open(w["fa"], w["gtf"], os.path.join(tmp, "ref"))
    s1, s2 = w["s1"], w["s2"]
    n, L = s1.shape
    lens = np.full(n, L, np.uint16)
    out = dict(n_pairs=n, read_len=L, kmer=h.kmer, reads_crc=zlib.crc32(s1.tobytes() + s2.tobytes()),
               chr_names=np.array(h.chr_names))
    mr = fresh_mreads(n, a.max_ed)
    more, rounds = 1, 0
    while more:
        more = h.next_contig()
        c = h.contig
        rounds += 1
        if not a.no_stage_dumps:
            for k, v in h.annotation().items():     # the reference's own tables for this contig
                out[f"c{c}_annot_{k}"] = v
            for nm, s in (("r1", s1), ("r2", s2)):
                sd = h.seed(s, lens)
                for k, v in sd.items():
                    out[f"c{c}_seed_{nm}_{k}"] = v
                cd = h.chain(s, lens)
                for k, v in cd.items():
                    out[f"c{c}_chain_{nm}_{k}"] = v
        if a.lookups:
            # GTFParser look-ups on random positions and on every position within 3 bases of an interval border
            ann = h.annotation()
            rng = np.random.default_rng(1000 + c)
            ref_len = h.index_pointers()[4]
            edges = np.concatenate([ann["iv_spos"][:ann["n_iv"]].astype(np.int64), ann["iv_epos"][:ann["n_iv"]].astype(np.int64)])
            edges = edges[edges < ref_len]
            near = (edges[:, None] + np.arange(-3, 4)[None, :]).ravel()
            pos = np.concatenate([rng.integers(1, ref_len - 400, size=a.lookups), near])
            pos = np.unique(pos[(pos >= 1) & (pos < ref_len - 400)]).astype(np.uint32)
            out[f"c{c}_lookup_pos"] = pos
            for tag, (ml, rl) in (("a", (h.kmer, 130)), ("b", (h.kmer, 30))):
                for k, v in h.annot_lookup(pos, ml, rl).items():
                    out[f"c{c}_lookup_{tag}_{k}"] = v
        if a.extend:
            for k, v in extend_problems(h, s1, s2, lens).items():
                out[f"c{c}_ext_{k}"] = v
        if a.circ:
            prob = circ_problems(h, L, a.circ, 4000 + c)
            for k, v in prob.items():
                out[f"c{c}_circ_p_{k}"] = v
            cres = h.circ_chain(prob)
            for k, v in cres.items():
                out[f"c{c}_circ_r_{k}"] = v
            ex = circ_exact_problems(h, prob, cres, 5000 + c)
            for k, v in ex.items():
                out[f"c{c}_circx_p_{k}"] = v
            for k, v in h.circ_exact(ex).items():
                out[f"c{c}_circx_r_{k}"] = v
        if a.circ_single:
            # stage 1 on this contig first (fresh MatchedReads: every back-splice pair of this contig is a candidate), then the
            # single-split path of stage 2 with every check_split_map call recorded
            mr1, _ = h.process_pe(s1, s2, lens, lens, fresh_mreads(n, a.max_ed))
            rec = h.circ_single(s1, s2, lens, lens, mr1)
            out[f"c{c}_circs_mr_chr"] = mr1
            import time
            t0 = time.time()
            real = h.circ_single_real(s1, s2, lens, lens, mr1)   # the unmodified call_circ on the same pairs
            out[f"c{c}_circreal_seconds"] = time.time() - t0
            for k, v in real.items():
                out[f"c{c}_circreal_{k}"] = v
            t0 = time.time()
            real2 = h.circ_call_real(s1, s2, lens, lens, mr1, 4)   # the unmodified call_circ on the double-split (CHI2BSJ) pairs
            out[f"c{c}_circreal2_seconds"] = time.time() - t0
            for k, v in real2.items():
                out[f"c{c}_circreal2_{k}"] = v
            for k, v in rec.items():
                out[f"c{c}_circs_{k}"] = v
            if len(rec["ret"]):
                # the same calls once more on the recorded inputs alone (check_split_map as a function of its arguments), timed
                import time
                t0 = time.time()
                again = h.circ_check(rec)
                out[f"c{c}_circs_check_seconds"] = time.time() - t0
                same = (again["ret"] == rec["ret"]).all() and (again["cr"] == rec["cr"]).all() and (again["r1"] == rec["out_r1"]).all() \
                    and (again["r2"] == rec["out_r2"]).all() and (again["pm"] == rec["out_pm"]).all()
                if not same:
                    raise SystemExit("check_split_map on the recorded inputs does not reproduce the recorded call")
                # the four-mate function on rows made from these calls (its single-BSJ arms, which the reads themselves rarely reach)
                p4 = circ_check4_problems(rec, 6000 + c)
                if len(p4["src"]):
                    r4 = h.circ_check4(p4)
                    for k, v in p4.items():
                        out[f"c{c}_circ4_{k}"] = v
                    out[f"c{c}_circ4_ret"] = r4["ret"]
                    out[f"c{c}_circ4_cr"] = r4["cr"]
        if a.se:
            out[f"c{c}_se_state"] = h.process_se(s1, lens)
        mr, st = h.process_pe(s1, s2, lens, lens, mr)
        out[f"c{c}_pe_mread"] = mr.copy()
        out[f"c{c}_pe_state"] = st
    out["n_contigs"] = rounds
    if a.align:
        probs = align_problems(77, a.align)
        for kind, (s, ns, t, nt, aux) in probs.items():
            r = h.align(kind, s, ns, t, nt, aux)
            for k, v in r.items():
                out[f"align{kind}_{k}"] = v
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    np.savez_compressed(a.out, **out)
    print("wrote", a.out, "contigs", rounds, "pairs", n)
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
