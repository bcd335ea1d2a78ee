// Stage 2 (ProcessCirc, src/process_circ.cpp), the split-read re-seeding and chaining step:
//   RegionalHashTable (src/hash_table.cpp:24-112)   a 4^ws-entry table of the ws-mers of ONE gene, positions relative to
//                                                   the gene start, rebuilt for every gene a candidate read overlaps
//   ProcessCirc::chaining (process_circ.cpp:678-737) look up every `step`-th ws-mer of the unmapped part of the read,
//                                                   chain_seeds_sorted_kbest2, keep the chains that miss the fewest k-mers
// Here the table is not per gene.  One index per contig lists every ws-mer position sorted by (hash, position)
// (CircIndex, built once on the device); the hits of a ws-mer inside a gene are then a contiguous run of that list,
// found by two bisections, and are chained in place -- the same ascending positions RegionalHashTable::add_loc collects.
#pragma once
#include "cm_chain.h"
#include "cm_align.h"
#include "cm_extend.h"
#include "cm_filter.h"

#define CM_CIRC_MAXHIT 1000   // MAXHIT, src/hash_table.cpp:6: a ws-mer with more hits in the gene has none

struct CircIndex {
    const uint32_t *pos;   // [n] 1-based contig positions of the ws-mers without N, sorted by (hash, position)
    const uint32_t *off;   // [4^ws + 1] run of every hash value
    int ws;
};

// RegionalHashTable::hash_val (hash_table.cpp:95-105): -1 when a base is not ACGT (either case)
CM_HD int cm_circ_hash(const uint8_t *s, int ws) {
    int val = 0;
    for (int i = 0; i < ws; i++) {
        const uint32_t u = s[i] & 0xDFu;
        int c;
        if (u == 'A') c = 0;
        else if (u == 'C') c = 1;
        else if (u == 'G') c = 2;
        else if (u == 'T') c = 3;
        else return -1;
        val = (val << 2) | c;
    }
    return val;
}

// first index in [lo, hi) with a[idx] >= x
CM_HD uint32_t cm_circ_lower(const uint32_t *a, uint32_t lo, uint32_t hi, uint32_t x) {
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (a[mid] < x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// The rows ProcessCirc::chaining builds (process_circ.cpp:685-708): one per `step`-th ws-mer of seq[qspos-1 .. qepos-1]
// that has no N.  row_pos[l] / row_cnt[l] = run of CircIndex::pos, qp[l] = read offset.  Returns kmer_cnt (= l there).
CM_HD int cm_circ_rows(const CircIndex &ix, const cm_params &p, const uint8_t *seq, uint32_t qspos, uint32_t qepos, uint32_t gene_start,
                       uint32_t gene_end, int step, int max_rows, uint32_t *row_pos, uint32_t *row_cnt, int32_t *qp, uint32_t &cells) {
    cells = 0;
    const int ws = ix.ws;
    if (qepos < qspos || (int) (qepos - qspos + 1) < ws) return 0;
    // ws-mers that lie inside the gene start at gene_start .. gene_end - ws + 1 (create_table(seq, 0, len): i <= len - ws)
    const bool any = gene_end >= gene_start && gene_end - gene_start + 1 >= (uint32_t) ws;
    const uint32_t p_lo = gene_start, p_hi = any ? gene_end - (uint32_t) ws + 1 : 0;
    int l = 0;
    for (uint32_t i = qspos - 1; i + ws <= qepos && l < max_rows; i += step) {
        const int hv = cm_circ_hash(seq + i, ws);
        if (hv < 0) continue;
        uint32_t lo = 0, cnt = 0;
        if (any) {
            const uint32_t a = ix.off[hv], b = ix.off[hv + 1];
            lo = cm_circ_lower(ix.pos, a, b, p_lo);
            const uint32_t hi = cm_circ_lower(ix.pos, lo, b, p_hi + 1);
            cnt = hi - lo;
        }
        if (cnt > CM_CIRC_MAXHIT) cnt = 0;                 // hash_table.cpp:75-78
        if (cnt > (uint32_t) p.seed_lim) cnt = 0;          // process_circ.cpp:706-707
        row_pos[l] = lo;
        row_cnt[l] = cnt;
        qp[l] = (int32_t) i;
        cells += cnt;
        l++;
    }
    return l;
}

// chain_seeds_sorted_kbest2's trim of the empty rows at the back (chain.cpp:351-356) and the row offsets of the dp arena
CM_HD int cm_circ_trim(int kmer_cnt, const uint32_t *row_cnt, uint32_t *row_off) {
    int T = kmer_cnt;
    while (T >= 1 && row_cnt[T - 1] == 0) T--;
    uint32_t c = 0;
    for (int t = 0; t < T; t++) { row_off[t] = c; c += row_cnt[t]; }
    return T;
}

// The cut after the chaining (process_circ.cpp:716-736): chains are in descending score order; the list ends at the first
// chain that misses more k-mers than the one before it.
CM_HD int cm_circ_cut(int kmer_cnt, int count, const uint32_t *chain_len) {
    int least = CM_INF;
    for (int j = 0; j < count; j++) {
        const int missing = kmer_cnt - (int) chain_len[j];
        if (missing > least) return j;
        least = missing;
    }
    return count;
}

// Apply the cut to a finished chain list and clear what lies behind it (the caller's arrays start zeroed, so the
// output is a function of the problem alone).
CM_HD void cm_circ_finish(int kmer_cnt, ChainOut &out) {
    const int before = *out.count, after = cm_circ_cut(kmer_cnt, before, out.len);
    for (int j = after; j < before; j++) {
        for (uint32_t f = 0; f < out.len[j]; f++) out.frags[j * out.F + f] = FragDev{0, 0};
        out.len[j] = 0;
        out.score[j] = 0.f;
    }
    *out.count = after;
}

// ---- ProcessCirc::find_exact_coord (process_circ.cpp:739-789) ---------------------------------------------------------
// same_transcript(segments, 3, common_tid) (src/utils.cpp:447-494): the first of the eight start / end combinations of the
// three mates whose exon intervals all exist and share a transcript.  The three-interval overload (:356-376) intersects
// the common transcripts of the first two intervals with the transcripts of the FIRST interval again -- the third is only
// tested for NULL -- so common_tid is the two-interval set of (seg0, seg1): TidSet{iv0, iv1}.
// The start look-ups are made on the mates themselves (find_exact_coord calls overlap_to_spos on them and they keep the result);
// the end look-ups happen on the copies same_transcript works on (segments.push_back) and are not kept.
CM_HD bool cm_circ_same_transcript3(const AnnotDev &a, MateD &m0, MateD &m1, MateD &m2, TidSet &ts) {
    cm_overlap_to_spos(a, m0); cm_overlap_to_spos(a, m1); cm_overlap_to_spos(a, m2);
    MateD c0m = m0, c1m = m1, c2m = m2;
    cm_overlap_to_epos(a, c0m); cm_overlap_to_epos(a, c1m); cm_overlap_to_epos(a, c2m);
    for (int c0 = 0; c0 < 2; c0++)
        for (int c1 = 0; c1 < 2; c1++)
            for (int c2 = 0; c2 < 2; c2++) {
                const int i0 = c0 ? c0m.exons_epos : c0m.exons_spos, i1 = c1 ? c1m.exons_epos : c1m.exons_spos,
                          i2 = c2 ? c2m.exons_epos : c2m.exons_spos;
                if (i0 < 0 || i1 < 0 || i2 < 0) continue;
                if (cm_same_transcript(a, i0, i1)) { ts.iv_a = i0; ts.iv_b = i1; return true; }
            }
    ts.iv_a = ts.iv_b = -1;
    return false;
}

// x.v->p.kmer must be the fragment length of the chain (window_size); r1 / r2 need spos and epos
// (contig coordinates) only; pm starts as MatchedMate().  qspos is 1-based as in the reference call.
CM_HDN bool cm_circ_exact(ExtCtx &x, MateD &r1, MateD &r2, MateD &pm, int dir, uint32_t qspos, const uint8_t *rseq, int rlen, int whole_len,
                          const ChainView &bc) {
    const cm_params &p = x.v->p;
    const AnnotDev &a = x.v->a;
    // set_mm (process_circ.cpp:1713-1752) = MatchedMate::set (src/common.cpp:154-161)
    pm.spos = bc.frags[0].rpos;
    pm.epos = bc.frags[bc.chain_len - 1].rpos + (uint32_t) p.kmer - 1;
    pm.qspos = qspos;
    pm.qepos = qspos + (uint32_t) rlen - 1;
    pm.matched_len = (pm.qepos + 1 >= pm.qspos) ? (pm.qepos - pm.qspos + 1) : 0;
    pm.dir = dir;
    --qspos;
    TidSet ts;
    if (!cm_circ_same_transcript3(a, r1, r2, pm, ts)) return false;
    pm.middle_ed = cm_calc_middle_ed(x, bc, p.max_ed, rseq, rlen);
    if (pm.middle_ed > p.max_ed) return false;
    pm.is_concord = false;
    int err = pm.middle_ed;
    pm.matched_len = (uint32_t) rlen;
    const bool lok = cm_extend_chain_left<true>(x, ts, bc, rseq + qspos, (int32_t) qspos, 0u /* MINLB */, pm, err);
    const bool rok = cm_extend_chain_right<true>(x, ts, bc, rseq, qspos == 0 ? rlen : whole_len, CM_MAXUB, pm, err);
    cm_update_match_mate_info(lok, rok, err, pm, p);
    return pm.type == CM_CONCRD;
}
