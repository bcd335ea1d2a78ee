// Stage 2, the single-split path end to end: the control flow of ProcessCirc::call_circ_single_split
// (src/process_circ.cpp:360-483) for a batch of read pairs, as rounds over the three batch entries
//   chaining (per pair and overlapping gene)  ->  find_exact_coord (chain by chain)  ->  check_split_map (for the CONCRD ones)
// Host code only (no CUDA): the same loop drives the device entries in libcm_gpu.so and the host build of the device functions
// in the CPU tests, through the CircEngine function table.
//
// The chains of one read cannot go through the entries all at once: the two mates of the pair keep what the previous
// check_split_map left on them (a trim by final_check, cached exon look-ups), remain_seq may have been pointed at the other
// read, and the loop ends at the first FR / RF or CR.  So round (g, j) handles chain j of gene g for every pair that is
// still going.
#pragma once
#include <cstdint>
#include <algorithm>
#include <cstring>
#include <vector>

#include "../../include/cm_gpu.h"

#define CM_CIRC_TOPCHAIN 10   // TOPCHAIN, src/process_circ.cpp:19
#define CM_CIRC_T_CHIBSJ 3    // CHIBSJ, src/common.h:62
#define CM_CIRC_T_CONCRD 0

struct CircEngine {
    void *self;
    int contig;
    int (*chain)(void *self, int contig, uint32_t n, const char *seqs, uint32_t stride, const uint32_t *qspos, const uint32_t *qepos,
                 const uint32_t *gene_start, const uint32_t *gene_end, int ws, int step, int F, int32_t *count, float *score, uint32_t *clen,
                 uint32_t *rpos, int32_t *qpos);
    int (*exact)(void *self, int contig, uint32_t n, const char *seqs, uint32_t stride, cm_circ_mate *r1, cm_circ_mate *r2, const int32_t *dir,
                 const uint32_t *qspos, const int32_t *rlen, const int32_t *whole_len, uint32_t F, const uint32_t *clen, const uint32_t *rpos,
                 const int32_t *qpos, int ws, cm_circ_mate *pm, int32_t *ret);
    int (*check)(void *self, int contig, uint32_t n, const char *remain, const char *fullmap, uint32_t stride, const int32_t *rlen,
                 const int32_t *flen, const uint8_t *r1_partial, cm_circ_mate *r1, cm_circ_mate *r2, cm_circ_mate *pm, int ws, int32_t *ret,
                 cm_circ_res *cr, uint8_t *remain_is_fullmap);
    // the double-split path (may be null for callers of circ_orch_single only): one (chain of R1's part, chain of R2's part) per row,
    // and the four-mate check_split_map
    int (*dbl)(void *self, int contig, uint32_t n, const char *r1seq, const char *r2seq, uint32_t stride, cm_circ_mate *r1, cm_circ_mate *r2,
               const uint32_t *qs1, const uint32_t *qe1, const uint32_t *qs2, const uint32_t *qe2, uint32_t F, const uint32_t *clen1,
               const uint32_t *rpos1, const int32_t *qpos1, const uint32_t *clen2, const uint32_t *rpos2, const int32_t *qpos2, int ws,
               cm_circ_mate *p1, cm_circ_mate *p2, int32_t *code);
    int (*check4)(void *self, int contig, uint32_t n, const char *r1seq, const char *r2seq, uint32_t stride, const int32_t *l1, const int32_t *l2,
                  const cm_circ_mate *r1, const cm_circ_mate *r2, const cm_circ_mate *p1, const cm_circ_mate *p2, int ws, int32_t *ret,
                  cm_circ_res *cr);
};

// MatchedMate(const MatchedRead &, r1_2, rlen, partial) (src/common.cpp:192-235); the three flags it leaves unset are 0
static inline void circ_orch_mate(const cm_mread &mr, int r1_2, int rlen, bool partial, cm_circ_mate &o) {
    memset(&o, 0, sizeof(o));
    cm_mate &m = o.m;
    m.type = mr.type;
    if (r1_2 == 1) {
        m.spos = mr.spos_r1; m.epos = mr.epos_r1; m.qspos = mr.qspos_r1; m.qepos = mr.qepos_r1;
        m.middle_ed = mr.ed_r1; m.matched_len = mr.mlen_r1; m.dir = mr.r1_forward ? 1 : -1;
    } else {
        m.spos = mr.spos_r2; m.epos = mr.epos_r2; m.qspos = mr.qspos_r2; m.qepos = mr.qepos_r2;
        m.middle_ed = mr.ed_r2; m.matched_len = mr.mlen_r2; m.dir = mr.r2_forward ? 1 : -1;
    }
    if (partial) {
        if ((m.qspos - 1) > ((uint32_t) rlen - m.qepos)) { m.sclen_left = 0; m.sclen_right = rlen - (int) m.qepos; }
        else { m.sclen_left = (int) m.qspos - 1; m.sclen_right = 0; }
    } else {
        m.sclen_left = (int) m.qspos - 1;
        m.sclen_right = rlen - (int) m.qepos;
    }
    o.exons_spos = o.exons_epos = -1;
    o.exon_ind_spos = o.exon_ind_epos = -1;
}

static inline void circ_orch_revcomp(const char *s, int len, char *rc) {
    for (int i = 0; i < len; i++) {
        char c = s[i];
        char r = 'N';
        if (c == 'A' || c == 'a') r = 'T';
        else if (c == 'C' || c == 'c') r = 'G';
        else if (c == 'G' || c == 'g') r = 'C';
        else if (c == 'T' || c == 't') r = 'A';
        rc[len - 1 - i] = r;
    }
}

struct CircSingleIn {
    uint32_t n;
    const char *seqs1, *seqs2;     // rows of `stride` bytes, as read
    uint32_t stride;
    const uint16_t *len1, *len2;
    const cm_mread *mr;            // stage 1's MatchedRead, positions in CONTIG coordinates (after chrloc2conloc)
    const uint32_t *gene_off;      // [n + 1]: genes of gtf_parser.get_gene_overlap(mm_r1.spos, false)->seg_list, in that order
    const uint32_t *gene_start, *gene_end;
    int ws, step, max_read_len, max_chain_len;   // max_chain_len: rows of the engine's chain lists (cm_params::max_chain_len)
    int also_type;                 // a second MatchedRead type to take besides CHIBSJ (call_circ_double_split calls
                                   // call_circ_single_split on CHI2BSJ pairs); -1: none
};

struct CircAttempt {               // one check_split_map call
    uint32_t pair, gene, chain;
    uint8_t r1_partial;
    cm_circ_mate in_r1, in_r2, in_pm, out_r1, out_r2, out_pm;
    int32_t ret;
    cm_circ_res cr;
};

// Runs the batch.  attempts: every check_split_map call in (pair, gene, chain) order -- the order the reference makes them in
// when it goes through the pairs one after the other.  best_type[i]: what call_circ_single_split ends with for pair i (NF 40
// when nothing was found, FR / RF, CR / NCR / MCR), -1 when the pair is not a single-split candidate of this contig;
// best_cr[i] is meaningful for CR / NCR / MCR (the CircRes that is pushed to circ_res).  Returns 0 or the engine's error code.
static inline int circ_orch_single(const CircEngine &E, const CircSingleIn &in, std::vector<CircAttempt> &attempts, int32_t *best_type,
                                   cm_circ_res *best_cr) {
    const uint32_t n = in.n, stride = in.stride;
    const int ws = in.ws, F = (in.max_read_len - ws) / in.step + 1;
    attempts.clear();
    struct Pair {
        bool going, r1_partial, remain_is_full;
        uint32_t qspos, qepos;
        int dir, rlen, whole_len, plen_remain, plen_full, best;
        cm_circ_mate r1, r2;
    };
    std::vector<Pair> P(n);
    // both orientations of the two reads of every candidate: remain / fullmap rows
    std::vector<char> remain((size_t) n * stride, 0), fullmap((size_t) n * stride, 0);
    uint32_t max_genes = 0;
    for (uint32_t i = 0; i < n; i++) {
        Pair &p = P[i];
        p.going = false;
        best_type[i] = -1;
        if (best_cr) memset(&best_cr[i], 0, sizeof(cm_circ_res));
        const cm_mread &mr = in.mr[i];
        if (mr.type != CM_CIRC_T_CHIBSJ && mr.type != in.also_type) continue;
        const int l1 = in.len1[i], l2 = in.len2[i];
        p.r1_partial = mr.mlen_r1 < mr.mlen_r2;
        const char *s1 = in.seqs1 + (size_t) i * stride, *s2 = in.seqs2 + (size_t) i * stride;
        char *rem = &remain[(size_t) i * stride], *ful = &fullmap[(size_t) i * stride];
        auto put = [&](char *dst, const char *src, int len, bool fwd) { if (fwd) memcpy(dst, src, len); else circ_orch_revcomp(src, len, dst); };
        if (p.r1_partial) { put(rem, s1, l1, mr.r1_forward); put(ful, s2, l2, mr.r2_forward); p.plen_remain = l1; p.plen_full = l2; }
        else { put(rem, s2, l2, mr.r2_forward); put(ful, s1, l1, mr.r1_forward); p.plen_remain = l2; p.plen_full = l1; }
        circ_orch_mate(mr, 1, l1, p.r1_partial, p.r1);
        circ_orch_mate(mr, 2, l2, !p.r1_partial, p.r2);
        const cm_mate &pm = p.r1_partial ? p.r1.m : p.r2.m;
        const uint32_t plen = (uint32_t) p.plen_remain;
        const bool right_side = (pm.qspos - 1) > (plen - pm.qepos);
        p.qspos = right_side ? 1 : pm.qepos + 1;
        p.qepos = right_side ? pm.qspos - 1 : plen;
        p.whole_len = (int) plen;
        p.rlen = (int) (p.qepos - p.qspos + 1);
        p.dir = (p.r1_partial ? mr.r1_forward : mr.r2_forward) ? 1 : -1;
        p.best = 40;   // NF
        p.remain_is_full = false;
        best_type[i] = 40;
        if (p.qepos < p.qspos || p.rlen < ws) continue;
        const uint32_t ng = in.gene_off[i + 1] - in.gene_off[i];
        if (ng == 0) continue;
        p.going = true;
        if (ng > max_genes) max_genes = ng;
    }
    std::vector<uint32_t> idx, sub, csub;
    for (uint32_t g = 0; g < max_genes; g++) {
        idx.clear();
        for (uint32_t i = 0; i < n; i++)
            if (P[i].going && g < in.gene_off[i + 1] - in.gene_off[i]) idx.push_back(i);
        if (idx.empty()) break;
        const uint32_t m = (uint32_t) idx.size();
        // chaining of (pair, gene g) on the read remain_seq points at now
        std::vector<char> cs((size_t) m * stride);
        std::vector<uint32_t> qs(m), qe(m), gs(m), ge(m);
        for (uint32_t k = 0; k < m; k++) {
            const uint32_t i = idx[k];
            memcpy(&cs[(size_t) k * stride], P[i].remain_is_full ? &fullmap[(size_t) i * stride] : &remain[(size_t) i * stride], stride);
            qs[k] = P[i].qspos; qe[k] = P[i].qepos;
            gs[k] = in.gene_start[in.gene_off[i] + g]; ge[k] = in.gene_end[in.gene_off[i] + g];
        }
        std::vector<int32_t> count(m);
        const size_t CH = (size_t) in.max_chain_len;
        std::vector<float> score(m * CH);
        std::vector<uint32_t> clen(m * CH), rpos(m * CH * F);
        std::vector<int32_t> qpos(m * CH * F);
        int r = E.chain(E.self, E.contig, m, cs.data(), stride, qs.data(), qe.data(), gs.data(), ge.data(), ws, in.step, F, count.data(), score.data(),
                        clen.data(), rpos.data(), qpos.data());
        if (r != 0) return r;
        for (int j = 0; j < CM_CIRC_TOPCHAIN; j++) {
            sub.clear();
            for (uint32_t k = 0; k < m; k++)
                if (P[idx[k]].going && j < (count[k] < CM_CIRC_TOPCHAIN ? count[k] : CM_CIRC_TOPCHAIN)) sub.push_back(k);
            if (sub.empty()) break;
            const uint32_t q = (uint32_t) sub.size();
            std::vector<char> xs((size_t) q * stride), xf((size_t) q * stride);
            std::vector<cm_circ_mate> r1(q), r2(q), pm(q);
            std::vector<int32_t> dir(q), rlen(q), wlen(q), ret(q), xq(q * (size_t) F), flen(q), remlen(q);
            std::vector<uint32_t> xqs(q), xcl(q), xr(q * (size_t) F);
            std::vector<uint8_t> r1p(q);
            for (uint32_t t = 0; t < q; t++) {
                const uint32_t k = sub[t], i = idx[k];
                // remain_seq as it is now (a check_split_map of an earlier chain may have pointed it at the other read)
                memcpy(&xs[(size_t) t * stride], P[i].remain_is_full ? &fullmap[(size_t) i * stride] : &remain[(size_t) i * stride], stride);
                r1[t] = P[i].r1; r2[t] = P[i].r2;
                dir[t] = P[i].dir; xqs[t] = P[i].qspos; rlen[t] = P[i].rlen; wlen[t] = P[i].whole_len;
                xcl[t] = clen[k * CH + j];
                memcpy(&xr[(size_t) t * F], &rpos[(k * CH + j) * F], (size_t) F * 4);
                memcpy(&xq[(size_t) t * F], &qpos[(k * CH + j) * F], (size_t) F * 4);
            }
            r = E.exact(E.self, E.contig, q, xs.data(), stride, r1.data(), r2.data(), dir.data(), xqs.data(), rlen.data(), wlen.data(), (uint32_t) F,
                        xcl.data(), xr.data(), xq.data(), ws, pm.data(), ret.data());
            if (r != 0) return r;
            for (uint32_t t = 0; t < q; t++) { P[idx[sub[t]]].r1 = r1[t]; P[idx[sub[t]]].r2 = r2[t]; }   // the look-ups find_exact_coord cached
            csub.clear();
            for (uint32_t t = 0; t < q; t++)
                if (pm[t].m.type == CM_CIRC_T_CONCRD) csub.push_back(t);
            if (csub.empty()) continue;
            const uint32_t c = (uint32_t) csub.size();
            std::vector<cm_circ_mate> c1(c), c2(c), cp(c);
            std::vector<int32_t> cret(c);
            std::vector<cm_circ_res> ccr(c);
            std::vector<uint8_t> sw(c);
            std::vector<char> crem((size_t) c * stride), cful((size_t) c * stride);
            std::vector<int32_t> crl(c), cfl(c);
            std::vector<uint8_t> cr1p(c);
            for (uint32_t u = 0; u < c; u++) {
                const uint32_t t = csub[u], i = idx[sub[t]];
                c1[u] = P[i].r1; c2[u] = P[i].r2; cp[u] = pm[t];
                // remain_seq / fullmap_seq as ProcessCirc holds them now; their lengths never change
                memcpy(&crem[(size_t) u * stride], &xs[(size_t) t * stride], stride);
                memcpy(&cful[(size_t) u * stride], &fullmap[(size_t) i * stride], stride);
                crl[u] = P[i].plen_remain; cfl[u] = P[i].plen_full;
                cr1p[u] = P[i].r1_partial;
            }
            std::vector<cm_circ_mate> i1 = c1, i2 = c2, ip = cp;
            r = E.check(E.self, E.contig, c, crem.data(), cful.data(), stride, crl.data(), cfl.data(), cr1p.data(), c1.data(), c2.data(), cp.data(), ws,
                        cret.data(), ccr.data(), sw.data());
            if (r != 0) return r;
            for (uint32_t u = 0; u < c; u++) {
                const uint32_t t = csub[u], i = idx[sub[t]];
                Pair &p = P[i];
                if (cret[u] < 0) { p.best = -2; p.going = false; continue; }   // a fixed-size list of the device code overflowed: the pair is the caller's
                CircAttempt a;
                a.pair = i; a.gene = g; a.chain = (uint32_t) j; a.r1_partial = p.r1_partial;
                a.in_r1 = i1[u]; a.in_r2 = i2[u]; a.in_pm = ip[u];
                a.out_r1 = c1[u]; a.out_r2 = c2[u]; a.out_pm = cp[u];
                a.ret = cret[u]; a.cr = ccr[u];
                attempts.push_back(a);
                p.r1 = c1[u]; p.r2 = c2[u];
                if (sw[u]) p.remain_is_full = true;
                const int type = cret[u];
                if (type < 20) { p.best = type; p.going = false; }                      // FR / RF: return
                else if (type <= 22 && type < p.best) {                                 // CR / NCR / MCR better than what we have
                    p.best = type;
                    if (best_cr) best_cr[i] = ccr[u];
                    if (type == 20) p.going = false;                                    // CR: push and return
                }
            }
        }
    }
    for (uint32_t i = 0; i < n; i++)
        if (best_type[i] != -1) best_type[i] = P[i].best;
    // (pair, gene, chain) order
    std::vector<CircAttempt> sorted;
    sorted.reserve(attempts.size());
    std::vector<uint32_t> order(attempts.size());
    for (size_t k = 0; k < order.size(); k++) order[k] = (uint32_t) k;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t x, uint32_t y) {
        const CircAttempt &a = attempts[x], &b = attempts[y];
        if (a.pair != b.pair) return a.pair < b.pair;
        if (a.gene != b.gene) return a.gene < b.gene;
        return a.chain < b.chain;
    });
    for (uint32_t k : order) sorted.push_back(attempts[k]);
    attempts.swap(sorted);
    return 0;
}

// ---- call_circ_double_split (src/process_circ.cpp:485-645) for a batch ---------------------------------------------------------
// Unlike the single-split loop, the attempts of one pair do not feed each other (the four-mate check_split_map works on copies),
// so every (gene, chain of R1's part, chain of R2's part) is evaluated in one batch and the function's rules -- where it calls
// call_circ_single_split on the same pair, where it stops -- are applied afterwards in its order.
struct CircEvent {
    uint32_t pair;
    uint8_t kind;                  // 0: call_circ_single_split(record1, record2) is called here; 1: a four-mate check_split_map call
    cm_circ_mate r1, r2, p1, p2;   // kind 1: mm_r1, mm_r2, r1_partial_mm, r2_partial_mm as print_split_mapping prints them
    int32_t ret;
    cm_circ_res cr;
};

// events: per pair in order.  push_type[i]: CR / NCR / MCR when the function pushes its own best_cr (push_cr[i], push_spos_of_r1[i] =
// mm_r1.spos for get_shift), else -1.  Pairs that are not CHI2BSJ get no events.
static inline int circ_orch_double(const CircEngine &E, const CircSingleIn &in, std::vector<CircEvent> &events, int32_t *push_type,
                                   cm_circ_res *push_cr) {
    const uint32_t n = in.n, stride = in.stride;
    const int ws = in.ws, F = (in.max_read_len - ws) / in.step + 1;
    const size_t CH = (size_t) in.max_chain_len;
    events.clear();
    struct Pair {
        bool cand, early_single;
        uint32_t qs1, qe1, qs2, qe2, first_prob, ng;
        int rl1, rl2, l1, l2;
        cm_circ_mate r1, r2;
    };
    std::vector<Pair> P(n);
    std::vector<char> o1((size_t) n * stride, 0), o2((size_t) n * stride, 0);   // the two reads as mapped (r?_remain_seq)
    // chaining problems: (pair, gene, mate)
    std::vector<uint32_t> pr_pair, pr_gene, pr_mate;
    for (uint32_t i = 0; i < n; i++) {
        Pair &p = P[i];
        p.cand = false;
        push_type[i] = -1;
        if (push_cr) memset(&push_cr[i], 0, sizeof(cm_circ_res));
        const cm_mread &mr = in.mr[i];
        if (mr.type != 4) continue;   // CHI2BSJ
        p.l1 = in.len1[i]; p.l2 = in.len2[i];
        const char *s1 = in.seqs1 + (size_t) i * stride, *s2 = in.seqs2 + (size_t) i * stride;
        if (mr.r1_forward) memcpy(&o1[(size_t) i * stride], s1, p.l1); else circ_orch_revcomp(s1, p.l1, &o1[(size_t) i * stride]);
        if (mr.r2_forward) memcpy(&o2[(size_t) i * stride], s2, p.l2); else circ_orch_revcomp(s2, p.l2, &o2[(size_t) i * stride]);
        const bool rs1 = (mr.qspos_r1 - 1) > ((uint32_t) p.l1 - mr.qepos_r1), rs2 = (mr.qspos_r2 - 1) > ((uint32_t) p.l2 - mr.qepos_r2);
        p.qs1 = rs1 ? 1 : mr.qepos_r1 + 1; p.qe1 = rs1 ? mr.qspos_r1 - 1 : (uint32_t) p.l1;
        p.qs2 = rs2 ? 1 : mr.qepos_r2 + 1; p.qe2 = rs2 ? mr.qspos_r2 - 1 : (uint32_t) p.l2;
        p.rl1 = (int) (p.qe1 - p.qs1 + 1); p.rl2 = (int) (p.qe2 - p.qs2 + 1);
        if (p.rl1 < ws && p.rl2 < ws) continue;                    // fully mapped: return
        p.early_single = p.rl1 < ws || p.rl2 < ws;                 // single split first -- and the function goes on
        p.cand = true;
        circ_orch_mate(mr, 1, p.l1, true, p.r1);
        circ_orch_mate(mr, 2, p.l2, true, p.r2);
        p.ng = in.gene_off[i + 1] - in.gene_off[i];
        p.first_prob = (uint32_t) pr_pair.size();
        for (uint32_t g = 0; g < p.ng; g++)
            for (uint32_t mt = 0; mt < 2; mt++) { pr_pair.push_back(i); pr_gene.push_back(g); pr_mate.push_back(mt); }
    }
    const uint32_t np = (uint32_t) pr_pair.size();
    std::vector<int32_t> count(np, 0);
    std::vector<uint32_t> clen(np * CH), rpos(np * CH * (size_t) F);
    std::vector<int32_t> qpos(np * CH * (size_t) F);
    if (np) {
        std::vector<char> cs((size_t) np * stride);
        std::vector<uint32_t> qs(np), qe(np), gs(np), ge(np);
        std::vector<float> score(np * CH);
        for (uint32_t k = 0; k < np; k++) {
            const uint32_t i = pr_pair[k];
            const Pair &p = P[i];
            memcpy(&cs[(size_t) k * stride], pr_mate[k] ? &o2[(size_t) i * stride] : &o1[(size_t) i * stride], stride);
            qs[k] = pr_mate[k] ? p.qs2 : p.qs1; qe[k] = pr_mate[k] ? p.qe2 : p.qe1;
            gs[k] = in.gene_start[in.gene_off[i] + pr_gene[k]]; ge[k] = in.gene_end[in.gene_off[i] + pr_gene[k]];
            if (qe[k] < qs[k]) { qs[k] = 1; qe[k] = 0; }            // nothing left of this mate: chaining gives no chain
        }
        int r = E.chain(E.self, E.contig, np, cs.data(), stride, qs.data(), qe.data(), gs.data(), ge.data(), ws, in.step, F, count.data(), score.data(),
                        clen.data(), rpos.data(), qpos.data());
        if (r != 0) return r;
    }
    // all chain pairs of the genes where both parts have chains
    struct Combo { uint32_t pair, gene, j, k, prob1; };
    std::vector<Combo> combos;
    for (uint32_t i = 0; i < n; i++) {
        const Pair &p = P[i];
        if (!p.cand) continue;
        for (uint32_t g = 0; g < p.ng; g++) {
            const uint32_t a = p.first_prob + 2 * g;
            const int c1 = count[a] < CM_CIRC_TOPCHAIN ? count[a] : CM_CIRC_TOPCHAIN, c2 = count[a + 1] < CM_CIRC_TOPCHAIN ? count[a + 1] : CM_CIRC_TOPCHAIN;
            if (count[a] <= 0 || count[a + 1] <= 0) continue;
            for (int j = 0; j < c1; j++)
                for (int k = 0; k < c2; k++) combos.push_back(Combo{i, g, (uint32_t) j, (uint32_t) k, a});
        }
    }
    const uint32_t nc = (uint32_t) combos.size();
    std::vector<cm_circ_mate> d1(nc), d2(nc), dp1(nc), dp2(nc);
    std::vector<int32_t> code(nc, 0), cret(nc, 0);
    std::vector<cm_circ_res> ccr(nc);
    std::vector<uint8_t> checked(nc, 0);
    if (nc) {
        std::vector<char> a1((size_t) nc * stride), a2((size_t) nc * stride);
        std::vector<uint32_t> qs1(nc), qe1(nc), qs2(nc), qe2(nc), cl1(nc), cl2(nc), rp1((size_t) nc * F), rp2((size_t) nc * F);
        std::vector<int32_t> qp1((size_t) nc * F), qp2((size_t) nc * F);
        for (uint32_t t = 0; t < nc; t++) {
            const Combo &c = combos[t];
            const Pair &p = P[c.pair];
            memcpy(&a1[(size_t) t * stride], &o1[(size_t) c.pair * stride], stride);
            memcpy(&a2[(size_t) t * stride], &o2[(size_t) c.pair * stride], stride);
            d1[t] = p.r1; d2[t] = p.r2;
            qs1[t] = p.qs1; qe1[t] = p.qe1; qs2[t] = p.qs2; qe2[t] = p.qe2;
            const size_t x1 = (size_t) c.prob1 * CH + c.j, x2 = ((size_t) c.prob1 + 1) * CH + c.k;
            cl1[t] = clen[x1]; cl2[t] = clen[x2];
            memcpy(&rp1[(size_t) t * F], &rpos[x1 * F], (size_t) F * 4); memcpy(&qp1[(size_t) t * F], &qpos[x1 * F], (size_t) F * 4);
            memcpy(&rp2[(size_t) t * F], &rpos[x2 * F], (size_t) F * 4); memcpy(&qp2[(size_t) t * F], &qpos[x2 * F], (size_t) F * 4);
        }
        int r = E.dbl(E.self, E.contig, nc, a1.data(), a2.data(), stride, d1.data(), d2.data(), qs1.data(), qe1.data(), qs2.data(), qe2.data(), (uint32_t) F,
                      cl1.data(), rp1.data(), qp1.data(), cl2.data(), rp2.data(), qp2.data(), ws, dp1.data(), dp2.data(), code.data());
        if (r != 0) return r;
        std::vector<uint32_t> sel;
        for (uint32_t t = 0; t < nc; t++)
            if (code[t] == 2 && dp1[t].m.type == CM_CIRC_T_CONCRD && dp2[t].m.type == CM_CIRC_T_CONCRD) sel.push_back(t);
        const uint32_t m = (uint32_t) sel.size();
        if (m) {
            std::vector<char> b1((size_t) m * stride), b2((size_t) m * stride);
            std::vector<int32_t> l1(m), l2(m), rr(m);
            std::vector<cm_circ_mate> m1(m), m2(m), q1(m), q2(m);
            std::vector<cm_circ_res> cr(m);
            for (uint32_t u = 0; u < m; u++) {
                const uint32_t t = sel[u];
                memcpy(&b1[(size_t) u * stride], &a1[(size_t) t * stride], stride);
                memcpy(&b2[(size_t) u * stride], &a2[(size_t) t * stride], stride);
                l1[u] = P[combos[t].pair].l1; l2[u] = P[combos[t].pair].l2;
                m1[u] = d1[t]; m2[u] = d2[t]; q1[u] = dp1[t]; q2[u] = dp2[t];
            }
            r = E.check4(E.self, E.contig, m, b1.data(), b2.data(), stride, l1.data(), l2.data(), m1.data(), m2.data(), q1.data(), q2.data(), ws, rr.data(),
                         cr.data());
            if (r != 0) return r;
            for (uint32_t u = 0; u < m; u++) { checked[sel[u]] = 1; cret[sel[u]] = rr[u]; ccr[sel[u]] = cr[u]; }   // rr < 0: a list overflowed
        }
    }
    // the function's own order and rules, pair by pair
    uint32_t t = 0;   // combos are in (pair, gene, j, k) order
    for (uint32_t i = 0; i < n; i++) {
        const Pair &p = P[i];
        if (!p.cand) continue;
        auto single_here = [&]() { CircEvent e; memset(&e, 0, sizeof(e)); e.pair = i; e.kind = 0; events.push_back(e); };
        if (p.early_single) single_here();
        int best = 40;
        bool returned = false;
        if (p.ng == 0) {                                            // gene not found: return
            while (t < nc && combos[t].pair == i) t++;
            continue;
        }
        for (uint32_t g = 0; g < p.ng && !returned; g++) {
            const uint32_t a = p.first_prob + 2 * g;
            if (count[a] <= 0 && count[a + 1] <= 0) continue;
            if (count[a] <= 0 || count[a + 1] <= 0) { single_here(); continue; }
            for (; t < nc && combos[t].pair == i && combos[t].gene == g; t++) {
                if (returned || !checked[t]) continue;
                if (cret[t] < 0) { push_type[i] = -2; returned = true; continue; }   // reached a row whose lists overflowed: the pair is the caller's
                CircEvent e;
                e.pair = i; e.kind = 1; e.r1 = d1[t]; e.r2 = d2[t]; e.p1 = dp1[t]; e.p2 = dp2[t]; e.ret = cret[t]; e.cr = ccr[t];
                events.push_back(e);
                const int type = cret[t];
                if (type < 20) { best = type; returned = true; }
                else if (type <= 22 && type < best) {
                    best = type;
                    if (push_cr) push_cr[i] = ccr[t];
                    if (type == 20) { push_type[i] = 20; returned = true; }
                }
            }
        }
        while (t < nc && combos[t].pair == i) t++;
        if (returned) continue;
        if (best >= 20 && best <= 22) push_type[i] = best;
        else single_here();
    }
    return 0;
}
