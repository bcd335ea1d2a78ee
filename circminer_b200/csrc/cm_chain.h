// Splice-aware seed chaining with k-best back-tracking:
// chain_seeds_sorted_kbest (src/chain.cpp:73-301) restated over compact seed rows.
// Only even GIMatchedKmer slots are ever filled (ll_step 2, src/match_read.cpp:275), so a
// "row" here is slot 2*t with qpos = t*kmer; positions are read in place from ContigDev::pos.
//
// The reference keeps a std::map<double, chain_cell_list> that receives one entry per strict
// improvement of a dp cell, at most maxChainLen per distinct score, in (ii desc, i asc) visit
// order.  Here every improvement is appended to a record list in the same visit order; the
// back-track then walks distinct scores in descending order and, inside one score, the first
// maxChainLen records in list order -- the same sequence the map iteration produces.
#pragma once
#include "cm_types.h"
#include "cm_annot.h"

struct ChainRec {
    double score;
    uint32_t loc;    // row << 16 | index in row
    uint32_t pad;
};

#define CM_NOPREV 0xFFFFFFFFu

#if defined(__CUDA_ARCH__)
#define CM_DADD(a, b) __dadd_rn((a), (b))
#define CM_DSUB(a, b) __dsub_rn((a), (b))
#define CM_DMUL(a, b) __dmul_rn((a), (b))
#else
#define CM_DADD(a, b) ((a) + (b))
#define CM_DSUB(a, b) ((a) - (b))
#define CM_DMUL(a, b) ((a) * (b))
#endif

// score_alpha / score_beta (chain.cpp:13-22): dp + 2e4*kmer - 0.1*|distr - distt|, evaluated left to right
CM_HD double cm_chain_score(double prev, int distr, int distt, int kmer) {
    int maxd = distr < distt ? distt : distr;
    int mind = distr < distt ? distr : distt;
    double alpha = CM_DMUL(2e4, (double) kmer);
    double beta = CM_DMUL(0.1, (double) (maxd - mind));
    return CM_DSUB(CM_DADD(prev, alpha), beta);
}

// Where the rows of a chaining problem come from.  Stage 1 (chain_seeds_sorted_kbest): row t sits at read offset t * kmer and
// its hits are ContigDev::pos entries.  Stage 2 (chain_seeds_sorted_kbest2, chain.cpp:310-539, called from
// ProcessCirc::chaining): explicit read offsets, window_size-mers, hit positions relative to the gene start.
struct ChainGeomMain {
    static constexpr int MAXROWS = CM_MAX_SEEDS;
    const uint32_t *pos;
    int kmer;
    CM_HD uint32_t at(uint32_t off) const { return pos[off]; }
    CM_HD int qpos(int t) const { return t * kmer; }
};
#define CM_CIRC_MAX_ROWS 128                 // (maxReadLength - window_size) / step + 1 rows at most
struct ChainGeomCirc {
    static constexpr int MAXROWS = CM_CIRC_MAX_ROWS;
    const uint32_t *pos;     // contig positions of the window_size-mers, ascending inside a row
    uint32_t sub;            // gene start: the reference chains gene-relative positions (RegionalHashTable::create_table(seq, 0, len))
    int kmer;
    const int32_t *qp;       // read offset of every row
    CM_HD uint32_t at(uint32_t off) const { return pos[off] - sub; }
    CM_HD int qpos(int t) const { return qp[t]; }
};

struct ChainOut {
    int32_t *count;        // -> chain count of this strand
    float *score;          // [C]
    uint32_t *len;         // [C]
    FragDev *frags;        // [C][F]
    int F;
};

// DP pass.  dp_score/dp_prev: this strand's cells, rows laid out back to back (row_off[t]).
// When recs != nullptr improvements are appended (up to rec_cap); the return value is the number of
// improvements seen (so a first call with recs == nullptr sizes the arena).  links counts candidates
// reaching chain.cpp:173.
template <class G>
CM_HD uint32_t cm_chain_dp_g(const G &g, const DevView &v, int seq_len, int T, const uint32_t *row_pos_off, const uint32_t *row_cnt,
                             const uint32_t *row_off, double *dp_score, uint32_t *dp_prev, ChainRec *recs, uint32_t rec_cap,
                             uint64_t *links) {
    const int kmer = g.kmer;
    uint32_t nrec = 0;
    uint64_t nlinks = 0;
    for (int t = 0; t < T; t++)
        for (uint32_t i = 0; i < row_cnt[t]; i++) {
            dp_score[row_off[t] + i] = (double) kmer;
            dp_prev[row_off[t] + i] = CM_NOPREV;
        }
    uint32_t lb_ind[G::MAXROWS];
    for (int t = T - 2; t >= 0; t--) {
        if (row_cnt[t] == 0) continue;
        uint32_t read_remain = (uint32_t) (seq_len - g.qpos(t) - kmer);
        for (int u = 0; u < T; u++) lb_ind[u] = 0;
        for (uint32_t i = 0; i < row_cnt[t]; i++) {
            uint32_t seg_start = g.at(row_pos_off[t] + i);
            uint32_t seg_end = seg_start + kmer - 1;
            uint32_t max_lpos_lim = CM_MAXUB, max_exon_end = 0;
            int ol_iv = -1;
            double best = dp_score[row_off[t] + i];
            uint32_t best_prev = dp_prev[row_off[t] + i];
            for (int u = t + 1; u < T; u++) {
                uint32_t cu = row_cnt[u];
                if (cu == 0 || lb_ind[u] >= cu) continue;
                const uint32_t pu0 = row_pos_off[u];
                auto pu = [&](uint32_t idx) { return g.at(pu0 + idx); };
                if (seg_start + (uint32_t) v.p.max_intron_len < pu(lb_ind[u])) continue;
                // first index >= lb_ind[u] with pu[idx] > seg_start (rows are ascending): gallop + bisect
                // instead of the reference's linear cursor walk (chain.cpp:153-156), same result
                if (pu(lb_ind[u]) <= seg_start) {
                    uint32_t lo = lb_ind[u], step = 1;
                    while (lo + step < cu && pu(lo + step) <= seg_start) { lo += step; step <<= 1; }
                    uint32_t hi = lo + step < cu ? lo + step : cu;   // pu(lo) <= seg_start < pu(hi) (or hi == cu)
                    while (hi - lo > 1) {
                        uint32_t mid = (lo + hi) >> 1;
                        if (pu(mid) <= seg_start) lo = mid; else hi = mid;
                    }
                    lb_ind[u] = hi;
                }
                if (lb_ind[u] >= cu) continue;
                if (max_lpos_lim == CM_MAXUB)
                    max_lpos_lim = cm_upper_bound(v.a, v.p, seg_start, (uint32_t) kmer, read_remain, max_exon_end, ol_iv);
                int distr = g.qpos(u) - g.qpos(t) - kmer;
                int read_dist = distr;
                uint32_t j = lb_ind[u];
                while (j < cu && pu(j) <= max_lpos_lim) {
                    nlinks++;
                    int genome_dist;
                    if (max_exon_end == 0 || (pu(j) + kmer - 1) <= max_exon_end) genome_dist = (int) (pu(j) - seg_end - 1);
                    else genome_dist = CM_INF;
                    int distt, trans_dist;
                    int gd = genome_dist - read_dist;
                    if ((gd < 0 ? -gd : gd) <= v.p.max_ed) {
                        distt = genome_dist;
                    } else if (cm_check_junction(v.a, v.p, seg_start, pu(j), ol_iv, kmer, read_dist, trans_dist)) {
                        distt = trans_dist;
                    } else {
                        j++;
                        continue;
                    }
                    double temp = cm_chain_score(dp_score[row_off[u] + j], distr, distt, kmer);
                    if (temp > best) {
                        best = temp;
                        best_prev = ((uint32_t) u << 16) | j;
                        if (recs && nrec < rec_cap) {
                            recs[nrec].score = temp;
                            recs[nrec].loc = ((uint32_t) t << 16) | i;
                            recs[nrec].pad = 0;
                        }
                        nrec++;
                    }
                    j++;
                }
            }
            dp_score[row_off[t] + i] = best;
            dp_prev[row_off[t] + i] = best_prev;
        }
    }
    if (links) *links += nlinks;
    return nrec;
}
CM_HDN uint32_t cm_chain_dp(const DevView &v, int seq_len, int T, const uint32_t *row_pos_off, const uint32_t *row_cnt,
                            const uint32_t *row_off, double *dp_score, uint32_t *dp_prev, ChainRec *recs, uint32_t rec_cap,
                            uint64_t *links) {
    const ChainGeomMain g{v.c.pos, v.p.kmer};
    return cm_chain_dp_g(g, v, seq_len, T, row_pos_off, row_cnt, row_off, dp_score, dp_prev, recs, rec_cap, links);
}

// Back-tracking (chain.cpp:234-300) over the recorded improvements.
// rpos_add: added to every reported position (the `shift` of chain_seeds_sorted_kbest2).  There the `repeats` set holds the
// SHIFTED positions of the emitted chains (chain.cpp:509-510) while the look-up key is the unshifted frags[].info (:494-495):
// kept as it is -- with a non-zero shift the filter only fires when a gene-relative position happens to equal a contig position.
template <class G>
CM_HD void cm_chain_backtrack_g(const G &g, const DevView &v, int T, const uint32_t *row_pos_off, const uint32_t *row_cnt,
                                const uint32_t *row_off, const double *dp_score, const uint32_t *dp_prev, const ChainRec *recs,
                                uint32_t nrec, ChainOut out, uint32_t rpos_add) {
    const int kmer = g.kmer;
    const uint32_t max_best = (uint32_t) v.p.max_chain_len;
    uint32_t best_count = 0;
    double best_score = (double) kmer;
    for (uint32_t r = 0; r < nrec; r++)
        if (recs[r].score > best_score) best_score = recs[r].score;
    double thresh = 0;
    bool first = true;
    uint64_t bloom[16];
    for (int q = 0; q < 16; q++) bloom[q] = 0;
    while (best_count < max_best) {
        // next distinct score below thresh
        bool found = false;
        double s = 0;
        for (uint32_t r = 0; r < nrec; r++) {
            double x = recs[r].score;
            if ((first || x < thresh) && (!found || x > s)) {
                s = x;
                found = true;
            }
        }
        if (!found) break;
        first = false;
        thresh = s;
        uint32_t in_group = 0;
        for (uint32_t r = 0; r < nrec && in_group < max_best; r++) {
            if (recs[r].score != s) continue;
            in_group++;
            if (best_count >= max_best) break;
            uint32_t loc = recs[r].loc;
            uint32_t spos = g.at(row_pos_off[loc >> 16] + (loc & 0xFFFF));
            if (s < best_score) {
                // `repeats`: rpos of every non-first fragment of the chains emitted so far.  A small Bloom
                // filter answers most queries; the exact scan runs only on a filter hit.
                bool rep = false;
                if ((bloom[(spos >> 6) & 15u] >> (spos & 63u)) & 1ull)
                for (uint32_t c = 0; c < best_count && !rep; c++)
                    for (uint32_t f = 1; f < out.len[c]; f++)
                        if (out.frags[c * out.F + f].rpos == spos) {
                            rep = true;
                            break;
                        }
                if (rep) continue;
            }
            uint32_t j = best_count++;
            uint32_t n = 0;
            uint32_t cur = loc;
            while (cur != CM_NOPREV) {
                uint32_t row = cur >> 16, ind = cur & 0xFFFF;
                out.frags[j * out.F + n].rpos = g.at(row_pos_off[row] + ind) + rpos_add;
                out.frags[j * out.F + n].qpos = (int32_t) g.qpos((int) row);
                if (n) { uint32_t rp = g.at(row_pos_off[row] + ind) + rpos_add; bloom[(rp >> 6) & 15u] |= 1ull << (rp & 63u); }
                n++;
                cur = dp_prev[row_off[row] + ind];
            }
            out.score[j] = (float) s;
            out.len[j] = n;
        }
    }
    if (best_count == 0) {   // chains of size 1 (chain.cpp:283-298)
        for (int t = T - 1; t >= 0; t--)
            for (uint32_t i = 0; i < row_cnt[t]; i++) {
                if (best_count >= max_best) break;
                uint32_t j = best_count++;
                out.frags[j * out.F].rpos = g.at(row_pos_off[t] + i) + rpos_add;
                out.frags[j * out.F].qpos = g.qpos(t);
                out.score[j] = (float) dp_score[row_off[t] + i];
                out.len[j] = 1;
            }
    }
    *out.count = (int32_t) best_count;
}
CM_HDN void cm_chain_backtrack(const DevView &v, int T, const uint32_t *row_pos_off, const uint32_t *row_cnt,
                               const uint32_t *row_off, const double *dp_score, const uint32_t *dp_prev, const ChainRec *recs,
                               uint32_t nrec, ChainOut out) {
    const ChainGeomMain g{v.c.pos, v.p.kmer};
    cm_chain_backtrack_g(g, v, T, row_pos_off, row_cnt, row_off, dp_score, dp_prev, recs, nrec, out, 0u);
}

// Rows of one strand from the seed stage output.  Returns T (rows after the trailing-empty trim,
// chain.cpp:112) and the number of dp cells; high_hits as FilterRead::get_best_chains counts them
// (filter.cpp:478-481).
CM_HD int cm_chain_rows(const DevView &v, const uint32_t *seed_off, const uint32_t *seed_cnt, uint32_t *row_pos_off,
                        uint32_t *row_cnt, uint32_t *row_off, uint32_t &cells, int &high_hits) {
    int T = 0;
    cells = 0;
    high_hits = 0;
    for (int t = 0; t < v.n_slots; t++) {
        uint32_t raw = seed_cnt[t];
        uint32_t kept = raw > (uint32_t) v.p.seed_lim ? 0 : raw;
        if (raw > (uint32_t) v.p.seed_lim) high_hits++;
        row_pos_off[t] = seed_off[t];
        row_cnt[t] = kept;
        row_off[t] = cells;
        cells += kept;
        if (kept) T = t + 1;
    }
    return T;
}
