// Warp-per-strand chaining for strands with many dp cells (repeat-rich reads): the same
// chain_seeds_sorted_kbest (src/chain.cpp:73-301) as cm_chain.h, with the hits of one row spread over
// the lanes of a warp.  Device only; cm_chain.h stays the single-thread statement (and the hostsim one).
//
// Why it is the same computation:
//  * Row t only reads dp cells of rows u > t, so all hits i of row t are independent of each other.
//  * The reference's cursor lb_ind[u] is stateful across i, but after hit i has been processed it always
//    equals L(i) = first index with pos_u > seg_start(i) (rows are ascending), or both are >= cnt_u:
//    it is either advanced to L(i), or left alone because pos_u[cursor] > seg_start(i) + maxIntronLen,
//    in which case the cursor already is L(i).  So lane i rebuilds the "stale" cursor the reference
//    tests first (chain.cpp:149) as L(i-1), taken from the neighbouring lane.
//  * score2chain receives one entry per strict improvement in (t desc, i asc, u asc, j asc) order; lanes
//    count their improvements, an exclusive warp scan gives each lane its slice of the record list, and
//    a second pass writes the records there -- the list is identical to the serial one.
#pragma once

template <class G>
__device__ __forceinline__ uint32_t cw_upper(const G &g, uint32_t base, uint32_t lo, uint32_t hi, uint32_t x) {
    // first index in [lo, hi) with row[idx] > x (hi when none); row[idx] = g.at(base + idx)
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (g.at(base + mid) <= x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// One hit (t, i): the u / j loops of chain.cpp:139-212.  WRITE = false counts the improvements,
// WRITE = true also stores them (up to `room` records) at out[0..].
template <bool WRITE, class G>
__device__ __forceinline__ uint32_t cw_eval(const G &g, const DevView &v, int T, int t, uint32_t i, uint32_t seg_start, uint32_t read_remain,
                                            const uint32_t *row_pos_off, const uint32_t *row_cnt, const uint32_t *row_off,
                                            const uint32_t *lbv, const uint32_t *stv, const double *dp_score, double &best,
                                            uint32_t &best_prev, ChainRec *out, uint32_t room, uint64_t &nlinks) {
    const int kmer = g.kmer;
    const uint32_t seg_end = seg_start + kmer - 1;
    uint32_t max_lpos_lim = CM_MAXUB, max_exon_end = 0, nimp = 0;
    int ol_iv = -1;
    for (int u = t + 1; u < T; u++) {
        const uint32_t cu = row_cnt[u];
        if (cu == 0 || stv[u] >= cu) continue;
        const uint32_t pu0 = row_pos_off[u];
        auto pu = [&](uint32_t idx) { return g.at(pu0 + idx); };
        if (seg_start + (uint32_t) v.p.max_intron_len < pu(stv[u])) continue;
        uint32_t j = lbv[u];
        if (j >= cu) continue;
        if (max_lpos_lim == CM_MAXUB)
            max_lpos_lim = cm_upper_bound(v.a, v.p, seg_start, (uint32_t) kmer, read_remain, max_exon_end, ol_iv);
        const int distr = g.qpos(u) - g.qpos(t) - kmer;
        while (j < cu && pu(j) <= max_lpos_lim) {
            nlinks++;
            int genome_dist;
            if (max_exon_end == 0 || (pu(j) + kmer - 1) <= max_exon_end) genome_dist = (int) (pu(j) - seg_end - 1);
            else genome_dist = CM_INF;
            int distt, trans_dist;
            int gd = genome_dist - distr;
            if ((gd < 0 ? -gd : gd) <= v.p.max_ed) {
                distt = genome_dist;
            } else if (cm_check_junction(v.a, v.p, seg_start, pu(j), ol_iv, kmer, distr, trans_dist)) {
                distt = trans_dist;
            } else {
                j++;
                continue;
            }
            double temp = cm_chain_score(dp_score[row_off[u] + j], distr, distt, kmer);
            if (temp > best) {
                best = temp;
                best_prev = ((uint32_t) u << 16) | j;
                if (WRITE && nimp < room) {
                    out[nimp].score = temp;
                    out[nimp].loc = ((uint32_t) t << 16) | i;
                    out[nimp].pad = 0;
                }
                nimp++;
            }
            j++;
        }
    }
    return nimp;
}

// DP pass of one strand by one warp.  Returns the number of improvement records (all lanes).
template <class G>
__device__ uint32_t cw_chain_dp(const G &g, const DevView &v, int seq_len, int T, const uint32_t *row_pos_off, const uint32_t *row_cnt,
                                const uint32_t *row_off, double *dp_score, uint32_t *dp_prev, ChainRec *recs, uint32_t rec_cap,
                                uint64_t &nlinks) {
    const unsigned FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u;
    const int kmer = g.kmer;
    for (int t = 0; t < T; t++)
        for (uint32_t i = lane; i < row_cnt[t]; i += 32) {
            dp_score[row_off[t] + i] = (double) kmer;
            dp_prev[row_off[t] + i] = CM_NOPREV;
        }
    uint32_t nrec = 0;
    uint32_t lbv[G::MAXROWS], stv[G::MAXROWS], carry[G::MAXROWS];
    for (int t = T - 2; t >= 0; t--) {
        const uint32_t ct = row_cnt[t];
        if (ct == 0) continue;
        __syncwarp();   // rows u > t are final and visible to every lane
        const uint32_t read_remain = (uint32_t) (seq_len - g.qpos(t) - kmer);
        for (int u = t + 1; u < T; u++) carry[u] = 0;
        for (uint32_t c0 = 0; c0 < ct; c0 += 32) {
            const uint32_t i = c0 + lane;
            const bool act = i < ct;
            const uint32_t seg_start = act ? g.at(row_pos_off[t] + i) : 0xFFFFFFFFu;
            // cursors of every later row, uniformly over the warp: L(i) and the stale L(i-1)
            for (int u = t + 1; u < T; u++) {
                const uint32_t cu = row_cnt[u];
                uint32_t lb = cu ? cw_upper(g, row_pos_off[u], 0, cu, seg_start) : 0;
                uint32_t st = __shfl_up_sync(FULL, lb, 1);
                if (lane == 0) st = carry[u];
                carry[u] = __shfl_sync(FULL, lb, 31);
                lbv[u] = lb;
                stv[u] = st;
            }
            double best = (double) kmer;
            uint32_t best_prev = CM_NOPREV;
            uint32_t nimp = 0;
            uint64_t nl = 0;
            if (act) nimp = cw_eval<false>(g, v, T, t, i, seg_start, read_remain, row_pos_off, row_cnt, row_off, lbv, stv, dp_score, best,
                                           best_prev, nullptr, 0, nl);
            nlinks += nl;
            // slice of the record list for this lane
            uint32_t incl = nimp;
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t y = __shfl_up_sync(FULL, incl, o);
                if (lane >= (uint32_t) o) incl += y;
            }
            const uint32_t excl = incl - nimp, total = __shfl_sync(FULL, incl, 31);
            if (act && nimp && recs) {
                const uint32_t at = nrec + excl;
                if (at < rec_cap) {
                    double b2 = (double) kmer;
                    uint32_t p2 = CM_NOPREV;
                    uint64_t nl2 = 0;
                    cw_eval<true>(g, v, T, t, i, seg_start, read_remain, row_pos_off, row_cnt, row_off, lbv, stv, dp_score, b2, p2,
                                  recs + at, rec_cap - at, nl2);
                }
            }
            nrec += total;
            if (act) {
                dp_score[row_off[t] + i] = best;
                dp_prev[row_off[t] + i] = best_prev;
            }
        }
    }
    __syncwarp();
    return nrec;
}

__device__ __forceinline__ double cw_max_double(double x) {
    for (int o = 16; o; o >>= 1) {
        double y = __shfl_xor_sync(0xffffffffu, x, o);
        x = y > x ? y : x;
    }
    return x;
}

// Back-tracking (chain.cpp:234-300) by one warp: the scans over the record list run 32 records at a time,
// the per-chain work (repeats test, walk along the prev pointers) stays sequential in list order.
// rpos_add: see cm_chain_backtrack_g (the repeats look-up key stays unshifted, the stored positions are shifted)
template <class G>
__device__ void cw_chain_backtrack(const G &g, const DevView &v, int T, const uint32_t *row_pos_off, const uint32_t *row_cnt, const uint32_t *row_off,
                                   const double *dp_score, const uint32_t *dp_prev, const ChainRec *recs, uint32_t nrec, ChainOut out,
                                   uint32_t rpos_add) {
    const unsigned FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u;
    const int kmer = g.kmer;
    const uint32_t max_best = (uint32_t) v.p.max_chain_len;
    const double NONE = -1e300;
    uint32_t best_count = 0;
    double best_score = NONE;
    for (uint32_t r = lane; r < nrec; r += 32) best_score = recs[r].score > best_score ? recs[r].score : best_score;
    best_score = cw_max_double(best_score);
    if (best_score < (double) kmer) best_score = (double) kmer;
    double thresh = 0;
    bool first = true;
    while (best_count < max_best) {
        double s = NONE;   // next distinct score below thresh
        for (uint32_t r = lane; r < nrec; r += 32) {
            double x = recs[r].score;
            if ((first || x < thresh) && x > s) s = x;
        }
        s = cw_max_double(s);
        if (s == NONE) break;
        first = false;
        thresh = s;
        uint32_t in_group = 0;
        bool done = false;
        for (uint32_t r0 = 0; r0 < nrec && !done; r0 += 32) {
            const uint32_t r = r0 + lane;
            const bool hit = r < nrec && recs[r].score == s;
            const uint32_t my_loc = hit ? recs[r].loc : 0;
            unsigned m = __ballot_sync(FULL, hit);
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1;
                if (in_group >= max_best) { done = true; break; }
                in_group++;
                if (best_count >= max_best) { done = true; break; }
                const uint32_t loc = __shfl_sync(FULL, my_loc, src);
                const uint32_t spos = g.at(row_pos_off[loc >> 16] + (loc & 0xFFFF));
                if (s < best_score) {
                    // `repeats`: lane c scans the non-first fragments of emitted chain c (at most 30 chains)
                    bool rep = false;
                    if (lane < best_count)
                        for (uint32_t f = 1; f < out.len[lane]; f++)
                            if (out.frags[lane * out.F + f].rpos == spos) { rep = true; break; }
                    if (__any_sync(FULL, rep)) continue;
                }
                const uint32_t j = best_count++;
                if (lane == 0) {
                    uint32_t n = 0, cur = loc;
                    while (cur != CM_NOPREV) {
                        uint32_t row = cur >> 16, ind = cur & 0xFFFF;
                        out.frags[j * out.F + n].rpos = g.at(row_pos_off[row] + ind) + rpos_add;
                        out.frags[j * out.F + n].qpos = (int32_t) g.qpos((int) row);
                        n++;
                        cur = dp_prev[row_off[row] + ind];
                    }
                    out.score[j] = (float) s;
                    out.len[j] = n;
                }
                __syncwarp();
            }
        }
    }
    if (best_count == 0 && lane == 0) {   // chains of size 1 (chain.cpp:283-298)
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
    best_count = __shfl_sync(FULL, best_count, 0);
    if (lane == 0) *out.count = (int32_t) best_count;
    __syncwarp();
}

// K2b': heavy strands (the first n_heavy entries of the work-signature order), one warp each
__global__ void __launch_bounds__(128) k_chain_warp(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, const uint64_t *__restrict__ cell_off, double *__restrict__ dp_score,
                                                    uint32_t *__restrict__ dp_prev, ChainRec *__restrict__ recs, uint32_t rec_mult,
                                                    uint32_t *__restrict__ ovf_strand, uint32_t *__restrict__ ovf_nrec,
                                                    uint32_t *__restrict__ n_ovf, const uint32_t *__restrict__ order, uint32_t n_heavy,
                                                    cm_counters *__restrict__ ctr) {
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (w >= n_heavy) return;
    const uint32_t s = order[w];
    const uint32_t *so = b.seed_off + (size_t) s * v.n_slots, *sc = b.seed_cnt + (size_t) s * v.n_slots;
    uint32_t row_pos_off[CM_MAX_SEEDS], row_cnt[CM_MAX_SEEDS], row_off[CM_MAX_SEEDS], cells;
    int hh;
    const int T = cm_chain_rows(v, so, sc, row_pos_off, row_cnt, row_off, cells, hh);
    if (lane == 0) {
        b.chain_hh[s] = hh;
        b.chain_cnt[s] = 0;
    }
    if (T <= 0) return;
    const uint64_t base = cell_off[s];
    const uint32_t cap = cm_rec_cap(cells, rec_mult);
    ChainRec *r = recs + cm_rec_base(base, s, rec_mult);
    uint64_t links = 0;
    const ChainGeomMain g{v.c.pos, v.p.kmer};
    const uint32_t nrec = cw_chain_dp(g, v, b.strand_len[s], T, row_pos_off, row_cnt, row_off, dp_score + base, dp_prev + base, r, cap, links);
    if (nrec > cap) {
        if (lane == 0) {
            uint32_t k = atomicAdd(n_ovf, 1u);
            ovf_strand[k] = s;
            ovf_nrec[k] = nrec;
        }
    } else {
        ChainOut out;
        out.count = b.chain_cnt + s;
        out.score = b.chain_score + (size_t) s * b.C;
        out.len = b.chain_len + (size_t) s * b.C;
        out.frags = b.chain_frags + (size_t) s * b.C * b.F;
        out.F = b.F;
        cw_chain_backtrack(g, v, T, row_pos_off, row_cnt, row_off, dp_score + base, dp_prev + base, r, nrec, out, 0u);
    }
    if (links) {
        cm_counters *c = ctr + (w & (CM_CTR_SLOTS - 1));
        atomicAdd((unsigned long long *) &c->chain_links, (unsigned long long) links);
    }
}
