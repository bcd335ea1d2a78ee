// Warp-per-pair classification for pairs with many candidate chain pairs (both mates in repeat families: up to
// 30 x 30 MatePairs).  One thread per pair makes such a pair the critical path of the whole kernel (about 5 ms of an
// 11 ms launch on C2).  Device only; cm_filter.h stays the single-thread statement (and the hostsim one).
//
// process_mates (src/filter.cpp:244-395) visits the MatePairs in (i, j) order and each visit may update the running
// best mapping `mr` (MatchedRead::update / go_for_update compare against it) or return at once.  Here the visits of
// one row i are evaluated *speculatively* by the lanes of a warp, each on a private copy of `mr` as it stood before
// the row's pending candidates, and committed in order: a lane's evaluation is exact as long as no earlier pending
// candidate changed `mr`; at the first one that did (or returned), its state is committed, everything after it is
// discarded and re-evaluated from the new state.  Updates are rare among the candidates of a repeat-rich pair, so
// nearly all of the work is done 30-wide.  The side results (min_ret, *_genic) are a minimum and a last-value, folded
// over the committed lanes.  The result is the serial one by construction.
#pragma once

static_assert(CM_MAX_CHAINS <= 32, "one lane per chain");

__device__ __forceinline__ void cw_bcast_mread(cm_mread &m, int src) {
    static_assert(sizeof(cm_mread) % 4 == 0, "cm_mread is a whole number of words");
    uint32_t *w = reinterpret_cast<uint32_t *>(&m);
#pragma unroll
    for (unsigned k = 0; k < sizeof(cm_mread) / 4; k++) w[k] = __shfl_sync(0xffffffffu, w[k], src);
}
__device__ __forceinline__ bool cw_same_mread(const cm_mread &a, const cm_mread &b) {
    const uint32_t *x = reinterpret_cast<const uint32_t *>(&a), *y = reinterpret_cast<const uint32_t *>(&b);
    bool same = true;
#pragma unroll
    for (unsigned k = 0; k < sizeof(cm_mread) / 4; k++) same = same && x[k] == y[k];
    return same;
}

// The unpaired chains of one side (src/filter.cpp:356-386), one lane per chain.  The reference extends them one after the
// other into ONE MatchedMate: the value extend_chain_both_sides returns does not depend on what the previous chain left
// in it (every field it reads it has written first), min_ret is a minimum over the chains, and the exon look-ups behind
// r*_genic happen only for the first unpaired chain (overlap_to_spos / _epos return at once when looked_up_* is set,
// src/utils.cpp:667-683) -- so the lanes are independent and the fold is exact.
__device__ __forceinline__ void cw_leftover_side(FilterCtx &f, const StrandView &sv, uint32_t paired, int dir, int &min_ret, bool &genic) {
    const unsigned FULL = 0xffffffffu;
    const int lane = (int) (threadIdx.x & 31u);
    const int n = sv.chains.count;
    const uint32_t todo = ~paired & (n >= 32 ? FULL : ((1u << n) - 1u));
    if (min_ret == CM_CONCRD || !todo) return;
    const AnnotDev &a = f.x.v->a;
    MateD mm;
    mate_init(mm, f.x.v->p);
    int ex = 0x7fffffff;
    if ((todo >> lane) & 1u) ex = cm_extend_chain_both_sides(f.x, cl_chain(sv.chains, lane), sv.seq, sv.seq_len, mm, dir);
    const int m = __reduce_min_sync(FULL, ex);
    min_ret = m < min_ret ? m : min_ret;
    const int first = __ffs(todo) - 1;
    int g = 0;
    if (lane == first) {
        cm_overlap_to_spos(a, mm);
        cm_overlap_to_epos(a, mm);
        g = (mm.exons_spos >= 0) || (mm.exons_epos >= 0);
    }
    genic = __shfl_sync(FULL, g, first) != 0;
}

// The rest of process_mates after the MatePair loop (cm_mates_leftovers), by one warp
__device__ int cw_mates_leftovers(FilterCtx &f, const StrandView &fw, const StrandView &bw, cm_mread &mr, uint32_t f_paired,
                                  uint32_t b_paired, MatesAcc acc) {
    if (mr.type == CM_CONCRD || mr.type == CM_DISCRD || mr.type == CM_CHIORF || mr.type == CM_CHIBSJ || mr.type == CM_CHI2BSJ)
        return mr.type;
    int min_ret1 = acc.min_ret1, min_ret2 = acc.min_ret2;
    bool r1_genic = acc.r1_genic, r2_genic = acc.r2_genic;
    cw_leftover_side(f, fw, f_paired, 1, min_ret1, r1_genic);
    cw_leftover_side(f, bw, b_paired, -1, min_ret2, r2_genic);
    int new_type = (((min_ret1 == CM_ORPHAN) && (min_ret2 == CM_CONCRD)) || ((min_ret1 == CM_CONCRD) && (min_ret2 == CM_ORPHAN)))
                       ? CM_OEANCH
                       : ((min_ret1 == CM_ORPHAN) || (min_ret2 == CM_ORPHAN))
                             ? CM_ORPHAN
                             : ((min_ret1 == CM_CONCRD) && (min_ret2 == CM_CONCRD) && (r1_genic && r2_genic))
                                   ? CM_CHIFUS
                                   : ((min_ret1 == CM_CONCRD) && (min_ret2 == CM_CONCRD)) ? CM_OEA2 : CM_CANDID;
    mr_update_type(mr, new_type);
    return mr.type;
}

// process_mates by one warp; every lane enters and leaves with the same mr and return value
__device__ int cw_process_mates(FilterCtx &f, const StrandView &fw, const StrandView &bw, cm_mread &mr, bool r1_forward) {
    const unsigned FULL = 0xffffffffu;
    const int lane = (int) (threadIdx.x & 31u);
    const AnnotDev &a = f.x.v->a;
    const cm_params &p = f.x.v->p;
    const int nf = fw.chains.count, nb = bw.chains.count;
    const int saved_type = mr.type;
    // lane c looks up the exon interval of forward chain c and of backward chain c; every lane keeps all of them
    int my_f = lane < nf ? cm_location_overlap_ne(a, cl_chain(fw.chains, lane).frags[0].rpos) : -1;
    int my_b = lane < nb ? cm_location_overlap_ne(a, cl_chain(bw.chains, lane).frags[0].rpos) : -1;
    int f_iv[CM_MAX_CHAINS], b_iv[CM_MAX_CHAINS];
    for (int c = 0; c < CM_MAX_CHAINS; c++) {
        f_iv[c] = __shfl_sync(FULL, my_f, c);
        b_iv[c] = __shfl_sync(FULL, my_b, c);
    }
    // pair_chains: lane i computes row i of the code matrix (2 bits per j)
    unsigned long long row = 0;
    uint32_t my_bp = 0;
    if (lane < nf) {
        ChainView ci = cl_chain(fw.chains, lane);
        for (int j = 0; j < nb; j++) {
            int c = cm_pair_code(a, p, ci, cl_chain(bw.chains, j), f_iv[lane], b_iv[j], saved_type);
            if (c) {
                row |= (unsigned long long) c << (2 * j);
                my_bp |= 1u << j;
            }
        }
    }
    const uint32_t f_paired = __ballot_sync(FULL, row != 0);
    const uint32_t b_paired = __reduce_or_sync(FULL, my_bp);
    MatesAcc acc;
    acc.min_ret1 = acc.min_ret2 = CM_ORPHAN;
    acc.r1_genic = acc.r2_genic = false;
    for (int i = 0; i < nf; i++) {
        const unsigned long long r = __shfl_sync(FULL, row, i);
        if (!r) continue;
        // candidates of this row in j order
        int ncand = 0;
        for (int j = 0; j < nb; j++) ncand += ((r >> (2 * j)) & 3ull) != 0;
        int start = 0;
        while (start < ncand) {
            const int k = start + lane;
            const bool act = k < ncand;
            cm_mread m2 = mr;
            MatesAcc a2 = acc;
            bool early = false;
            if (act) {
                int j = 0, seen = -1, c = 0;
                for (; j < nb; j++) {
                    c = (int) ((r >> (2 * j)) & 3ull);
                    if (c && ++seen == k) break;
                }
                early = cm_mates_candidate(f, fw, bw, i, j, c - 1, f_iv[i], b_iv[j], m2, r1_forward, a2);
            }
            const bool changed = act && (early || !cw_same_mread(m2, mr));
            const unsigned chm = __ballot_sync(FULL, changed);
            const int nact = ncand - start < 32 ? ncand - start : 32;
            const int upto = chm ? __ffs(chm) - 1 : nact - 1;   // lanes 0..upto are committed
            const int m1 = __reduce_min_sync(FULL, lane <= upto ? a2.min_ret1 : 0x7fffffff);
            const int m2r = __reduce_min_sync(FULL, lane <= upto ? a2.min_ret2 : 0x7fffffff);
            acc.min_ret1 = m1 < acc.min_ret1 ? m1 : acc.min_ret1;
            acc.min_ret2 = m2r < acc.min_ret2 ? m2r : acc.min_ret2;
            acc.r1_genic = __shfl_sync(FULL, (int) a2.r1_genic, upto) != 0;
            acc.r2_genic = __shfl_sync(FULL, (int) a2.r2_genic, upto) != 0;
            const bool ret_now = __shfl_sync(FULL, (int) early, upto) != 0;
            if (chm) {
                mr = m2;
                cw_bcast_mread(mr, upto);
            }
            if (ret_now) return CM_CONCRD;
            start += upto + 1;
        }
    }
    return cw_mates_leftovers(f, fw, bw, mr, f_paired, b_paired, acc);
}

// FilterRead::process_read, PE (src/filter.cpp:124-241), by one warp
__device__ int cw_process_read_pe(FilterCtx &f, const PairView &pv, cm_mread &mr) {
    const cm_params &p = f.x.v->p;
    const int nf1 = pv.s[0].chains.count, nb1 = pv.s[1].chains.count, nf2 = pv.s[2].chains.count, nb2 = pv.s[3].chains.count;
    if (nf1 + nb1 + nf2 + nb2 <= 0) {
        int t = ((pv.high_hits[0] + pv.high_hits[1] > 0) && (pv.high_hits[2] + pv.high_hits[3] > 0)) ? CM_NOPROC_MANYHIT : CM_NOPROC_NOMATCH;
        mr_update_type(mr, t);
        return t;
    }
    if ((nf1 + nb1 <= 0) || (nf2 + nb2 <= 0)) {
        mr_update_type(mr, CM_OEANCH);
        return CM_OEANCH;
    }
    float fc1 = nf1 > 0 ? pv.s[0].chains.score[0] : 0.f;
    float bc1 = nb1 > 0 ? pv.s[1].chains.score[0] : 0.f;
    float fc2 = nf2 > 0 ? pv.s[2].chains.score[0] : 0.f;
    float bc2 = nb2 > 0 ? pv.s[3].chains.score[0] : 0.f;
    const bool first = __fadd_rn(fc1, bc2) >= __fadd_rn(fc2, bc1);
    for (int k = 0; k < 2; k++) {
        bool attempt_a = (k == 0) == first;
        int r = attempt_a ? cw_process_mates(f, pv.s[0], pv.s[3], mr, true) : cw_process_mates(f, pv.s[2], pv.s[1], mr, false);
        if (p.scan_level == 0 && r == CM_CONCRD) return CM_CONCRD;
    }
    return mr.type;
}

// K3': heavy pairs (the first n_heavy entries of the pair order), one warp each
__global__ void __launch_bounds__(128) k_process_pe_warp(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, const cm_mread *__restrict__ in, cm_mread *__restrict__ out,
                                                         int32_t *__restrict__ state, uint32_t n_heavy, const uint32_t *__restrict__ order,
                                                         cm_counters *__restrict__ ctr) {
    const uint32_t w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (w >= n_heavy) return;
    const uint32_t i = order[w];
    FilterCtx f;
    ext_ctx_init(f.x, &v);
    f.contig_id = v.contig_id;
    PairView pv;
    for (int k = 0; k < 4; k++) {
        pv.s[k] = cm_strand_view(b, 4 * i + k);
        pv.high_hits[k] = b.chain_hh[4 * i + k];
    }
    cm_mread mr = in[i];
    const int r = cw_process_read_pe(f, pv, mr);
    if (lane == 0) {
        state[i] = r;
        out[i] = mr;
    }
    // every lane did its own candidates and leftover chains
    stats_flush(f.x.st, lane == 0 ? 1u : 0u, ctr);
}
