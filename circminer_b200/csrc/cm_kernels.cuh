// Chaining / classification kernels and the batch pipeline.  Included by cm_gpu.cu after the context.
#pragma once

// ------------------------------------------------------------------------------------------
// K2a: dp cells per strand (sum of kept hit-list lengths) -> exclusive scan gives arena offsets
// ------------------------------------------------------------------------------------------
__global__ void k_cell_count(const __grid_constant__ DevView v, const uint32_t *__restrict__ seed_cnt, uint32_t n_strands, uint32_t *__restrict__ cells) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_strands) return;
    uint32_t c = 0;
    for (int t = 0; t < v.n_slots; t++) {
        uint32_t raw = seed_cnt[(size_t) s * v.n_slots + t];
        c += raw > (uint32_t) v.p.seed_lim ? 0 : raw;
    }
    cells[s] = c;
}

CM_HD uint64_t cm_rec_base(uint64_t cell_off, uint32_t strand, uint32_t mult) { return (uint64_t) mult * cell_off + 8ull * strand; }
CM_HD uint32_t cm_rec_cap(uint32_t cells, uint32_t mult) { return mult * cells + 8u; }

// ------------------------------------------------------------------------------------------
// K2b: chaining, one thread per strand.  Strands whose improvement list outgrows its slice of the
// record arena are queued for k_chain_big (exact two-pass sizing) instead of being truncated.
// ------------------------------------------------------------------------------------------
#ifndef CM_CHAIN_LOCAL
#define CM_CHAIN_LOCAL 1
#endif
#ifndef CM_CHAIN_LOCAL_CELLS
#define CM_CHAIN_LOCAL_CELLS 64
#endif
#ifndef CM_CHAIN_LOCAL_RECS
#define CM_CHAIN_LOCAL_RECS 32
#endif
#ifndef CM_CHAIN_MINB
#define CM_CHAIN_MINB 6   // <= 80 registers: k_chain is latency-bound, more resident warps pay (25.95 vs 26.06 ms per step)
#endif
#ifndef CM_PE_MINB
#define CM_PE_MINB 3
#endif
#ifndef CM_PE_BLOCK
#define CM_PE_BLOCK 128   // threads per block of the per-pair kernel
#endif
__global__ void __launch_bounds__(128, CM_CHAIN_MINB) k_chain(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, const uint64_t *__restrict__ cell_off, double *__restrict__ dp_score,
                                               uint32_t *__restrict__ dp_prev, ChainRec *__restrict__ recs, uint32_t rec_mult,
                                               uint32_t *__restrict__ ovf_strand, uint32_t *__restrict__ ovf_nrec,
                                               uint32_t *__restrict__ n_ovf, const uint32_t *__restrict__ order, uint32_t n_units,
                                               cm_counters *__restrict__ ctr) {
    uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t links = 0;
    if (slot < n_units) {
        const uint32_t s = order[slot];   // strands with the same row shape sit in the same warp
        const uint32_t *so = b.seed_off + (size_t) s * v.n_slots, *sc = b.seed_cnt + (size_t) s * v.n_slots;
        uint32_t row_pos_off[CM_MAX_SEEDS], row_cnt[CM_MAX_SEEDS], row_off[CM_MAX_SEEDS], cells;
        int hh;
        int T = cm_chain_rows(v, so, sc, row_pos_off, row_cnt, row_off, cells, hh);
        b.chain_hh[s] = hh;
        b.chain_cnt[s] = 0;
        if (T > 0) {
            uint64_t base = cell_off[s];
            uint32_t cap = cm_rec_cap(cells, rec_mult);
            ChainRec *r = recs + cm_rec_base(base, s, rec_mult);
#if CM_CHAIN_LOCAL_RECS
            // same for the improvement records of a strand whose slice is small (unique hits: 7 cells -> 29 records)
            ChainRec lr[CM_CHAIN_LOCAL_RECS];
            if (cap <= CM_CHAIN_LOCAL_RECS) r = lr;
#endif
#if CM_CHAIN_LOCAL
            // dp cells of a light strand (<= CM_CHAIN_LOCAL_CELLS; the heavy ones go to k_chain_warp) live in local memory:
            // the hardware interleaves local words across the lanes of a warp, so the per-thread rows that were scattered
            // 128-byte lines in the global arena become coalesced, L1-resident accesses.  The arena is still what
            // k_chain_big and the warp kernel use.
            double ls[CM_CHAIN_LOCAL_CELLS];
            uint32_t lp[CM_CHAIN_LOCAL_CELLS];
            const bool loc = cells <= CM_CHAIN_LOCAL_CELLS;
            double *ds = loc ? ls : dp_score + base;
            uint32_t *dpv = loc ? lp : dp_prev + base;
#else
            double *ds = dp_score + base;
            uint32_t *dpv = dp_prev + base;
#endif
            uint32_t nrec = cm_chain_dp(v, b.strand_len[s], T, row_pos_off, row_cnt, row_off, ds, dpv, r, cap, &links);
            if (nrec > cap) {
                uint32_t k = atomicAdd(n_ovf, 1u);
                ovf_strand[k] = s;
                ovf_nrec[k] = nrec;
            } else {
                ChainOut out;
                out.count = b.chain_cnt + s;
                out.score = b.chain_score + (size_t) s * b.C;
                out.len = b.chain_len + (size_t) s * b.C;
                out.frags = b.chain_frags + (size_t) s * b.C * b.F;
                out.F = b.F;
                cm_chain_backtrack(v, T, row_pos_off, row_cnt, row_off, ds, dpv, r, nrec, out);
            }
        }
    }
    if (links) {
        cm_counters *c = ctr + ((blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) & (CM_CTR_SLOTS - 1));
        atomicAdd((unsigned long long *) &c->chain_links, (unsigned long long) links);
    }
}

// overflow path: record arena sized exactly from the counts k_chain produced
__global__ void __launch_bounds__(128) k_chain_big(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, const uint64_t *__restrict__ cell_off, double *__restrict__ dp_score,
                                                   uint32_t *__restrict__ dp_prev, const uint32_t *__restrict__ ovf_strand,
                                                   const uint64_t *__restrict__ ovf_off, const uint32_t *__restrict__ ovf_nrec, uint32_t n_ovf,
                                                   ChainRec *__restrict__ recs) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_ovf) return;
    uint32_t s = ovf_strand[k];
    const uint32_t *so = b.seed_off + (size_t) s * v.n_slots, *sc = b.seed_cnt + (size_t) s * v.n_slots;
    uint32_t row_pos_off[CM_MAX_SEEDS], row_cnt[CM_MAX_SEEDS], row_off[CM_MAX_SEEDS], cells;
    int hh;
    int T = cm_chain_rows(v, so, sc, row_pos_off, row_cnt, row_off, cells, hh);
    uint64_t base = cell_off[s];
    ChainRec *r = recs + ovf_off[k];
    uint32_t nrec = cm_chain_dp(v, b.strand_len[s], T, row_pos_off, row_cnt, row_off, dp_score + base, dp_prev + base, r, ovf_nrec[k], nullptr);
    ChainOut out;
    out.count = b.chain_cnt + s;
    out.score = b.chain_score + (size_t) s * b.C;
    out.len = b.chain_len + (size_t) s * b.C;
    out.frags = b.chain_frags + (size_t) s * b.C * b.F;
    out.F = b.F;
    cm_chain_backtrack(v, T, row_pos_off, row_cnt, row_off, dp_score + base, dp_prev + base, r, nrec, out);
}

#include "cm_chain_warp.cuh"

// ------------------------------------------------------------------------------------------
// K3: extension + classification, one thread per pair (PE) / read (SE)
// ------------------------------------------------------------------------------------------
// Work counters.  No block barrier here: the threads of a warp leave the per-pair code at unrelated points and
// an aligned bar.sync reached by a diverged warp is undefined (it lost contributions on sm_100a).  Each thread
// adds its non-zero counts into one of CM_CTR_SLOTS copies of cm_counters (summed by cm_gpu_get_counters),
// chosen by warp, so same-address atomics stay rare.
__device__ __forceinline__ void stats_flush(const AlignStats &st, uint32_t units, cm_counters *ctr) {
    cm_counters *c = ctr + ((blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) & (CM_CTR_SLOTS - 1));
    if (st.xdrop_cells) atomicAdd((unsigned long long *) &c->xdrop_cells, (unsigned long long) st.xdrop_cells);
    if (st.edit_cells) atomicAdd((unsigned long long *) &c->edit_cells, (unsigned long long) st.edit_cells);
    if (st.xdrop_calls) atomicAdd((unsigned long long *) &c->xdrop_calls, (unsigned long long) st.xdrop_calls);
    if (st.edit_calls) atomicAdd((unsigned long long *) &c->edit_calls, (unsigned long long) st.edit_calls);
    if (units) atomicAdd((unsigned int *) &c->pairs, units);   // low word of the slot; 2^32 units per slot between resets
}

__global__ void __launch_bounds__(CM_PE_BLOCK, CM_PE_MINB) k_process_pe(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, const cm_mread *__restrict__ in, cm_mread *__restrict__ out,
                                                    int32_t *__restrict__ state, uint32_t n_pairs, const uint32_t *__restrict__ order,
                                                    cm_counters *__restrict__ ctr) {
    uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    AlignStats st = {0, 0, 0, 0};
    const bool live = slot < n_pairs;
    if (live) {
        const uint32_t i = order[slot];   // pairs with the same control-flow signature sit in the same warp
        cm_mread mr = in[i];
        state[i] = cm_pair_process(v, b, i, mr, st);
        out[i] = mr;
    }
    stats_flush(st, live ? 1u : 0u, ctr);
}

__global__ void __launch_bounds__(128) k_process_se(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, int32_t *__restrict__ state, uint32_t n_reads,
                                                    cm_counters *__restrict__ ctr) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    AlignStats st = {0, 0, 0, 0};
    if (i < n_reads) state[i] = cm_read_process_se(v, b, i, st);
    stats_flush(st, i < n_reads ? 1u : 0u, ctr);
}

#include "cm_pair_warp.cuh"

// stage-level parity entry: chains of the forward strand of each row, in the API layout [n][C][n_slots]
__global__ void k_chain_export(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, uint32_t n_rows, int FO, int32_t *__restrict__ chain_count, int32_t *__restrict__ high_hits,
                               float *__restrict__ score, uint32_t *__restrict__ chain_len, uint32_t *__restrict__ frag_rpos,
                               int32_t *__restrict__ frag_qpos) {
    uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= n_rows * (uint32_t) b.C) return;
    uint32_t row = gid / b.C, k = gid % b.C, s = 2 * row;
    if (k == 0) {
        chain_count[row] = b.chain_cnt[s];
        high_hits[row] = b.chain_hh[s];
    }
    bool live = (int) k < b.chain_cnt[s];
    score[gid] = live ? b.chain_score[(size_t) s * b.C + k] : 0.f;
    uint32_t L = live ? b.chain_len[(size_t) s * b.C + k] : 0;
    chain_len[gid] = L;
    for (int f = 0; f < FO; f++) {
        bool lf = (uint32_t) f < L && f < b.F;
        const FragDev *fr = b.chain_frags + ((size_t) s * b.C + k) * b.F + f;
        frag_rpos[(size_t) gid * FO + f] = lf ? fr->rpos : 0;
        frag_qpos[(size_t) gid * FO + f] = lf ? fr->qpos : 0;
    }
}

// alignment kinds of cm_gpu_align_batch, one thread per problem
__global__ void k_align(cm_params p, int kind, uint32_t n, const uint8_t *__restrict__ s, const uint16_t *__restrict__ n_s,
                        const uint8_t *__restrict__ t, const uint16_t *__restrict__ n_t, uint32_t stride, const int32_t *__restrict__ aux,
                        int32_t *__restrict__ out_ed, int32_t *__restrict__ out_sclen, int32_t *__restrict__ out_indel,
                        int32_t *__restrict__ out_score) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t *si = s + (size_t) i * stride, *ti = t + (size_t) i * stride;
    int ed = 0, sc = 0, indel = 0, score = 0;
    if (kind == 0 || kind == 1) ed = cm_drop_alignment_sc(si, n_s[i], ti, n_t[i], kind == 1, p, sc, indel, score, nullptr);
    else if (kind == 2 || kind == 3) ed = cm_local_alignment(si, n_s[i], ti, n_t[i], kind == 3, p, indel, score, nullptr);
    else ed = cm_one_side_banded(si, n_s[i], ti, n_t[i], aux[i], nullptr);
    out_ed[i] = ed; out_sclen[i] = sc; out_indel[i] = indel; out_score[i] = score;
}

// ------------------------------------------------------------------------------------------
// Work signatures.  Thread-per-unit kernels diverge on data-dependent control flow; sorting the units
// by a cheap signature first puts units that take the same path into the same warp.  Results do not
// depend on the order (every unit writes its own slots).
// ------------------------------------------------------------------------------------------
// strand: heavy first (inverted cell count), then the mask of non-empty rows
__global__ void k_strand_key(const __grid_constant__ DevView v, const uint32_t *__restrict__ seed_cnt, uint32_t n_strands, uint32_t *__restrict__ key,
                             uint32_t *__restrict__ idx, uint32_t heavy_th, uint32_t *__restrict__ n_heavy) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_strands) return;
    uint32_t cells = 0, mask = 0;
    for (int t = 0; t < v.n_slots; t++) {
        uint32_t raw = seed_cnt[(size_t) s * v.n_slots + t];
        uint32_t kept = raw > (uint32_t) v.p.seed_lim ? 0 : raw;
        cells += kept;
        if (kept) mask |= 1u << t;
    }
    uint32_t cq = cells > 255 ? 255 : cells;
    key[s] = ((255u - cq) << 24) | (mask & 0xFFFFFFu);
    idx[s] = s;
    // strands with more than heavy_th (< 255) cells sort to the front and go to the warp-per-strand kernel
    unsigned hm = __ballot_sync(__activemask(), cells > heavy_th);
    if (hm && (threadIdx.x & 31u) == (unsigned) (__ffs(hm) - 1)) atomicAdd(n_heavy, (uint32_t) __popc(hm));
}

// pair: chain counts of the four strands, attempt order, and the shape of the leading chains
__global__ void k_pair_key(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, uint32_t n_pairs, unsigned long long *__restrict__ key, uint32_t *__restrict__ idx,
                           uint32_t heavy_th, uint32_t *__restrict__ n_heavy, int sig_mode) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pairs) return;
    unsigned long long k = 0;
    int cnt[4];
    float sc[4];
    for (int q = 0; q < 4; q++) {
        cnt[q] = b.chain_cnt[4 * i + q];
        sc[q] = cnt[q] > 0 ? b.chain_score[(size_t) (4 * i + q) * b.C] : 0.f;
        int c = cnt[q] > 7 ? 7 : cnt[q];
        k = (k << 3) | (unsigned) c;
    }
    bool first = __fadd_rn(sc[0], sc[3]) >= __fadd_rn(sc[2], sc[1]);
    k = (k << 1) | (first ? 1u : 0u);
    // the two strands of the first attempt: (R1 fwd, R2 rc) or (R2 fwd, R1 rc)
    int sa = first ? 0 : 2, sb = first ? 3 : 1;
    for (int q = 0; q < 2; q++) {
        uint32_t s = 4 * i + (q ? sb : sa);
        unsigned sig = 0;
        if (b.chain_cnt[s] > 0) {
            uint32_t L = b.chain_len[(size_t) s * b.C];
            const FragDev *fr = b.chain_frags + (size_t) s * b.C * b.F;
            unsigned q0 = (unsigned) (fr[0].qpos / v.p.kmer) & 31u, q1 = (unsigned) (fr[L - 1].qpos / v.p.kmer) & 31u;
            sig = (q0 << 10) | (q1 << 5) | (L & 31u);
        }
        k = (k << 15) | sig;
    }
    // how far the transcript walk of the leading chains' ends went (EndSpec::n): ends that leave their exon take
    // the long way through extend_left / extend_right
    if (sig_mode) {
        unsigned w = 0;
        for (int q = 0; q < 2; q++) {
            uint32_t s = 4 * i + (q ? sb : sa);
            for (int e = 0; e < 2; e++) {
                unsigned nn = 0;
                // EndSpec::n is written only for ends that had something to align (ChainAux::has_l / has_r)
                if (b.chain_cnt[s] > 0 && b.spec && b.aux && b.chain_base && b.spec_k > 0) {
                    const ChainAux &ax = b.aux[b.chain_base[s]];
                    if (e ? ax.has_r : ax.has_l) nn = b.spec[2 * (size_t) b.chain_base[s] + e].n;
                }
                if (sig_mode == 1) nn = nn ? 1u : 0u;
                w = (w << 2) | (nn > 3u ? 3u : nn);
            }
        }
        k = (k << 8) | w;
    }
    // pairs heavier than heavy_th (candidate chain pairs of both attempts + chains to extend as leftovers) sort to the
    // front: one warp each
    const bool heavy = (uint32_t) (cnt[0] * cnt[3] + cnt[2] * cnt[1] + cnt[0] + cnt[1] + cnt[2] + cnt[3]) > heavy_th;
    if (heavy) k |= 1ull << (sig_mode ? 51 : 43);
    key[i] = k;
    idx[i] = i;
    unsigned hm = __ballot_sync(__activemask(), heavy);
    if (hm && (threadIdx.x & 31u) == (unsigned) (__ffs(hm) - 1)) atomicAdd(n_heavy, (uint32_t) __popc(hm));
}

// ------------------------------------------------------------------------------------------
// K2c: dense pre-alignment.  The per-pair code is control-flow heavy and its threads diverge; the DP
// work it would do is mostly determined by the chains alone (gap edit distances, genome-text X-drop of
// the two read ends), so it is done here first, one thread per (chain, kind) with all lanes of a warp
// on the same kind, and handed to k_process_* as a memo (ChainAux / DropHint).
// ------------------------------------------------------------------------------------------
__global__ void k_spec_count(const __grid_constant__ BatchDev b, uint32_t *__restrict__ cnt) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= b.n_strands) return;
    int n = b.chain_cnt[s];
    cnt[s] = (uint32_t) (n < b.spec_k ? n : b.spec_k);
}

// Work items of the pre-alignment: (chain, kind) with a cost key = read bases the DP has to cover.  Items are
// sorted by descending cost, so the lanes of a warp run DPs of the same kind and about the same length and the
// zero-cost items (nothing to align: ChainAux stays zeroed, has_* = 0) fall to the end, where whole warps exit.
__global__ void k_chain_list(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, uint32_t n_chains, uint8_t *__restrict__ key, uint32_t *__restrict__ item) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= b.n_strands) return;
    uint32_t base = b.chain_base[s];
    int n = b.chain_cnt[s];
    if (n > b.spec_k) n = b.spec_k;
    const int k = v.p.kmer, len = b.strand_len[s];
    for (int c = 0; c < n; c++) {
        const FragDev *fr = b.chain_frags + ((size_t) s * b.C + c) * b.F;
        const uint32_t L = b.chain_len[(size_t) s * b.C + c];
        int gap = 0;
        for (uint32_t i = 0; i + 1 < L; i++) {
            int g = fr[i + 1].qpos - fr[i].qpos - k;
            if (g > 0) gap += g;
        }
        int ml = L ? fr[0].qpos : 0, mr = L ? len - (fr[L - 1].qpos + k) : 0;
        const uint32_t id = base + (uint32_t) c;
        const int cost[3] = {gap, ml, mr};
        for (int kind = 0; kind < 3; kind++) {
            int q = cost[kind] < 0 ? 0 : (cost[kind] > 255 ? 255 : cost[kind]);
            key[(size_t) kind * n_chains + id] = (uint8_t) q;
            item[(size_t) kind * n_chains + id] = ((uint32_t) kind << 30) | id;
        }
    }
}

// chain id -> (strand << 5 | chain)
__global__ void k_chain_ids(const __grid_constant__ BatchDev b, uint32_t *__restrict__ list) {
    uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= b.n_strands) return;
    uint32_t base = b.chain_base[s];
    int n = b.chain_cnt[s];
    if (n > b.spec_k) n = b.spec_k;
    for (int c = 0; c < n; c++) list[base + c] = (s << 5) | (uint32_t) c;
}

__global__ void __launch_bounds__(128) k_chain_spec(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, const uint32_t *__restrict__ list, const uint8_t *__restrict__ key,
                                                    const uint32_t *__restrict__ item, uint32_t n_chains, uint32_t *__restrict__ walk_q,
                                                    uint32_t *__restrict__ n_walk, cm_counters *__restrict__ ctr) {
    uint64_t gid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    ExtCtx x;
    ext_ctx_init(x, &v);
    if (gid >= 3ull * n_chains || key[gid] == 0) return;
    const uint32_t it = item[gid];
    const int kind = (int) (it >> 30);
    const uint32_t id = it & 0x3FFFFFFFu;
    uint32_t e = list[id], s = e >> 5, c = e & 31u;
    ChainView ch;
    ch.frags = b.chain_frags + ((size_t) s * b.C + c) * b.F;
    ch.chain_len = b.chain_len[(size_t) s * b.C + c];
    ch.aux = nullptr;
    ch.spec = nullptr;
    const bool walk = cm_chain_spec(x, ch, b.strands + (size_t) s * b.stride, b.strand_len[s], kind, b.aux[id],
                                    b.spec ? b.spec + 2 * (size_t) id : nullptr);
    if (walk) {   // queue the end for k_chain_walk; the lanes that are here together share one atomic
        const unsigned act = __activemask(), lane = threadIdx.x & 31u;
        const int leader = __ffs(act) - 1;
        uint32_t base = 0;
        if ((int) lane == leader) base = atomicAdd(n_walk, (uint32_t) __popc(act));
        base = __shfl_sync(act, base, leader);
        walk_q[base + __popc(act & ((1u << lane) - 1u))] = it;
    }
    stats_flush(x.st, 0u, ctr);
}

// transcript walks of the queued chain ends (roughly in the cost order k_chain_spec met them)
__global__ void __launch_bounds__(128) k_chain_walk(const __grid_constant__ DevView v, const __grid_constant__ BatchDev b, const uint32_t *__restrict__ list, const uint32_t *__restrict__ walk_q,
                                                    const uint32_t *__restrict__ n_walk, cm_counters *__restrict__ ctr) {
    uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= *n_walk) return;
    ExtCtx x;
    ext_ctx_init(x, &v);
    const uint32_t it = walk_q[gid];
    const int kind = (int) (it >> 30);
    const uint32_t id = it & 0x3FFFFFFFu;
    uint32_t e = list[id], s = e >> 5, c = e & 31u;
    ChainView ch;
    ch.frags = b.chain_frags + ((size_t) s * b.C + c) * b.F;
    ch.chain_len = b.chain_len[(size_t) s * b.C + c];
    ch.aux = nullptr;
    ch.spec = nullptr;
    cm_chain_walk(x, ch, b.strands + (size_t) s * b.stride, b.strand_len[s], kind == 2, b.aux[id], b.spec + 2 * (size_t) id);
    stats_flush(x.st, 0u, ctr);
}

// ------------------------------------------------------------------------------------------
// Rounds on the device.  The reference maps one packed contig per "round" (src/circminer.cpp:229-308): a pair that
// is final in round c is printed, every other pair is written to the remain FASTQs with its MatchedRead in the
// header (src/filter.cpp:413-455) and parsed back in round c + 1 (src/fastq_parser.cpp:203-269).  With every
// contig resident the same state machine runs here without files: k_round_commit applies the skip rule
// (src/circminer.cpp:386-390) to the results of one contig, k_round_compact gathers the survivors' rows, seed keys
// and carried MatchedRead (after the header's lossy round trip) into the batch of the next contig.
// ------------------------------------------------------------------------------------------
// one thread per pair of the current batch; ids == nullptr: the batch is the caller's (pair k is input pair k)
__global__ void k_round_commit(const __grid_constant__ cm_params p, uint32_t n_cur, const uint32_t *__restrict__ ids,
                               const cm_mread *__restrict__ out, const int32_t *__restrict__ state,
                               const uint16_t *__restrict__ strand_len, int is_last, int contig_id,
                               cm_mread *__restrict__ final_mread, int32_t *__restrict__ final_state,
                               int32_t *__restrict__ final_round, uint8_t *__restrict__ survive,
                               cm_mread *__restrict__ snap_mread, int32_t *__restrict__ snap_state) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_cur) return;
    const uint32_t id = ids ? ids[k] : k;
    const cm_mread m = out[k];
    const int st = state[k];
    const bool skip = cm_skip_rule(p, st, m, (int) strand_len[4 * (size_t) k] + (int) strand_len[4 * (size_t) k + 2]);
    if (snap_mread) snap_mread[id] = m;
    if (snap_state) snap_state[id] = st;
    const bool fin = skip || is_last;
    if (fin) {
        final_mread[id] = m;
        final_state[id] = st;
        if (final_round) final_round[id] = contig_id;
    }
    survive[k] = fin ? 0 : 1;
}

// one warp per surviving pair: rows of the four strands, their lengths and seed keys, the carried MatchedRead
__global__ void __launch_bounds__(256) k_round_compact(const __grid_constant__ cm_params p, uint32_t n_surv, const uint32_t *__restrict__ surv,
                                const uint32_t *__restrict__ ids, uint32_t stride, int F,
                                const uint8_t *__restrict__ strands, const uint16_t *__restrict__ slen, const int2 *__restrict__ key,
                                const cm_mread *__restrict__ out,
                                uint8_t *__restrict__ strands_n, uint16_t *__restrict__ slen_n, int2 *__restrict__ key_n,
                                cm_mread *__restrict__ in_n, uint32_t *__restrict__ ids_n) {
    const uint32_t j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (j >= n_surv) return;
    const uint32_t k = surv[j];
    const uint4 *src = (const uint4 *) (strands + (size_t) 4 * k * stride);
    uint4 *dst = (uint4 *) (strands_n + (size_t) 4 * j * stride);
    for (uint32_t w = lane; w < stride / 4; w += 32) dst[w] = src[w];   // 4 rows of stride bytes = stride / 4 uint4
    const int2 *ks = key + (size_t) 4 * k * F;
    int2 *kd = key_n + (size_t) 4 * j * F;
    for (int w = (int) lane; w < 4 * F; w += 32) kd[w] = ks[w];
    if (lane < 4) slen_n[4 * (size_t) j + lane] = slen[4 * (size_t) k + lane];
    if (lane == 0) {
        in_n[j] = cm_header_round_trip(out[k], p);
        ids_n[j] = ids ? ids[k] : k;
    }
}

// ------------------------------------------------------------------------------------------
// Stage-level parity entry for the annotation look-ups the chaining / extension / classification code makes
// (GTFParser::get_upper_bound src/gene_annotation.h:121, get_location_overlap_ind src/gene_annotation.cpp:555,
// get_shift :451): one thread per position.
// ------------------------------------------------------------------------------------------
__global__ void k_annot_lookup(const __grid_constant__ DevView v, uint32_t n, const uint32_t *__restrict__ pos, uint32_t mlen, uint32_t rlen,
                               uint32_t *__restrict__ ub, uint32_t *__restrict__ max_end, int32_t *__restrict__ ol_iv,
                               int32_t *__restrict__ loc_iv, int32_t *__restrict__ loc_ind, int32_t *__restrict__ shift_idx,
                               uint8_t *__restrict__ intronic) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t p = pos[i];
    uint32_t me = 0;
    int ol = -1;
    ub[i] = cm_upper_bound(v.a, v.p, p, mlen, rlen, me, ol);
    max_end[i] = me;
    ol_iv[i] = ol;
    int ind = -1;
    loc_iv[i] = cm_location_overlap_ind(v.a, p, ind);
    loc_ind[i] = ind;
    shift_idx[i] = cm_get_shift(v.a, p);
    intronic[i] = cm_bit(v.a.intronic, v.a.n_bits, p) ? 1 : 0;
}

// ------------------------------------------------------------------------------------------
// Stage-level parity entry for the extension (TransExtension::extend_both_mates src/extend.cpp:37-125,
// extend_chain_both_sides :131-213): one thread per problem, chains given by the caller, every DP inline.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void cm_mate_export(const MateD &m, cm_mate &o) {
    o.spos = m.spos; o.epos = m.epos; o.qspos = m.qspos; o.qepos = m.qepos; o.matched_len = m.matched_len;
    o.left_ed = m.left_ed; o.middle_ed = m.middle_ed; o.right_ed = m.right_ed; o.sclen_left = m.sclen_left; o.sclen_right = m.sclen_right;
    o.dir = m.dir; o.type = m.type;
    o.is_concord = m.is_concord; o.left_ok = m.left_ok; o.right_ok = m.right_ok; o.pad_ = 0;
}
__global__ void __launch_bounds__(128) k_extend(const __grid_constant__ DevView v, int mode, uint32_t n, const uint8_t *__restrict__ lseq,
                                                const uint8_t *__restrict__ rseq, uint32_t stride, const uint16_t *__restrict__ llen,
                                                const uint16_t *__restrict__ rlen, int F, const uint32_t *__restrict__ lcl,
                                                const FragDev *__restrict__ lfr, const uint32_t *__restrict__ rcl, const FragDev *__restrict__ rfr,
                                                const int32_t *__restrict__ iv_a, const int32_t *__restrict__ iv_b, const int32_t *__restrict__ dir,
                                                cm_mate *__restrict__ lmm, cm_mate *__restrict__ rmm, int32_t *__restrict__ ret) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    ExtCtx x;
    ext_ctx_init(x, &v);
    ChainView lc, rc;
    lc.frags = lfr + (size_t) i * F; lc.chain_len = lcl[i]; lc.aux = nullptr; lc.spec = nullptr;
    MateD a, b;
    mate_init(a, v.p);
    mate_init(b, v.p);
    if (mode == 0) {
        rc.frags = rfr + (size_t) i * F; rc.chain_len = rcl[i]; rc.aux = nullptr; rc.spec = nullptr;
        TidSet ts;
        ts.iv_a = iv_a[i];
        ts.iv_b = iv_b[i];
        ret[i] = cm_extend_both_mates(x, lc, rc, ts, lseq + (size_t) i * stride, rseq + (size_t) i * stride, 1, 1, llen[i], rlen[i], a, b) ? 1 : 0;
        cm_mate_export(b, rmm[i]);
    } else {
        ret[i] = cm_extend_chain_both_sides(x, lc, lseq + (size_t) i * stride, llen[i], a, dir[i]);
    }
    cm_mate_export(a, lmm[i]);
}
