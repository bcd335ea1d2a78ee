// Stage 2 on the device, first piece: the per-contig ws-mer index and ProcessCirc::chaining as a batch entry
// (cm_circ.h has the device functions).  Included by cm_gpu.cu.
#pragma once
#include "cm_circ.h"
#include "cm_circ_split.h"

// key of the ws-mer starting at 1-based position P = tid + 1: its hash, or 4^ws when it holds a non-ACGT base or runs
// past the end of the contig (sorted to the back and dropped)
__global__ void k_circ_keys(const uint8_t *genome, uint32_t len, int ws, uint32_t *keys, uint32_t *vals) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= len) return;
    const uint32_t P = t + 1;
    int hv = -1;
    if (P + (uint32_t) ws - 1 <= len) hv = cm_circ_hash(genome + P, ws);
    keys[t] = hv < 0 ? (1u << (2 * ws)) : (uint32_t) hv;
    vals[t] = P;
}

// off[h] = first index of the sorted keys with key >= h, h = 0 .. 4^ws
__global__ void k_circ_offsets(const uint32_t *keys, uint32_t n, uint32_t n_off, uint32_t *off) {
    const uint32_t h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= n_off) return;
    off[h] = cm_circ_lower(keys, 0, n, h);
}

static int build_circ_index(cm_gpu_ctx *ctx, int contig_id, int ws) {
    if (ws < 4 || ws > 12) return fail(ctx, CM_ERR_ARG, "window size %d outside 4..12", ws);
    ContigHost *ch = get_contig(ctx, contig_id, false);
    if (!ch || !ch->loaded) return fail(ctx, CM_ERR_STATE, "contig %d not loaded", contig_id);
    if (ch->circ_ws == ws && ch->circ_pos) return CM_OK;
    if (ctx->parent) return fail(ctx, CM_ERR_STATE, "the ws-mer index of contig %d is built through the root context (cm_gpu_build_circ_index)", contig_id);
    if (ch->circ_pos) return fail(ctx, CM_ERR_STATE, "contig %d already has a ws-mer index for window size %d", contig_id, ch->circ_ws);
    cudaStream_t st = ctx->stream;
    const uint32_t len = ch->d.ref_len;
    std::vector<void *> tmp;
    auto cleanup = [&]() { for (void *q : tmp) cudaFree(q); };
    uint32_t *k0 = nullptr, *k1 = nullptr, *v0 = nullptr, *v1 = nullptr, *off = nullptr;
    uint8_t *d_tmp = nullptr;
    int r;
    if ((r = dev_alloc(ctx, tmp, &k0, (size_t) len)) != CM_OK || (r = dev_alloc(ctx, tmp, &k1, (size_t) len)) != CM_OK ||
        (r = dev_alloc(ctx, tmp, &v0, (size_t) len)) != CM_OK || (r = dev_alloc(ctx, ch->owned, &v1, (size_t) len)) != CM_OK ||
        (r = dev_alloc(ctx, ch->owned, &off, ((size_t) 1 << (2 * ws)) + 2)) != CM_OK) { cleanup(); return r; }
    k_circ_keys<<<grid_for(len, 256), 256, 0, st>>>(ch->d.genome, len, ws, k0, v0);
    ctx->launches++;
    size_t tb = 0;
    // stable sort on the 2 ws + 1 key bits: positions stay ascending inside a hash value
    cub::DeviceRadixSort::SortPairs(nullptr, tb, k0, k1, v0, v1, (int64_t) len, 0, 2 * ws + 1, st);
    if ((r = dev_alloc(ctx, tmp, &d_tmp, tb + 16)) != CM_OK) { cleanup(); return r; }
    cub::DeviceRadixSort::SortPairs(d_tmp, tb, k0, k1, v0, v1, (int64_t) len, 0, 2 * ws + 1, st);
    const uint32_t n_off = (1u << (2 * ws)) + 1;
    k_circ_offsets<<<grid_for(n_off, 256), 256, 0, st>>>(k1, len, n_off, off);
    ctx->launches += 2;
    cudaError_t e = cudaStreamSynchronize(st);
    cleanup();
    if (e != cudaSuccess || (e = cudaGetLastError()) != cudaSuccess) return fail(ctx, CM_ERR_CUDA, "ws-mer index build: %s", cudaGetErrorString(e));
    ch->circ_pos = v1;
    ch->circ_off = off;
    ch->circ_ws = ws;
    return CM_OK;
}

extern "C" int cm_gpu_build_circ_index(cm_gpu_ctx *ctx, int contig_id, int window_size) {
    if (!ctx) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    return build_circ_index(ctx, contig_id, window_size);
}

struct CircBatch {
    uint32_t n;
    const uint8_t *seqs;
    uint32_t stride;
    const uint32_t *qspos, *qepos, *gene_start, *gene_end;
    int step, R;                    // R = row slots per problem
    uint32_t *row_pos, *row_cnt;    // [n][R]
    int32_t *qp;                    // [n][R]
    int32_t *kmer_cnt;              // [n]
    uint32_t *cells;                // [n]
};

// rows of every problem (thread per problem)
__global__ void k_circ_rows(const __grid_constant__ DevView v, const CircIndex ix, const __grid_constant__ CircBatch b) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= b.n) return;
    uint32_t cells = 0;
    const size_t o = (size_t) i * b.R;
    b.kmer_cnt[i] = cm_circ_rows(ix, v.p, b.seqs + (size_t) i * b.stride, b.qspos[i], b.qepos[i], b.gene_start[i], b.gene_end[i], b.step, b.R,
                                 b.row_pos + o, b.row_cnt + o, b.qp + o, cells);
    b.cells[i] = cells;
}

// chain_seeds_sorted_kbest2 + the cut of ProcessCirc::chaining, one WARP per problem (cm_chain_warp.cuh over the stage-2 row
// geometry: the hits of a row are spread over the lanes).  A thread per problem is not enough here: a read that falls into a
// tandem repeat of a gene has rows of up to seedLim hits that all lie within reach of each other (8-mers every 2 bp), about
// 10^8 link evaluations for one problem -- seconds on one thread, and the whole batch waits for it.
// list == nullptr: every problem, with the record slice rec_base[k] .. + rec_cap[k]; a problem whose improvements outgrow
// its slice is appended to ovf (with the number it needs) and left for a second launch over that list with exact slices.
__global__ void __launch_bounds__(128) k_circ_chain(const __grid_constant__ DevView v, const CircIndex ix, const __grid_constant__ CircBatch b,
                                                    const uint32_t *list, uint32_t n_list, const uint64_t *cell_base, double *dp_score,
                                                    uint32_t *dp_prev, const uint64_t *rec_base, const uint32_t *rec_cap, ChainRec *recs,
                                                    uint32_t *ovf, uint32_t *ovf_nrec, uint32_t *n_ovf, int F, int32_t *out_count,
                                                    float *out_score, uint32_t *out_len, FragDev *out_frags) {
    const uint32_t k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (k >= n_list) return;
    const uint32_t i = list ? list[k] : k;
    const int C = v.p.max_chain_len;
    const size_t o = (size_t) i * b.R;
    const uint32_t *row_pos = b.row_pos + o, *row_cnt = b.row_cnt + o;
    uint32_t row_off[CM_CIRC_MAX_ROWS];
    const int kmer_cnt = b.kmer_cnt[i];
    const int T = cm_circ_trim(kmer_cnt, row_cnt, row_off);
    ChainOut out;
    out.count = out_count + i;
    out.score = out_score + (size_t) i * C;
    out.len = out_len + (size_t) i * C;
    out.frags = out_frags + (size_t) i * C * F;
    out.F = F;
    if (T <= 0) {
        if (lane == 0) *out.count = 0;
        return;
    }
    const ChainGeomCirc g{ix.pos, b.gene_start[i], ix.ws, b.qp + o};
    double *ds = dp_score + cell_base[i];
    uint32_t *dpv = dp_prev + cell_base[i];
    ChainRec *r = recs + rec_base[k];
    const uint32_t cap = rec_cap[k];
    uint64_t links = 0;
    const uint32_t nrec = cw_chain_dp(g, v, (int) b.qepos[i], T, row_pos, row_cnt, row_off, ds, dpv, r, cap, links);
    if (nrec > cap) {
        if (lane == 0) {
            const uint32_t s = atomicAdd(n_ovf, 1u);
            ovf[s] = i;
            ovf_nrec[s] = nrec;
            *out.count = -1;
        }
        return;
    }
    cw_chain_backtrack(g, v, T, row_pos, row_cnt, row_off, ds, dpv, r, nrec, out, b.gene_start[i]);
    if (lane == 0) cm_circ_finish(kmer_cnt, out);
}

// fragments as two planes (the layout of cm_gpu_chain_batch's outputs)
__global__ void k_circ_split_frags(const FragDev *__restrict__ fr, size_t n, uint32_t *__restrict__ rpos, int32_t *__restrict__ qpos) {
    const size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const FragDev f = fr[i];
    rpos[i] = f.rpos;
    qpos[i] = f.qpos;
}

static int circ_scan_u64(const std::vector<uint64_t> &in, std::vector<uint64_t> &out) {
    out.resize(in.size() + 1);
    uint64_t s = 0;
    for (size_t i = 0; i < in.size(); i++) { out[i] = s; s += in[i]; }
    out[in.size()] = s;
    return 0;
}

extern "C" int cm_gpu_circ_chain_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride, const uint32_t *qspos,
                                       const uint32_t *qepos, const uint32_t *gene_start, const uint32_t *gene_end, int window_size, int step,
                                       int F, int32_t *chain_count, int32_t *kmer_cnt, float *score, uint32_t *chain_len,
                                       uint32_t *frag_rpos, int32_t *frag_qpos) {
    if (!ctx) return CM_ERR_ARG;
    if (n == 0) return CM_OK;
    if (!seqs || !qspos || !qepos || !gene_start || !gene_end || !chain_count || !score || !chain_len || !frag_rpos || !frag_qpos)
        return fail(ctx, CM_ERR_ARG, "null argument");
    if (step < 1) return fail(ctx, CM_ERR_ARG, "step < 1");
    if (F < 1 || F > CM_CIRC_MAX_ROWS) return fail(ctx, CM_ERR_ARG, "F = %d outside 1..%d", F, CM_CIRC_MAX_ROWS);
    for (uint32_t i = 0; i < n; i++) {
        if (qspos[i] < 1 || qepos[i] > stride) return fail(ctx, CM_ERR_ARG, "problem %u: [qspos, qepos] outside the row", i);
        if (qepos[i] >= qspos[i] && (int) (qepos[i] - qspos[i] + 1) >= window_size &&
            (int) ((qepos[i] - qspos[i] + 1 - window_size) / step + 1) > F)
            return fail(ctx, CM_ERR_ARG, "problem %u: more than F = %d k-mers", i, F);
    }
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig_id, true, v);
    if (r != CM_OK) return r;
    if ((r = build_circ_index(ctx, contig_id, window_size)) != CM_OK) return r;
    ContigHost *ch = get_contig(ctx, contig_id, false);
    const CircIndex ix{ch->circ_pos, ch->circ_off, ch->circ_ws};
    cudaStream_t st = ctx->stream;
    std::vector<void *> tmp;
    auto cleanup = [&]() { for (void *q : tmp) cudaFree(q); };
    const bool verbose = getenv("CM_GPU_VERBOSE") != nullptr;
    auto now = []() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t_ph[6] = {now(), 0, 0, 0, 0, 0};
#define CIRC_TRY(x) do { if ((r = (x)) != CM_OK) { cleanup(); return r; } } while (0)
#define CIRC_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { cleanup(); return fail(ctx, CM_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e_)); } } while (0)
    const int C = ctx->p.max_chain_len, R = F;
    uint8_t *d_seq = nullptr;
    uint32_t *d_in = nullptr, *d_row_pos = nullptr, *d_row_cnt = nullptr, *d_cells = nullptr;
    int32_t *d_qp = nullptr, *d_kc = nullptr;
    CIRC_TRY(dev_alloc(ctx, tmp, &d_seq, (size_t) n * stride + 64));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_in, (size_t) n * 4));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_row_pos, (size_t) n * R));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_row_cnt, (size_t) n * R));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_qp, (size_t) n * R));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_kc, (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_cells, (size_t) n));
    CIRC_CUDA(cudaMemcpyAsync(d_seq, seqs, (size_t) n * stride, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_in, qspos, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_in + n, qepos, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_in + 2 * (size_t) n, gene_start, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_in + 3 * (size_t) n, gene_end, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CircBatch b{n, d_seq, stride, d_in, d_in + n, d_in + 2 * (size_t) n, d_in + 3 * (size_t) n, step, R, d_row_pos, d_row_cnt, d_qp, d_kc, d_cells};
    k_circ_rows<<<grid_for(n, 128), 128, 0, st>>>(v, ix, b);
    ctx->launches++;
    CIRC_CUDA(cudaGetLastError());
    std::vector<uint32_t> h_cells(n);
    CIRC_CUDA(cudaMemcpyAsync(h_cells.data(), d_cells, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaStreamSynchronize(st));
    t_ph[1] = now();
    // arenas: dp cells back to back; records 3 per cell + 8 per problem (as the stage-1 kernel), exact on the second pass
    std::vector<uint64_t> cells64(h_cells.begin(), h_cells.end()), cell_base, recn(n), rec_base;
    std::vector<uint32_t> rec_cap(n);
    for (uint32_t i = 0; i < n; i++) { rec_cap[i] = 3 * h_cells[i] + 8; recn[i] = rec_cap[i]; }
    circ_scan_u64(cells64, cell_base);
    circ_scan_u64(recn, rec_base);
    uint64_t *d_cell_base = nullptr, *d_rec_base = nullptr;
    uint32_t *d_rec_cap = nullptr, *d_prev = nullptr, *d_ovf = nullptr;
    double *d_score = nullptr;
    ChainRec *d_recs = nullptr;
    CIRC_TRY(dev_alloc(ctx, tmp, &d_cell_base, (size_t) n + 1));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_rec_base, (size_t) n + 1));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_rec_cap, (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_score, (size_t) cell_base[n] + 1));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_prev, (size_t) cell_base[n] + 1));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_recs, (size_t) rec_base[n] + 1));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_ovf, 2 * (size_t) n + 2));
    CIRC_CUDA(cudaMemcpyAsync(d_cell_base, cell_base.data(), ((size_t) n + 1) * 8, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_rec_base, rec_base.data(), ((size_t) n + 1) * 8, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_rec_cap, rec_cap.data(), (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemsetAsync(d_ovf + 2 * (size_t) n, 0, 8, st));
    int32_t *d_count = nullptr;
    float *d_oscore = nullptr;
    uint32_t *d_olen = nullptr;
    FragDev *d_frags = nullptr;
    CIRC_TRY(dev_alloc(ctx, tmp, &d_count, (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_oscore, (size_t) n * C));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_olen, (size_t) n * C));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_frags, (size_t) n * C * F));
    CIRC_CUDA(cudaMemsetAsync(d_oscore, 0, (size_t) n * C * 4, st));
    CIRC_CUDA(cudaMemsetAsync(d_olen, 0, (size_t) n * C * 4, st));
    CIRC_CUDA(cudaMemsetAsync(d_frags, 0, (size_t) n * C * F * sizeof(FragDev), st));
    uint32_t *d_novf = d_ovf + 2 * (size_t) n;
    k_circ_chain<<<grid_for((size_t) n * 32, 128), 128, 0, st>>>(v, ix, b, nullptr, n, d_cell_base, d_score, d_prev, d_rec_base, d_rec_cap, d_recs, d_ovf, d_ovf + n,
                                                  d_novf, F, d_count, d_oscore, d_olen, d_frags);
    ctx->launches++;
    CIRC_CUDA(cudaGetLastError());
    uint32_t n_ovf = 0;
    t_ph[2] = now();
    CIRC_CUDA(cudaMemcpyAsync(&n_ovf, d_novf, 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaStreamSynchronize(st));
    t_ph[3] = now();
    if (n_ovf) {
        std::vector<uint32_t> ovf(n_ovf), need(n_ovf);
        CIRC_CUDA(cudaMemcpy(ovf.data(), d_ovf, (size_t) n_ovf * 4, cudaMemcpyDeviceToHost));
        CIRC_CUDA(cudaMemcpy(need.data(), d_ovf + n, (size_t) n_ovf * 4, cudaMemcpyDeviceToHost));
        std::vector<uint64_t> need64(need.begin(), need.end()), base2;
        circ_scan_u64(need64, base2);
        uint64_t *d_base2 = nullptr;
        uint32_t *d_cap2 = nullptr, *d_list = nullptr, *d_ovf2 = nullptr;
        ChainRec *d_recs2 = nullptr;
        CIRC_TRY(dev_alloc(ctx, tmp, &d_base2, (size_t) n_ovf + 1));
        CIRC_TRY(dev_alloc(ctx, tmp, &d_cap2, (size_t) n_ovf));
        CIRC_TRY(dev_alloc(ctx, tmp, &d_list, (size_t) n_ovf));
        CIRC_TRY(dev_alloc(ctx, tmp, &d_recs2, (size_t) base2[n_ovf] + 1));
        CIRC_TRY(dev_alloc(ctx, tmp, &d_ovf2, 2 * (size_t) n_ovf + 2));
        CIRC_CUDA(cudaMemcpyAsync(d_base2, base2.data(), ((size_t) n_ovf + 1) * 8, cudaMemcpyHostToDevice, st));
        CIRC_CUDA(cudaMemcpyAsync(d_cap2, need.data(), (size_t) n_ovf * 4, cudaMemcpyHostToDevice, st));
        CIRC_CUDA(cudaMemcpyAsync(d_list, ovf.data(), (size_t) n_ovf * 4, cudaMemcpyHostToDevice, st));
        CIRC_CUDA(cudaMemsetAsync(d_ovf2 + 2 * (size_t) n_ovf, 0, 8, st));
        k_circ_chain<<<grid_for((size_t) n_ovf * 32, 128), 128, 0, st>>>(v, ix, b, d_list, n_ovf, d_cell_base, d_score, d_prev, d_base2, d_cap2, d_recs2, d_ovf2,
                                                      d_ovf2 + n_ovf, d_ovf2 + 2 * (size_t) n_ovf, F, d_count, d_oscore, d_olen, d_frags);
        ctx->launches++;
        CIRC_CUDA(cudaGetLastError());
    }
    if (verbose) { cudaStreamSynchronize(st); }
    t_ph[4] = now();
    // fragments come back as two planes, like cm_gpu_chain_batch
    const size_t nf = (size_t) n * C * F;
    uint32_t *d_rp = nullptr;
    int32_t *d_qp2 = nullptr;
    CIRC_TRY(dev_alloc(ctx, tmp, &d_rp, nf));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_qp2, nf));
    k_circ_split_frags<<<grid_for(nf, 256), 256, 0, st>>>(d_frags, nf, d_rp, d_qp2);
    ctx->launches++;
    CIRC_CUDA(cudaGetLastError());
    CIRC_CUDA(cudaMemcpyAsync(chain_count, d_count, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    if (kmer_cnt) CIRC_CUDA(cudaMemcpyAsync(kmer_cnt, d_kc, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(score, d_oscore, (size_t) n * C * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(chain_len, d_olen, (size_t) n * C * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(frag_rpos, d_rp, nf * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(frag_qpos, d_qp2, nf * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaStreamSynchronize(st));
    cleanup();
    t_ph[5] = now();
    if (verbose)
        fprintf(stderr, "cm_gpu: circ chaining of %u problems: rows %.3f s, arenas + uploads %.3f s (%.1f M cells), chain kernel %.3f s, overflow pass (%u) %.3f s, "
                "results to the host %.3f s\n", n, t_ph[1] - t_ph[0], t_ph[2] - t_ph[1], (double) cell_base[n] / 1e6, t_ph[3] - t_ph[2], n_ovf,
                t_ph[4] - t_ph[3], t_ph[5] - t_ph[4]);
#undef CIRC_TRY
#undef CIRC_CUDA
    return CM_OK;
}

// ---- ProcessCirc::find_exact_coord, one thread per problem -------------------------------------------------------------
struct CircExactIn {
    uint32_t n;
    const uint8_t *seqs;
    uint32_t stride;
    const uint32_t *r1_spos, *r1_epos, *r2_spos, *r2_epos, *qspos;
    const int32_t *dir, *rlen, *whole_len;
    int F;
    const uint32_t *chain_len;
    const FragDev *frags;
};

__global__ void __launch_bounds__(128) k_circ_exact(const __grid_constant__ DevView v, const __grid_constant__ CircExactIn in, cm_mate *pmm,
                                                    int32_t *ret) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= in.n) return;
    ExtCtx x;
    ext_ctx_init(x, &v);
    ChainView bc;
    bc.frags = in.frags + (size_t) i * in.F;
    bc.chain_len = in.chain_len[i];
    bc.aux = nullptr;
    bc.spec = nullptr;
    MateD a, b, pm;
    mate_init(a, v.p);
    mate_init(b, v.p);
    mate_init(pm, v.p);
    a.spos = in.r1_spos[i]; a.epos = in.r1_epos[i];
    b.spos = in.r2_spos[i]; b.epos = in.r2_epos[i];
    ret[i] = cm_circ_exact(x, a, b, pm, in.dir[i], in.qspos[i], in.seqs + (size_t) i * in.stride, in.rlen[i], in.whole_len[i], bc) ? 1 : 0;
    cm_mate_export(pm, pmm[i]);
}

extern "C" int cm_gpu_circ_exact_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride, const uint32_t *r1_spos,
                                       const uint32_t *r1_epos, const uint32_t *r2_spos, const uint32_t *r2_epos, const int32_t *dir,
                                       const uint32_t *qspos, const int32_t *rlen, const int32_t *whole_len, uint32_t frag_slots,
                                       const uint32_t *chain_len, const uint32_t *frag_rpos, const int32_t *frag_qpos, int window_size,
                                       cm_mate *pmm, int32_t *ret) {
    if (!ctx) return CM_ERR_ARG;
    if (n == 0) return CM_OK;
    if (!seqs || !r1_spos || !r1_epos || !r2_spos || !r2_epos || !dir || !qspos || !rlen || !whole_len || !chain_len || !frag_rpos || !frag_qpos ||
        !pmm || !ret)
        return fail(ctx, CM_ERR_ARG, "null argument");
    if (window_size < 4 || window_size > 12 || frag_slots < 1) return fail(ctx, CM_ERR_ARG, "window size outside 4..12 or no fragment slots");
    for (uint32_t i = 0; i < n; i++) {
        if (chain_len[i] < 1 || chain_len[i] > frag_slots) return fail(ctx, CM_ERR_ARG, "problem %u: chain of %u fragments (slots: %u)", i, chain_len[i], frag_slots);
        if (qspos[i] < 1 || rlen[i] < 1 || qspos[i] - 1 + (uint32_t) rlen[i] > stride || whole_len[i] < 0 || (uint32_t) whole_len[i] > stride)
            return fail(ctx, CM_ERR_ARG, "problem %u: read range outside the row", i);
    }
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig_id, true, v);
    if (r != CM_OK) return r;
    v.p.kmer = window_size;   // fragment length of the chains; the extension code reads nothing else from it
    cudaStream_t st = ctx->stream;
    std::vector<void *> tmp;
    auto cleanup = [&]() { for (void *q : tmp) cudaFree(q); };
#define CIRC_TRY(x) do { if ((r = (x)) != CM_OK) { cleanup(); return r; } } while (0)
#define CIRC_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { cleanup(); return fail(ctx, CM_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e_)); } } while (0)
    const size_t F = frag_slots;
    uint8_t *d_seq = nullptr;
    uint32_t *d_u = nullptr;
    FragDev *d_fr = nullptr;
    cm_mate *d_mm = nullptr;
    int32_t *d_ret = nullptr;
    CIRC_TRY(dev_alloc(ctx, tmp, &d_seq, (size_t) n * stride + 64));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_u, (size_t) n * 9));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_fr, (size_t) n * F));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_mm, (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_ret, (size_t) n));
    std::vector<FragDev> h_fr((size_t) n * F);
    for (size_t q = 0; q < h_fr.size(); q++) { h_fr[q].rpos = frag_rpos[q]; h_fr[q].qpos = frag_qpos[q]; }
    const void *srcs[9] = {r1_spos, r1_epos, r2_spos, r2_epos, qspos, dir, rlen, whole_len, chain_len};
    CIRC_CUDA(cudaMemcpyAsync(d_seq, seqs, (size_t) n * stride, cudaMemcpyHostToDevice, st));
    for (int k = 0; k < 9; k++) CIRC_CUDA(cudaMemcpyAsync(d_u + (size_t) k * n, srcs[k], (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_fr, h_fr.data(), h_fr.size() * sizeof(FragDev), cudaMemcpyHostToDevice, st));
    CircExactIn in{n, d_seq, stride, d_u, d_u + n, d_u + 2 * (size_t) n, d_u + 3 * (size_t) n, d_u + 4 * (size_t) n,
                   (const int32_t *) (d_u + 5 * (size_t) n), (const int32_t *) (d_u + 6 * (size_t) n), (const int32_t *) (d_u + 7 * (size_t) n),
                   (int) F, d_u + 8 * (size_t) n, d_fr};
    k_circ_exact<<<grid_for(n, 128), 128, 0, st>>>(v, in, d_mm, d_ret);
    ctx->launches++;
    CIRC_CUDA(cudaGetLastError());
    CIRC_CUDA(cudaMemcpyAsync(pmm, d_mm, (size_t) n * sizeof(cm_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(ret, d_ret, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaStreamSynchronize(st));
    cleanup();
#undef CIRC_TRY
#undef CIRC_CUDA
    return CM_OK;
}

// ---- ProcessCirc::check_split_map (three mates) and everything under it, one thread per problem ------------------------------
__global__ void __launch_bounds__(64) k_circ_check(const __grid_constant__ DevView v, uint32_t n, const uint8_t *__restrict__ remain_seq,
                                                   const uint8_t *__restrict__ fullmap_seq, uint32_t stride, const int32_t *__restrict__ remain_len,
                                                   const int32_t *__restrict__ fullmap_len, const uint8_t *__restrict__ r1_partial,
                                                   cm_circ_mate *r1, cm_circ_mate *r2, cm_circ_mate *pm, int32_t *__restrict__ ret, cm_circ_res *cr,
                                                   uint8_t *__restrict__ remain_is_fullmap) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    ret[i] = cm_circ_check_row(v, remain_seq + (size_t) i * stride, remain_len[i], fullmap_seq + (size_t) i * stride, fullmap_len[i],
                               r1_partial[i] != 0, r1[i], r2[i], pm[i], cr[i], remain_is_fullmap + i);
}

static int circ_check_impl(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *remain_seq, const char *fullmap_seq, uint32_t stride,
                           const int32_t *remain_len, const int32_t *fullmap_len, const uint8_t *r1_partial, cm_circ_mate *mm_r1,
                           cm_circ_mate *mm_r2, cm_circ_mate *partial_mm, int window_size, int32_t *ret, cm_circ_res *cr, uint8_t *remain_is_fullmap) {
    if (!ctx) return CM_ERR_ARG;
    if (n == 0) return CM_OK;
    if (!remain_seq || !fullmap_seq || !remain_len || !fullmap_len || !r1_partial || !mm_r1 || !mm_r2 || !partial_mm || !ret || !cr)
        return fail(ctx, CM_ERR_ARG, "null argument");
    if (window_size < 4 || window_size > 12) return fail(ctx, CM_ERR_ARG, "window size outside 4..12");
    for (uint32_t i = 0; i < n; i++)
        if (remain_len[i] < 0 || fullmap_len[i] < 0 || (uint32_t) remain_len[i] > stride || (uint32_t) fullmap_len[i] > stride)
            return fail(ctx, CM_ERR_ARG, "problem %u: read longer than the row", i);
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig_id, true, v);
    if (r != CM_OK) return r;
    v.p.kmer = window_size;
    cudaStream_t st = ctx->stream;
    std::vector<void *> tmp;
    auto cleanup = [&]() { for (void *q : tmp) cudaFree(q); };
#define CIRC_TRY(x) do { if ((r = (x)) != CM_OK) { cleanup(); return r; } } while (0)
#define CIRC_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { cleanup(); return fail(ctx, CM_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e_)); } } while (0)
    uint8_t *d_rs = nullptr, *d_fs = nullptr, *d_rp = nullptr;
    int32_t *d_len = nullptr, *d_ret = nullptr;
    cm_circ_mate *d_m = nullptr;
    cm_circ_res *d_cr = nullptr;
    const size_t bytes = (size_t) n * stride;
    CIRC_TRY(dev_alloc(ctx, tmp, &d_rs, bytes + 64));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_fs, bytes + 64));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_rp, 2 * (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_len, 2 * (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_ret, (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_m, 3 * (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_cr, (size_t) n));
    CIRC_CUDA(cudaMemcpyAsync(d_rs, remain_seq, bytes, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_fs, fullmap_seq, bytes, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_rp, r1_partial, (size_t) n, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_len, remain_len, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_len + n, fullmap_len, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_m, mm_r1, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_m + n, mm_r2, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_m + 2 * (size_t) n, partial_mm, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyHostToDevice, st));
    k_circ_check<<<grid_for(n, 64), 64, 0, st>>>(v, n, d_rs, d_fs, stride, d_len, d_len + n, d_rp, d_m, d_m + n, d_m + 2 * (size_t) n, d_ret, d_cr, d_rp + n);
    ctx->launches++;
    CIRC_CUDA(cudaGetLastError());
    CIRC_CUDA(cudaMemcpyAsync(mm_r1, d_m, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(mm_r2, d_m + n, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(partial_mm, d_m + 2 * (size_t) n, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(ret, d_ret, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(cr, d_cr, (size_t) n * sizeof(cm_circ_res), cudaMemcpyDeviceToHost, st));
    if (remain_is_fullmap) CIRC_CUDA(cudaMemcpyAsync(remain_is_fullmap, d_rp + n, (size_t) n, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaStreamSynchronize(st));
    cleanup();
#undef CIRC_TRY
#undef CIRC_CUDA
    return CM_OK;
}
extern "C" int cm_gpu_circ_check_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *remain_seq, const char *fullmap_seq, uint32_t stride,
                                       const int32_t *remain_len, const int32_t *fullmap_len, const uint8_t *r1_partial, cm_circ_mate *mm_r1,
                                       cm_circ_mate *mm_r2, cm_circ_mate *partial_mm, int window_size, int32_t *ret, cm_circ_res *cr) {
    return circ_check_impl(ctx, contig_id, n, remain_seq, fullmap_seq, stride, remain_len, fullmap_len, r1_partial, mm_r1, mm_r2, partial_mm,
                           window_size, ret, cr, nullptr);
}

// ---- find_exact_coord with the pair's mates carried from chain to chain (cm_circ_exact_row) ----------------------------------
__global__ void __launch_bounds__(128) k_circ_exact_mates(const __grid_constant__ DevView v, const __grid_constant__ CircExactIn in, cm_circ_mate *r1,
                                                          cm_circ_mate *r2, cm_circ_mate *pm, int32_t *ret) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= in.n) return;
    ChainView bc;
    bc.frags = in.frags + (size_t) i * in.F;
    bc.chain_len = in.chain_len[i];
    bc.aux = nullptr;
    bc.spec = nullptr;
    ret[i] = cm_circ_exact_row(v, r1[i], r2[i], pm[i], in.dir[i], in.qspos[i], in.seqs + (size_t) i * in.stride, in.rlen[i], in.whole_len[i], bc);
}

extern "C" int cm_gpu_circ_exact_mates_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride, cm_circ_mate *mm_r1,
                                             cm_circ_mate *mm_r2, const int32_t *dir, const uint32_t *qspos, const int32_t *rlen,
                                             const int32_t *whole_len, uint32_t frag_slots, const uint32_t *chain_len, const uint32_t *frag_rpos,
                                             const int32_t *frag_qpos, int window_size, cm_circ_mate *partial_mm, int32_t *ret) {
    if (!ctx) return CM_ERR_ARG;
    if (n == 0) return CM_OK;
    if (!seqs || !mm_r1 || !mm_r2 || !dir || !qspos || !rlen || !whole_len || !chain_len || !frag_rpos || !frag_qpos || !partial_mm || !ret)
        return fail(ctx, CM_ERR_ARG, "null argument");
    if (window_size < 4 || window_size > 12 || frag_slots < 1) return fail(ctx, CM_ERR_ARG, "window size outside 4..12 or no fragment slots");
    for (uint32_t i = 0; i < n; i++) {
        if (chain_len[i] < 1 || chain_len[i] > frag_slots) return fail(ctx, CM_ERR_ARG, "problem %u: chain of %u fragments (slots: %u)", i, chain_len[i], frag_slots);
        if (qspos[i] < 1 || rlen[i] < 1 || qspos[i] - 1 + (uint32_t) rlen[i] > stride || whole_len[i] < 0 || (uint32_t) whole_len[i] > stride)
            return fail(ctx, CM_ERR_ARG, "problem %u: read range outside the row", i);
    }
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig_id, true, v);
    if (r != CM_OK) return r;
    v.p.kmer = window_size;
    cudaStream_t st = ctx->stream;
    std::vector<void *> tmp;
    auto cleanup = [&]() { for (void *q : tmp) cudaFree(q); };
#define CIRC_TRY(x) do { if ((r = (x)) != CM_OK) { cleanup(); return r; } } while (0)
#define CIRC_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { cleanup(); return fail(ctx, CM_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e_)); } } while (0)
    const size_t F = frag_slots;
    uint8_t *d_seq = nullptr;
    uint32_t *d_u = nullptr;
    FragDev *d_fr = nullptr;
    cm_circ_mate *d_mm = nullptr;
    int32_t *d_ret = nullptr;
    CIRC_TRY(dev_alloc(ctx, tmp, &d_seq, (size_t) n * stride + 64));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_u, (size_t) n * 5));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_fr, (size_t) n * F));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_mm, 3 * (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_ret, (size_t) n));
    std::vector<FragDev> h_fr((size_t) n * F);
    for (size_t q = 0; q < h_fr.size(); q++) { h_fr[q].rpos = frag_rpos[q]; h_fr[q].qpos = frag_qpos[q]; }
    const void *srcs[5] = {qspos, dir, rlen, whole_len, chain_len};
    CIRC_CUDA(cudaMemcpyAsync(d_seq, seqs, (size_t) n * stride, cudaMemcpyHostToDevice, st));
    for (int k = 0; k < 5; k++) CIRC_CUDA(cudaMemcpyAsync(d_u + (size_t) k * n, srcs[k], (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_fr, h_fr.data(), h_fr.size() * sizeof(FragDev), cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_mm, mm_r1, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_mm + n, mm_r2, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyHostToDevice, st));
    CircExactIn in{n, d_seq, stride, nullptr, nullptr, nullptr, nullptr, d_u, (const int32_t *) (d_u + n), (const int32_t *) (d_u + 2 * (size_t) n),
                   (const int32_t *) (d_u + 3 * (size_t) n), (int) F, d_u + 4 * (size_t) n, d_fr};
    k_circ_exact_mates<<<grid_for(n, 128), 128, 0, st>>>(v, in, d_mm, d_mm + n, d_mm + 2 * (size_t) n, d_ret);
    ctx->launches++;
    CIRC_CUDA(cudaGetLastError());
    CIRC_CUDA(cudaMemcpyAsync(mm_r1, d_mm, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(mm_r2, d_mm + n, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(partial_mm, d_mm + 2 * (size_t) n, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(ret, d_ret, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaStreamSynchronize(st));
    cleanup();
#undef CIRC_TRY
#undef CIRC_CUDA
    return CM_OK;
}

// ---- call_circ_single_split for a batch (cm_circ_orch.h over the three device entries) ---------------------------------------
#include "cm_circ_orch.h"
static int gpu_eng_chain(void *self, int contig, uint32_t n, const char *seqs, uint32_t stride, const uint32_t *qspos, const uint32_t *qepos,
                         const uint32_t *gs, const uint32_t *ge, int ws, int step, int F, int32_t *count, float *score, uint32_t *clen, uint32_t *rpos,
                         int32_t *qpos) {
    return cm_gpu_circ_chain_batch((cm_gpu_ctx *) self, contig, n, seqs, stride, qspos, qepos, gs, ge, ws, step, F, count, nullptr, score, clen, rpos, qpos);
}
static int gpu_eng_exact(void *self, int contig, uint32_t n, const char *seqs, uint32_t stride, cm_circ_mate *r1, cm_circ_mate *r2, const int32_t *dir,
                         const uint32_t *qspos, const int32_t *rlen, const int32_t *wl, uint32_t F, const uint32_t *clen, const uint32_t *rpos,
                         const int32_t *qpos, int ws, cm_circ_mate *pm, int32_t *ret) {
    return cm_gpu_circ_exact_mates_batch((cm_gpu_ctx *) self, contig, n, seqs, stride, r1, r2, dir, qspos, rlen, wl, F, clen, rpos, qpos, ws, pm, ret);
}
static int gpu_eng_check(void *self, int contig, uint32_t n, const char *rem, const char *ful, uint32_t stride, const int32_t *rl, const int32_t *fl,
                         const uint8_t *r1p, cm_circ_mate *r1, cm_circ_mate *r2, cm_circ_mate *pm, int ws, int32_t *ret, cm_circ_res *cr, uint8_t *sw) {
    return circ_check_impl((cm_gpu_ctx *) self, contig, n, rem, ful, stride, rl, fl, r1p, r1, r2, pm, ws, ret, cr, sw);
}

extern "C" int cm_gpu_circ_single_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride,
                                        const uint16_t *len1, const uint16_t *len2, const cm_mread *mr, const uint32_t *gene_off,
                                        const uint32_t *gene_start, const uint32_t *gene_end, int window_size, int step, uint32_t attempt_cap,
                                        uint32_t *n_attempts, uint32_t *att_pair, uint8_t *att_r1_partial, cm_circ_mate *att_in_r1,
                                        cm_circ_mate *att_in_r2, cm_circ_mate *att_in_pm, int32_t *att_ret, cm_circ_res *att_cr,
                                        cm_circ_mate *att_out_r1, cm_circ_mate *att_out_r2, cm_circ_mate *att_out_pm, int32_t *best_type,
                                        cm_circ_res *best_cr) {
    if (!ctx) return CM_ERR_ARG;
    if (n_attempts) *n_attempts = 0;
    if (n == 0) return CM_OK;
    if (!seqs1 || !seqs2 || !len1 || !len2 || !mr || !gene_off || !gene_start || !gene_end || !n_attempts || !best_type)
        return fail(ctx, CM_ERR_ARG, "null argument");
    if (step < 1 || window_size < 4 || window_size > 12) return fail(ctx, CM_ERR_ARG, "window size outside 4..12 or step < 1");
    for (uint32_t i = 0; i < n; i++)
        if (len1[i] > stride || len2[i] > stride || gene_off[i + 1] < gene_off[i]) return fail(ctx, CM_ERR_ARG, "pair %u: read longer than the row or gene list out of order", i);
    const CircEngine E{ctx, contig_id, gpu_eng_chain, gpu_eng_exact, gpu_eng_check};
    const CircSingleIn in{n, seqs1, seqs2, stride, len1, len2, mr, gene_off, gene_start, gene_end, window_size, step, ctx->p.max_read_len,
                          ctx->p.max_chain_len, 4};
    std::vector<CircAttempt> att;
    const int r = circ_orch_single(E, in, att, best_type, best_cr);
    if (r == -100) return fail(ctx, CM_ERR_STATE, "a fixed-size list of the stage-2 classification overflowed (CM_CIRC_TID_CAP / CM_CIRC_JUNC_CAP)");
    if (r != 0) return r;
    *n_attempts = (uint32_t) att.size();
    for (size_t k = 0; k < att.size() && k < attempt_cap; k++) {
        if (att_pair) att_pair[k] = att[k].pair;
        if (att_r1_partial) att_r1_partial[k] = att[k].r1_partial;
        if (att_in_r1) att_in_r1[k] = att[k].in_r1;
        if (att_in_r2) att_in_r2[k] = att[k].in_r2;
        if (att_in_pm) att_in_pm[k] = att[k].in_pm;
        if (att_ret) att_ret[k] = att[k].ret;
        if (att_cr) att_cr[k] = att[k].cr;
        if (att_out_r1) att_out_r1[k] = att[k].out_r1;
        if (att_out_r2) att_out_r2[k] = att[k].out_r2;
        if (att_out_pm) att_out_pm[k] = att[k].out_pm;
    }
    return CM_OK;
}

// ---- call_circ_double_split for a batch (circ_orch_double) -------------------------------------------------------------------
struct CircDoubleIn {
    uint32_t n;
    const uint8_t *r1seq, *r2seq;
    uint32_t stride;
    const uint32_t *qs1, *qe1, *qs2, *qe2, *clen1, *clen2;
    int F;
    const FragDev *fr1, *fr2;
};
__global__ void __launch_bounds__(128) k_circ_double(const __grid_constant__ DevView v, const __grid_constant__ CircDoubleIn in, cm_circ_mate *r1,
                                                     cm_circ_mate *r2, cm_circ_mate *p1, cm_circ_mate *p2, int32_t *code) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= in.n) return;
    ChainView c1, c2;
    c1.frags = in.fr1 + (size_t) i * in.F; c1.chain_len = in.clen1[i]; c1.aux = nullptr; c1.spec = nullptr;
    c2.frags = in.fr2 + (size_t) i * in.F; c2.chain_len = in.clen2[i]; c2.aux = nullptr; c2.spec = nullptr;
    code[i] = cm_circ_double_row(v, r1[i], r2[i], in.qs1[i], in.qe1[i], in.qs2[i], in.qe2[i], in.r1seq + (size_t) i * in.stride,
                                 in.r2seq + (size_t) i * in.stride, c1, c2, p1[i], p2[i]);
}
__global__ void __launch_bounds__(64) k_circ_check4(const __grid_constant__ DevView v, uint32_t n, const uint8_t *__restrict__ r1seq,
                                                    const uint8_t *__restrict__ r2seq, uint32_t stride, const int32_t *__restrict__ l1,
                                                    const int32_t *__restrict__ l2, const uint8_t *__restrict__ fullmap_sel,
                                                    const cm_circ_mate *__restrict__ m, int32_t *__restrict__ ret, cm_circ_res *cr) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint8_t *a = r1seq + (size_t) i * stride, *b = r2seq + (size_t) i * stride;
    const int sel = fullmap_sel ? fullmap_sel[i] : 0;   // what ProcessCirc::fullmap_seq holds: nothing, R1 or R2
    ret[i] = cm_circ_check4_row(v, a, l1[i], b, l2[i], sel == 1 ? a : sel == 2 ? b : nullptr, sel == 1 ? l1[i] : sel == 2 ? l2[i] : 0, m[i], m[n + i],
                                m[2 * (size_t) n + i], m[3 * (size_t) n + i], cr[i]);
}

#define CIRC_TRY(x) do { if ((r = (x)) != CM_OK) { cleanup(); return r; } } while (0)
#define CIRC_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { cleanup(); return fail(ctx, CM_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e_)); } } while (0)
static int gpu_eng_dbl(void *self, int contig, uint32_t n, const char *r1seq, const char *r2seq, uint32_t stride, cm_circ_mate *r1, cm_circ_mate *r2,
                       const uint32_t *qs1, const uint32_t *qe1, const uint32_t *qs2, const uint32_t *qe2, uint32_t F, const uint32_t *clen1,
                       const uint32_t *rpos1, const int32_t *qpos1, const uint32_t *clen2, const uint32_t *rpos2, const int32_t *qpos2, int ws,
                       cm_circ_mate *p1, cm_circ_mate *p2, int32_t *code) {
    cm_gpu_ctx *ctx = (cm_gpu_ctx *) self;
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig, true, v);
    if (r != CM_OK) return r;
    v.p.kmer = ws;
    cudaStream_t st = ctx->stream;
    std::vector<void *> tmp;
    auto cleanup = [&]() { for (void *q : tmp) cudaFree(q); };
    uint8_t *d_s1 = nullptr, *d_s2 = nullptr;
    uint32_t *d_u = nullptr;
    FragDev *d_f = nullptr;
    cm_circ_mate *d_m = nullptr;
    int32_t *d_code = nullptr;
    const size_t bytes = (size_t) n * stride;
    CIRC_TRY(dev_alloc(ctx, tmp, &d_s1, bytes + 64));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_s2, bytes + 64));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_u, 6 * (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_f, 2 * (size_t) n * F));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_m, 4 * (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_code, (size_t) n));
    std::vector<FragDev> hf(2 * (size_t) n * F);
    for (size_t q = 0; q < (size_t) n * F; q++) {
        hf[q].rpos = rpos1[q]; hf[q].qpos = qpos1[q];
        hf[(size_t) n * F + q].rpos = rpos2[q]; hf[(size_t) n * F + q].qpos = qpos2[q];
    }
    const void *srcs[6] = {qs1, qe1, qs2, qe2, clen1, clen2};
    CIRC_CUDA(cudaMemcpyAsync(d_s1, r1seq, bytes, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_s2, r2seq, bytes, cudaMemcpyHostToDevice, st));
    for (int k = 0; k < 6; k++) CIRC_CUDA(cudaMemcpyAsync(d_u + (size_t) k * n, srcs[k], (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_f, hf.data(), hf.size() * sizeof(FragDev), cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_m, r1, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_m + n, r2, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyHostToDevice, st));
    CircDoubleIn in{n, d_s1, d_s2, stride, d_u, d_u + n, d_u + 2 * (size_t) n, d_u + 3 * (size_t) n, d_u + 4 * (size_t) n, d_u + 5 * (size_t) n, (int) F,
                    d_f, d_f + (size_t) n * F};
    k_circ_double<<<grid_for(n, 128), 128, 0, st>>>(v, in, d_m, d_m + n, d_m + 2 * (size_t) n, d_m + 3 * (size_t) n, d_code);
    ctx->launches++;
    CIRC_CUDA(cudaGetLastError());
    CIRC_CUDA(cudaMemcpyAsync(r1, d_m, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(r2, d_m + n, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(p1, d_m + 2 * (size_t) n, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(p2, d_m + 3 * (size_t) n, (size_t) n * sizeof(cm_circ_mate), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(code, d_code, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaStreamSynchronize(st));
    cleanup();
    return CM_OK;
}
static int circ_check4_impl(cm_gpu_ctx *ctx, int contig, uint32_t n, const char *r1seq, const char *r2seq, uint32_t stride, const int32_t *l1,
                            const int32_t *l2, const uint8_t *fullmap_sel, const cm_circ_mate *r1, const cm_circ_mate *r2, const cm_circ_mate *p1,
                            const cm_circ_mate *p2, int ws, int32_t *ret, cm_circ_res *cr) {
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig, true, v);
    if (r != CM_OK) return r;
    v.p.kmer = ws;
    cudaStream_t st = ctx->stream;
    std::vector<void *> tmp;
    auto cleanup = [&]() { for (void *q : tmp) cudaFree(q); };
    uint8_t *d_s1 = nullptr, *d_s2 = nullptr;
    int32_t *d_l = nullptr, *d_ret = nullptr;
    cm_circ_mate *d_m = nullptr;
    cm_circ_res *d_cr = nullptr;
    const size_t bytes = (size_t) n * stride;
    CIRC_TRY(dev_alloc(ctx, tmp, &d_s1, bytes + 64));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_s2, bytes + 64));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_l, 2 * (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_ret, (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_m, 4 * (size_t) n));
    CIRC_TRY(dev_alloc(ctx, tmp, &d_cr, (size_t) n));
    uint8_t *d_sel = nullptr;
    if (fullmap_sel) {
        CIRC_TRY(dev_alloc(ctx, tmp, &d_sel, (size_t) n));
        CIRC_CUDA(cudaMemcpyAsync(d_sel, fullmap_sel, (size_t) n, cudaMemcpyHostToDevice, st));
    }
    CIRC_CUDA(cudaMemcpyAsync(d_s1, r1seq, bytes, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_s2, r2seq, bytes, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_l, l1, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    CIRC_CUDA(cudaMemcpyAsync(d_l + n, l2, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    const cm_circ_mate *src[4] = {r1, r2, p1, p2};
    for (int k = 0; k < 4; k++) CIRC_CUDA(cudaMemcpyAsync(d_m + (size_t) k * n, src[k], (size_t) n * sizeof(cm_circ_mate), cudaMemcpyHostToDevice, st));
    k_circ_check4<<<grid_for(n, 64), 64, 0, st>>>(v, n, d_s1, d_s2, stride, d_l, d_l + n, d_sel, d_m, d_ret, d_cr);
    ctx->launches++;
    CIRC_CUDA(cudaGetLastError());
    CIRC_CUDA(cudaMemcpyAsync(ret, d_ret, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaMemcpyAsync(cr, d_cr, (size_t) n * sizeof(cm_circ_res), cudaMemcpyDeviceToHost, st));
    CIRC_CUDA(cudaStreamSynchronize(st));
    cleanup();
    return CM_OK;
}
static int gpu_eng_check4(void *self, int contig, uint32_t n, const char *r1seq, const char *r2seq, uint32_t stride, const int32_t *l1, const int32_t *l2,
                          const cm_circ_mate *r1, const cm_circ_mate *r2, const cm_circ_mate *p1, const cm_circ_mate *p2, int ws, int32_t *ret,
                          cm_circ_res *cr) {
    return circ_check4_impl((cm_gpu_ctx *) self, contig, n, r1seq, r2seq, stride, l1, l2, nullptr, r1, r2, p1, p2, ws, ret, cr);
}
#undef CIRC_TRY
#undef CIRC_CUDA

extern "C" int cm_gpu_circ_check4_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *r1_seq, const char *r2_seq, uint32_t stride,
                                        const int32_t *r1_len, const int32_t *r2_len, const uint8_t *fullmap_sel, const cm_circ_mate *mm_r1_1,
                                        const cm_circ_mate *mm_r2_1, const cm_circ_mate *mm_r1_2, const cm_circ_mate *mm_r2_2, int window_size,
                                        int32_t *ret, cm_circ_res *cr) {
    if (!ctx) return CM_ERR_ARG;
    if (n == 0) return CM_OK;
    if (!r1_seq || !r2_seq || !r1_len || !r2_len || !mm_r1_1 || !mm_r2_1 || !mm_r1_2 || !mm_r2_2 || !ret || !cr) return fail(ctx, CM_ERR_ARG, "null argument");
    if (window_size < 4 || window_size > 12) return fail(ctx, CM_ERR_ARG, "window size outside 4..12");
    for (uint32_t i = 0; i < n; i++)
        if (r1_len[i] < 0 || r2_len[i] < 0 || (uint32_t) r1_len[i] > stride || (uint32_t) r2_len[i] > stride || (fullmap_sel && fullmap_sel[i] > 2))
            return fail(ctx, CM_ERR_ARG, "row %u: read longer than the row, or fullmap_sel outside 0..2", i);
    return circ_check4_impl(ctx, contig_id, n, r1_seq, r2_seq, stride, r1_len, r2_len, fullmap_sel, mm_r1_1, mm_r2_1, mm_r1_2, mm_r2_2, window_size, ret, cr);
}

extern "C" int cm_gpu_circ_double_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride,
                                        const uint16_t *len1, const uint16_t *len2, const cm_mread *mr, const uint32_t *gene_off,
                                        const uint32_t *gene_start, const uint32_t *gene_end, int window_size, int step, uint32_t event_cap,
                                        uint32_t *n_events, uint32_t *ev_pair, uint8_t *ev_kind, cm_circ_mate *ev_r1, cm_circ_mate *ev_r2,
                                        cm_circ_mate *ev_p1, cm_circ_mate *ev_p2, int32_t *ev_ret, cm_circ_res *ev_cr, int32_t *push_type,
                                        cm_circ_res *push_cr) {
    if (!ctx) return CM_ERR_ARG;
    if (n_events) *n_events = 0;
    if (n == 0) return CM_OK;
    if (!seqs1 || !seqs2 || !len1 || !len2 || !mr || !gene_off || !gene_start || !gene_end || !n_events || !push_type)
        return fail(ctx, CM_ERR_ARG, "null argument");
    if (step < 1 || window_size < 4 || window_size > 12) return fail(ctx, CM_ERR_ARG, "window size outside 4..12 or step < 1");
    for (uint32_t i = 0; i < n; i++)
        if (len1[i] > stride || len2[i] > stride || gene_off[i + 1] < gene_off[i]) return fail(ctx, CM_ERR_ARG, "pair %u: read longer than the row or gene list out of order", i);
    const CircEngine E{ctx, contig_id, gpu_eng_chain, gpu_eng_exact, gpu_eng_check, gpu_eng_dbl, gpu_eng_check4};
    const CircSingleIn in{n, seqs1, seqs2, stride, len1, len2, mr, gene_off, gene_start, gene_end, window_size, step, ctx->p.max_read_len,
                          ctx->p.max_chain_len, 4};
    std::vector<CircEvent> ev;
    const int r = circ_orch_double(E, in, ev, push_type, push_cr);
    if (r == -100) return fail(ctx, CM_ERR_STATE, "a fixed-size list of the stage-2 classification overflowed (CM_CIRC_TID_CAP / CM_CIRC_JUNC_CAP)");
    if (r != 0) return r;
    *n_events = (uint32_t) ev.size();
    for (size_t k = 0; k < ev.size() && k < event_cap; k++) {
        if (ev_pair) ev_pair[k] = ev[k].pair;
        if (ev_kind) ev_kind[k] = ev[k].kind;
        if (ev_r1) ev_r1[k] = ev[k].r1;
        if (ev_r2) ev_r2[k] = ev[k].r2;
        if (ev_p1) ev_p1[k] = ev[k].p1;
        if (ev_p2) ev_p2[k] = ev[k].p2;
        if (ev_ret) ev_ret[k] = ev[k].ret;
        if (ev_cr) ev_cr[k] = ev[k].cr;
    }
    return CM_OK;
}
