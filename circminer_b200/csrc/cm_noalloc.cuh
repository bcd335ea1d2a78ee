// cm_gpu_align_batch_noalloc: the alignment-kernel batch entry (cm_gpu_align_batch) for callers that may not allocate --
// code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
// Included by cm_gpu.cu (same translation unit as the array below).
//
// cm_gpu_align_batch stages its inputs and results in the context's growable scratch buffers (ENSURE -> cudaMalloc).  Here
// the staging area is ONE file-scope device array of CM_NOALLOC_SCRATCH_BYTES (4 MiB): every GPU that loads the library
// holds it until the process exits.  The entry point calls no cudaMalloc / cudaFree / allocating library routine; it
// rejects batches that do not fit.  Every call on a device shares the array and finds the previous call's contents in it:
// each call overwrites its input region and clears its result region before the kernel runs, and calls that overlap on
// the GPU (two contexts on one device, two host threads) race on it -- the caller serialises them.
#pragma once

#define CM_NOALLOC_SCRATCH_BYTES (4u << 20)
#define CM_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_noalloc_scratch[CM_NOALLOC_SCRATCH_BYTES];

// Device address of the array, once per device.  In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a
// device address: the address a kernel can use comes from cudaGetSymbolAddress with that device current.
static unsigned char *g_noalloc_ptr[CM_NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mu;

static int noalloc_scratch(cm_gpu_ctx *ctx, unsigned char **out) {
    if (ctx->device < 0 || ctx->device >= CM_NOALLOC_MAX_DEVICES) return fail(ctx, CM_ERR_ARG, "device %d beyond the %d the scratch table holds", ctx->device, CM_NOALLOC_MAX_DEVICES);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[ctx->device]) {
        void *p = nullptr;
        CUDA_TRY(ctx, cudaSetDevice(ctx->device));
        CUDA_TRY(ctx, cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail(ctx, CM_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        g_noalloc_ptr[ctx->device] = (unsigned char *) p;
    }
    *out = g_noalloc_ptr[ctx->device];
    return CM_OK;
}

static inline size_t noalloc_up(size_t x) { return (x + 255) & ~(size_t) 255; }

extern "C" size_t cm_gpu_noalloc_scratch_bytes(void) { return CM_NOALLOC_SCRATCH_BYTES; }

extern "C" int cm_gpu_align_batch_noalloc(cm_gpu_ctx *ctx, int kind, uint32_t n, const char *s, const uint16_t *n_s, const char *t,
                                          const uint16_t *n_t, uint32_t stride, const int32_t *aux, int32_t *out_ed, int32_t *out_sclen,
                                          int32_t *out_indel, int32_t *out_score) {
    if (!ctx) return CM_ERR_ARG;
    if (kind < 0 || kind > 4) return fail(ctx, CM_ERR_ARG, "alignment kind must be 0..4");
    if (n == 0) return CM_OK;
    if (!s || !t || !n_s || !n_t || !out_ed || !out_sclen || !out_indel || !out_score || (kind == 4 && !aux))
        return fail(ctx, CM_ERR_ARG, "null argument");
    if (stride == 0) return fail(ctx, CM_ERR_ARG, "stride is 0");
    // layout of the scratch array: s rows | t rows | n_s | n_t | aux | results (4 x n int32), each region 256-byte aligned
    const size_t bytes = (size_t) n * stride;
    const size_t o_s = 0, o_t = o_s + noalloc_up(bytes + 64), o_ns = o_t + noalloc_up(bytes + 64), o_nt = o_ns + noalloc_up((size_t) n * 2),
                 o_aux = o_nt + noalloc_up((size_t) n * 2), o_res = o_aux + noalloc_up((size_t) n * 4), need = o_res + noalloc_up((size_t) n * 16);
    if (need > CM_NOALLOC_SCRATCH_BYTES)
        return fail(ctx, CM_ERR_ARG, "batch of %u problems x stride %u needs %zu bytes of scratch, the fixed array holds %u: split the batch "
                    "(or use cm_gpu_align_batch, which allocates)", n, stride, need, CM_NOALLOC_SCRATCH_BYTES);
    for (uint32_t i = 0; i < n; i++) {
        if (n_s[i] > stride || n_t[i] > stride) return fail(ctx, CM_ERR_ARG, "problem %u longer than stride", i);
        if (kind == 4 && aux[i] > CM_MAX_BAND) return fail(ctx, CM_ERR_ARG, "problem %u: band %d > %d", i, aux[i], CM_MAX_BAND);
        bool full = (kind == 2 || kind == 3) ? (n_s[i] <= 2 * ctx->p.band_width || n_t[i] <= ctx->p.band_width)
                                             : (kind == 4 ? (aux[i] < 0 || n_s[i] <= aux[i]) : false);
        if (full && n_s[i] > CM_FULLDP_MAX) return fail(ctx, CM_ERR_ARG, "problem %u: unbanded s longer than %d", i, CM_FULLDP_MAX);
    }
    CUDA_TRY(ctx, cudaSetDevice(ctx->device));
    unsigned char *base = nullptr;
    int r = noalloc_scratch(ctx, &base);
    if (r != CM_OK) return r;
    cudaStream_t st = ctx->stream;
    CUDA_TRY(ctx, cudaMemcpyAsync(base + o_s, s, bytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(base + o_t, t, bytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(base + o_ns, n_s, (size_t) n * 2, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(base + o_nt, n_t, (size_t) n * 2, cudaMemcpyHostToDevice, st));
    if (aux) CUDA_TRY(ctx, cudaMemcpyAsync(base + o_aux, aux, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    else CUDA_TRY(ctx, cudaMemsetAsync(base + o_aux, 0, (size_t) n * 4, st));
    CUDA_TRY(ctx, cudaMemsetAsync(base + o_res, 0, (size_t) n * 16, st));   // the last call's results do not show through
    int32_t *o = (int32_t *) (base + o_res);
    k_align<<<grid_for(n, 128), 128, 0, st>>>(ctx->p, kind, n, base + o_s, (const uint16_t *) (base + o_ns), base + o_t,
                                              (const uint16_t *) (base + o_nt), stride, (const int32_t *) (base + o_aux), o, o + n,
                                              o + 2 * (size_t) n, o + 3 * (size_t) n);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(out_ed, o, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(out_sclen, o + n, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(out_indel, o + 2 * (size_t) n, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(out_score, o + 3 * (size_t) n, (size_t) n * 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return CM_OK;
}
