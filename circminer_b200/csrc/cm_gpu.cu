// B200 (sm_100a) implementation of the C-ABI in include/cm_gpu.h.
// Host runtime (context, HBM-resident index/annotation, batch staging) + kernels.
// No CPU fallback: every entry point needs a CUDA device and fails with CM_ERR_CUDA otherwise.
#include <cuda_runtime.h>
#include <cub/cub.cuh>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>
#include <thread>
#include <atomic>
#include <mutex>
#include <chrono>

#include "cm_pipeline.h"

#define CM_CTR_SLOTS 256   // copies of cm_counters on the device (power of two), summed on read
#define CM_MAX_ROUNDS 64   // packed contigs one cm_gpu_map_rounds_pe call may visit

// ------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
};

struct ContigHost {
    bool loaded = false;
    ContigDev d{};
    bool has_annot = false;
    AnnotDev a{};
    std::vector<void *> owned;        // device allocations of this contig
    std::vector<void *> owned_annot;
    uint32_t n_big = 0;               // buckets with a coarse checksum table
    // stage 2 (cm_circ.cuh): every ws-mer position of the contig sorted by (hash, position); in `owned`
    uint32_t *circ_pos = nullptr, *circ_off = nullptr;
    int circ_ws = 0;
};

struct cm_gpu_ctx {
    int device = 0;
    cm_params p{};
    int max_seg_cnt = 0, n_slots = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[7]{};   // [0..3] stage boundaries, [4..6] around k_chain_spec / k_chain_walk
    std::string err;
    std::vector<ContigHost> contigs;
    cm_gpu_ctx *parent = nullptr;     // cm_gpu_create_shared: index and annotation live in (and are owned by) the parent
    std::atomic<int> n_children{0};   // shared contexts alive on this root: resident contigs cannot be replaced meanwhile
    cudaStream_t stream2 = nullptr;   // side stream: kernels that run beside the main one (heavy strands)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cm_counters *d_counters = nullptr;
    uint64_t launches = 0;
    float stage_ms[5] = {0, 0, 0, 0, 0};
    bool align_timed = false;   // ev[4..6] were recorded by the last chain stage

    // staged batch (device resident)
    uint32_t n_pairs = 0;      // pairs (PE) or reads (SE)
    uint32_t n_strands = 0;    // 4 * pairs (PE) / 2 * reads (SE)
    uint32_t stride = 0;       // device row stride of strand rows
    bool staged_pe = false;
    DevBuf in_seq1, in_seq2, in_len1, in_len2, in_mread;  // raw uploads
    DevBuf strands, strand_len;                           // [n_strands][stride] fwd/rc rows, lengths
    DevBuf seed_key;                                      // [n_strands][batch_F] (hv, cv) of every seed slot
    DevBuf seed_off, seed_cnt;                            // [n_strands][batch_F]
    DevBuf cell_cnt, cell_off;                            // dp cells per strand, exclusive prefix (u64)
    DevBuf dp_score, dp_prev, recs;                       // chaining arenas
    DevBuf ovf_strand, ovf_nrec, ovf_n, ovf_off, ovf_recs;  // strands whose record list outgrew its slice
    DevBuf chain_cnt, chain_hh, chain_score, chain_len, chain_frags;   // chains per strand
    int batch_F = 1;           // fragment slots per chain for the staged batch
    uint32_t rec_mult = 3;     // record-arena slice = rec_mult * cells + 8 (CM_GPU_REC_MULT overrides; 0 forces the overflow path)
    uint32_t last_overflow = 0;
    DevBuf out_mread, out_state;
    DevBuf cub_tmp;
    DevBuf chain_spec;                                    // transcript-walk DPs of every pre-aligned chain end
    DevBuf chain_base, chain_ids, chain_aux;              // compact chain list + pre-alignment results
    uint32_t last_chains = 0;
    bool sync_each = false;    // CM_GPU_SYNC_EACH=1: synchronise after every launch and name the kernel that failed
    uint32_t heavy_th = 64;    // CM_GPU_HEAVY_TH: strands with more dp cells than this are chained by one warp each (< 255)
    uint32_t last_heavy = 0;
    // CM_GPU_HEAVY_PAIR_TH: pairs whose weight (candidate chain pairs of both attempts + chains of the four strands, i.e.
    // the MatePairs process_mates may extend plus the unpaired chains it extends afterwards) exceeds this get one warp
    // each: candidates and leftover chains run one per lane.  On a GRCh38-scale batch a tenth of the pairs carry 30 chains
    // per strand and, one thread each, were 2/3 of the step (extend stage 236 -> 81 ms per million pairs).
    uint32_t heavy_pair_th = 40;
    int pair_sig = 2;
    bool heavy_pair_forced = false;
    uint32_t last_heavy_pairs = 0;
    int spec_k = CM_MAX_CHAINS;            // CM_GPU_SPEC_K: leading chains per strand handed to the dense pre-alignment kernel
    bool no_walk_spec = false; // CM_GPU_NO_WALK_SPEC=1: pre-align the genome-text ends only (A-B)
    bool no_spec = false;      // CM_GPU_NO_SPEC=1: every DP inline in the per-pair kernel (debug / A-B)
    DevBuf spec_key[2], spec_item[2];                     // pre-alignment work items, before / after the cost sort
    DevBuf key32[2], key64[2], perm[2];                   // work-signature sort of strands / pairs
    DevBuf scratch[8];
    // rounds on the device (cm_gpu_map_rounds_pe): two compacted batches used in turn, per-pair final results
    DevBuf rb_strands[2], rb_slen[2], rb_key[2], rb_in[2], rb_ids[2];
    DevBuf fin_mread, fin_state, fin_round, surv_flag, surv_idx, surv_n, snap_mread, snap_state, round_out, round_state;
    cm_mread *out_view = nullptr;              // where launch_process writes when a round runs in chunks (else out_mread / out_state)
    int32_t *state_view = nullptr;
    uint32_t chunk_pairs = 1u << 20;           // CM_GPU_CHUNK: pairs per pipeline pass inside cm_gpu_map_rounds_pe
    cudaEvent_t ev_r[2]{};
    float rounds_ms[6] = {0, 0, 0, 0, 0, 0};   // stage sums over the rounds, [5] the whole run
    float rounds_align_ms[2] = {0, 0};
    uint32_t round_pairs[CM_MAX_ROUNDS] = {0};
    int rounds_run = 0;
    uint32_t rounds_n = 0;                     // pairs of the batch the final arrays describe
};

static thread_local std::string g_create_err;   // last cm_gpu_create failure of the calling thread

static int fail(cm_gpu_ctx *ctx, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf; else g_create_err = buf;
    return code;
}

#define CUDA_TRY(ctx, call)                                                                         \
    do {                                                                                            \
        cudaError_t e_ = (call);                                                                    \
        if (e_ != cudaSuccess)                                                                      \
            return fail(ctx, (e_ == cudaErrorMemoryAllocation) ? CM_ERR_ALLOC : CM_ERR_CUDA,        \
                        "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_));         \
    } while (0)

static int ensure(cm_gpu_ctx *ctx, DevBuf &b, size_t bytes) {
    if (bytes <= b.cap) return CM_OK;
    if (b.p) CUDA_TRY(ctx, cudaFree(b.p));
    b.p = nullptr;
    b.cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    CUDA_TRY(ctx, cudaMalloc(&b.p, want));
    b.cap = want;
    return CM_OK;
}
#define ENSURE(ctx, buf, bytes)                      \
    do {                                             \
        int r_ = ensure(ctx, buf, bytes);            \
        if (r_ != CM_OK) return r_;                  \
    } while (0)

static void release(DevBuf &b) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
}

static int dbg_sync(cm_gpu_ctx *ctx, const char *what);
#define DBG_SYNC(ctx, what)                                  \
    do {                                                     \
        int r_ = dbg_sync(ctx, what);                        \
        if (r_ != CM_OK) return r_;                          \
    } while (0)

static inline unsigned grid_for(size_t n, unsigned block) { return (unsigned) ((n + block - 1) / block); }

static int dbg_sync(cm_gpu_ctx *ctx, const char *what) {
    if (!ctx->sync_each) return CM_OK;
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return fail(ctx, CM_ERR_CUDA, "after %s: %s", what, cudaGetErrorString(e));
    return CM_OK;
}

extern "C" void cm_params_default(cm_params *p) {
    p->kmer = 20; p->max_read_len = 300; p->max_ed = 4; p->max_sc = 7; p->band_width = 3; p->seed_lim = 500;
    p->max_tlen = 500; p->max_intron_len = 2000000; p->max_chain_len = 30; p->scan_level = 0;
    p->mat_sc = 1; p->mis_sc = -3; p->ind_sc = -3; p->xdrop = 8;
}

extern "C" void cm_mread_init(cm_mread *m, const cm_params *p) {
    memset(m, 0, sizeof *m);
    m->r1_forward = 1; m->r2_forward = 1;
    m->ed_r1 = p->max_ed + 1; m->ed_r2 = p->max_ed + 1;
    m->type = CM_NOPROC_NOMATCH; m->tlen = CM_INF;
    m->chr_r1 = -1; m->chr_r2 = -1;
}

extern "C" const char *cm_gpu_last_error(const cm_gpu_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

__global__ void k_init_prep_lut();   // tables of the read staging kernel, below

static void destroy_ctx_resources(cm_gpu_ctx *ctx) {
    if (ctx->d_counters) cudaFree(ctx->d_counters);
    for (auto &ev : ctx->ev) if (ev) cudaEventDestroy(ev);
    for (auto &ev : ctx->ev_r) if (ev) cudaEventDestroy(ev);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    ctx->d_counters = nullptr;
    ctx->stream = ctx->stream2 = nullptr;
}

static int create_ctx(int device, const cm_params *params, cm_gpu_ctx **out, cm_gpu_ctx *shared_of);

extern "C" int cm_gpu_create(int device, const cm_params *params, cm_gpu_ctx **out) { return create_ctx(device, params, out, nullptr); }

static int create_ctx(int device, const cm_params *params, cm_gpu_ctx **out, cm_gpu_ctx *shared_of) {
    if (!params || !out) return fail(nullptr, CM_ERR_ARG, "null argument");
    *out = nullptr;
    const cm_params &p = *params;
    if (p.kmer < CM_WINDOW || p.kmer > CM_WINDOW + 8)
        return fail(nullptr, CM_ERR_ARG, "kmer must be in [14..22] (commandline_parser.cpp:243)");
    if (p.band_width < 0 || p.band_width > CM_MAX_BAND) return fail(nullptr, CM_ERR_ARG, "band_width must be in [0..%d]", CM_MAX_BAND);
    if (p.max_chain_len < 1 || p.max_chain_len > CM_MAX_CHAINS) return fail(nullptr, CM_ERR_ARG, "max_chain_len must be in [1..%d]", CM_MAX_CHAINS);
    if (p.max_read_len < p.kmer || p.max_read_len > 600) return fail(nullptr, CM_ERR_ARG, "max_read_len must be in [kmer..600]");
    if ((cm_max_seg_cnt(p) + 1) / 2 > CM_MAX_SEEDS) return fail(nullptr, CM_ERR_ARG, "too many seeds per read");
    if (p.seed_lim < 1 || p.seed_lim > 65535) return fail(nullptr, CM_ERR_ARG, "seed_lim must be in [1..65535]");
    if (p.max_ed < 0 || p.max_sc < 0 || p.max_ed > 64 || p.max_sc > 64) return fail(nullptr, CM_ERR_ARG, "max_ed/max_sc out of range");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0)
        return fail(nullptr, CM_ERR_CUDA, "no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(nullptr, CM_ERR_ARG, "device %d out of range (%d devices)", device, ndev);
    cm_gpu_ctx *ctx = new cm_gpu_ctx;
    ctx->device = device;
    ctx->p = p;
    ctx->max_seg_cnt = cm_max_seg_cnt(p);
    ctx->n_slots = (ctx->max_seg_cnt + 1) / 2;
    if (cudaSetDevice(device) != cudaSuccess) {
        delete ctx;
        return fail(nullptr, CM_ERR_CUDA, "cannot initialise device %d", device);
    }
    bool ok = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking) == cudaSuccess;
    for (auto &ev : ctx->ev) ok = ok && cudaEventCreate(&ev) == cudaSuccess;
    for (auto &ev : ctx->ev_r) ok = ok && cudaEventCreate(&ev) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming) == cudaSuccess;
    ok = ok && cudaMalloc(&ctx->d_counters, CM_CTR_SLOTS * sizeof(cm_counters)) == cudaSuccess;
    if (ok && !shared_of) {   // staging tables: written once per root context (a shared context reads its parent's)
        k_init_prep_lut<<<1, 256, 0, ctx->stream>>>();
        ok = cudaGetLastError() == cudaSuccess;
    }
    ok = ok && cudaMemsetAsync(ctx->d_counters, 0, CM_CTR_SLOTS * sizeof(cm_counters), ctx->stream) == cudaSuccess;
    ok = ok && cudaStreamSynchronize(ctx->stream) == cudaSuccess;
    if (!ok) {
        const cudaError_t ce = cudaGetLastError();
        destroy_ctx_resources(ctx);
        delete ctx;
        return fail(nullptr, CM_ERR_CUDA, "cannot create the context's streams / events / counters on device %d (%s)", device, cudaGetErrorString(ce));
    }
    if (const char *e = getenv("CM_GPU_REC_MULT")) ctx->rec_mult = (uint32_t) atoi(e);
    if (const char *e = getenv("CM_GPU_HEAVY_TH")) { int t = atoi(e); ctx->heavy_th = (uint32_t) (t < 0 ? 0 : (t > 254 ? 254 : t)); }
    if (const char *e = getenv("CM_GPU_PAIR_SIG")) ctx->pair_sig = atoi(e);
    if (const char *e = getenv("CM_GPU_HEAVY_PAIR_TH")) { ctx->heavy_pair_th = (uint32_t) atol(e); ctx->heavy_pair_forced = true; }
    if (const char *e = getenv("CM_GPU_SPEC_K")) ctx->spec_k = atoi(e) < 1 ? 1 : atoi(e);
    if (const char *e = getenv("CM_GPU_NO_WALK_SPEC")) ctx->no_walk_spec = atoi(e) != 0;
    if (const char *e = getenv("CM_GPU_NO_SPEC")) ctx->no_spec = atoi(e) != 0;
    if (const char *e = getenv("CM_GPU_SYNC_EACH")) ctx->sync_each = atoi(e) != 0;
    if (const char *e = getenv("CM_GPU_CHUNK")) { long c = atol(e); ctx->chunk_pairs = (uint32_t) (c < 1024 ? 1024 : c); }
    *out = ctx;
    return CM_OK;
}

// A second context on the same device that shares the parent's resident contigs (index, genome, annotation; read-only
// on this path) and owns its own stream and batch buffers: two host threads alternating batches over the two contexts
// overlap one batch's host<->device copies with the other's kernels.
extern "C" int cm_gpu_create_shared(cm_gpu_ctx *parent, cm_gpu_ctx **out) {
    if (!parent || !out) return CM_ERR_ARG;
    if (parent->parent) return fail(parent, CM_ERR_ARG, "the parent of a shared context must be a root context");
    int r = create_ctx(parent->device, &parent->p, out, parent);
    if (r != CM_OK) return r;
    (*out)->parent = parent;
    parent->n_children++;
    return CM_OK;
}

static void free_contig(ContigHost &c) {
    for (void *p : c.owned) cudaFree(p);
    for (void *p : c.owned_annot) cudaFree(p);
    c = ContigHost();
}
// the index loaders replace the index only: an annotation uploaded before (or kept across a reload) stays
static void free_index(ContigHost &c) {
    for (void *p : c.owned) cudaFree(p);
    c.owned.clear();
    c.d = ContigDev();
    c.loaded = false;
    c.n_big = 0;
    c.circ_pos = c.circ_off = nullptr;
    c.circ_ws = 0;
}
// a contig that shared contexts may be reading cannot be replaced under them
static int check_reload(cm_gpu_ctx *ctx, ContigHost *ch, int contig_id) {
    if ((ch->loaded || ch->has_annot) && ctx->n_children.load() > 0)
        return fail(ctx, CM_ERR_STATE, "contig %d is resident and %d shared context(s) may be using it: destroy them before reloading", contig_id,
                    ctx->n_children.load());
    return CM_OK;
}

extern "C" int cm_gpu_destroy(cm_gpu_ctx *ctx) {
    if (!ctx) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (auto &c : ctx->contigs) free_contig(c);
    DevBuf *bufs[] = {&ctx->in_seq1, &ctx->in_seq2, &ctx->in_len1, &ctx->in_len2, &ctx->in_mread, &ctx->strands,
                      &ctx->strand_len, &ctx->seed_key, &ctx->seed_off, &ctx->seed_cnt, &ctx->cell_cnt, &ctx->cell_off,
                      &ctx->dp_score, &ctx->dp_prev, &ctx->recs, &ctx->ovf_strand, &ctx->ovf_nrec, &ctx->ovf_n, &ctx->ovf_off,
                      &ctx->ovf_recs, &ctx->chain_cnt, &ctx->chain_hh, &ctx->chain_score, &ctx->chain_len,
                      &ctx->chain_frags, &ctx->out_mread, &ctx->out_state, &ctx->cub_tmp};
    for (DevBuf *b : bufs) release(*b);
    for (auto &b : ctx->scratch) release(b);
    release(ctx->chain_spec); release(ctx->chain_base); release(ctx->chain_ids); release(ctx->chain_aux);
    for (int q = 0; q < 2; q++) { release(ctx->spec_key[q]); release(ctx->spec_item[q]); }
    for (int q = 0; q < 2; q++) { release(ctx->key32[q]); release(ctx->key64[q]); release(ctx->perm[q]); }
    for (int q = 0; q < 2; q++) { release(ctx->rb_strands[q]); release(ctx->rb_slen[q]); release(ctx->rb_key[q]); release(ctx->rb_in[q]); release(ctx->rb_ids[q]); }
    DevBuf *rbufs[] = {&ctx->fin_mread, &ctx->fin_state, &ctx->fin_round, &ctx->surv_flag, &ctx->surv_idx, &ctx->surv_n, &ctx->snap_mread, &ctx->snap_state,
                       &ctx->round_out, &ctx->round_state};
    for (DevBuf *b : rbufs) release(*b);
    if (ctx->parent) ctx->parent->n_children--;
    destroy_ctx_resources(ctx);
    delete ctx;
    return CM_OK;
}

extern "C" int cm_gpu_get_counters(cm_gpu_ctx *ctx, cm_counters *out, int reset) {
    if (!ctx || !out) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    static_assert(sizeof(cm_counters) % 8 == 0, "cm_counters is an array of u64");
    std::vector<uint64_t> slots(CM_CTR_SLOTS * sizeof(cm_counters) / 8);
    CUDA_TRY(ctx, cudaMemcpy(slots.data(), ctx->d_counters, slots.size() * 8, cudaMemcpyDeviceToHost));
    uint64_t *o = (uint64_t *) out;
    const size_t nf = sizeof(cm_counters) / 8;
    for (size_t f = 0; f < nf; f++) o[f] = 0;
    for (size_t k = 0; k < slots.size(); k++) o[k % nf] += slots[k];
    out->kernel_launches = ctx->launches;
    if (reset) {
        // on the context's own (non-blocking) stream: a legacy-stream memset is not ordered against later launches
        CUDA_TRY(ctx, cudaMemsetAsync(ctx->d_counters, 0, CM_CTR_SLOTS * sizeof(cm_counters), ctx->stream));
        CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
        ctx->launches = 0;
    }
    return CM_OK;
}

extern "C" int cm_gpu_last_stage_ms(cm_gpu_ctx *ctx, float out[5]) {
    if (!ctx || !out) return CM_ERR_ARG;
    for (int i = 0; i < 5; i++) out[i] = ctx->stage_ms[i];
    return CM_OK;
}

extern "C" int cm_gpu_device_count(void) {
    int n = 0;
    return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}

static void preload_kernels();   // end of this file: the kernels are defined below

extern "C" int cm_gpu_warmup(int device) {
    if (cudaSetDevice(device) != cudaSuccess) return CM_ERR_CUDA;
    if (cudaFree(nullptr) != cudaSuccess) return CM_ERR_CUDA;
    preload_kernels();   // CUDA loads a kernel's code at its first launch: take that out of the first batch too
    return CM_OK;
}

extern "C" int cm_gpu_host_alloc(size_t bytes, void **out) {
    if (!out) return CM_ERR_ARG;
    *out = nullptr;
    return cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocPortable) == cudaSuccess ? CM_OK : CM_ERR_ALLOC;
}

extern "C" int cm_gpu_host_free(void *p) {
    return !p || cudaFreeHost(p) == cudaSuccess ? CM_OK : CM_ERR_CUDA;
}

extern "C" int cm_gpu_last_align_ms(cm_gpu_ctx *ctx, float out[2]) {
    if (!ctx || !out) return CM_ERR_ARG;
    out[0] = out[1] = 0.f;
    if (!ctx->align_timed) return CM_OK;
    if (cudaEventSynchronize(ctx->ev[6]) != cudaSuccess) return CM_ERR_CUDA;
    cudaEventElapsedTime(&out[0], ctx->ev[4], ctx->ev[5]);
    cudaEventElapsedTime(&out[1], ctx->ev[5], ctx->ev[6]);
    return CM_OK;
}

static ContigHost *get_contig(cm_gpu_ctx *ctx, int contig_id, bool create) {
    if (contig_id < 0 || contig_id > 4095) return nullptr;
    if (ctx->parent) return create ? nullptr : get_contig(ctx->parent, contig_id, false);   // shared contexts only read
    if ((size_t) contig_id >= ctx->contigs.size()) {
        if (!create) return nullptr;
        ctx->contigs.resize(contig_id + 1);
    }
    return &ctx->contigs[contig_id];
}

struct ContigHost;
static int build_dirx(cm_gpu_ctx *ctx, ContigHost *ch);

template <typename T>
static int dev_alloc(cm_gpu_ctx *ctx, std::vector<void *> &owned, T **out, size_t count) {
    void *p = nullptr;
    CUDA_TRY(ctx, cudaMalloc(&p, (count ? count : 1) * sizeof(T)));
    owned.push_back(p);
    *out = (T *) p;
    return CM_OK;
}

template <typename T>
static int dev_upload(cm_gpu_ctx *ctx, std::vector<void *> &owned, const T **out, const T *src, size_t count) {
    T *p = nullptr;
    int r = dev_alloc(ctx, owned, &p, count);
    if (r != CM_OK) return r;
    if (count) CUDA_TRY(ctx, cudaMemcpy(p, src, count * sizeof(T), cudaMemcpyHostToDevice));
    *out = p;
    return CM_OK;
}

// ------------------------------------------------------------------------------------------
// index: upload of the reference's in-memory layout
// ------------------------------------------------------------------------------------------
#pragma pack(push, 1)
struct GeneralIndexPacked {   // lib/mrsfast/Common.h:105-111
    uint16_t checksum;
    uint32_t info;
};
#pragma pack(pop)

// 3-bit packed genome (21 bases / word, MSB first; Common.c:174-196) -> ASCII at genome[1..len]
__global__ void k_unpack_cref(const uint64_t *__restrict__ cref, uint32_t len, uint8_t *__restrict__ genome) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;   // 0-based base index
    if (i >= len) return;
    uint64_t w = cref[i / 21];
    int v = (int) ((w >> (60 - 3 * (i % 21))) & 7);
    const char lut[8] = {'A', 'C', 'G', 'T', 'N', 'N', 'N', 'N'};   // GenomeSeeder::init, match_read.cpp:15-23
    genome[i + 1] = lut[v];
}

// host-side helpers of the loader: uninitialised buffers (no 1 GiB zero fill) and a fork-join over host threads
template <class T> struct HostBuf {
    T *p = nullptr;
    size_t n = 0;
    explicit HostBuf(size_t n_) : p((T *) malloc((n_ ? n_ : 1) * sizeof(T))), n(n_) {}
    ~HostBuf() { free(p); }
    HostBuf(const HostBuf &) = delete;
    HostBuf &operator=(const HostBuf &) = delete;
};
template <class F> static void host_parallel_for(int T, F &&fn) {
    std::vector<std::thread> th;
    for (int t = 1; t < T; t++) th.emplace_back([&fn, t] { fn(t); });
    fn(0);
    for (auto &x : th) x.join();
}

extern "C" int cm_gpu_load_contig(cm_gpu_ctx *ctx, int contig_id, const void *const *bucket_lists, const void *table_mem,
                                  uint64_t table_mem_entries, const uint64_t *cref_gen, uint32_t ref_gen_len) {
    if (!ctx) return CM_ERR_ARG;
    if (!bucket_lists || !table_mem || !cref_gen) return fail(ctx, CM_ERR_ARG, "null argument");
    cudaSetDevice(ctx->device);
    ContigHost *ch = get_contig(ctx, contig_id, true);
    if (!ch) return fail(ctx, CM_ERR_ARG, ctx->parent ? "contig %d: a shared context cannot load (load into its parent)" : "bad contig id %d", contig_id);
    { int rr = check_reload(ctx, ch, contig_id); if (rr != CM_OK) return rr; }
    free_index(*ch);
    const GeneralIndexPacked *mem = (const GeneralIndexPacked *) table_mem;
    // directory + SoA split on the host (setHashTablePointers order, HashTable.c:1074), by all host threads: the
    // 2 GiB pointer table is cut into slices, every slice counts its entries, a prefix over the slices gives each
    // one its first entry, then the slices fill their part of the directory and of the two arrays.
    const int T = (int) std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 64u);
    HostBuf<uint32_t> dir((size_t) CM_NBUCKETS + 1);
    if (!dir.p) return fail(ctx, CM_ERR_ALLOC, "host memory for the directory");
    std::vector<uint64_t> slice_total(T + 1, 0);
    std::atomic<uint32_t> bad_bucket(0xFFFFFFFFu);
    host_parallel_for(T, [&](int t) {
        const uint32_t h0 = (uint32_t) ((uint64_t) CM_NBUCKETS * t / T), h1 = (uint32_t) ((uint64_t) CM_NBUCKETS * (t + 1) / T);
        uint64_t sum = 0;
        for (uint32_t hv = h0; hv < h1; hv++) {
            const GeneralIndexPacked *l = (const GeneralIndexPacked *) bucket_lists[hv];
            if (!l) continue;
            if (l < mem || (uint64_t) (l - mem) + l[0].info + 1 > table_mem_entries) { bad_bucket.store(hv); return; }
            sum += l[0].info;
        }
        slice_total[t + 1] = sum;
    });
    if (bad_bucket.load() != 0xFFFFFFFFu) return fail(ctx, CM_ERR_ARG, "bucket %u points outside table_mem", bad_bucket.load());
    for (int t = 0; t < T; t++) slice_total[t + 1] += slice_total[t];
    const uint64_t total = slice_total[T];
    if (total > 0xFFFFFFFFull) return fail(ctx, CM_ERR_ARG, "more than 2^32 index entries");
    dir.p[CM_NBUCKETS] = (uint32_t) total;
    HostBuf<uint16_t> cks(total);
    HostBuf<uint32_t> pos(total);
    if (total && (!cks.p || !pos.p)) return fail(ctx, CM_ERR_ALLOC, "host memory for the index entries");
    host_parallel_for(T, [&](int t) {
        const uint32_t h0 = (uint32_t) ((uint64_t) CM_NBUCKETS * t / T), h1 = (uint32_t) ((uint64_t) CM_NBUCKETS * (t + 1) / T);
        uint32_t o = (uint32_t) slice_total[t];
        for (uint32_t hv = h0; hv < h1; hv++) {
            dir.p[hv] = o;
            const GeneralIndexPacked *l = (const GeneralIndexPacked *) bucket_lists[hv];
            if (!l) continue;
            const uint32_t n = l[0].info;
            for (uint32_t j = 0; j < n; j++) {
                cks.p[o + j] = l[1 + j].checksum;
                pos.p[o + j] = l[1 + j].info;
            }
            o += n;
        }
    });
    int r;
    if ((r = dev_upload(ctx, ch->owned, &ch->d.dir, dir.p, dir.n)) != CM_OK) return r;
    if ((r = dev_upload(ctx, ch->owned, &ch->d.cks, cks.p, cks.n)) != CM_OK) return r;
    if ((r = dev_upload(ctx, ch->owned, &ch->d.pos, pos.p, pos.n)) != CM_OK) return r;
    size_t nwords = (size_t) ref_gen_len / 21 + 1;
    const uint64_t *d_cref = nullptr;
    std::vector<void *> tmp;
    if ((r = dev_upload(ctx, tmp, &d_cref, cref_gen, (ref_gen_len + 20) / 21)) != CM_OK) return r;
    (void) nwords;
    uint8_t *genome = nullptr;
    if ((r = dev_alloc(ctx, ch->owned, &genome, (size_t) ref_gen_len + 64)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaMemsetAsync(genome, 'N', (size_t) ref_gen_len + 64, ctx->stream));
    if (ref_gen_len) {
        k_unpack_cref<<<grid_for(ref_gen_len, 256), 256, 0, ctx->stream>>>(d_cref, ref_gen_len, genome);
        ctx->launches++;
    }
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    for (void *p : tmp) cudaFree(p);
    ch->d.genome = genome;
    ch->d.n_entries = total;
    ch->d.ref_len = ref_gen_len;
    if ((r = build_dirx(ctx, ch)) != CM_OK) return r;
    ch->loaded = true;
    return CM_OK;
}

// ------------------------------------------------------------------------------------------
// index: device-side build (count -> scan -> scatter/sort), HashTable.c:769-839 semantics
// ------------------------------------------------------------------------------------------
// key = hv (28 bits) << 16 | checksum (<= 16 bits); one thread per k-mer start (0-based i, loc = i+1)
__global__ void k_kmer_keys(const uint8_t *__restrict__ genome /*1-based*/, uint32_t len, int kmer,
                            uint64_t *__restrict__ keys, uint32_t *__restrict__ locs, uint8_t *__restrict__ valid) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= len) return;
    uint8_t ok = 0;
    uint64_t key = 0;
    if (i + kmer <= len) {
        ok = 1;
        for (int j = 0; j < kmer; j++) {
            int v = cm_nhval(genome[1 + i + j]);
            if (v < 0) { ok = 0; break; }
            key = (key << 2) | (uint64_t) v;
        }
        if (ok) {
            int cs = kmer - CM_WINDOW;
            uint64_t hv = key >> (2 * cs);
            uint64_t cv = key & ((1ull << (2 * cs)) - 1);
            key = (hv << 16) | cv;
        }
    }
    keys[i] = key;
    locs[i] = i + 1;
    valid[i] = ok;
}

__global__ void k_bucket_hist(const uint64_t *__restrict__ keys, uint64_t n, uint32_t *__restrict__ dir) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    atomicAdd(&dir[(uint32_t) (keys[i] >> 16)], 1u);
}

__global__ void k_split_keys(const uint64_t *__restrict__ keys, uint64_t n, uint16_t *__restrict__ cks) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    cks[i] = (uint16_t) (keys[i] & 0xFFFF);
}

__global__ void k_ascii_to_genome(const char *__restrict__ seq, uint32_t len, uint8_t *__restrict__ genome) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= len) return;
    uint8_t c = (uint8_t) seq[i];
    // nVal (Common.c:229-233): upper-case ACGT kept, everything else is N
    genome[i + 1] = (c == 'A' || c == 'C' || c == 'G' || c == 'T') ? c : 'N';
}

// Device-side index build from the contig text already at genome[1..len] (count -> select -> stable radix sort on
// (hv, checksum) -> histogram / scan -> fused directory); what HashTable.c:769-839 + Sort.c:110-176 produce.
static int finish_build(cm_gpu_ctx *ctx, ContigHost *ch, uint8_t *genome, uint32_t len) {
    cudaStream_t st = ctx->stream;
    int r;
    std::vector<void *> tmp;
    auto cleanup = [&]() { for (void *p : tmp) cudaFree(p); };
    const bool verbose = getenv("CM_GPU_VERBOSE") != nullptr;
    auto now = [] { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; };
    const double t0 = now();
    uint64_t *keys = nullptr, *keys_sel = nullptr, *keys_alt = nullptr;
    uint32_t *locs = nullptr, *locs_sel = nullptr, *locs_alt = nullptr, *dirc = nullptr, *d_nsel = nullptr;
    uint8_t *valid = nullptr;
    if ((r = dev_alloc(ctx, tmp, &keys, (size_t) len)) != CM_OK || (r = dev_alloc(ctx, tmp, &locs, (size_t) len)) != CM_OK ||
        (r = dev_alloc(ctx, tmp, &valid, (size_t) len)) != CM_OK || (r = dev_alloc(ctx, tmp, &keys_sel, (size_t) len)) != CM_OK ||
        (r = dev_alloc(ctx, tmp, &locs_sel, (size_t) len)) != CM_OK || (r = dev_alloc(ctx, tmp, &d_nsel, 2)) != CM_OK ||
        (r = dev_alloc(ctx, ch->owned, &dirc, (size_t) CM_NBUCKETS + 1)) != CM_OK) { cleanup(); return r; }
    if (len) {
        k_kmer_keys<<<grid_for(len, 256), 256, 0, st>>>(genome, len, ctx->p.kmer, keys, locs, valid);
        ctx->launches++;
    }
    // compact valid k-mers (ascending position)
    size_t tb1 = 0, tb2 = 0, tb3 = 0, tb4 = 0;
    cub::DeviceSelect::Flagged(nullptr, tb1, keys, valid, keys_sel, d_nsel, (int64_t) len, st);
    cub::DeviceSelect::Flagged(nullptr, tb2, locs, valid, locs_sel, d_nsel, (int64_t) len, st);
    cub::DoubleBuffer<uint64_t> kb(keys_sel, keys);
    cub::DoubleBuffer<uint32_t> vb(locs_sel, locs);
    cub::DeviceRadixSort::SortPairs(nullptr, tb3, kb, vb, (int64_t) len, 0, 44, st);
    cub::DeviceScan::ExclusiveSum(nullptr, tb4, dirc, dirc, (int64_t) CM_NBUCKETS + 1, st);
    size_t tb = tb1 > tb2 ? tb1 : tb2;
    if (tb3 > tb) tb = tb3;
    if (tb4 > tb) tb = tb4;
    void *d_tmp = nullptr;
    if ((r = dev_alloc(ctx, tmp, (uint8_t **) &d_tmp, tb + 16)) != CM_OK) { cleanup(); return r; }
    const double t1 = now();
    uint32_t nsel = 0;
    if (len) {
        cub::DeviceSelect::Flagged(d_tmp, tb, keys, valid, keys_sel, d_nsel, (int64_t) len, st);
        cub::DeviceSelect::Flagged(d_tmp, tb, locs, valid, locs_sel, d_nsel, (int64_t) len, st);
        CUDA_TRY(ctx, cudaMemcpyAsync(&nsel, d_nsel, sizeof nsel, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
    }
    // stable LSD radix sort over (hv, checksum): positions stay ascending inside equal keys,
    // i.e. the (checksum, info) total order of Sort.c:110-176 inside each bucket
    if (nsel) {
        cub::DeviceRadixSort::SortPairs(d_tmp, tb, kb, vb, (int64_t) nsel, 0, 44, st);
    }
    CUDA_TRY(ctx, cudaMemsetAsync(dirc, 0, ((size_t) CM_NBUCKETS + 1) * sizeof(uint32_t), st));
    uint16_t *cks = nullptr;
    uint32_t *pos = nullptr;
    if ((r = dev_alloc(ctx, ch->owned, &cks, (size_t) nsel)) != CM_OK || (r = dev_alloc(ctx, ch->owned, &pos, (size_t) nsel)) != CM_OK) {
        cleanup();
        return r;
    }
    if (nsel) {
        k_bucket_hist<<<grid_for(nsel, 256), 256, 0, st>>>(kb.Current(), nsel, dirc);
        k_split_keys<<<grid_for(nsel, 256), 256, 0, st>>>(kb.Current(), nsel, cks);
        CUDA_TRY(ctx, cudaMemcpyAsync(pos, vb.Current(), (size_t) nsel * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
        ctx->launches += 2;
    }
    cub::DeviceScan::ExclusiveSum(d_tmp, tb, dirc, dirc, (int64_t) CM_NBUCKETS + 1, st);
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    CUDA_TRY(ctx, cudaGetLastError());
    const double t2 = now();
    cleanup();
    const double t3 = now();
    ch->d.dir = dirc;
    ch->d.cks = cks;
    ch->d.pos = pos;
    ch->d.genome = genome;
    ch->d.n_entries = nsel;
    ch->d.ref_len = len;
    if ((r = build_dirx(ctx, ch)) != CM_OK) return r;
    if (verbose)
        fprintf(stderr, "cm_gpu: index build of %u bp: allocations %.3f s, keys / select / sort / directory %.3f s, frees %.3f s, "
                "fused directory %.3f s\n", len, t1 - t0, t2 - t1, t3 - t2, now() - t3);
    ch->loaded = true;
    return CM_OK;
}

extern "C" int cm_gpu_build_contig(cm_gpu_ctx *ctx, int contig_id, const char *contig_seq, uint32_t ref_gen_len) {
    if (!ctx) return CM_ERR_ARG;
    if (!contig_seq) return fail(ctx, CM_ERR_ARG, "null argument");
    cudaSetDevice(ctx->device);
    ContigHost *ch = get_contig(ctx, contig_id, true);
    if (!ch) return fail(ctx, CM_ERR_ARG, ctx->parent ? "contig %d: a shared context cannot load (load into its parent)" : "bad contig id %d", contig_id);
    { int rr = check_reload(ctx, ch, contig_id); if (rr != CM_OK) return rr; }
    free_index(*ch);
    cudaStream_t st = ctx->stream;
    int r;
    const uint32_t len = ref_gen_len;
    uint8_t *genome = nullptr;
    if ((r = dev_alloc(ctx, ch->owned, &genome, (size_t) len + 64)) != CM_OK) return r;
    std::vector<void *> tmp;
    char *d_seq = nullptr;
    if ((r = dev_alloc(ctx, tmp, &d_seq, (size_t) len)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaMemcpyAsync(d_seq, contig_seq, len, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemsetAsync(genome, 'N', (size_t) len + 64, st));
    if (len) {
        k_ascii_to_genome<<<grid_for(len, 256), 256, 0, st>>>(d_seq, len, genome);
        ctx->launches++;
    }
    r = finish_build(ctx, ch, genome, len);
    cudaStreamSynchronize(st);
    for (void *p : tmp) cudaFree(p);
    return r;
}

// Same build from the reference's 3-bit packed contig (_ih_crefGen, 21 bases per word, Common.c:174-196): 0.4 bytes per
// base over PCIe instead of the 2 GiB pointer table + 6-byte entries that cm_gpu_load_contig converts and uploads.
extern "C" int cm_gpu_build_contig_packed(cm_gpu_ctx *ctx, int contig_id, const uint64_t *cref_gen, uint32_t ref_gen_len) {
    if (!ctx) return CM_ERR_ARG;
    if (!cref_gen) return fail(ctx, CM_ERR_ARG, "null argument");
    cudaSetDevice(ctx->device);
    ContigHost *ch = get_contig(ctx, contig_id, true);
    if (!ch) return fail(ctx, CM_ERR_ARG, ctx->parent ? "contig %d: a shared context cannot load (load into its parent)" : "bad contig id %d", contig_id);
    { int rr = check_reload(ctx, ch, contig_id); if (rr != CM_OK) return rr; }
    free_index(*ch);
    cudaStream_t st = ctx->stream;
    int r;
    const uint32_t len = ref_gen_len;
    uint8_t *genome = nullptr;
    if ((r = dev_alloc(ctx, ch->owned, &genome, (size_t) len + 64)) != CM_OK) return r;
    std::vector<void *> tmp;
    uint64_t *d_cref = nullptr;
    const size_t nwords = ((size_t) len + 20) / 21;
    if ((r = dev_alloc(ctx, tmp, &d_cref, nwords + 1)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaMemcpyAsync(d_cref, cref_gen, nwords * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemsetAsync(genome, 'N', (size_t) len + 64, st));
    if (len) {
        k_unpack_cref<<<grid_for(len, 256), 256, 0, st>>>(d_cref, len, genome);
        ctx->launches++;
    }
    r = finish_build(ctx, ch, genome, len);
    cudaStreamSynchronize(st);
    for (void *p : tmp) cudaFree(p);
    return r;
}

extern "C" int cm_gpu_contig_stats(cm_gpu_ctx *ctx, int contig_id, uint64_t *n_entries, uint32_t *ref_gen_len) {
    if (!ctx) return CM_ERR_ARG;
    ContigHost *ch = get_contig(ctx, contig_id, false);
    if (!ch || !ch->loaded) return fail(ctx, CM_ERR_STATE, "contig %d not loaded", contig_id);
    if (n_entries) *n_entries = ch->d.n_entries;
    if (ref_gen_len) *ref_gen_len = ch->d.ref_len;
    return CM_OK;
}

extern "C" int cm_gpu_read_bucket(cm_gpu_ctx *ctx, int contig_id, uint32_t hv, uint32_t cap, uint16_t *checksum,
                                  uint32_t *pos, uint32_t *n) {
    if (!ctx || !n) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    ContigHost *ch = get_contig(ctx, contig_id, false);
    if (!ch || !ch->loaded) return fail(ctx, CM_ERR_STATE, "contig %d not loaded", contig_id);
    if (hv >= CM_NBUCKETS) return fail(ctx, CM_ERR_ARG, "hv out of range");
    uint32_t d[2];
    CUDA_TRY(ctx, cudaMemcpy(d, ch->d.dir + hv, sizeof d, cudaMemcpyDeviceToHost));
    uint32_t cnt = d[1] - d[0];
    *n = cnt;
    uint32_t m = cnt < cap ? cnt : cap;
    if (m && checksum) CUDA_TRY(ctx, cudaMemcpy(checksum, ch->d.cks + d[0], m * sizeof(uint16_t), cudaMemcpyDeviceToHost));
    if (m && pos) CUDA_TRY(ctx, cudaMemcpy(pos, ch->d.pos + d[0], m * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    return CM_OK;
}

// ------------------------------------------------------------------------------------------
// annotation upload
// ------------------------------------------------------------------------------------------
extern "C" int cm_gpu_load_annotation(cm_gpu_ctx *ctx, int contig_id, const cm_annot_flat *A) {
    if (!ctx) return CM_ERR_ARG;
    if (!A) return fail(ctx, CM_ERR_ARG, "null argument");
    cudaSetDevice(ctx->device);
    ContigHost *ch = get_contig(ctx, contig_id, true);
    if (!ch) return fail(ctx, CM_ERR_ARG, ctx->parent ? "contig %d: a shared context cannot load (load into its parent)" : "bad contig id %d", contig_id);
    { int rr = check_reload(ctx, ch, contig_id); if (rr != CM_OK && ch->has_annot) return rr; }
    for (void *p : ch->owned_annot) cudaFree(p);
    ch->owned_annot.clear();
    ch->has_annot = false;
    if (A->n_iv == 0) return fail(ctx, CM_ERR_ARG, "annotation needs at least the dummy interval (gene_annotation.cpp:368-382)");
    if (A->n_chr == 0) return fail(ctx, CM_ERR_ARG, "annotation needs the contig's chromosome table");
    AnnotDev &a = ch->a;
    a = AnnotDev();
    int r;
    auto &own = ch->owned_annot;
#define UP(field, count)                                                                  \
    if ((r = dev_upload(ctx, own, &a.field, A->field, (size_t) (count))) != CM_OK) return r;
    a.n_iv = A->n_iv; a.n_seg = A->n_seg; a.n_trans = A->n_trans; a.n_genes = A->n_genes; a.n_chr = A->n_chr;
    a.n_bits = A->n_bits;
    UP(iv_spos, A->n_iv) UP(iv_epos, A->n_iv) UP(iv_max_end, A->n_iv) UP(iv_min_end, A->n_iv)
    UP(iv_max_next_exon, A->n_iv) UP(iv_seg_off, A->n_iv + 1)
    UP(seg_start, A->n_seg) UP(seg_end, A->n_seg) UP(seg_next_exon_beg, A->n_seg) UP(seg_gene_id, A->n_seg)
    UP(seg_tid_off, A->n_seg + 1) UP(tid, A->n_tid)
    UP(trans_start_ind, A->n_trans) UP(t2s_off, A->n_trans + 1)
    {
        uint32_t n_t2s = 0;
        if (A->n_trans) n_t2s = A->t2s_off[A->n_trans];
        UP(t2s, n_t2s)
    }
    UP(gene_start, A->n_genes) UP(gene_end, A->n_genes) UP(chr_shift, A->n_chr) UP(chr_id, A->n_chr)
    UP(near_border, (A->n_bits + 63) / 64) UP(intronic, (A->n_bits + 63) / 64)
#undef UP
    if (!getenv("CM_GPU_NO_IVLUT")) {   // radix index over the interval starts (cm_iv_search)
        const uint32_t shift = 10;
        const uint32_t top = A->iv_spos[A->n_iv - 1];
        const uint32_t nb = (top >> shift) + 2;
        std::vector<uint32_t> lut((size_t) nb + 1);
        uint32_t k = 0;
        for (uint32_t b = 0; b <= nb; b++) {
            const uint64_t lim = (uint64_t) b << shift;
            while (k < A->n_iv && (uint64_t) A->iv_spos[k] < lim) k++;
            lut[b] = k;
        }
        if ((r = dev_upload(ctx, own, &a.iv_lut, lut.data(), lut.size())) != CM_OK) return r;
        a.iv_lut_n = nb;
        a.iv_lut_shift = shift;
    }
    ch->has_annot = true;
    return CM_OK;
}

// fused directory: one thread per bucket; large buckets take a row of the coarse table (numbered by an atomic counter)
__global__ void k_build_dirx(const uint32_t *__restrict__ dir, const uint16_t *__restrict__ cks, DirX *__restrict__ dirx, uint32_t big_th,
                             uint32_t *__restrict__ n_big, uint32_t *__restrict__ big_hv, uint32_t big_cap) {
    uint32_t hv = blockIdx.x * blockDim.x + threadIdx.x;
    if (hv >= CM_NBUCKETS) return;
    uint32_t lo = dir[hv], hi = dir[hv + 1];
    DirX e;
    e.off = lo;
    e.cnt = hi - lo;
    for (uint32_t k = 0; k < CM_DIRX_INLINE; k++) e.c[k] = lo + k < hi ? cks[lo + k] : (uint16_t) 0xFFFFu;
    if (big_hv && e.cnt >= big_th) {
        const uint32_t id = atomicAdd(n_big, 1u);
        if (id < big_cap) big_hv[id] = hv;
        e.c[0] = (uint16_t) (id & 0xFFFFu);
        e.c[1] = (uint16_t) (id >> 16);
    }
    uint4 *dst = reinterpret_cast<uint4 *>(dirx + hv);
    const uint32_t *src = reinterpret_cast<const uint32_t *>(&e);
    for (unsigned q = 0; q < sizeof(DirX) / 16; q++) dst[q] = make_uint4(src[4 * q], src[4 * q + 1], src[4 * q + 2], src[4 * q + 3]);
}
// number of buckets of >= big_th entries
__global__ void k_count_big(const uint32_t *__restrict__ dir, uint32_t big_th, uint32_t *__restrict__ n_big) {
    uint32_t hv = blockIdx.x * blockDim.x + threadIdx.x;
    const bool big = hv < CM_NBUCKETS && dir[hv + 1] - dir[hv] >= big_th;
    const unsigned m = __ballot_sync(0xffffffffu, big);
    if (m && (threadIdx.x & 31u) == 0) atomicAdd(n_big, (uint32_t) __popc(m));
}
// coarse table of one large bucket per warp: row[t] = first entry with (cks >> shift) >= t, row[2^bits] = cnt
__global__ void k_build_big_tab(const uint32_t *__restrict__ dir, const uint16_t *__restrict__ cks, const uint32_t *__restrict__ big_hv,
                                uint32_t n_big, int bits, int shift, uint32_t *__restrict__ tab) {
    const uint32_t wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (wid >= n_big) return;
    const uint32_t hv = big_hv[wid], lo = dir[hv], cnt = dir[hv + 1] - lo;
    const uint32_t nt = 1u << bits;
    uint32_t *row = tab + (size_t) wid * (nt + 1);
    for (uint32_t t = lane; t <= nt; t += 32) {
        uint32_t a = 0, b = cnt;
        if (t == nt) a = cnt;
        else
            while (a < b) {
                const uint32_t mid = (a + b) >> 1;
                if ((uint32_t) (cks[lo + mid] >> shift) < t) a = mid + 1; else b = mid;
            }
        row[t] = a;
    }
}
static int build_dirx(cm_gpu_ctx *ctx, ContigHost *ch) {
    ch->d.dirx = nullptr;
    ch->d.big_tab = nullptr;
    ch->d.big_th = 0xFFFFFFFFu;
    ch->d.big_bits = 0;
    ch->d.big_shift = 0;
    if (getenv("CM_GPU_NO_DIRX")) return CM_OK;
    cudaStream_t st = ctx->stream;
    DirX *dx = nullptr;
    int r = dev_alloc(ctx, ch->owned, &dx, (size_t) CM_NBUCKETS);
    if (r != CM_OK) return r;
    const int cs_bits = 2 * (ctx->p.kmer - CM_WINDOW);
    uint32_t big_th = CM_BIG_TH_DEFAULT;
    if (const char *e = getenv("CM_GPU_BIG_TH")) { long v = atol(e); big_th = v <= CM_DIRX_INLINE ? CM_DIRX_INLINE + 1 : (uint32_t) v; }
    uint32_t n_big = 0, *d_nbig = nullptr, *big_hv = nullptr, *tab = nullptr;
    std::vector<void *> tmp;
    if (cs_bits > 0) {   // with no checksum (k = 14) every entry of a bucket matches: nothing to search
        if ((r = dev_alloc(ctx, tmp, &d_nbig, 2)) != CM_OK) return r;
        CUDA_TRY(ctx, cudaMemsetAsync(d_nbig, 0, 8, st));
        k_count_big<<<grid_for(CM_NBUCKETS, 256), 256, 0, st>>>(ch->d.dir, big_th, d_nbig);
        CUDA_TRY(ctx, cudaMemcpyAsync(&n_big, d_nbig, 4, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
        ctx->launches++;
    }
    const int bits = cs_bits < 10 ? cs_bits : 10;
    if (n_big) {
        if ((r = dev_alloc(ctx, tmp, &big_hv, (size_t) n_big)) != CM_OK || (r = dev_alloc(ctx, ch->owned, &tab, (size_t) n_big * ((1u << bits) + 1))) != CM_OK) {
            for (void *p : tmp) cudaFree(p);
            return r;
        }
        CUDA_TRY(ctx, cudaMemsetAsync(d_nbig, 0, 4, st));
    }
    k_build_dirx<<<grid_for(CM_NBUCKETS, 256), 256, 0, st>>>(ch->d.dir, ch->d.cks, dx, big_th, d_nbig, n_big ? big_hv : nullptr, n_big);
    ctx->launches++;
    if (n_big) {
        k_build_big_tab<<<grid_for((size_t) n_big * 32, 256), 256, 0, st>>>(ch->d.dir, ch->d.cks, big_hv, n_big, bits, cs_bits - bits, tab);
        ctx->launches++;
    }
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    for (void *p : tmp) cudaFree(p);
    ch->d.dirx = dx;
    if (n_big) {
        ch->d.big_tab = tab;
        ch->d.big_th = big_th;
        ch->d.big_bits = bits;
        ch->d.big_shift = cs_bits - bits;
    }
    ch->n_big = n_big;
    return CM_OK;
}

// ------------------------------------------------------------------------------------------
// batch staging
// ------------------------------------------------------------------------------------------
// strand rows: PE pair i -> strands 4i (R1 fwd), 4i+1 (R1 rc), 4i+2 (R2 fwd), 4i+3 (R2 rc)
//              SE read i -> strands 2i (fwd), 2i+1 (rc)
// One warp per input read.  Besides the two ASCII rows it emits the (hv, cv) pair of every seed slot
// (hashVal / checkSumVal, lib/mrsfast/Common.c:199-225) so that the gather kernel never touches read
// bytes: hv = -1 / cv = -1 on a non-ACGT base, hv = -2 when the slot holds no seed (q + k > len).
#define CM_PREP_BLOCK 256
// byte -> complement ([0..255]) and byte -> 2-bit code, 0x80 when not ACGT ([256..511]); written once per context
__device__ uint8_t g_prep_lut[512];
__global__ void k_init_prep_lut() {
    int v = cm_nhval((uint8_t) threadIdx.x);
    g_prep_lut[threadIdx.x] = cm_comp((uint8_t) threadIdx.x);
    g_prep_lut[256 + threadIdx.x] = v < 0 ? 0x80 : (uint8_t) v;
}
__global__ void __launch_bounds__(CM_PREP_BLOCK) k_make_strands(const char *__restrict__ s1, const char *__restrict__ s2, const uint16_t *__restrict__ l1,
                               const uint16_t *__restrict__ l2, uint32_t n, uint32_t in_stride, uint32_t stride,
                               uint8_t *__restrict__ strands, uint16_t *__restrict__ slen, int kmer, int F,
                               int2 *__restrict__ seed_key) {
    __shared__ uint8_t lut_comp[256], lut_nh[256];
    extern __shared__ __align__(16) uint8_t warp_buf[];   // per warp: raw read (+ 4 bytes of alignment slack), forward codes, reverse-complement codes
    lut_comp[threadIdx.x] = g_prep_lut[threadIdx.x];
    lut_nh[threadIdx.x] = g_prep_lut[256 + threadIdx.x];
    __syncthreads();
    uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    uint32_t n_reads = s2 ? 2 * n : n;
    if (warp >= n_reads) return;
    uint32_t i = s2 ? warp >> 1 : warp;
    bool second = s2 && (warp & 1);
    const uint8_t *src = (const uint8_t *) (second ? s2 : s1) + (size_t) i * in_stride;
    const uint32_t len = second ? l2[i] : l1[i];
    uint8_t *wb = warp_buf + (size_t) (threadIdx.x >> 5) * (3 * stride + 16);
    // the read as aligned words (rows of the upload are in_stride bytes apart: any alignment); the upload buffers carry
    // 16 bytes of slack behind the last row
    const uint32_t mis = (uint32_t) ((uintptr_t) src & 3u);
    const uint32_t *src_w = (const uint32_t *) (src - mis);
    for (uint32_t w = lane; 4 * w < mis + len; w += 32) ((uint32_t *) wb)[w] = src_w[w];
    const uint8_t *raw = wb + mis;
    uint8_t *cf = wb + stride + 16, *cr = cf + stride;
    __syncwarp();
    uint32_t *fw = (uint32_t *) (strands + (size_t) (2 * warp) * stride), *rc = fw + stride / 4;
    for (uint32_t w = lane; w < stride / 4; w += 32) {
        uint32_t wf = 0, wr = 0, kf = 0, kr = 0;
#pragma unroll
        for (uint32_t k = 0; k < 4; k++) {
            const uint32_t j = 4 * w + k;
            const uint8_t c = j < len ? raw[j] : 0, r = j < len ? lut_comp[raw[len - 1 - j]] : 0;
            wf |= (uint32_t) c << (8 * k);
            wr |= (uint32_t) r << (8 * k);
            kf |= (uint32_t) lut_nh[c] << (8 * k);
            kr |= (uint32_t) lut_nh[r] << (8 * k);
        }
        fw[w] = wf;
        rc[w] = wr;
        ((uint32_t *) cf)[w] = kf;
        ((uint32_t *) cr)[w] = kr;
    }
    if (lane == 0) {
        slen[2 * warp] = (uint16_t) len;
        slen[2 * warp + 1] = (uint16_t) len;
    }
    __syncwarp();
    for (int idx = (int) lane; idx < 2 * F; idx += 32) {
        int str = idx / F, t = idx - str * F;
        int q = t * kmer;
        int hv = -2, cv = 0;
        if (q + kmer <= (int) len) {
            const uint8_t *code = (str ? cr : cf) + q;
            uint32_t bad_h = 0, bad_c = 0;
            hv = 0;
            for (int x = 0; x < CM_WINDOW; x++) { uint32_t v = code[x]; bad_h |= v; hv = (hv << 2) | (int) (v & 3u); }
            for (int x = CM_WINDOW; x < kmer; x++) { uint32_t v = code[x]; bad_c |= v; cv = (cv << 2) | (int) (v & 3u); }
            if (bad_h & 0x80u) hv = -1;
            if (bad_c & 0x80u) cv = -1;
        }
        seed_key[((size_t) 2 * warp + str) * F + t] = make_int2(hv, cv);
    }
}

static int stage_reads(cm_gpu_ctx *ctx, uint32_t n, const char *s1, const char *s2, uint32_t in_stride,
                       const uint16_t *l1, const uint16_t *l2, const cm_mread *in) {
    if (n == 0) { ctx->n_pairs = 0; ctx->n_strands = 0; return CM_OK; }
    if (!s1 || !l1 || (s2 && !l2)) return fail(ctx, CM_ERR_ARG, "null argument");
    if (in_stride == 0) return fail(ctx, CM_ERR_ARG, "stride is 0");
    uint32_t max_len = 0;
    for (uint32_t i = 0; i < n; i++) {
        if (l1[i] > max_len) max_len = l1[i];
        if (s2 && l2[i] > max_len) max_len = l2[i];
        if (l1[i] > in_stride || (s2 && l2[i] > in_stride)) return fail(ctx, CM_ERR_ARG, "read %u longer than stride", i);
        if (l1[i] > ctx->p.max_read_len || (s2 && l2[i] > ctx->p.max_read_len))
            return fail(ctx, CM_ERR_ARG, "read %u longer than max_read_len", i);
        if (l1[i] == 0 || (s2 && l2[i] == 0)) return fail(ctx, CM_ERR_ARG, "read %u is empty (fastq_parser.cpp:157 would underflow)", i);
    }
    cudaStream_t st = ctx->stream;
    uint32_t stride = (max_len + 15) / 16 * 16;   // strand rows as long as this batch's longest read, not max_read_len
    uint32_t n_strands = (s2 ? 4u : 2u) * n;
    ENSURE(ctx, ctx->in_seq1, (size_t) n * in_stride + 16);
    ENSURE(ctx, ctx->in_len1, (size_t) n * 2);
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->in_seq1.p, s1, (size_t) n * in_stride, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->in_len1.p, l1, (size_t) n * 2, cudaMemcpyHostToDevice, st));
    if (s2) {
        ENSURE(ctx, ctx->in_seq2, (size_t) n * in_stride + 16);
        ENSURE(ctx, ctx->in_len2, (size_t) n * 2);
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->in_seq2.p, s2, (size_t) n * in_stride, cudaMemcpyHostToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->in_len2.p, l2, (size_t) n * 2, cudaMemcpyHostToDevice, st));
    }
    ENSURE(ctx, ctx->in_mread, (size_t) n * sizeof(cm_mread));
    if (in) CUDA_TRY(ctx, cudaMemcpyAsync(ctx->in_mread.p, in, (size_t) n * sizeof(cm_mread), cudaMemcpyHostToDevice, st));
    ENSURE(ctx, ctx->strands, (size_t) n_strands * stride);
    ENSURE(ctx, ctx->strand_len, (size_t) n_strands * 2);
    ctx->n_pairs = n;
    ctx->n_strands = n_strands;
    ctx->stride = stride;
    ctx->staged_pe = s2 != nullptr;
    ctx->batch_F = cm_seed_count((int) max_len, ctx->p.kmer);
    if (ctx->batch_F < 1) ctx->batch_F = 1;
    return CM_OK;
}

static int launch_prep(cm_gpu_ctx *ctx, uint32_t in_stride) {
    if (!ctx->n_pairs) return CM_OK;
    uint32_t n_reads = ctx->staged_pe ? 2 * ctx->n_pairs : ctx->n_pairs;
    ENSURE(ctx, ctx->seed_key, (size_t) ctx->n_strands * ctx->batch_F * sizeof(int2));
    k_make_strands<<<grid_for((size_t) n_reads * 32, CM_PREP_BLOCK), CM_PREP_BLOCK, (size_t) (CM_PREP_BLOCK / 32) * (3 * ctx->stride + 16), ctx->stream>>>(
        (const char *) ctx->in_seq1.p, ctx->staged_pe ? (const char *) ctx->in_seq2.p : nullptr,
        (const uint16_t *) ctx->in_len1.p, (const uint16_t *) ctx->in_len2.p, ctx->n_pairs, in_stride, ctx->stride,
        (uint8_t *) ctx->strands.p, (uint16_t *) ctx->strand_len.p, ctx->p.kmer, ctx->batch_F, (int2 *) ctx->seed_key.p);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    DBG_SYNC(ctx, "k_make_strands");
    return CM_OK;
}

// ------------------------------------------------------------------------------------------
// K1: seed lookup.  One thread per (strand, seed slot): 8 coalesced bytes of key in, then the dependent
// random gathers directory (dir[hv], dir[hv+1]) -> checksum run (binary search).  Positions are NOT
// copied: chaining reads them in place through (off, cnt).
// ------------------------------------------------------------------------------------------
#ifndef CM_SEED_ILP
#define CM_SEED_ILP 1   // measured on C3 (28 M seeds): 1.62 ms at 1 and 2, 1.68 at 4, 1.95 at 8 -- the kernel is not short of loads in flight
#endif
__global__ void __launch_bounds__(256) k_seed(const __grid_constant__ DevView v, const int2 *__restrict__ seed_key, uint64_t total,
                                              uint32_t *__restrict__ seed_off, uint32_t *__restrict__ seed_cnt,
                                              cm_counters *__restrict__ ctr) {
    __shared__ uint32_t sh[4];   // looked, sectors, hits, warps done
    if (threadIdx.x < 4) sh[threadIdx.x] = 0;
    __syncthreads();             // at the start every warp arrives at once
    // CM_SEED_ILP seeds per thread, a block apart (coalesced key reads): all their directory entries are requested before the
    // first one is used, so that a thread keeps several random DRAM reads in flight
    const uint64_t g0 = (uint64_t) blockIdx.x * (blockDim.x * CM_SEED_ILP) + threadIdx.x;
    uint32_t sectors = 0, hits = 0, looked = 0;
    int2 key[CM_SEED_ILP];
    uint32_t w[CM_SEED_ILP][CM_DIRX_WORDS];
#pragma unroll
    for (int u = 0; u < CM_SEED_ILP; u++) {
        const uint64_t gid = g0 + (uint64_t) u * blockDim.x;
        key[u] = gid < total ? seed_key[gid] : make_int2(-2, 0);
    }
    if (v.c.dirx) {
#pragma unroll
        for (int u = 0; u < CM_SEED_ILP; u++) cm_dirx_load(v.c, (key[u].x < 0 || key[u].y < 0) ? -1 : key[u].x, w[u]);
    }
#pragma unroll
    for (int u = 0; u < CM_SEED_ILP; u++) {
        const uint64_t gid = g0 + (uint64_t) u * blockDim.x;
        if (gid >= total) continue;
        SeedHit h;
        h.off = 0;
        h.cnt = 0;
        if (key[u].x != -2) {
            uint32_t sec = 0;
            h = v.c.dirx ? cm_seed_resolve_x(v.c, w[u], key[u].x, key[u].y, &sec) : cm_seed_lookup(v.c, key[u].x, key[u].y, &sec);
            looked++;
            const uint32_t kept = h.cnt > (uint32_t) v.p.seed_lim ? 0 : h.cnt;
            hits += kept;
            sectors += sec + (6 * kept) / 32;
        }
        seed_off[gid] = h.off;
        seed_cnt[gid] = h.cnt;
    }
    // Work counters (roofline numerators), block-aggregated without a barrier behind the look-ups: every
    // warp adds its sums to shared memory and the last warp to arrive adds the block's totals to one of CM_CTR_SLOTS
    // copies.  A __syncthreads here made every warp of a block wait for the one with the longest chain of dependent probes
    // (a third of the kernel's stall cycles in the ncu capture); per-warp global atomics on three addresses bounded it before.
    for (int o = 16; o; o >>= 1) {
        sectors += __shfl_down_sync(0xffffffffu, sectors, o);
        hits += __shfl_down_sync(0xffffffffu, hits, o);
        looked += __shfl_down_sync(0xffffffffu, looked, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (looked) atomicAdd(&sh[0], looked);
        if (sectors) atomicAdd(&sh[1], sectors);
        if (hits) atomicAdd(&sh[2], hits);
        __threadfence_block();
        if (atomicAdd(&sh[3], 1u) == (blockDim.x >> 5) - 1) {
            __threadfence_block();
            cm_counters *c = ctr + (blockIdx.x & (CM_CTR_SLOTS - 1));
            const uint32_t a = atomicAdd(&sh[0], 0u), b = atomicAdd(&sh[1], 0u), d = atomicAdd(&sh[2], 0u);
            if (a) atomicAdd((unsigned long long *) &c->seeds, (unsigned long long) a);
            if (b) atomicAdd((unsigned long long *) &c->seed_sectors, (unsigned long long) b);
            if (d) atomicAdd((unsigned long long *) &c->seed_hits, (unsigned long long) d);
        }
    }
}

static int make_view(cm_gpu_ctx *ctx, int contig_id, bool need_annot, DevView &v) {
    ContigHost *ch = get_contig(ctx, contig_id, false);
    if (!ch || !ch->loaded) return fail(ctx, CM_ERR_STATE, "contig %d not loaded (call cm_gpu_load_contig / cm_gpu_build_contig)", contig_id);
    if (need_annot && !ch->has_annot) return fail(ctx, CM_ERR_STATE, "contig %d has no annotation (call cm_gpu_load_annotation)", contig_id);
    v.p = ctx->p;
    v.c = ch->d;
    v.a = ch->has_annot ? ch->a : AnnotDev();
    v.contig_id = contig_id;
    v.max_seg_cnt = ctx->max_seg_cnt;
    v.n_slots = ctx->batch_F;   // seed / chain arrays of the staged batch use the batch's own slot count
    return CM_OK;
}

static int launch_seed(cm_gpu_ctx *ctx, const DevView &v) {
    if (!ctx->n_strands) return CM_OK;
    size_t total = (size_t) ctx->n_strands * ctx->batch_F;
    ENSURE(ctx, ctx->seed_off, total * 4);
    ENSURE(ctx, ctx->seed_cnt, total * 4);
    k_seed<<<grid_for(total, 256 * CM_SEED_ILP), 256, 0, ctx->stream>>>(v, (const int2 *) ctx->seed_key.p, (uint64_t) total,
                                                         (uint32_t *) ctx->seed_off.p, (uint32_t *) ctx->seed_cnt.p, ctx->d_counters);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    DBG_SYNC(ctx, "k_seed");
    return CM_OK;
}

// per-slot summary for the stage-level parity entry
__global__ void k_seed_summary(const __grid_constant__ DevView v, const uint16_t *__restrict__ slen, uint32_t n_strands, const uint32_t *__restrict__ seed_off,
                               const uint32_t *__restrict__ seed_cnt, int32_t *__restrict__ qpos, uint32_t *__restrict__ frag_count,
                               uint8_t *__restrict__ has_frags, uint32_t *__restrict__ first_info, uint32_t *__restrict__ last_info,
                               unsigned long long *__restrict__ sum_info) {
    const int S = v.max_seg_cnt;
    uint64_t gid = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (uint64_t) n_strands * S) return;
    uint32_t strand = (uint32_t) (gid / S), s = (uint32_t) (gid % S);
    int32_t q = 0;
    uint32_t fc = 0, fi = 0, li = 0;
    uint8_t hf = 0;
    unsigned long long sum = 0;
    if ((s & 1) == 0) {
        uint32_t slot = s >> 1;
        int len = slen[strand];
        if ((int) slot < v.n_slots && (int) (slot * v.p.kmer + v.p.kmer) <= len) {
            q = (int32_t) (slot * v.p.kmer);
            uint32_t cnt = seed_cnt[(size_t) strand * v.n_slots + slot], off = seed_off[(size_t) strand * v.n_slots + slot];
            hf = cnt > 0;
            fc = cnt > (uint32_t) v.p.seed_lim ? 0 : cnt;
            if (fc) {
                fi = v.c.pos[off];
                li = v.c.pos[off + fc - 1];
                for (uint32_t j = 0; j < fc; j++) sum += v.c.pos[off + j];
            }
        }
    }
    qpos[gid] = q;
    frag_count[gid] = fc;
    has_frags[gid] = hf;
    first_info[gid] = fi;
    last_info[gid] = li;
    sum_info[gid] = sum;
}

template <typename T>
static int fetch(cm_gpu_ctx *ctx, T *host, const void *dev, size_t count) {
    if (!host || !count) return CM_OK;
    CUDA_TRY(ctx, cudaMemcpyAsync(host, dev, count * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
    return CM_OK;
}

extern "C" int cm_gpu_seed_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride,
                                 const uint16_t *len, int32_t *qpos, uint32_t *frag_count, uint8_t *has_frags,
                                 uint32_t *first_info, uint32_t *last_info, uint64_t *sum_info) {
    if (!ctx) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig_id, false, v);
    if (r != CM_OK) return r;
    if (n == 0) return CM_OK;
    // each input row is ONE strand here: stage as SE and keep only the forward rows
    if ((r = stage_reads(ctx, n, seqs, nullptr, stride, len, nullptr, nullptr)) != CM_OK) return r;
    if ((r = make_view(ctx, contig_id, false, v)) != CM_OK) return r;
    if ((r = launch_prep(ctx, stride)) != CM_OK) return r;
    if ((r = launch_seed(ctx, v)) != CM_OK) return r;
    size_t total = (size_t) ctx->n_strands * ctx->max_seg_cnt;
    ENSURE(ctx, ctx->scratch[0], total * 4);
    ENSURE(ctx, ctx->scratch[1], total * 4);
    ENSURE(ctx, ctx->scratch[2], total);
    ENSURE(ctx, ctx->scratch[3], total * 4);
    ENSURE(ctx, ctx->scratch[4], total * 4);
    ENSURE(ctx, ctx->scratch[5], total * 8);
    k_seed_summary<<<grid_for(total, 256), 256, 0, ctx->stream>>>(
        v, (const uint16_t *) ctx->strand_len.p, ctx->n_strands, (const uint32_t *) ctx->seed_off.p, (const uint32_t *) ctx->seed_cnt.p,
        (int32_t *) ctx->scratch[0].p, (uint32_t *) ctx->scratch[1].p, (uint8_t *) ctx->scratch[2].p, (uint32_t *) ctx->scratch[3].p,
        (uint32_t *) ctx->scratch[4].p, (unsigned long long *) ctx->scratch[5].p);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    // rows 2i are the forward strands
    const int S = ctx->max_seg_cnt;
    auto pull = [&](void *host, const void *dev, size_t esz) -> int {
        if (!host) return CM_OK;
        CUDA_TRY(ctx, cudaMemcpy2DAsync(host, (size_t) S * esz, dev, (size_t) 2 * S * esz, (size_t) S * esz, n, cudaMemcpyDeviceToHost,
                                        ctx->stream));
        return CM_OK;
    };
    if ((r = pull(qpos, ctx->scratch[0].p, 4)) != CM_OK) return r;
    if ((r = pull(frag_count, ctx->scratch[1].p, 4)) != CM_OK) return r;
    if ((r = pull(has_frags, ctx->scratch[2].p, 1)) != CM_OK) return r;
    if ((r = pull(first_info, ctx->scratch[3].p, 4)) != CM_OK) return r;
    if ((r = pull(last_info, ctx->scratch[4].p, 4)) != CM_OK) return r;
    if ((r = pull(sum_info, ctx->scratch[5].p, 8)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return CM_OK;
}

#include "cm_kernels.cuh"
#include "cm_host_pipeline.cuh"
#include "cm_index_io.cuh"
#include "cm_noalloc.cuh"
#include "cm_annot_io.cuh"
#include "cm_circ.cuh"

// the hot-path kernels, loaded on the current device (cudaFuncGetAttributes forces the lazy load)
static void preload_kernels() {
    cudaFuncAttributes fa;
    const void *ks[] = {(const void *) k_make_strands, (const void *) k_seed, (const void *) k_cell_count, (const void *) k_strand_key,
                        (const void *) k_chain, (const void *) k_chain_warp, (const void *) k_chain_big, (const void *) k_spec_count,
                        (const void *) k_chain_ids, (const void *) k_chain_list, (const void *) k_chain_spec, (const void *) k_chain_walk,
                        (const void *) k_pair_key, (const void *) k_process_pe, (const void *) k_process_pe_warp, (const void *) k_process_se,
                        (const void *) k_round_commit, (const void *) k_round_compact, (const void *) k_unpack_cref, (const void *) k_kmer_keys,
                        (const void *) k_bucket_hist, (const void *) k_split_keys, (const void *) k_build_dirx, (const void *) k_count_big,
                        (const void *) k_build_big_tab, (const void *) k_init_prep_lut};
    for (const void *k : ks) cudaFuncGetAttributes(&fa, k);
    cudaGetLastError();
}
