// chaining / alignment / extension / classification kernels and the fused batch entries
#pragma once
extern "C" int cm_gpu_chain_batch(cm_gpu_ctx *ctx, int, uint32_t, const char *, uint32_t, const uint16_t *, int32_t *, int32_t *,
                                  float *, uint32_t *, uint32_t *, int32_t *) { return fail(ctx, CM_ERR_STATE, "not built yet"); }
extern "C" int cm_gpu_align_batch(cm_gpu_ctx *ctx, int, uint32_t, const char *, const uint16_t *, const char *, const uint16_t *,
                                  uint32_t, const int32_t *, int32_t *, int32_t *, int32_t *, int32_t *) { return fail(ctx, CM_ERR_STATE, "not built yet"); }
extern "C" int cm_gpu_process_batch_pe(cm_gpu_ctx *ctx, int, uint32_t, const char *, const char *, uint32_t, const uint16_t *,
                                       const uint16_t *, const cm_mread *, cm_mread *, int32_t *) { return fail(ctx, CM_ERR_STATE, "not built yet"); }
extern "C" int cm_gpu_process_batch_se(cm_gpu_ctx *ctx, int, uint32_t, const char *, uint32_t, const uint16_t *, int32_t *) { return fail(ctx, CM_ERR_STATE, "not built yet"); }
extern "C" int cm_gpu_stage_batch_pe(cm_gpu_ctx *ctx, uint32_t n, const char *s1, const char *s2, uint32_t stride, const uint16_t *l1,
                                     const uint16_t *l2, const cm_mread *in) {
    if (!ctx) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    int r = stage_reads(ctx, n, s1, s2, stride, l1, l2, in);
    if (r != CM_OK) return r;
    if ((r = launch_prep(ctx, stride)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return CM_OK;
}
extern "C" int cm_gpu_run_staged_pe(cm_gpu_ctx *ctx, int contig_id, int upto_stage) {
    if (!ctx) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig_id, upto_stage > 1, v);
    if (r != CM_OK) return r;
    cudaEventRecord(ctx->ev[0], ctx->stream);
    if ((r = launch_seed(ctx, v)) != CM_OK) return r;
    cudaEventRecord(ctx->ev[1], ctx->stream);
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    cudaEventElapsedTime(&ctx->stage_ms[1], ctx->ev[0], ctx->ev[1]);
    ctx->stage_ms[4] = ctx->stage_ms[1];
    return CM_OK;
}
extern "C" int cm_gpu_fetch_staged_pe(cm_gpu_ctx *ctx, cm_mread *, int32_t *) { return fail(ctx, CM_ERR_STATE, "not built yet"); }
