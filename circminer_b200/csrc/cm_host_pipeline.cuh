// Host side of the batch pipeline: arena sizing, launches, stage timing, C-ABI batch entries.
#pragma once

static int make_batch(cm_gpu_ctx *ctx, BatchDev &b) {
    const size_t ns = ctx->n_strands;
    const int C = ctx->p.max_chain_len, F = ctx->batch_F;
    ENSURE(ctx, ctx->seed_off, ns * F * 4);
    ENSURE(ctx, ctx->seed_cnt, ns * F * 4);
    ENSURE(ctx, ctx->chain_cnt, ns * 4);
    ENSURE(ctx, ctx->chain_hh, ns * 4);
    ENSURE(ctx, ctx->chain_score, ns * C * 4);
    ENSURE(ctx, ctx->chain_len, ns * C * 4);
    ENSURE(ctx, ctx->chain_frags, ns * C * F * sizeof(FragDev));
    b.strands = (const uint8_t *) ctx->strands.p;
    b.strand_len = (const uint16_t *) ctx->strand_len.p;
    b.n_strands = (uint32_t) ns;
    b.stride = ctx->stride;
    b.F = F;
    b.C = C;
    b.seed_off = (uint32_t *) ctx->seed_off.p;
    b.seed_cnt = (uint32_t *) ctx->seed_cnt.p;
    b.chain_cnt = (int32_t *) ctx->chain_cnt.p;
    b.chain_hh = (int32_t *) ctx->chain_hh.p;
    b.chain_score = (float *) ctx->chain_score.p;
    b.chain_len = (uint32_t *) ctx->chain_len.p;
    b.chain_frags = (FragDev *) ctx->chain_frags.p;
    b.chain_base = nullptr;
    b.aux = nullptr;
    b.spec = nullptr;
    b.spec_k = 0;
    return CM_OK;
}

static int launch_chain(cm_gpu_ctx *ctx, const DevView &v, BatchDev &b) {
    if (!ctx->n_strands) return CM_OK;
    cudaStream_t st = ctx->stream;
    const uint32_t ns = ctx->n_strands;
    ENSURE(ctx, ctx->cell_cnt, (size_t) ns * 4);
    ENSURE(ctx, ctx->cell_off, ((size_t) ns + 1) * 8);
    k_cell_count<<<grid_for(ns, 256), 256, 0, st>>>(v, b.seed_cnt, ns, (uint32_t *) ctx->cell_cnt.p);
    ctx->launches++;
    DBG_SYNC(ctx, "k_cell_count");
    size_t tb = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tb, (const uint32_t *) ctx->cell_cnt.p, (uint64_t *) ctx->cell_off.p, (int) ns, st);
    ENSURE(ctx, ctx->cub_tmp, tb + 16);
    cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tb, (const uint32_t *) ctx->cell_cnt.p, (uint64_t *) ctx->cell_off.p, (int) ns, st);
    ctx->launches++;
    // strands in work-signature order (heavy first) + the number of heavy ones
    ENSURE(ctx, ctx->key32[0], (size_t) ns * 4);
    ENSURE(ctx, ctx->key32[1], (size_t) ns * 4);
    ENSURE(ctx, ctx->perm[0], (size_t) ns * 4);
    ENSURE(ctx, ctx->perm[1], (size_t) ns * 4);
    ENSURE(ctx, ctx->ovf_n, 8);
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->ovf_n.p, 0, 8, st));   // [0] overflow queue length, [1] heavy strands
    k_strand_key<<<grid_for(ns, 256), 256, 0, st>>>(v, b.seed_cnt, ns, (uint32_t *) ctx->key32[0].p, (uint32_t *) ctx->perm[0].p,
                                                    ctx->heavy_th, (uint32_t *) ctx->ovf_n.p + 1);
    {
        size_t tb2 = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, tb2, (const uint32_t *) ctx->key32[0].p, (uint32_t *) ctx->key32[1].p,
                                        (const uint32_t *) ctx->perm[0].p, (uint32_t *) ctx->perm[1].p, (int) ns, 0, 32, st);
        ENSURE(ctx, ctx->cub_tmp, tb2 + 16);
        cub::DeviceRadixSort::SortPairs(ctx->cub_tmp.p, tb2, (const uint32_t *) ctx->key32[0].p, (uint32_t *) ctx->key32[1].p,
                                        (const uint32_t *) ctx->perm[0].p, (uint32_t *) ctx->perm[1].p, (int) ns, 0, 32, st);
    }
    ctx->launches += 2;
    // total cells = off[ns-1] + cnt[ns-1]
    uint64_t last_off = 0;
    uint32_t last_cnt = 0, n_heavy = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&last_off, (uint64_t *) ctx->cell_off.p + (ns - 1), 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(&last_cnt, (uint32_t *) ctx->cell_cnt.p + (ns - 1), 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(&n_heavy, (uint32_t *) ctx->ovf_n.p + 1, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    DBG_SYNC(ctx, "k_strand_key + sort");
    ctx->last_heavy = n_heavy;
    const uint64_t cells = last_off + last_cnt;
    ENSURE(ctx, ctx->dp_score, (size_t) (cells + 1) * 8);
    ENSURE(ctx, ctx->dp_prev, (size_t) (cells + 1) * 4);
    const uint32_t mult = ctx->rec_mult;
    ENSURE(ctx, ctx->recs, (size_t) ((uint64_t) mult * cells + 8ull * ns + 1) * sizeof(ChainRec));
    ENSURE(ctx, ctx->ovf_strand, (size_t) ns * 4);
    ENSURE(ctx, ctx->ovf_nrec, (size_t) ns * 4);
    if (n_heavy) {
        // heavy and light strands are independent: the warp kernel runs beside k_chain on a second stream
        cudaEventRecord(ctx->ev_fork, st);
        cudaStreamWaitEvent(ctx->stream2, ctx->ev_fork, 0);
        k_chain_warp<<<grid_for((size_t) n_heavy * 32, 128), 128, 0, ctx->stream2>>>(v, b, (const uint64_t *) ctx->cell_off.p, (double *) ctx->dp_score.p,
                                                                          (uint32_t *) ctx->dp_prev.p, (ChainRec *) ctx->recs.p, mult,
                                                                          (uint32_t *) ctx->ovf_strand.p, (uint32_t *) ctx->ovf_nrec.p,
                                                                          (uint32_t *) ctx->ovf_n.p, (const uint32_t *) ctx->perm[1].p, n_heavy,
                                                                          ctx->d_counters);
        cudaEventRecord(ctx->ev_join, ctx->stream2);
        ctx->launches++;
    }
    if (ns > n_heavy)
        k_chain<<<grid_for(ns - n_heavy, 128), 128, 0, st>>>(v, b, (const uint64_t *) ctx->cell_off.p, (double *) ctx->dp_score.p,
                                                             (uint32_t *) ctx->dp_prev.p, (ChainRec *) ctx->recs.p, mult,
                                                             (uint32_t *) ctx->ovf_strand.p, (uint32_t *) ctx->ovf_nrec.p, (uint32_t *) ctx->ovf_n.p,
                                                             (const uint32_t *) ctx->perm[1].p + n_heavy, ns - n_heavy, ctx->d_counters);
    ctx->launches++;
    if (n_heavy) cudaStreamWaitEvent(st, ctx->ev_join, 0);
    CUDA_TRY(ctx, cudaGetLastError());
    DBG_SYNC(ctx, "k_chain");
    uint32_t n_ovf = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&n_ovf, ctx->ovf_n.p, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    ctx->last_overflow = n_ovf;
    if (n_ovf) {
        std::vector<uint32_t> nrec(n_ovf);
        std::vector<uint64_t> off(n_ovf);
        CUDA_TRY(ctx, cudaMemcpy(nrec.data(), ctx->ovf_nrec.p, (size_t) n_ovf * 4, cudaMemcpyDeviceToHost));
        uint64_t tot = 0;
        for (uint32_t k = 0; k < n_ovf; k++) { off[k] = tot; tot += nrec[k]; }
        ENSURE(ctx, ctx->ovf_off, (size_t) n_ovf * 8);
        ENSURE(ctx, ctx->ovf_recs, (size_t) (tot + 1) * sizeof(ChainRec));
        CUDA_TRY(ctx, cudaMemcpyAsync(ctx->ovf_off.p, off.data(), (size_t) n_ovf * 8, cudaMemcpyHostToDevice, st));
        k_chain_big<<<grid_for(n_ovf, 128), 128, 0, st>>>(v, b, (const uint64_t *) ctx->cell_off.p, (double *) ctx->dp_score.p,
                                                         (uint32_t *) ctx->dp_prev.p, (const uint32_t *) ctx->ovf_strand.p,
                                                         (const uint64_t *) ctx->ovf_off.p, (const uint32_t *) ctx->ovf_nrec.p, n_ovf,
                                                         (ChainRec *) ctx->ovf_recs.p);
        ctx->launches++;
        CUDA_TRY(ctx, cudaGetLastError());
        CUDA_TRY(ctx, cudaStreamSynchronize(st));   // off[] is host memory
    }
    // dense pre-alignment of every chain (memo for the per-pair kernel)
    ENSURE(ctx, ctx->chain_base, ((size_t) ns + 1) * 4);
    ENSURE(ctx, ctx->key32[0], (size_t) ns * 4);
    b.spec_k = ctx->spec_k;
    uint32_t *spec_cnt = (uint32_t *) ctx->key32[0].p;   // the strand sort is done with its keys
    k_spec_count<<<grid_for(ns, 256), 256, 0, st>>>(b, spec_cnt);
    {
        size_t tb3 = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, tb3, (const uint32_t *) spec_cnt, (uint32_t *) ctx->chain_base.p, (int) ns, st);
        ENSURE(ctx, ctx->cub_tmp, tb3 + 16);
        cub::DeviceScan::ExclusiveSum(ctx->cub_tmp.p, tb3, (const uint32_t *) spec_cnt, (uint32_t *) ctx->chain_base.p, (int) ns, st);
        ctx->launches += 2;
    }
    ctx->align_timed = false;
    uint32_t last_base = 0, last_n = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&last_base, (uint32_t *) ctx->chain_base.p + (ns - 1), 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(&last_n, spec_cnt + (ns - 1), 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const uint32_t n_chains = last_base + (uint32_t) last_n;
    ctx->last_chains = n_chains;
    b.chain_base = (const uint32_t *) ctx->chain_base.p;
    if (n_chains && !ctx->no_spec) {
        ENSURE(ctx, ctx->chain_ids, (size_t) n_chains * 4);
        ENSURE(ctx, ctx->chain_aux, (size_t) n_chains * sizeof(ChainAux));
        b.aux = (ChainAux *) ctx->chain_aux.p;
        if (!ctx->no_walk_spec) {
            ENSURE(ctx, ctx->chain_spec, (size_t) n_chains * 2 * sizeof(EndSpec));
            b.spec = (EndSpec *) ctx->chain_spec.p;
        }
        CUDA_TRY(ctx, cudaMemsetAsync(ctx->chain_aux.p, 0, (size_t) n_chains * sizeof(ChainAux), st));
        const size_t n_items = 3ull * n_chains;
        ENSURE(ctx, ctx->spec_key[0], n_items);
        ENSURE(ctx, ctx->spec_key[1], n_items);
        ENSURE(ctx, ctx->spec_item[0], n_items * 4);
        ENSURE(ctx, ctx->spec_item[1], n_items * 4);
        k_chain_ids<<<grid_for(ns, 256), 256, 0, st>>>(b, (uint32_t *) ctx->chain_ids.p);
        k_chain_list<<<grid_for(ns, 256), 256, 0, st>>>(v, b, n_chains, (uint8_t *) ctx->spec_key[0].p, (uint32_t *) ctx->spec_item[0].p);
        DBG_SYNC(ctx, "k_chain_list");
        {
            size_t tb4 = 0;
            cub::DeviceRadixSort::SortPairsDescending(nullptr, tb4, (const uint8_t *) ctx->spec_key[0].p, (uint8_t *) ctx->spec_key[1].p,
                                                      (const uint32_t *) ctx->spec_item[0].p, (uint32_t *) ctx->spec_item[1].p, (int) n_items, 0, 8, st);
            ENSURE(ctx, ctx->cub_tmp, tb4 + 16);
            cub::DeviceRadixSort::SortPairsDescending(ctx->cub_tmp.p, tb4, (const uint8_t *) ctx->spec_key[0].p, (uint8_t *) ctx->spec_key[1].p,
                                                      (const uint32_t *) ctx->spec_item[0].p, (uint32_t *) ctx->spec_item[1].p, (int) n_items, 0, 8, st);
        }
        // spec_item[0] is free again after the sort: it becomes the walk queue (at most 2 ends per chain)
        uint32_t *walk_q = (uint32_t *) ctx->spec_item[0].p, *n_walk = (uint32_t *) ctx->ovf_n.p + 1;
        CUDA_TRY(ctx, cudaMemsetAsync(n_walk, 0, 4, st));
        cudaEventRecord(ctx->ev[4], st);
        k_chain_spec<<<grid_for(n_items, 128), 128, 0, st>>>(v, b, (const uint32_t *) ctx->chain_ids.p, (const uint8_t *) ctx->spec_key[1].p,
                                                            (const uint32_t *) ctx->spec_item[1].p, n_chains, walk_q, n_walk, ctx->d_counters);
        cudaEventRecord(ctx->ev[5], st);
        if (b.spec) {
            k_chain_walk<<<grid_for(2ull * n_chains, 128), 128, 0, st>>>(v, b, (const uint32_t *) ctx->chain_ids.p, walk_q, n_walk, ctx->d_counters);
            ctx->launches++;
        }
        cudaEventRecord(ctx->ev[6], st);
        ctx->align_timed = true;
        ctx->launches += 4;
        CUDA_TRY(ctx, cudaGetLastError());
        DBG_SYNC(ctx, "k_chain_spec");
    }
    return CM_OK;
}

static int launch_process(cm_gpu_ctx *ctx, const DevView &v, const BatchDev &b) {
    if (!ctx->n_pairs) return CM_OK;
    const uint32_t n = ctx->n_pairs;
    int32_t *out_state = ctx->state_view;
    cm_mread *out_mread = ctx->out_view;
    if (!out_state) {
        ENSURE(ctx, ctx->out_state, (size_t) n * 4);
        out_state = (int32_t *) ctx->out_state.p;
    }
    if (ctx->staged_pe) {
        if (!out_mread) {
            ENSURE(ctx, ctx->out_mread, (size_t) n * sizeof(cm_mread));
            out_mread = (cm_mread *) ctx->out_mread.p;
        }
        ENSURE(ctx, ctx->key64[0], (size_t) n * 8);
        ENSURE(ctx, ctx->key64[1], (size_t) n * 8);
        ENSURE(ctx, ctx->perm[0], (size_t) n * 4);
        ENSURE(ctx, ctx->perm[1], (size_t) n * 4);
        cudaStream_t st = ctx->stream;
        ENSURE(ctx, ctx->ovf_n, 8);
        uint32_t *d_heavy = (uint32_t *) ctx->ovf_n.p + 1;
        CUDA_TRY(ctx, cudaMemsetAsync(d_heavy, 0, 4, st));
        const bool split_heavy = ctx->heavy_pair_th != 0xFFFFFFFFu;
        k_pair_key<<<grid_for(n, 256), 256, 0, st>>>(v, b, n, (unsigned long long *) ctx->key64[0].p, (uint32_t *) ctx->perm[0].p,
                                                     split_heavy ? ctx->heavy_pair_th : 0xFFFFFFFFu, d_heavy, ctx->pair_sig);
        size_t tb = 0;
        cub::DeviceRadixSort::SortPairsDescending(nullptr, tb, (const unsigned long long *) ctx->key64[0].p, (unsigned long long *) ctx->key64[1].p,
                                        (const uint32_t *) ctx->perm[0].p, (uint32_t *) ctx->perm[1].p, (int) n, 0, ctx->pair_sig ? 52 : 44, st);
        ENSURE(ctx, ctx->cub_tmp, tb + 16);
        cub::DeviceRadixSort::SortPairsDescending(ctx->cub_tmp.p, tb, (const unsigned long long *) ctx->key64[0].p, (unsigned long long *) ctx->key64[1].p,
                                        (const uint32_t *) ctx->perm[0].p, (uint32_t *) ctx->perm[1].p, (int) n, 0, ctx->pair_sig ? 52 : 44, st);
        ctx->launches += 2;
        uint32_t n_heavy = 0;
        if (split_heavy) {   // the only host round trip of this stage; none when every pair takes the thread kernel
            CUDA_TRY(ctx, cudaMemcpyAsync(&n_heavy, d_heavy, 4, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(ctx, cudaStreamSynchronize(st));
        }
        DBG_SYNC(ctx, "k_pair_key + sort");
        ctx->last_heavy_pairs = n_heavy;
        if (n_heavy) {   // heavy pairs beside the light ones, on the side stream
            cudaEventRecord(ctx->ev_fork, st);
            cudaStreamWaitEvent(ctx->stream2, ctx->ev_fork, 0);
            k_process_pe_warp<<<grid_for((size_t) n_heavy * 32, 128), 128, 0, ctx->stream2>>>(
                v, b, (const cm_mread *) ctx->in_mread.p, out_mread, out_state, n_heavy,
                (const uint32_t *) ctx->perm[1].p, ctx->d_counters);
            cudaEventRecord(ctx->ev_join, ctx->stream2);
            ctx->launches++;
        }
        if (n > n_heavy)
            k_process_pe<<<grid_for(n - n_heavy, CM_PE_BLOCK), CM_PE_BLOCK, 0, st>>>(v, b, (const cm_mread *) ctx->in_mread.p, out_mread,
                                                      out_state, n - n_heavy, (const uint32_t *) ctx->perm[1].p + n_heavy, ctx->d_counters);
        if (n_heavy) cudaStreamWaitEvent(st, ctx->ev_join, 0);
    } else {
        k_process_se<<<grid_for(n, 128), 128, 0, ctx->stream>>>(v, b, out_state, n, ctx->d_counters);
    }
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    DBG_SYNC(ctx, "k_process");
    return CM_OK;
}

// seed -> chain -> extend/classify on the staged batch; stage times from events on the context's stream
static int run_pipeline(cm_gpu_ctx *ctx, int contig_id, int upto) {
    DevView v;
    int r = make_view(ctx, contig_id, upto > 1, v);
    if (r != CM_OK) return r;
    for (float &m : ctx->stage_ms) m = 0.f;
    if (!ctx->n_strands) return CM_OK;
    BatchDev b;
    if ((r = make_batch(ctx, b)) != CM_OK) return r;
    cudaStream_t st = ctx->stream;
    cudaEventRecord(ctx->ev[0], st);
    if ((r = launch_seed(ctx, v)) != CM_OK) return r;
    cudaEventRecord(ctx->ev[1], st);
    if (upto >= 2 && (r = launch_chain(ctx, v, b)) != CM_OK) return r;
    cudaEventRecord(ctx->ev[2], st);
    if (upto >= 3 && (r = launch_process(ctx, v, b)) != CM_OK) return r;
    cudaEventRecord(ctx->ev[3], st);
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ctx->stage_ms[1], ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&ctx->stage_ms[2], ctx->ev[1], ctx->ev[2]);
    cudaEventElapsedTime(&ctx->stage_ms[3], ctx->ev[2], ctx->ev[3]);
    cudaEventElapsedTime(&ctx->stage_ms[4], ctx->ev[0], ctx->ev[3]);
    return CM_OK;
}

// ---- C-ABI batch entries ---------------------------------------------------------------
extern "C" int cm_gpu_chain_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride, const uint16_t *len,
                                  int32_t *chain_count, int32_t *high_hits, float *score, uint32_t *chain_len, uint32_t *frag_rpos,
                                  int32_t *frag_qpos) {
    if (!ctx) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig_id, true, v);
    if (r != CM_OK) return r;
    if (n == 0) return CM_OK;
    // each input row is one strand: stage as SE reads, export the forward rows
    if ((r = stage_reads(ctx, n, seqs, nullptr, stride, len, nullptr, nullptr)) != CM_OK) return r;
    if ((r = make_view(ctx, contig_id, true, v)) != CM_OK) return r;
    if ((r = launch_prep(ctx, stride)) != CM_OK) return r;
    if ((r = run_pipeline(ctx, contig_id, 2)) != CM_OK) return r;
    BatchDev b;
    if ((r = make_batch(ctx, b)) != CM_OK) return r;
    const size_t C = ctx->p.max_chain_len, FO = ctx->n_slots, rows = n;
    ENSURE(ctx, ctx->scratch[0], rows * 4);
    ENSURE(ctx, ctx->scratch[1], rows * 4);
    ENSURE(ctx, ctx->scratch[2], rows * C * 4);
    ENSURE(ctx, ctx->scratch[3], rows * C * 4);
    ENSURE(ctx, ctx->scratch[4], rows * C * FO * 4);
    ENSURE(ctx, ctx->scratch[5], rows * C * FO * 4);
    k_chain_export<<<grid_for(rows * C, 256), 256, 0, ctx->stream>>>(v, b, n, (int) FO, (int32_t *) ctx->scratch[0].p, (int32_t *) ctx->scratch[1].p,
                                                                    (float *) ctx->scratch[2].p, (uint32_t *) ctx->scratch[3].p,
                                                                    (uint32_t *) ctx->scratch[4].p, (int32_t *) ctx->scratch[5].p);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    if ((r = fetch(ctx, chain_count, ctx->scratch[0].p, rows)) != CM_OK) return r;
    if ((r = fetch(ctx, high_hits, ctx->scratch[1].p, rows)) != CM_OK) return r;
    if ((r = fetch(ctx, score, ctx->scratch[2].p, rows * C)) != CM_OK) return r;
    if ((r = fetch(ctx, chain_len, ctx->scratch[3].p, rows * C)) != CM_OK) return r;
    if ((r = fetch(ctx, frag_rpos, ctx->scratch[4].p, rows * C * FO)) != CM_OK) return r;
    if ((r = fetch(ctx, frag_qpos, ctx->scratch[5].p, rows * C * FO)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return CM_OK;
}

extern "C" int cm_gpu_align_batch(cm_gpu_ctx *ctx, int kind, uint32_t n, const char *s, const uint16_t *n_s, const char *t,
                                  const uint16_t *n_t, uint32_t stride, const int32_t *aux, int32_t *out_ed, int32_t *out_sclen,
                                  int32_t *out_indel, int32_t *out_score) {
    if (!ctx) return CM_ERR_ARG;
    if (kind < 0 || kind > 4) return fail(ctx, CM_ERR_ARG, "alignment kind must be 0..4");
    if (n == 0) return CM_OK;
    if (!s || !t || !n_s || !n_t || !out_ed || !out_sclen || !out_indel || !out_score || (kind == 4 && !aux))
        return fail(ctx, CM_ERR_ARG, "null argument");
    for (uint32_t i = 0; i < n; i++) {
        if (n_s[i] > stride || n_t[i] > stride) return fail(ctx, CM_ERR_ARG, "problem %u longer than stride", i);
        if (kind == 4 && aux[i] > CM_MAX_BAND) return fail(ctx, CM_ERR_ARG, "problem %u: band %d > %d", i, aux[i], CM_MAX_BAND);
        // full-DP fallbacks keep one column of s in local memory
        bool full = (kind == 2 || kind == 3) ? (n_s[i] <= 2 * ctx->p.band_width || n_t[i] <= ctx->p.band_width)
                                             : (kind == 4 ? (aux[i] < 0 || n_s[i] <= aux[i]) : false);
        if (full && n_s[i] > CM_FULLDP_MAX) return fail(ctx, CM_ERR_ARG, "problem %u: unbanded s longer than %d", i, CM_FULLDP_MAX);
    }
    cudaSetDevice(ctx->device);
    cudaStream_t st = ctx->stream;
    const size_t bytes = (size_t) n * stride;
    ENSURE(ctx, ctx->scratch[0], bytes + 64);
    ENSURE(ctx, ctx->scratch[1], bytes + 64);
    ENSURE(ctx, ctx->scratch[2], (size_t) n * 2);
    ENSURE(ctx, ctx->scratch[3], (size_t) n * 2);
    ENSURE(ctx, ctx->scratch[4], (size_t) n * 4);
    ENSURE(ctx, ctx->scratch[5], (size_t) n * 16);
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->scratch[0].p, s, bytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->scratch[1].p, t, bytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->scratch[2].p, n_s, (size_t) n * 2, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->scratch[3].p, n_t, (size_t) n * 2, cudaMemcpyHostToDevice, st));
    if (aux) CUDA_TRY(ctx, cudaMemcpyAsync(ctx->scratch[4].p, aux, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    int32_t *o = (int32_t *) ctx->scratch[5].p;
    k_align<<<grid_for(n, 128), 128, 0, st>>>(ctx->p, kind, n, (const uint8_t *) ctx->scratch[0].p, (const uint16_t *) ctx->scratch[2].p,
                                              (const uint8_t *) ctx->scratch[1].p, (const uint16_t *) ctx->scratch[3].p, stride,
                                              (const int32_t *) ctx->scratch[4].p, o, o + n, o + 2 * (size_t) n, o + 3 * (size_t) n);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    int r;
    if ((r = fetch(ctx, out_ed, o, n)) != CM_OK) return r;
    if ((r = fetch(ctx, out_sclen, o + n, n)) != CM_OK) return r;
    if ((r = fetch(ctx, out_indel, o + 2 * (size_t) n, n)) != CM_OK) return r;
    if ((r = fetch(ctx, out_score, o + 3 * (size_t) n, n)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return CM_OK;
}

extern "C" int cm_gpu_stage_batch_pe(cm_gpu_ctx *ctx, uint32_t n, const char *s1, const char *s2, uint32_t stride, const uint16_t *l1,
                                     const uint16_t *l2, const cm_mread *in) {
    if (!ctx) return CM_ERR_ARG;
    if (n && (!s2 || !in)) return fail(ctx, CM_ERR_ARG, "null argument");
    cudaSetDevice(ctx->device);
    int r = stage_reads(ctx, n, s1, s2, stride, l1, l2, in);
    if (r != CM_OK) return r;
    if ((r = launch_prep(ctx, stride)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return CM_OK;
}

extern "C" int cm_gpu_run_staged_pe(cm_gpu_ctx *ctx, int contig_id, int upto_stage) {
    if (!ctx) return CM_ERR_ARG;
    if (upto_stage < 1 || upto_stage > 3) return fail(ctx, CM_ERR_ARG, "upto_stage must be 1..3");
    cudaSetDevice(ctx->device);
    return run_pipeline(ctx, contig_id, upto_stage);
}

extern "C" int cm_gpu_fetch_staged_pe(cm_gpu_ctx *ctx, cm_mread *out, int32_t *state) {
    if (!ctx) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    if (!ctx->n_pairs) return CM_OK;
    if (!ctx->out_state.p || (ctx->staged_pe && !ctx->out_mread.p)) return fail(ctx, CM_ERR_STATE, "no results: run stage 3 first");
    int r;
    if (ctx->staged_pe && (r = fetch(ctx, out, ctx->out_mread.p, ctx->n_pairs)) != CM_OK) return r;
    if ((r = fetch(ctx, state, ctx->out_state.p, ctx->n_pairs)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return CM_OK;
}

extern "C" int cm_gpu_process_batch_pe(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride,
                                       const uint16_t *len1, const uint16_t *len2, const cm_mread *in, cm_mread *out, int32_t *state) {
    if (!ctx) return CM_ERR_ARG;
    if (n && (!out || !state)) return fail(ctx, CM_ERR_ARG, "null argument");
    int r = cm_gpu_stage_batch_pe(ctx, n, seqs1, seqs2, stride, len1, len2, in);
    if (r != CM_OK) return r;
    if ((r = run_pipeline(ctx, contig_id, 3)) != CM_OK) return r;
    return cm_gpu_fetch_staged_pe(ctx, out, state);
}

extern "C" int cm_gpu_process_batch_se(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride, const uint16_t *len,
                                       int32_t *state) {
    if (!ctx) return CM_ERR_ARG;
    if (n && !state) return fail(ctx, CM_ERR_ARG, "null argument");
    cudaSetDevice(ctx->device);
    int r = stage_reads(ctx, n, seqs, nullptr, stride, len, nullptr, nullptr);
    if (r != CM_OK) return r;
    if ((r = launch_prep(ctx, stride)) != CM_OK) return r;
    if ((r = run_pipeline(ctx, contig_id, 3)) != CM_OK) return r;
    return cm_gpu_fetch_staged_pe(ctx, nullptr, state);
}

// ---- rounds on the device ----------------------------------------------------------------
// The staged batch visits contigs first .. first + n_contigs - 1 in order; after each contig the pairs the
// reference would print are committed and the rest is gathered into the next contig's batch (k_round_commit /
// k_round_compact).  The caller's staged batch stays intact (the compacted batches live in their own buffers).
struct StagedBatchGuard {
    cm_gpu_ctx *ctx;
    DevBuf strands, slen, key, in;
    uint32_t n_pairs, n_strands;
    explicit StagedBatchGuard(cm_gpu_ctx *c)
        : ctx(c), strands(c->strands), slen(c->strand_len), key(c->seed_key), in(c->in_mread), n_pairs(c->n_pairs), n_strands(c->n_strands) {}
    ~StagedBatchGuard() {
        ctx->strands = strands; ctx->strand_len = slen; ctx->seed_key = key; ctx->in_mread = in;
        ctx->n_pairs = n_pairs; ctx->n_strands = n_strands;
    }
};

static int run_rounds(cm_gpu_ctx *ctx, int first, int n_contigs, bool want_snap) {
    if (n_contigs < 1 || n_contigs > CM_MAX_ROUNDS || first < 0) return fail(ctx, CM_ERR_ARG, "rounds: contigs [%d, %d) out of range", first, first + n_contigs);
    for (float &m : ctx->rounds_ms) m = 0.f;
    ctx->rounds_align_ms[0] = ctx->rounds_align_ms[1] = 0.f;
    for (auto &x : ctx->round_pairs) x = 0;
    ctx->rounds_run = 0;
    ctx->rounds_n = ctx->n_pairs;
    if (!ctx->n_pairs) return CM_OK;
    if (!ctx->staged_pe) return fail(ctx, CM_ERR_STATE, "rounds need a staged paired-end batch");
    for (int r = 0; r < n_contigs; r++) {   // fail before any work when a contig is missing
        DevView v;
        int rc = make_view(ctx, first + r, true, v);
        if (rc != CM_OK) return rc;
    }
    const uint32_t n = ctx->n_pairs;
    cudaStream_t st = ctx->stream;
    ENSURE(ctx, ctx->fin_mread, (size_t) n * sizeof(cm_mread));
    ENSURE(ctx, ctx->fin_state, (size_t) n * 4);
    ENSURE(ctx, ctx->fin_round, (size_t) n * 4);
    ENSURE(ctx, ctx->surv_flag, (size_t) n);
    ENSURE(ctx, ctx->surv_idx, (size_t) n * 4);
    ENSURE(ctx, ctx->surv_n, 8);
    if (want_snap) {
        ENSURE(ctx, ctx->snap_mread, (size_t) n_contigs * n * sizeof(cm_mread));
        ENSURE(ctx, ctx->snap_state, (size_t) n_contigs * n * 4);
        CUDA_TRY(ctx, cudaMemsetAsync(ctx->snap_mread.p, 0, (size_t) n_contigs * n * sizeof(cm_mread), st));
        CUDA_TRY(ctx, cudaMemsetAsync(ctx->snap_state.p, 0xFF, (size_t) n_contigs * n * 4, st));   // -1: the pair did not visit the contig
    }
    StagedBatchGuard guard(ctx);
    const uint32_t *ids = nullptr;
    int cur = -1;   // -1: the caller's staged batch, 0 / 1: compacted set in use
    const uint32_t chunk = ctx->chunk_pairs;
    ENSURE(ctx, ctx->round_out, (size_t) n * sizeof(cm_mread));
    ENSURE(ctx, ctx->round_state, (size_t) n * 4);
    cudaEventRecord(ctx->ev_r[0], st);
    for (int r = 0; r < n_contigs; r++) {
        const int c = first + r;
        const bool last = r == n_contigs - 1;
        const uint32_t n_cur = ctx->n_pairs;
        // the round's batch in chunks of at most chunk_pairs pairs: the arenas of the chaining and pre-alignment stages
        // are sized by the chunk, not by the batch (a repeat-rich GRCh38 batch needs ~30 GB per million pairs)
        {
            const DevBuf r_strands = ctx->strands, r_slen = ctx->strand_len, r_key = ctx->seed_key, r_in = ctx->in_mread;
            int rc = CM_OK;
            for (uint32_t start = 0; start < n_cur && rc == CM_OK; start += chunk) {
                const uint32_t mcur = n_cur - start < chunk ? n_cur - start : chunk;
                ctx->strands.p = (uint8_t *) r_strands.p + (size_t) 4 * start * ctx->stride;
                ctx->strand_len.p = (uint16_t *) r_slen.p + (size_t) 4 * start;
                ctx->seed_key.p = (int2 *) r_key.p + (size_t) 4 * start * ctx->batch_F;
                ctx->in_mread.p = (cm_mread *) r_in.p + start;
                ctx->out_view = (cm_mread *) ctx->round_out.p + start;
                ctx->state_view = (int32_t *) ctx->round_state.p + start;
                ctx->n_pairs = mcur;
                ctx->n_strands = 4 * mcur;
                rc = run_pipeline(ctx, c, 3);
                if (rc == CM_OK) {
                    for (int k = 1; k < 5; k++) ctx->rounds_ms[k] += ctx->stage_ms[k];
                    float am[2];
                    if (cm_gpu_last_align_ms(ctx, am) == CM_OK) { ctx->rounds_align_ms[0] += am[0]; ctx->rounds_align_ms[1] += am[1]; }
                }
            }
            ctx->strands = r_strands; ctx->strand_len = r_slen; ctx->seed_key = r_key; ctx->in_mread = r_in;
            ctx->out_view = nullptr; ctx->state_view = nullptr;
            ctx->n_pairs = n_cur;
            ctx->n_strands = 4 * n_cur;
            if (rc != CM_OK) return rc;
        }
        ctx->round_pairs[r] = n_cur;
        ctx->rounds_run = r + 1;
        k_round_commit<<<grid_for(n_cur, 256), 256, 0, st>>>(
            ctx->p, n_cur, ids, (const cm_mread *) ctx->round_out.p, (const int32_t *) ctx->round_state.p, (const uint16_t *) ctx->strand_len.p,
            last ? 1 : 0, c, (cm_mread *) ctx->fin_mread.p, (int32_t *) ctx->fin_state.p, (int32_t *) ctx->fin_round.p,
            (uint8_t *) ctx->surv_flag.p, want_snap ? (cm_mread *) ctx->snap_mread.p + (size_t) r * n : nullptr,
            want_snap ? (int32_t *) ctx->snap_state.p + (size_t) r * n : nullptr);
        ctx->launches++;
        CUDA_TRY(ctx, cudaGetLastError());
        if (last) break;
        cub::CountingInputIterator<uint32_t> it(0);
        size_t tb = 0;
        cub::DeviceSelect::Flagged(nullptr, tb, it, (const uint8_t *) ctx->surv_flag.p, (uint32_t *) ctx->surv_idx.p, (uint32_t *) ctx->surv_n.p, (int) n_cur, st);
        ENSURE(ctx, ctx->cub_tmp, tb + 16);
        cub::DeviceSelect::Flagged(ctx->cub_tmp.p, tb, it, (const uint8_t *) ctx->surv_flag.p, (uint32_t *) ctx->surv_idx.p, (uint32_t *) ctx->surv_n.p, (int) n_cur, st);
        ctx->launches++;
        uint32_t n_surv = 0;
        CUDA_TRY(ctx, cudaMemcpyAsync(&n_surv, ctx->surv_n.p, 4, cudaMemcpyDeviceToHost, st));
        CUDA_TRY(ctx, cudaStreamSynchronize(st));
        if (!n_surv) break;
        const int nx = cur == 0 ? 1 : 0;
        const int F = ctx->batch_F;
        ENSURE(ctx, ctx->rb_strands[nx], (size_t) n_surv * 4 * ctx->stride);
        ENSURE(ctx, ctx->rb_slen[nx], (size_t) n_surv * 4 * 2);
        ENSURE(ctx, ctx->rb_key[nx], (size_t) n_surv * 4 * F * sizeof(int2));
        ENSURE(ctx, ctx->rb_in[nx], (size_t) n_surv * sizeof(cm_mread));
        ENSURE(ctx, ctx->rb_ids[nx], (size_t) n_surv * 4);
        k_round_compact<<<grid_for((size_t) n_surv * 32, 256), 256, 0, st>>>(
            ctx->p, n_surv, (const uint32_t *) ctx->surv_idx.p, ids, ctx->stride, F, (const uint8_t *) ctx->strands.p,
            (const uint16_t *) ctx->strand_len.p, (const int2 *) ctx->seed_key.p, (const cm_mread *) ctx->round_out.p,
            (uint8_t *) ctx->rb_strands[nx].p, (uint16_t *) ctx->rb_slen[nx].p, (int2 *) ctx->rb_key[nx].p, (cm_mread *) ctx->rb_in[nx].p,
            (uint32_t *) ctx->rb_ids[nx].p);
        ctx->launches++;
        CUDA_TRY(ctx, cudaGetLastError());
        DBG_SYNC(ctx, "k_round_compact");
        ctx->strands = ctx->rb_strands[nx];
        ctx->strand_len = ctx->rb_slen[nx];
        ctx->seed_key = ctx->rb_key[nx];
        ctx->in_mread = ctx->rb_in[nx];
        ids = (const uint32_t *) ctx->rb_ids[nx].p;
        ctx->n_pairs = n_surv;
        ctx->n_strands = 4 * n_surv;
        cur = nx;
    }
    cudaEventRecord(ctx->ev_r[1], st);
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ctx->rounds_ms[5], ctx->ev_r[0], ctx->ev_r[1]);
    return CM_OK;
}

extern "C" int cm_gpu_run_staged_rounds_pe(cm_gpu_ctx *ctx, int first_contig, int n_contigs) {
    if (!ctx) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    return run_rounds(ctx, first_contig, n_contigs, false);
}

extern "C" int cm_gpu_fetch_rounds_pe(cm_gpu_ctx *ctx, cm_mread *out, int32_t *state, int32_t *final_round) {
    if (!ctx) return CM_ERR_ARG;
    cudaSetDevice(ctx->device);
    const uint32_t n = ctx->rounds_n;
    if (!n) return CM_OK;
    if (!ctx->rounds_run) return fail(ctx, CM_ERR_STATE, "no results: run the rounds first");
    int r;
    if ((r = fetch(ctx, out, ctx->fin_mread.p, n)) != CM_OK) return r;
    if ((r = fetch(ctx, state, ctx->fin_state.p, n)) != CM_OK) return r;
    if ((r = fetch(ctx, final_round, ctx->fin_round.p, n)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return CM_OK;
}

extern "C" int cm_gpu_last_rounds_ms(cm_gpu_ctx *ctx, float out[8], uint32_t *pairs_per_round, int cap) {
    if (!ctx || !out) return CM_ERR_ARG;
    for (int i = 0; i < 6; i++) out[i] = ctx->rounds_ms[i];
    out[6] = ctx->rounds_align_ms[0];
    out[7] = ctx->rounds_align_ms[1];
    if (pairs_per_round)
        for (int r = 0; r < cap; r++) pairs_per_round[r] = r < ctx->rounds_run ? ctx->round_pairs[r] : 0;
    return ctx->rounds_run;
}

extern "C" int cm_gpu_map_rounds_pe(cm_gpu_ctx *ctx, int first_contig, int n_contigs, uint32_t n, const char *seqs1, const char *seqs2,
                                    uint32_t stride, const uint16_t *len1, const uint16_t *len2, const cm_mread *in, cm_mread *out,
                                    int32_t *state, int32_t *final_round, cm_mread *round_out, int32_t *round_state) {
    if (!ctx) return CM_ERR_ARG;
    if (n && (!out || !state)) return fail(ctx, CM_ERR_ARG, "null argument");
    int r = cm_gpu_stage_batch_pe(ctx, n, seqs1, seqs2, stride, len1, len2, in);
    if (r != CM_OK) return r;
    const bool snap = round_out || round_state;
    if ((r = run_rounds(ctx, first_contig, n_contigs, snap)) != CM_OK) return r;
    if (!n) return CM_OK;
    if ((r = fetch(ctx, out, ctx->fin_mread.p, n)) != CM_OK) return r;
    if ((r = fetch(ctx, state, ctx->fin_state.p, n)) != CM_OK) return r;
    if ((r = fetch(ctx, final_round, ctx->fin_round.p, n)) != CM_OK) return r;
    if ((r = fetch(ctx, round_out, ctx->snap_mread.p, (size_t) n_contigs * n)) != CM_OK) return r;
    if ((r = fetch(ctx, round_state, ctx->snap_state.p, (size_t) n_contigs * n)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return CM_OK;
}

extern "C" int cm_gpu_annot_lookup_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const uint32_t *pos, uint32_t mlen, uint32_t rlen,
                                         uint32_t *ub, uint32_t *max_end, int32_t *ol_iv, int32_t *loc_iv, int32_t *loc_ind,
                                         int32_t *shift_idx, uint8_t *intronic) {
    if (!ctx) return CM_ERR_ARG;
    if (n && (!pos || !ub || !max_end || !ol_iv || !loc_iv || !loc_ind || !shift_idx || !intronic)) return fail(ctx, CM_ERR_ARG, "null argument");
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig_id, true, v);
    if (r != CM_OK) return r;
    if (!n) return CM_OK;
    cudaStream_t st = ctx->stream;
    for (int k = 0; k < 8; k++) ENSURE(ctx, ctx->scratch[k], (size_t) n * 4);
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->scratch[0].p, pos, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    k_annot_lookup<<<grid_for(n, 256), 256, 0, st>>>(v, n, (const uint32_t *) ctx->scratch[0].p, mlen, rlen, (uint32_t *) ctx->scratch[1].p,
                                                    (uint32_t *) ctx->scratch[2].p, (int32_t *) ctx->scratch[3].p, (int32_t *) ctx->scratch[4].p,
                                                    (int32_t *) ctx->scratch[5].p, (int32_t *) ctx->scratch[6].p, (uint8_t *) ctx->scratch[7].p);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    if ((r = fetch(ctx, ub, ctx->scratch[1].p, n)) != CM_OK || (r = fetch(ctx, max_end, ctx->scratch[2].p, n)) != CM_OK ||
        (r = fetch(ctx, ol_iv, ctx->scratch[3].p, n)) != CM_OK || (r = fetch(ctx, loc_iv, ctx->scratch[4].p, n)) != CM_OK ||
        (r = fetch(ctx, loc_ind, ctx->scratch[5].p, n)) != CM_OK || (r = fetch(ctx, shift_idx, ctx->scratch[6].p, n)) != CM_OK ||
        (r = fetch(ctx, intronic, ctx->scratch[7].p, n)) != CM_OK)
        return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return CM_OK;
}

extern "C" int cm_gpu_extend_batch(cm_gpu_ctx *ctx, int contig_id, int mode, uint32_t n, const char *lseq, const char *rseq, uint32_t stride,
                                   const uint16_t *llen, const uint16_t *rlen, uint32_t frag_slots, const uint32_t *l_chain_len,
                                   const uint32_t *l_rpos, const int32_t *l_qpos, const uint32_t *r_chain_len, const uint32_t *r_rpos,
                                   const int32_t *r_qpos, const int32_t *iv_a, const int32_t *iv_b, const int32_t *dir, cm_mate *lmm,
                                   cm_mate *rmm, int32_t *ret) {
    if (!ctx) return CM_ERR_ARG;
    if (mode != 0 && mode != 1) return fail(ctx, CM_ERR_ARG, "mode must be 0 (extend_both_mates) or 1 (extend_chain_both_sides)");
    if (!n) return CM_OK;
    const bool both = mode == 0;
    if (!lseq || !llen || !l_chain_len || !l_rpos || !l_qpos || !lmm || !ret || !frag_slots || frag_slots > CM_MAX_SEEDS ||
        (both && (!rseq || !rlen || !r_chain_len || !r_rpos || !r_qpos || !iv_a || !iv_b || !rmm)) || (!both && !dir))
        return fail(ctx, CM_ERR_ARG, "null or out-of-range argument");
    const uint32_t F = frag_slots;
    for (uint32_t i = 0; i < n; i++) {
        if (llen[i] > stride || (both && rlen[i] > stride) || llen[i] > ctx->p.max_read_len) return fail(ctx, CM_ERR_ARG, "problem %u: read longer than stride", i);
        if (l_chain_len[i] > F || (both && r_chain_len[i] > F)) return fail(ctx, CM_ERR_ARG, "problem %u: chain longer than frag_slots", i);
    }
    cudaSetDevice(ctx->device);
    DevView v;
    int r = make_view(ctx, contig_id, true, v);
    if (r != CM_OK) return r;
    cudaStream_t st = ctx->stream;
    // frags as FragDev {rpos, qpos} pairs
    std::vector<FragDev> hl((size_t) n * F), hr(both ? (size_t) n * F : 0);
    for (size_t k = 0; k < (size_t) n * F; k++) {
        hl[k].rpos = l_rpos[k]; hl[k].qpos = l_qpos[k];
        if (both) { hr[k].rpos = r_rpos[k]; hr[k].qpos = r_qpos[k]; }
    }
    const size_t sb = (size_t) n * stride, fb = (size_t) n * F * sizeof(FragDev);
    DevBuf *S = ctx->scratch;
    ENSURE(ctx, S[0], 2 * sb + 64);
    ENSURE(ctx, S[1], (size_t) n * 4);                    // llen | rlen (u16 each)
    ENSURE(ctx, S[2], 2 * fb);
    ENSURE(ctx, S[3], (size_t) n * 8);                    // chain lengths l | r
    ENSURE(ctx, S[4], (size_t) n * 12);                   // iv_a | iv_b | dir
    ENSURE(ctx, S[5], (size_t) n * 2 * sizeof(cm_mate));
    ENSURE(ctx, S[6], (size_t) n * 4);
    uint8_t *d_seq = (uint8_t *) S[0].p;
    uint16_t *d_len = (uint16_t *) S[1].p;
    FragDev *d_fr = (FragDev *) S[2].p;
    uint32_t *d_cl = (uint32_t *) S[3].p;
    int32_t *d_iv = (int32_t *) S[4].p;
    cm_mate *d_mm = (cm_mate *) S[5].p;
    CUDA_TRY(ctx, cudaMemsetAsync(S[4].p, 0, (size_t) n * 12, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_seq, lseq, sb, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_len, llen, (size_t) n * 2, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_fr, hl.data(), fb, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_cl, l_chain_len, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    if (both) {
        CUDA_TRY(ctx, cudaMemcpyAsync(d_seq + sb, rseq, sb, cudaMemcpyHostToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(d_len + n, rlen, (size_t) n * 2, cudaMemcpyHostToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(d_fr + (size_t) n * F, hr.data(), fb, cudaMemcpyHostToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(d_cl + n, r_chain_len, (size_t) n * 4, cudaMemcpyHostToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(d_iv, iv_a, (size_t) n * 4, cudaMemcpyHostToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(d_iv + n, iv_b, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    } else {
        CUDA_TRY(ctx, cudaMemcpyAsync(d_iv + 2 * (size_t) n, dir, (size_t) n * 4, cudaMemcpyHostToDevice, st));
    }
    k_extend<<<grid_for(n, 128), 128, 0, st>>>(v, mode, n, d_seq, d_seq + sb, stride, d_len, d_len + n, (int) F, d_cl, d_fr, d_cl + n,
                                              d_fr + (size_t) n * F, d_iv, d_iv + n, d_iv + 2 * (size_t) n, d_mm, d_mm + n, (int32_t *) S[6].p);
    ctx->launches++;
    CUDA_TRY(ctx, cudaGetLastError());
    if ((r = fetch(ctx, lmm, d_mm, n)) != CM_OK || (both && (r = fetch(ctx, rmm, d_mm + n, n)) != CM_OK) || (r = fetch(ctx, ret, S[6].p, n)) != CM_OK) return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    return CM_OK;
}
