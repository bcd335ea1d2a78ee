// TEST INFRASTRUCTURE ONLY -- never part of the product.
// The drop-in shim (integration/gpu_map_reads.cpp + host_feed.h) compiled into one translation unit with a stand-in
// for the C-ABI calls it makes, so that the HOST side of the drop-in -- block reads, parallel FASTQ parse, header
// tokens -> cm_mread, PAM / remain-FASTQ text, write order -- can be checked byte for byte against the reference on a
// box without a GPU.  The stand-in answers cm_gpu_process_batch_pe with the reference's own FilterRead::process_read
// (src/filter.cpp:124), one pair at a time, on the scratch of mapping thread 0; nothing here touches libcm_gpu.so.
// Built by oracle/Makefile into oracle/_ref/circminer_feedtest; used by tests/test_host_feed.py.
#define CM_SHIM_KEEP_REF_TABLE 1   // the stand-in below maps with the reference's own table
#include "../../integration/gpu_map_reads.cpp"

namespace {
FilterArgs *mock_scratch() { return g_fa; }
char mock_comp(char c) {
    switch (c) {
    case 'A': case 'a': return 'T';
    case 'C': case 'c': return 'G';
    case 'G': case 'g': return 'C';
    case 'T': case 't': return 'A';
    case 'N': case 'n': return 'N';
    default: return 0;   // FASTQParser::comp is zero elsewhere (src/fastq_parser.cpp:127-139)
    }
}
}   // namespace

extern "C" {
void cm_params_default(cm_params *p) {
    p->kmer = 20; p->max_read_len = 300; p->max_ed = 4; p->max_sc = 7; p->band_width = 3; p->seed_lim = 500;
    p->max_tlen = 500; p->max_intron_len = 2000000; p->max_chain_len = 30; p->scan_level = 0;
    p->mat_sc = 1; p->mis_sc = -3; p->ind_sc = -3; p->xdrop = 8;
}
void cm_mread_init(cm_mread *m, const cm_params *p) {
    memset(m, 0, sizeof *m);
    m->r1_forward = 1; m->r2_forward = 1;
    m->ed_r1 = p->max_ed + 1; m->ed_r2 = p->max_ed + 1;
    m->type = NOPROC_NOMATCH; m->tlen = (int32_t) INF;
    m->chr_r1 = -1; m->chr_r2 = -1;
}
int cm_gpu_warmup(int) { return CM_OK; }
int cm_gpu_device_count(void) { return 1; }
int cm_gpu_create(int, const cm_params *, cm_gpu_ctx **out) { *out = (cm_gpu_ctx *) &g_params; return CM_OK; }
int cm_gpu_create_shared(cm_gpu_ctx *parent, cm_gpu_ctx **out) { *out = parent; return CM_OK; }
const char *cm_gpu_last_error(const cm_gpu_ctx *) { return "mock"; }
int cm_gpu_load_contig(cm_gpu_ctx *, int, const void *const *, const void *, uint64_t, const uint64_t *, uint32_t) { return CM_OK; }
int cm_gpu_build_contig_packed(cm_gpu_ctx *, int, const uint64_t *, uint32_t) { return CM_OK; }
int cm_gpu_load_annotation(cm_gpu_ctx *, int, const cm_annot_flat *) { return CM_OK; }
int cm_annot_file_write(const char *, const cm_annot_flat *, uint32_t, uint64_t) { return CM_ERR_ARG; }
int cm_annot_file_read(const char *, cm_annot_flat *, uint32_t *, uint64_t *, void **) { return CM_ERR_ARG; }
void cm_annot_file_free(void *) {}
// stage 2 has no stand-in here: the host-feed tests run the binary with CM_GPU_STAGE2=0 (the reference's own do_process)
int cm_gpu_circ_single_batch(cm_gpu_ctx *, int, uint32_t, const char *, const char *, uint32_t, const uint16_t *, const uint16_t *, const cm_mread *,
                             const uint32_t *, const uint32_t *, const uint32_t *, int, int, uint32_t, uint32_t *, uint32_t *, uint8_t *, cm_circ_mate *,
                             cm_circ_mate *, cm_circ_mate *, int32_t *, cm_circ_res *, cm_circ_mate *, cm_circ_mate *, cm_circ_mate *, int32_t *,
                             cm_circ_res *) { return CM_ERR_STATE; }
int cm_gpu_circ_double_batch(cm_gpu_ctx *, int, uint32_t, const char *, const char *, uint32_t, const uint16_t *, const uint16_t *, const cm_mread *,
                             const uint32_t *, const uint32_t *, const uint32_t *, int, int, uint32_t, uint32_t *, uint32_t *, uint8_t *, cm_circ_mate *,
                             cm_circ_mate *, cm_circ_mate *, cm_circ_mate *, int32_t *, cm_circ_res *, int32_t *, cm_circ_res *) { return CM_ERR_STATE; }
int cm_gpu_host_alloc(size_t bytes, void **out) { *out = malloc(bytes ? bytes : 1); return *out ? CM_OK : CM_ERR_ALLOC; }
int cm_gpu_host_free(void *p) { free(p); return CM_OK; }

int cm_gpu_process_batch_pe(cm_gpu_ctx *, int, uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride,
                            const uint16_t *len1, const uint16_t *len2, const cm_mread *in, cm_mread *out, int32_t *state) {
    static std::mutex mu;   // the shim runs one host thread per context; the reference's scratch here is thread 0's
    std::lock_guard<std::mutex> lk(mu);
    static Record *r1 = nullptr, *r2 = nullptr;
    if (!r1) { r1 = new Record; r1->init(); r2 = new Record; r2->init(); }
    FilterArgs *fa = mock_scratch();
    for (uint32_t i = 0; i < n; i++) {
        Record *rr[2] = {r1, r2};
        const char *ss[2] = {seqs1 + (size_t) i * stride, seqs2 + (size_t) i * stride};
        const int ll[2] = {len1[i], len2[i]};
        for (int k = 0; k < 2; k++) {
            memcpy(rr[k]->seq, ss[k], ll[k]); rr[k]->seq[ll[k]] = 0;
            for (int j = 0; j < ll[k]; j++) rr[k]->rcseq[ll[k] - j - 1] = mock_comp(ss[k][j]);
            rr[k]->rcseq[ll[k]] = 0;
            rr[k]->seq_len = ll[k];
            strcpy(rr[k]->rname, "r");
        }
        from_pod(&in[i], r1->mr);
        state[i] = filter_read.process_read(fa->id, r1, r2, fa->kmer_size, fa->fl, fa->bl, *(fa->fbc_r1), *(fa->bbc_r1),
                                            *(fa->fbc_r2), *(fa->bbc_r2));
        to_pod(r1->mr, &out[i]);
    }
    return CM_OK;
}
}
