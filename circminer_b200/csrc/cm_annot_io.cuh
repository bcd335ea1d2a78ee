// The flattened annotation of one packed contig (cm_annot_flat) as a file: built once per (GTF, packed reference, -l) instead
// of on every run.  The reference has no such artefact: GTFParser::load_gtf (src/gene_annotation.cpp:191-399)
// parses the GTF and sets the two 1.3-Gbit masks bit by bit in every process.  Host-only code; included by cm_gpu.cu.
//
//   file := "CMANNOT1", u32 version (1), u32 max_read_len, u64 tag (caller's: e.g. size / mtime of the GTF), u32 counts[7]
//           (n_iv n_seg n_tid n_trans n_genes n_chr n_t2s), u64 n_bits, then the arrays in cm_annot_flat order, each padded to
//           8 bytes; the masks last.
#pragma once

struct AnnotFileHeader {
    char magic[8];
    uint32_t version, max_read_len;
    uint64_t tag;
    uint32_t n_iv, n_seg, n_tid, n_trans, n_genes, n_chr, n_t2s, pad_;
    uint64_t n_bits;
};

struct AnnotFileArray { const void *p; size_t bytes; };

static void annot_arrays(const cm_annot_flat &A, uint32_t n_t2s, AnnotFileArray out[21]) {
    const size_t nw = (size_t) ((A.n_bits + 63) / 64);
    const AnnotFileArray a[21] = {
        {A.iv_spos, 4ull * A.n_iv}, {A.iv_epos, 4ull * A.n_iv}, {A.iv_max_end, 4ull * A.n_iv}, {A.iv_min_end, 4ull * A.n_iv},
        {A.iv_max_next_exon, 4ull * A.n_iv}, {A.iv_seg_off, 4ull * (A.n_iv + 1)},
        {A.seg_start, 4ull * A.n_seg}, {A.seg_end, 4ull * A.n_seg}, {A.seg_next_exon_beg, 4ull * A.n_seg}, {A.seg_gene_id, 4ull * A.n_seg},
        {A.seg_tid_off, 4ull * (A.n_seg + 1)}, {A.tid, 4ull * A.n_tid},
        {A.trans_start_ind, 4ull * A.n_trans}, {A.t2s_off, 4ull * (A.n_trans + 1)}, {A.t2s, (size_t) n_t2s},
        {A.gene_start, 4ull * A.n_genes}, {A.gene_end, 4ull * A.n_genes}, {A.chr_shift, 4ull * A.n_chr}, {A.chr_id, 4ull * A.n_chr},
        {A.near_border, 8ull * nw}, {A.intronic, 8ull * nw}};
    for (int i = 0; i < 21; i++) out[i] = a[i];
}

extern "C" int cm_annot_file_write(const char *path, const cm_annot_flat *A, uint32_t max_read_len, uint64_t tag) {
    if (!path || !A || !A->t2s_off) return CM_ERR_ARG;
    FILE *f = fopen(path, "wb");
    if (!f) return CM_ERR_ARG;
    AnnotFileHeader h;
    memset(&h, 0, sizeof h);
    memcpy(h.magic, "CMANNOT1", 8);
    h.version = 1; h.max_read_len = max_read_len; h.tag = tag;
    h.n_iv = A->n_iv; h.n_seg = A->n_seg; h.n_tid = A->n_tid; h.n_trans = A->n_trans; h.n_genes = A->n_genes; h.n_chr = A->n_chr;
    h.n_t2s = A->n_trans ? A->t2s_off[A->n_trans] : 0;
    h.n_bits = A->n_bits;
    AnnotFileArray arr[21];
    annot_arrays(*A, h.n_t2s, arr);
    bool ok = fwrite(&h, sizeof h, 1, f) == 1;
    const char zero[8] = {0};
    for (int i = 0; i < 21 && ok; i++) {
        if (arr[i].bytes) ok = arr[i].p && fwrite(arr[i].p, 1, arr[i].bytes, f) == arr[i].bytes;
        const size_t pad = (8 - arr[i].bytes % 8) % 8;
        if (ok && pad) ok = fwrite(zero, 1, pad, f) == pad;
    }
    ok = (fclose(f) == 0) && ok;
    if (!ok) remove(path);
    return ok ? CM_OK : CM_ERR_ARG;
}

// Reads the file into ONE malloc'ed block (*storage, release with cm_annot_file_free) and points `out` into it.
extern "C" int cm_annot_file_read(const char *path, cm_annot_flat *out, uint32_t *max_read_len, uint64_t *tag, void **storage) {
    if (!path || !out || !storage) return CM_ERR_ARG;
    *storage = nullptr;
    FILE *f = fopen(path, "rb");
    if (!f) return CM_ERR_ARG;
    AnnotFileHeader h;
    if (fread(&h, sizeof h, 1, f) != 1 || memcmp(h.magic, "CMANNOT1", 8) || h.version != 1) { fclose(f); return CM_ERR_ARG; }
    cm_annot_flat A;
    memset(&A, 0, sizeof A);
    A.n_iv = h.n_iv; A.n_seg = h.n_seg; A.n_tid = h.n_tid; A.n_trans = h.n_trans; A.n_genes = h.n_genes; A.n_chr = h.n_chr; A.n_bits = h.n_bits;
    AnnotFileArray arr[21];
    annot_arrays(A, h.n_t2s, arr);   // sizes only (pointers are null)
    size_t total = 0;
    for (int i = 0; i < 21; i++) total += (arr[i].bytes + 7) / 8 * 8;
    char *blk = (char *) malloc(total ? total : 8);
    if (!blk) { fclose(f); return CM_ERR_ALLOC; }
    if (total && fread(blk, 1, total, f) != total) { free(blk); fclose(f); return CM_ERR_ARG; }
    fclose(f);
    const void *ptr[21];
    size_t o = 0;
    for (int i = 0; i < 21; i++) { ptr[i] = blk + o; o += (arr[i].bytes + 7) / 8 * 8; }
    A.iv_spos = (const uint32_t *) ptr[0]; A.iv_epos = (const uint32_t *) ptr[1]; A.iv_max_end = (const uint32_t *) ptr[2];
    A.iv_min_end = (const uint32_t *) ptr[3]; A.iv_max_next_exon = (const uint32_t *) ptr[4]; A.iv_seg_off = (const uint32_t *) ptr[5];
    A.seg_start = (const uint32_t *) ptr[6]; A.seg_end = (const uint32_t *) ptr[7]; A.seg_next_exon_beg = (const uint32_t *) ptr[8];
    A.seg_gene_id = (const uint32_t *) ptr[9]; A.seg_tid_off = (const uint32_t *) ptr[10]; A.tid = (const uint32_t *) ptr[11];
    A.trans_start_ind = (const int32_t *) ptr[12]; A.t2s_off = (const uint32_t *) ptr[13]; A.t2s = (const uint8_t *) ptr[14];
    A.gene_start = (const uint32_t *) ptr[15]; A.gene_end = (const uint32_t *) ptr[16]; A.chr_shift = (const uint32_t *) ptr[17];
    A.chr_id = (const int32_t *) ptr[18]; A.near_border = (const uint64_t *) ptr[19]; A.intronic = (const uint64_t *) ptr[20];
    // the CSR offsets must be consistent with the counts before anything indexes with them
    if (A.iv_seg_off[A.n_iv] != A.n_seg || A.seg_tid_off[A.n_seg] != A.n_tid || (A.n_trans && A.t2s_off[A.n_trans] != h.n_t2s)) { free(blk); return CM_ERR_ARG; }
    *out = A;
    if (max_read_len) *max_read_len = h.max_read_len;
    if (tag) *tag = h.tag;
    *storage = blk;
    return CM_OK;
}

extern "C" void cm_annot_file_free(void *storage) { free(storage); }

// cm_gpu_load_annotation from a file written by cm_annot_file_write; refuses a file made for another -l or tag
extern "C" int cm_gpu_load_annotation_file(cm_gpu_ctx *ctx, int contig_id, const char *path, uint64_t expect_tag) {
    if (!ctx) return CM_ERR_ARG;
    cm_annot_flat A;
    uint32_t mrl = 0;
    uint64_t tag = 0;
    void *st = nullptr;
    int r = cm_annot_file_read(path, &A, &mrl, &tag, &st);
    if (r != CM_OK) return fail(ctx, r, "cannot read annotation file %s", path ? path : "(null)");
    if ((int) mrl != ctx->p.max_read_len || tag != expect_tag) {
        free(st);
        return fail(ctx, CM_ERR_STATE, "annotation file %s was made for max_read_len %u, tag %llu (context: %d, expected tag %llu)", path, mrl,
                    (unsigned long long) tag, ctx->p.max_read_len, (unsigned long long) expect_tag);
    }
    r = cm_gpu_load_annotation(ctx, contig_id, &A);
    free(st);
    return r;
}
