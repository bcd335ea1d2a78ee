// TEST INFRASTRUCTURE ONLY -- never part of the product.
// The drop-in shim (integration/gpu_map_reads.cpp + host_feed.h) compiled into one translation unit with a stand-in
// for the C-ABI calls it makes, so that the HOST side of the drop-in -- block reads, parallel FASTQ parse, header
// tokens -> cm_mread, PAM / remain-FASTQ text, write order -- can be checked byte for byte against the reference on a
// box without a GPU.  The stand-in answers cm_gpu_process_batch_pe with the reference's own FilterRead::process_read
// (src/filter.cpp:124), one pair at a time, on the scratch of mapping thread 0; nothing here touches libcm_gpu.so.
// With CM_MOCK_HOSTSIM=<tests/hostsim/libhostsim*.so> it also answers the two stage-2 batch entries, through the g++ build of the
// device headers and the same host orchestration the library uses (cm_circ_orch.h), so integration/gpu_circ_detect.cpp runs too;
// with CM_MOCK_STAGE1=hostsim as well, the mapping batches go through that build instead of the reference's process_read: the whole
// product logic (shim + device code) inside the reference's driver, on the CPU.
// Built by oracle/Makefile into oracle/_ref/circminer_feedtest; used by tests/test_host_feed.py.
#define CM_SHIM_KEEP_REF_TABLE 1   // the stand-in below maps with the reference's own table
#include "../../integration/gpu_map_reads.cpp"
#include <dlfcn.h>
#include <map>

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

// the host build of the device headers, loaded at run time (it is its own translation unit: no name meets the reference's)
struct HostSimLib {
    void *(*create)(const cm_params *) = nullptr;
    int (*set_text)(void *, int, const uint8_t *, uint32_t) = nullptr;
    int (*load_annotation)(void *, int, const cm_annot_flat *) = nullptr;
    int (*circ_single)(void *, int, const uint32_t *, const uint32_t *, uint32_t, const char *, const char *, uint32_t, const uint16_t *, const uint16_t *,
                       const cm_mread *, const uint32_t *, const uint32_t *, const uint32_t *, int, int, uint32_t, uint32_t *, uint8_t *, cm_circ_mate *,
                       cm_circ_mate *, cm_circ_mate *, int32_t *, cm_circ_res *, cm_circ_mate *, cm_circ_mate *, cm_circ_mate *, int32_t *,
                       cm_circ_res *) = nullptr;
    int (*circ_double)(void *, int, const uint32_t *, const uint32_t *, uint32_t, const char *, const char *, uint32_t, const uint16_t *, const uint16_t *,
                       const cm_mread *, const uint32_t *, const uint32_t *, const uint32_t *, int, int, uint32_t, uint32_t *, uint8_t *, cm_circ_mate *,
                       cm_circ_mate *, cm_circ_mate *, cm_circ_mate *, int32_t *, cm_circ_res *, int32_t *, cm_circ_res *) = nullptr;
    int (*load_contig)(void *, int, const void *const *, const void *, const uint8_t *, uint32_t) = nullptr;
    int (*process_batch_pe)(void *, int, uint32_t, const char *, const char *, uint32_t, const uint16_t *, const uint16_t *, const cm_mread *, cm_mread *,
                            int32_t *, cm_counters *) = nullptr;
    void *ctx = nullptr;
    int state = 0;   // 0 not tried, 1 loaded, -1 not asked for
    bool stage1 = false;   // CM_MOCK_STAGE1=hostsim: the mapping batches go through the host build as well
} g_hs;
bool mock_hs_load() {
    if (g_hs.state) return g_hs.state > 0;
    const char *so = getenv("CM_MOCK_HOSTSIM");
    g_hs.state = -1;
    if (!so || !*so) return false;
    void *h = dlopen(so, RTLD_NOW | RTLD_LOCAL);
    if (!h) { fprintf(stderr, "feed_mock: %s\n", dlerror()); exit(1); }
    *(void **) &g_hs.create = dlsym(h, "hs_create");
    *(void **) &g_hs.set_text = dlsym(h, "hs_set_text");
    *(void **) &g_hs.load_annotation = dlsym(h, "hs_load_annotation");
    *(void **) &g_hs.circ_single = dlsym(h, "hs_circ_single");
    *(void **) &g_hs.circ_double = dlsym(h, "hs_circ_double");
    *(void **) &g_hs.load_contig = dlsym(h, "hs_load_contig");
    *(void **) &g_hs.process_batch_pe = dlsym(h, "hs_process_batch_pe");
    const char *s1 = getenv("CM_MOCK_STAGE1");
    g_hs.stage1 = s1 && !strcmp(s1, "hostsim") && g_hs.load_contig && g_hs.process_batch_pe;
    fprintf(stderr, "feed_mock: stage 2 through %s%s\n", so, g_hs.stage1 ? ", the mapping batches too" : "");
    if (!g_hs.create || !g_hs.set_text || !g_hs.load_annotation || !g_hs.circ_single || !g_hs.circ_double) { fprintf(stderr, "feed_mock: symbols missing in %s\n", so); exit(1); }
    g_hs.state = 1;
    return true;
}
void *mock_hs_ctx() {
    if (!g_hs.ctx) g_hs.ctx = g_hs.create(&g_params);
    return g_hs.ctx;
}
// an output array of the C-ABI that the caller may leave NULL: the host build writes all of them
template <class T> struct MockBuf {
    std::vector<T> own;
    T *p;
    MockBuf(T *given, size_t n) : p(given) { if (!given) { own.resize(n ? n : 1); p = own.data(); } }
};
struct MockContig {
    std::vector<char> text;
    const void *loaded_for = nullptr;
    std::vector<uint32_t> pos, off;     // ws-mer index of the contig: positions (1-based) sorted by (hash, position), run offsets
    int ws = 0;
    cm_annot_flat flat;
    std::vector<std::vector<uint8_t>> store;
    template <class T> const T *copy(const T *src, size_t n) {
        store.emplace_back((const uint8_t *) src, (const uint8_t *) src + (n ? n : 1) * sizeof(T));
        return (const T *) store.back().data();
    }
    void keep(const cm_annot_flat *A) {
        store.clear();
        flat = *A;
        flat.iv_spos = copy(A->iv_spos, A->n_iv); flat.iv_epos = copy(A->iv_epos, A->n_iv); flat.iv_max_end = copy(A->iv_max_end, A->n_iv);
        flat.iv_min_end = copy(A->iv_min_end, A->n_iv); flat.iv_max_next_exon = copy(A->iv_max_next_exon, A->n_iv);
        flat.iv_seg_off = copy(A->iv_seg_off, (size_t) A->n_iv + 1);
        flat.seg_start = copy(A->seg_start, A->n_seg); flat.seg_end = copy(A->seg_end, A->n_seg);
        flat.seg_next_exon_beg = copy(A->seg_next_exon_beg, A->n_seg); flat.seg_gene_id = copy(A->seg_gene_id, A->n_seg);
        flat.seg_tid_off = copy(A->seg_tid_off, (size_t) A->n_seg + 1);
        flat.tid = copy(A->tid, A->n_tid);
        flat.trans_start_ind = copy(A->trans_start_ind, A->n_trans); flat.t2s_off = copy(A->t2s_off, (size_t) A->n_trans + 1);
        flat.t2s = copy(A->t2s, A->n_trans ? A->t2s_off[A->n_trans] : 0);
        flat.gene_start = copy(A->gene_start, A->n_genes); flat.gene_end = copy(A->gene_end, A->n_genes);
        flat.chr_shift = copy(A->chr_shift, A->n_chr); flat.chr_id = copy(A->chr_id, A->n_chr);
        flat.near_border = copy(A->near_border, (size_t) ((A->n_bits + 63) / 64)); flat.intronic = copy(A->intronic, (size_t) ((A->n_bits + 63) / 64));
    }
    void circ_index(int w) {
        if (ws == w && !off.empty()) return;
        ws = w;
        const size_t nb = (size_t) 1 << (2 * w), L = text.size();
        off.assign(nb + 1, 0);
        std::vector<uint32_t> key(L, 0xffffffffu);
        auto code = [](char ch) { switch (ch) { case 'A': case 'a': return 0; case 'C': case 'c': return 1; case 'G': case 'g': return 2; case 'T': case 't': return 3; default: return 4; } };
        uint32_t h = 0;
        int good = 0;
        for (size_t i = 0; i < L; i++) {
            const int v = code(text[i]);
            if (v == 4) { good = 0; h = 0; continue; }
            h = ((h << 2) | (uint32_t) v) & (uint32_t) (nb - 1);
            if (++good >= w) { key[i + 1 - w] = h; off[h + 1]++; }
        }
        for (size_t b = 0; b < nb; b++) off[b + 1] += off[b];
        pos.assign(off[nb] ? off[nb] : 1, 0);
        std::vector<uint32_t> at(off.begin(), off.end() - 1);
        for (size_t i = 0; i < L; i++)
            if (key[i] != 0xffffffffu) pos[at[key[i]]++] = (uint32_t) i + 1;
    }
};
std::map<int, MockContig> g_mock_contigs;
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
// ---- stage 2 over the host build of the device headers (tests/hostsim), when CM_MOCK_HOSTSIM names its library -------------------
// With it the stand-in also answers cm_gpu_circ_single_batch / cm_gpu_circ_double_batch, so the drop-in's do_process
// (integration/gpu_circ_detect.cpp) runs on the CPU and its candidates.pam / circ_report can be compared with the reference's.
// Without it those two entries fail and the tests run the binary with CM_GPU_STAGE2=0.
std::mutex g_mock_mu;   // the shim builds the contig on all its devices at once; here they are one host-build context
int cm_gpu_build_contig_packed(cm_gpu_ctx *, int c, const uint64_t *, uint32_t len) {
    std::lock_guard<std::mutex> lk(g_mock_mu);
    if (!mock_hs_load()) return CM_OK;
    if (g_mock_contigs.count(c) && g_mock_contigs[c].text.size() == len && g_mock_contigs[c].loaded_for == _ih_crefGen) return CM_OK;   // another device's copy
    std::vector<char> text((size_t) len + 2);
    if (!genome_seeder.pac2char_otf(1, (int) len, text.data())) return CM_ERR_STATE;   // the contig the reference has just loaded
    MockContig &m = g_mock_contigs[c];
    m.text.assign(text.begin(), text.begin() + len);
    m.loaded_for = _ih_crefGen;
    m.pos.clear(); m.off.clear(); m.ws = 0;
    if (g_hs.stage1)   // with the k-mer table the reference has just loaded (what cm_gpu_load_contig takes)
        return g_hs.load_contig(mock_hs_ctx(), c, (const void *const *) _ih_hashTable, _ih_hashTableMem, (const uint8_t *) m.text.data(), len) == 0 ? CM_OK : CM_ERR_STATE;
    return g_hs.set_text(mock_hs_ctx(), c, (const uint8_t *) m.text.data(), len) == 0 ? CM_OK : CM_ERR_STATE;
}
int cm_gpu_load_annotation(cm_gpu_ctx *, int c, const cm_annot_flat *A) {
    std::lock_guard<std::mutex> lk(g_mock_mu);
    // the two bit vectors as the shim hands them over (copied by words from the std::bitsets) against the bitsets read bit by bit
    for (uint64_t k = 0; k < A->n_bits; k++) {
        const bool nb = (size_t) c < near_border_bs.size() && near_border_bs[c][k], in = (size_t) c < intronic_bs.size() && intronic_bs[c][k];
        if (((A->near_border[k >> 6] >> (k & 63)) & 1) != (uint64_t) nb || ((A->intronic[k >> 6] >> (k & 63)) & 1) != (uint64_t) in) {
            fprintf(stderr, "feed_mock: annotation bit %llu of contig %d differs from the reference's bitset\n", (unsigned long long) k, c);
            exit(1);
        }
    }
    for (uint64_t k = A->n_bits; k < ((A->n_bits + 63) / 64) * 64; k++)
        if (((A->near_border[k >> 6] >> (k & 63)) & 1) || ((A->intronic[k >> 6] >> (k & 63)) & 1)) { fprintf(stderr, "feed_mock: annotation bits past n_bits set\n"); exit(1); }
    if (!mock_hs_load()) return CM_OK;
    MockContig &m = g_mock_contigs[c];
    m.keep(A);                                       // the host build keeps pointers; the shim's arrays die after this call
    return g_hs.load_annotation(mock_hs_ctx(), c, &m.flat) == 0 ? CM_OK : CM_ERR_STATE;
}
int cm_annot_file_write(const char *, const cm_annot_flat *, uint32_t, uint64_t) { return CM_ERR_ARG; }
int cm_annot_file_read(const char *, cm_annot_flat *, uint32_t *, uint64_t *, void **) { return CM_ERR_ARG; }
void cm_annot_file_free(void *) {}
int cm_gpu_circ_single_batch(cm_gpu_ctx *, int c, uint32_t n, const char *s1, const char *s2, uint32_t stride, const uint16_t *l1, const uint16_t *l2,
                             const cm_mread *mr, const uint32_t *goff, const uint32_t *gs, const uint32_t *ge, int ws, int step, uint32_t cap,
                             uint32_t *n_att, uint32_t *att_pair, uint8_t *att_r1p, cm_circ_mate *in_r1, cm_circ_mate *in_r2, cm_circ_mate *in_pm,
                             int32_t *ret, cm_circ_res *cr, cm_circ_mate *out_r1, cm_circ_mate *out_r2, cm_circ_mate *out_pm, int32_t *best_type,
                             cm_circ_res *best_cr) {
    std::lock_guard<std::mutex> lk(g_mock_mu);
    if (!mock_hs_load()) return CM_ERR_STATE;
    if (n_att) *n_att = 0;
    if (n == 0) return CM_OK;
    MockContig &m = g_mock_contigs[c];
    m.circ_index(ws);
    MockBuf<uint32_t> b_pair(att_pair, cap);
    MockBuf<uint8_t> b_r1p(att_r1p, cap);
    MockBuf<cm_circ_mate> b_i1(in_r1, cap), b_i2(in_r2, cap), b_ip(in_pm, cap), b_o1(out_r1, cap), b_o2(out_r2, cap), b_op(out_pm, cap);
    MockBuf<int32_t> b_ret(ret, cap);
    MockBuf<cm_circ_res> b_cr(cr, cap);
    const int r = g_hs.circ_single(mock_hs_ctx(), c, m.pos.data(), m.off.data(), n, s1, s2, stride, l1, l2, mr, goff, gs, ge, ws, step, cap, b_pair.p, b_r1p.p,
                                   b_i1.p, b_i2.p, b_ip.p, b_ret.p, b_cr.p, b_o1.p, b_o2.p, b_op.p, best_type, best_cr);
    if (r < 0) return CM_ERR_STATE;
    *n_att = (uint32_t) r;
    return CM_OK;
}
int cm_gpu_circ_double_batch(cm_gpu_ctx *, int c, uint32_t n, const char *s1, const char *s2, uint32_t stride, const uint16_t *l1, const uint16_t *l2,
                             const cm_mread *mr, const uint32_t *goff, const uint32_t *gs, const uint32_t *ge, int ws, int step, uint32_t cap,
                             uint32_t *n_ev, uint32_t *ev_pair, uint8_t *ev_kind, cm_circ_mate *ev_r1, cm_circ_mate *ev_r2, cm_circ_mate *ev_p1,
                             cm_circ_mate *ev_p2, int32_t *ev_ret, cm_circ_res *ev_cr, int32_t *push_type, cm_circ_res *push_cr) {
    std::lock_guard<std::mutex> lk(g_mock_mu);
    if (!mock_hs_load()) return CM_ERR_STATE;
    if (n_ev) *n_ev = 0;
    if (n == 0) return CM_OK;
    MockContig &m = g_mock_contigs[c];
    m.circ_index(ws);
    MockBuf<uint32_t> b_pair(ev_pair, cap);
    MockBuf<uint8_t> b_kind(ev_kind, cap);
    MockBuf<cm_circ_mate> b_r1(ev_r1, cap), b_r2(ev_r2, cap), b_p1(ev_p1, cap), b_p2(ev_p2, cap);
    MockBuf<int32_t> b_ret(ev_ret, cap);
    MockBuf<cm_circ_res> b_cr(ev_cr, cap);
    const int r = g_hs.circ_double(mock_hs_ctx(), c, m.pos.data(), m.off.data(), n, s1, s2, stride, l1, l2, mr, goff, gs, ge, ws, step, cap, b_pair.p, b_kind.p,
                                   b_r1.p, b_r2.p, b_p1.p, b_p2.p, b_ret.p, b_cr.p, push_type, push_cr);
    if (r < 0) return CM_ERR_STATE;
    *n_ev = (uint32_t) r;
    return CM_OK;
}
int cm_gpu_host_alloc(size_t bytes, void **out) { *out = malloc(bytes ? bytes : 1); return *out ? CM_OK : CM_ERR_ALLOC; }
int cm_gpu_host_free(void *p) { free(p); return CM_OK; }

int cm_gpu_process_batch_pe(cm_gpu_ctx *, int contig, uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride,
                            const uint16_t *len1, const uint16_t *len2, const cm_mread *in, cm_mread *out, int32_t *state) {
    if (mock_hs_load() && g_hs.stage1)   // the device code's host build instead of the reference's process_read
        return g_hs.process_batch_pe(mock_hs_ctx(), contig, n, seqs1, seqs2, stride, len1, len2, in, out, state, nullptr) == 0 ? CM_OK : CM_ERR_STATE;
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
