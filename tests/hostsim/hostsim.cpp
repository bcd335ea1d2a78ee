// TEST INFRASTRUCTURE ONLY -- never loaded by the product path (circminer_b200/).
//
// g++ build of the very same CM_HD functions the CUDA kernels run (circminer_b200/csrc/cm_*.h), so
// that the per-read logic can be compared with the reference (oracle/_ref) in the GPU-less build
// container.  It is a debugging aid for bit-exactness, not a CPU fallback: the C-ABI library
// libcm_gpu.so has no host path and fails without a device.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../circminer_b200/csrc/cm_pipeline.h"
#include "../../circminer_b200/csrc/cm_circ.h"
#define CM_CIRC_TRACE 1
#include "../../circminer_b200/csrc/cm_circ_split.h"
#include "../../circminer_b200/csrc/cm_circ_orch.h"

#pragma pack(push, 1)
struct GeneralIndexPacked {
    uint16_t checksum;
    uint32_t info;
};
#pragma pack(pop)

struct HsContig {
    std::vector<uint32_t> dir, pos;
    std::vector<uint16_t> cks;
    std::vector<uint8_t> genome;
    ContigDev d{};
    AnnotDev a{};
    bool has_annot = false;
};

struct HsCtx {
    cm_params p;
    int max_seg_cnt, n_slots;
    std::vector<HsContig *> contigs;
};

static uint8_t hs_comp(uint8_t c) { return cm_comp(c); }

extern "C" {

HsCtx *hs_create(const cm_params *p) {
    HsCtx *c = new HsCtx;
    c->p = *p;
    c->max_seg_cnt = cm_max_seg_cnt(*p);
    c->n_slots = (c->max_seg_cnt + 1) / 2;
    return c;
}

void hs_destroy(HsCtx *c) {
    for (auto *x : c->contigs) delete x;
    delete c;
}

// same sources as cm_gpu_load_contig, plus the contig text as ASCII (genome_ascii[0] = base 1)
int hs_load_contig(HsCtx *c, int contig_id, const void *const *bucket_lists, const void *table_mem, const uint8_t *genome_ascii,
                   uint32_t ref_len) {
    if ((size_t) contig_id >= c->contigs.size()) c->contigs.resize(contig_id + 1, nullptr);
    delete c->contigs[contig_id];
    HsContig *h = new HsContig;
    c->contigs[contig_id] = h;
    h->dir.resize((size_t) CM_NBUCKETS + 1);
    uint64_t total = 0;
    for (uint32_t hv = 0; hv < CM_NBUCKETS; hv++) {
        h->dir[hv] = (uint32_t) total;
        const GeneralIndexPacked *l = (const GeneralIndexPacked *) bucket_lists[hv];
        if (l) total += l[0].info;
    }
    h->dir[CM_NBUCKETS] = (uint32_t) total;
    h->cks.resize(total);
    h->pos.resize(total);
    for (uint32_t hv = 0; hv < CM_NBUCKETS; hv++) {
        const GeneralIndexPacked *l = (const GeneralIndexPacked *) bucket_lists[hv];
        if (!l) continue;
        for (uint32_t j = 0; j < l[0].info; j++) {
            h->cks[h->dir[hv] + j] = l[1 + j].checksum;
            h->pos[h->dir[hv] + j] = l[1 + j].info;
        }
    }
    (void) table_mem;
    h->genome.assign((size_t) ref_len + 64, 'N');
    memcpy(h->genome.data() + 1, genome_ascii, ref_len);
    h->d.dir = h->dir.data();
    h->d.cks = h->cks.data();
    h->d.pos = h->pos.data();
    h->d.genome = h->genome.data();
    h->d.n_entries = total;
    h->d.ref_len = ref_len;
    return 0;
}

// index given directly as the SoA arrays (built by the caller from the contig text)
int hs_set_index(HsCtx *c, int contig_id, const uint32_t *dir, const uint16_t *cks, const uint32_t *pos, uint64_t n_entries,
                 const uint8_t *genome_ascii, uint32_t ref_len) {
    if ((size_t) contig_id >= c->contigs.size()) c->contigs.resize(contig_id + 1, nullptr);
    delete c->contigs[contig_id];
    HsContig *h = new HsContig;
    c->contigs[contig_id] = h;
    h->dir.assign(dir, dir + (size_t) CM_NBUCKETS + 1);
    h->cks.assign(cks, cks + n_entries);
    h->pos.assign(pos, pos + n_entries);
    h->genome.assign((size_t) ref_len + 64, 'N');
    memcpy(h->genome.data() + 1, genome_ascii, ref_len);
    h->d.dir = h->dir.data(); h->d.cks = h->cks.data(); h->d.pos = h->pos.data(); h->d.genome = h->genome.data();
    h->d.n_entries = n_entries; h->d.ref_len = ref_len;
    return 0;
}

// contig text alone, no k-mer index: enough for stage 2 (tests/hostfeed/feed_mock.cpp)
int hs_set_text(HsCtx *c, int contig_id, const uint8_t *genome_ascii, uint32_t ref_len) {
    if ((size_t) contig_id >= c->contigs.size()) c->contigs.resize(contig_id + 1, nullptr);
    delete c->contigs[contig_id];
    HsContig *h = new HsContig;
    c->contigs[contig_id] = h;
    h->genome.assign((size_t) ref_len + 64, 'N');
    memcpy(h->genome.data() + 1, genome_ascii, ref_len);
    h->d.dir = nullptr; h->d.cks = nullptr; h->d.pos = nullptr; h->d.genome = h->genome.data();
    h->d.n_entries = 0; h->d.ref_len = ref_len;
    return 0;
}

// Host build of the fused directory + coarse tables of the large buckets (what build_dirx does on the device), then
// cm_seed_lookup_x against the plain cm_seed_lookup on every stored (bucket, checksum) of a sample of entries, their
// neighbours checksum +- 1, and random keys.  Returns the number of look-ups that differ (0 = the two are the same map).
long hs_seed_x_check(HsCtx *c, int contig_id, uint32_t big_th, uint32_t n_probe, uint64_t seed) {
    HsContig *h = c->contigs[contig_id];
    const int cs_bits = 2 * (c->p.kmer - CM_WINDOW);
    const int bits = cs_bits < 10 ? cs_bits : 10;
    DirX *dx = (DirX *) calloc((size_t) CM_NBUCKETS, sizeof(DirX));   // untouched pages stay unmapped: only non-empty buckets are written
    std::vector<uint32_t> tab;
    uint32_t n_big = 0;
    for (uint32_t hv = 0; hv < CM_NBUCKETS; hv++) {
        const uint32_t lo = h->dir[hv], hi = h->dir[hv + 1];
        if (hi == lo) continue;
        DirX e;
        e.off = lo;
        e.cnt = hi - lo;
        for (uint32_t k = 0; k < CM_DIRX_INLINE; k++) e.c[k] = lo + k < hi ? h->cks[lo + k] : (uint16_t) 0xFFFFu;
        if (cs_bits > 0 && e.cnt >= big_th) {
            const uint32_t id = n_big++, nt = 1u << bits;
            e.c[0] = (uint16_t) (id & 0xFFFFu);
            e.c[1] = (uint16_t) (id >> 16);
            for (uint32_t t = 0; t <= nt; t++) {
                uint32_t a = 0;
                if (t == nt) a = e.cnt;
                else while (a < e.cnt && (uint32_t) (h->cks[lo + a] >> (cs_bits - bits)) < t) a++;
                tab.push_back(a);
            }
        }
        dx[hv] = e;
    }
    ContigDev d = h->d;
    d.dirx = dx;
    d.big_tab = n_big ? tab.data() : nullptr;
    d.big_th = n_big ? big_th : 0xFFFFFFFFu;
    d.big_bits = bits;
    d.big_shift = cs_bits - bits;
    ContigDev plain = h->d;
    plain.dirx = nullptr;
    long bad = 0;
    uint64_t x = seed * 0x9E3779B97F4A7C15ull + 1;
    auto rnd = [&]() { x ^= x << 13; x ^= x >> 7; x ^= x << 17; return x; };
    auto probe = [&](int hv, int cv) {
        uint32_t s1, s2;
        SeedHit a = cm_seed_lookup(plain, hv, cv, &s1), b = cm_seed_lookup_x(d, hv, cv, &s2);
        if (a.cnt != b.cnt || (a.cnt && a.off != b.off) || s1 != s2) bad++;
    };
    const int cmax = cs_bits ? (1 << cs_bits) : 1;
    // every large bucket in full, then samples
    for (uint32_t hv = 0; hv < CM_NBUCKETS; hv++)
        if (dx[hv].cnt >= big_th)
            for (uint32_t j = 0; j < dx[hv].cnt; j++) {
                const int cv = h->cks[dx[hv].off + j];
                probe((int) hv, cv);
                if (cv + 1 < cmax) probe((int) hv, cv + 1);
                if (cv > 0) probe((int) hv, cv - 1);
            }
    for (uint32_t k = 0; k < n_probe && h->d.n_entries; k++) {
        const uint64_t e = rnd() % h->d.n_entries;
        // bucket of entry e
        uint32_t a = 0, b = CM_NBUCKETS;
        while (a < b) { uint32_t mid = (a + b) >> 1; if (h->dir[mid + 1] <= e) a = mid + 1; else b = mid; }
        const int cv = h->cks[e];
        probe((int) a, cv);
        if (cv + 1 < cmax) probe((int) a, cv + 1);
        if (cv > 0) probe((int) a, cv - 1);
        probe((int) (rnd() % CM_NBUCKETS), (int) (rnd() % cmax));
    }
    probe(-1, 0);
    probe(0, -1);
    free(dx);
    return bad + (n_big == 0 && big_th < 50 ? 1000000 : 0);   // a small threshold must produce large buckets on the test data
}

// the arrays of `A` must outlive the context (the Python wrapper keeps them)
int hs_load_annotation(HsCtx *c, int contig_id, const cm_annot_flat *A) {
    HsContig *h = c->contigs[contig_id];
    AnnotDev &a = h->a;
    a.n_iv = A->n_iv; a.iv_spos = A->iv_spos; a.iv_epos = A->iv_epos; a.iv_max_end = A->iv_max_end;
    a.iv_min_end = A->iv_min_end; a.iv_max_next_exon = A->iv_max_next_exon; a.iv_seg_off = A->iv_seg_off;
    a.n_seg = A->n_seg; a.seg_start = A->seg_start; a.seg_end = A->seg_end; a.seg_next_exon_beg = A->seg_next_exon_beg;
    a.seg_gene_id = A->seg_gene_id; a.seg_tid_off = A->seg_tid_off; a.tid = A->tid;
    a.n_trans = A->n_trans; a.trans_start_ind = A->trans_start_ind; a.t2s_off = A->t2s_off; a.t2s = A->t2s;
    a.n_genes = A->n_genes; a.gene_start = A->gene_start; a.gene_end = A->gene_end;
    a.n_chr = A->n_chr; a.chr_shift = A->chr_shift; a.chr_id = A->chr_id;
    a.n_bits = A->n_bits; a.near_border = A->near_border; a.intronic = A->intronic;
    h->has_annot = true;
    return 0;
}

static DevView make_view(HsCtx *c, int contig_id) {
    DevView v;
    v.p = c->p;
    v.c = c->contigs[contig_id]->d;
    v.a = c->contigs[contig_id]->a;
    v.contig_id = contig_id;
    v.max_seg_cnt = c->max_seg_cnt;
    v.n_slots = c->n_slots;
    return v;
}

// host mirror of the device batch: strands + seeds + chains
struct HsBatch {
    std::vector<uint8_t> strands;
    std::vector<uint16_t> slen;
    std::vector<uint32_t> seed_off, seed_cnt, chain_len;
    std::vector<int32_t> chain_cnt, chain_hh;
    std::vector<float> chain_score;
    std::vector<FragDev> chain_frags;
    std::vector<uint32_t> chain_base;
    std::vector<ChainAux> aux;
    std::vector<EndSpec> spec;
    BatchDev b{};
};

static void build_batch(HsCtx *c, const DevView &v, HsBatch &B, uint32_t n_reads, const char *const *srcs, const uint16_t *const *lens,
                        int n_src, uint32_t in_stride, uint64_t *links) {
    // read r of source k -> strands 2*(r*n_src+k) (fwd) and +1 (rc)
    uint32_t stride = (uint32_t) ((c->p.max_read_len + 15) / 16 * 16);
    uint32_t ns = 2 * n_reads * n_src;
    B.strands.assign((size_t) ns * stride, 0);
    B.slen.resize(ns);
    for (uint32_t r = 0; r < n_reads; r++)
        for (int k = 0; k < n_src; k++) {
            uint32_t s0 = 2 * (r * n_src + k);
            const uint8_t *src = (const uint8_t *) srcs[k] + (size_t) r * in_stride;
            uint32_t len = lens[k][r];
            uint8_t *fw = &B.strands[(size_t) s0 * stride], *rc = fw + stride;
            for (uint32_t j = 0; j < len; j++) {
                fw[j] = src[j];
                rc[j] = hs_comp(src[len - 1 - j]);
            }
            B.slen[s0] = B.slen[s0 + 1] = (uint16_t) len;
        }
    int C = c->p.max_chain_len, F = c->n_slots;
    B.seed_off.assign((size_t) ns * c->n_slots, 0);
    B.seed_cnt.assign((size_t) ns * c->n_slots, 0);
    B.chain_cnt.assign(ns, 0);
    B.chain_hh.assign(ns, 0);
    B.chain_score.assign((size_t) ns * C, 0.f);
    B.chain_len.assign((size_t) ns * C, 0);
    B.chain_frags.assign((size_t) ns * C * F, FragDev{0, 0});
    BatchDev &b = B.b;
    b.strands = B.strands.data(); b.strand_len = B.slen.data(); b.n_strands = ns; b.stride = stride; b.F = F; b.C = C;
    b.seed_off = B.seed_off.data(); b.seed_cnt = B.seed_cnt.data(); b.chain_cnt = B.chain_cnt.data(); b.chain_hh = B.chain_hh.data();
    b.chain_score = B.chain_score.data(); b.chain_len = B.chain_len.data(); b.chain_frags = B.chain_frags.data();
    std::vector<double> dps;
    std::vector<uint32_t> dpp;
    std::vector<ChainRec> recs;
    for (uint32_t s = 0; s < ns; s++) {
        uint32_t *so = &B.seed_off[(size_t) s * c->n_slots], *sc = &B.seed_cnt[(size_t) s * c->n_slots];
        cm_strand_seed(v, b.strands + (size_t) s * stride, B.slen[s], so, sc);
        uint32_t row_pos_off[CM_MAX_SEEDS], row_cnt[CM_MAX_SEEDS], row_off[CM_MAX_SEEDS], cells;
        int hh;
        int T = cm_chain_rows(v, so, sc, row_pos_off, row_cnt, row_off, cells, hh);
        B.chain_hh[s] = hh;
        if (T <= 0) continue;
        dps.resize(cells + 1);
        dpp.resize(cells + 1);
        uint32_t nrec = cm_chain_dp(v, B.slen[s], T, row_pos_off, row_cnt, row_off, dps.data(), dpp.data(), nullptr, 0, nullptr);
        recs.resize(nrec + 1);
        cm_chain_dp(v, B.slen[s], T, row_pos_off, row_cnt, row_off, dps.data(), dpp.data(), recs.data(), nrec, links);
        ChainOut out;
        out.count = &B.chain_cnt[s];
        out.score = &B.chain_score[(size_t) s * C];
        out.len = &B.chain_len[(size_t) s * C];
        out.frags = &B.chain_frags[(size_t) s * C * F];
        out.F = F;
        cm_chain_backtrack(v, T, row_pos_off, row_cnt, row_off, dps.data(), dpp.data(), recs.data(), nrec, out);
    }
    // dense pre-alignment of every chain (k_chain_spec on the device); HS_NO_SPEC=1 keeps every DP inline
    b.chain_base = nullptr;
    b.aux = nullptr;
    b.spec = nullptr;
    b.spec_k = getenv("HS_SPEC_K") ? atoi(getenv("HS_SPEC_K")) : CM_MAX_CHAINS;
    if (!getenv("HS_NO_SPEC")) {
        B.chain_base.assign(ns + 1, 0);
        for (uint32_t s = 0; s < ns; s++) B.chain_base[s + 1] = B.chain_base[s] + (uint32_t) (B.chain_cnt[s] < b.spec_k ? B.chain_cnt[s] : b.spec_k);
        B.aux.assign(B.chain_base[ns] + 1, ChainAux{});
        B.spec.assign(2 * (size_t) B.chain_base[ns] + 2, EndSpec{});
        const bool walk = !getenv("HS_NO_WALK_SPEC");
        ExtCtx x;
        ext_ctx_init(x, &v);
        for (uint32_t s = 0; s < ns; s++) {
            ChainListView l = cm_chain_list(b, s);
            for (int k = 0; k < l.count && k < b.spec_k; k++) {
                ChainView ch = cl_chain(l, k);
                EndSpec *sp = walk ? &B.spec[2 * (size_t) (B.chain_base[s] + k)] : nullptr;
                ChainAux &ax = B.aux[B.chain_base[s] + k];
                for (int kind = 0; kind < 3; kind++)
                    if (cm_chain_spec(x, ch, b.strands + (size_t) s * stride, B.slen[s], kind, ax, sp))
                        cm_chain_walk(x, ch, b.strands + (size_t) s * stride, B.slen[s], kind == 2, ax, sp);
            }
        }
        b.chain_base = B.chain_base.data();
        b.aux = B.aux.data();
        b.spec = walk ? B.spec.data() : nullptr;
    }
}

int hs_chain_batch(HsCtx *c, int contig_id, uint32_t n, const char *seqs, uint32_t stride, const uint16_t *len, int32_t *chain_count,
                   int32_t *high_hits, float *score, uint32_t *chain_len, uint32_t *frag_rpos, int32_t *frag_qpos) {
    DevView v = make_view(c, contig_id);
    HsBatch B;
    const char *srcs[1] = {seqs};
    const uint16_t *lens[1] = {len};
    build_batch(c, v, B, n, srcs, lens, 1, stride, nullptr);
    int C = c->p.max_chain_len, F = c->n_slots;
    for (uint32_t i = 0; i < n; i++) {
        uint32_t s = 2 * i;   // forward strand of row i
        chain_count[i] = B.chain_cnt[s];
        high_hits[i] = B.chain_hh[s];
        for (int k = 0; k < C; k++) {
            bool live = k < B.chain_cnt[s];
            score[(size_t) i * C + k] = live ? B.chain_score[(size_t) s * C + k] : 0.f;
            uint32_t L = live ? B.chain_len[(size_t) s * C + k] : 0;
            chain_len[(size_t) i * C + k] = L;
            for (int f = 0; f < F; f++) {
                bool lf = (uint32_t) f < L;
                frag_rpos[((size_t) i * C + k) * F + f] = lf ? B.chain_frags[((size_t) s * C + k) * F + f].rpos : 0;
                frag_qpos[((size_t) i * C + k) * F + f] = lf ? B.chain_frags[((size_t) s * C + k) * F + f].qpos : 0;
            }
        }
    }
    return 0;
}

// ProcessCirc::chaining per row through the device functions of cm_circ.h; circ_pos / circ_off = the contig's ws-mer index
// (positions sorted by (hash, position), run offsets), made by the caller
int hs_circ_chain(HsCtx *c, int contig_id, const uint32_t *circ_pos, const uint32_t *circ_off, uint32_t n, const char *seqs, uint32_t stride,
                  const uint32_t *qspos, const uint32_t *qepos, const uint32_t *gene_start, const uint32_t *gene_end, int ws, int step, int F,
                  int32_t *chain_count, int32_t *kmer_cnt, float *score, uint32_t *chain_len, uint32_t *frag_rpos, int32_t *frag_qpos) {
    DevView v = make_view(c, contig_id);
    const CircIndex ix{circ_pos, circ_off, ws};
    const int C = c->p.max_chain_len;
    std::vector<double> dps;
    std::vector<uint32_t> dpp;
    std::vector<ChainRec> recs;
    std::vector<FragDev> frags((size_t) C * F);
    for (uint32_t i = 0; i < n; i++) {
        uint32_t row_pos[CM_CIRC_MAX_ROWS], row_cnt[CM_CIRC_MAX_ROWS], row_off[CM_CIRC_MAX_ROWS], cells;
        int32_t qp[CM_CIRC_MAX_ROWS];
        const int kc = cm_circ_rows(ix, c->p, (const uint8_t *) seqs + (size_t) i * stride, qspos[i], qepos[i], gene_start[i], gene_end[i], step, F,
                                    row_pos, row_cnt, qp, cells);
        if (kmer_cnt) kmer_cnt[i] = kc;
        const int T = cm_circ_trim(kc, row_cnt, row_off);
        std::fill(frags.begin(), frags.end(), FragDev{0, 0});
        ChainOut out;
        int32_t cnt = 0;
        out.count = &cnt;
        out.score = score + (size_t) i * C;
        out.len = chain_len + (size_t) i * C;
        out.frags = frags.data();
        out.F = F;
        for (int k = 0; k < C; k++) { out.score[k] = 0.f; out.len[k] = 0; }
        if (T > 0) {
            const ChainGeomCirc g{ix.pos, gene_start[i], ws, qp};
            dps.resize(cells + 1);
            dpp.resize(cells + 1);
            uint32_t nrec = cm_chain_dp_g(g, v, (int) qepos[i], T, row_pos, row_cnt, row_off, dps.data(), dpp.data(), (ChainRec *) nullptr, 0u, nullptr);
            recs.resize(nrec + 1);
            cm_chain_dp_g(g, v, (int) qepos[i], T, row_pos, row_cnt, row_off, dps.data(), dpp.data(), recs.data(), nrec, nullptr);
            cm_chain_backtrack_g(g, v, T, row_pos, row_cnt, row_off, dps.data(), dpp.data(), recs.data(), nrec, out, gene_start[i]);
            cm_circ_finish(kc, out);
        }
        chain_count[i] = cnt;
        for (size_t q = 0; q < frags.size(); q++) {
            frag_rpos[(size_t) i * C * F + q] = frags[q].rpos;
            frag_qpos[(size_t) i * C * F + q] = frags[q].qpos;
        }
    }
    return 0;
}

// ProcessCirc::find_exact_coord per row through cm_circ_exact (same arguments as cm_gpu_circ_exact_batch)
int hs_circ_exact(HsCtx *c, int contig_id, uint32_t n, const char *seqs, uint32_t stride, const uint32_t *r1_spos, const uint32_t *r1_epos,
                  const uint32_t *r2_spos, const uint32_t *r2_epos, const int32_t *dir, const uint32_t *qspos, const int32_t *rlen,
                  const int32_t *whole_len, uint32_t F, const uint32_t *chain_len, const uint32_t *frag_rpos, const int32_t *frag_qpos, int ws,
                  cm_mate *pmm, int32_t *ret) {
    DevView v = make_view(c, contig_id);
    v.p.kmer = ws;
    std::vector<FragDev> fr(F);
    for (uint32_t i = 0; i < n; i++) {
        ExtCtx x;
        ext_ctx_init(x, &v);
        for (uint32_t f = 0; f < F; f++) fr[f] = FragDev{frag_rpos[(size_t) i * F + f], frag_qpos[(size_t) i * F + f]};
        ChainView bc;
        bc.frags = fr.data(); bc.chain_len = chain_len[i]; bc.aux = nullptr; bc.spec = nullptr;
        MateD a, b, pm;
        mate_init(a, v.p); mate_init(b, v.p); mate_init(pm, v.p);
        a.spos = r1_spos[i]; a.epos = r1_epos[i];
        b.spos = r2_spos[i]; b.epos = r2_epos[i];
        ret[i] = cm_circ_exact(x, a, b, pm, dir[i], qspos[i], (const uint8_t *) seqs + (size_t) i * stride, rlen[i], whole_len[i], bc) ? 1 : 0;
        cm_mate &o = pmm[i];
        o.spos = pm.spos; o.epos = pm.epos; o.qspos = pm.qspos; o.qepos = pm.qepos; o.matched_len = pm.matched_len;
        o.left_ed = pm.left_ed; o.middle_ed = pm.middle_ed; o.right_ed = pm.right_ed; o.sclen_left = pm.sclen_left; o.sclen_right = pm.sclen_right;
        o.dir = pm.dir; o.type = pm.type;
        o.is_concord = pm.is_concord; o.left_ok = pm.left_ok; o.right_ok = pm.right_ok; o.pad_ = 0;
    }
    return 0;
}

// ProcessCirc::check_split_map (three mates) per row through cm_circ_check_row (arguments of cm_gpu_circ_check_batch)
int hs_circ_check(HsCtx *c, int contig_id, uint32_t n, const char *remain_seq, const char *fullmap_seq, uint32_t stride, const int32_t *remain_len,
                  const int32_t *fullmap_len, const uint8_t *r1_partial, cm_circ_mate *mm_r1, cm_circ_mate *mm_r2, cm_circ_mate *partial_mm, int ws,
                  int32_t *ret, cm_circ_res *cr) {
    DevView v = make_view(c, contig_id);
    v.p.kmer = ws;
    for (uint32_t i = 0; i < n; i++)
        ret[i] = cm_circ_check_row(v, (const uint8_t *) remain_seq + (size_t) i * stride, remain_len[i], (const uint8_t *) fullmap_seq + (size_t) i * stride,
                                   fullmap_len[i], r1_partial[i] != 0, mm_r1[i], mm_r2[i], partial_mm[i], cr[i]);
    return 0;
}

// find_exact_coord with the pair's mates carried from chain to chain (arguments of cm_gpu_circ_exact_mates_batch)
int hs_circ_exact_mates(HsCtx *c, int contig_id, uint32_t n, const char *seqs, uint32_t stride, cm_circ_mate *r1, cm_circ_mate *r2, const int32_t *dir,
                        const uint32_t *qspos, const int32_t *rlen, const int32_t *whole_len, uint32_t F, const uint32_t *chain_len,
                        const uint32_t *frag_rpos, const int32_t *frag_qpos, int ws, cm_circ_mate *pm, int32_t *ret) {
    DevView v = make_view(c, contig_id);
    v.p.kmer = ws;
    std::vector<FragDev> fr(F);
    for (uint32_t i = 0; i < n; i++) {
        for (uint32_t f = 0; f < F; f++) fr[f] = FragDev{frag_rpos[(size_t) i * F + f], frag_qpos[(size_t) i * F + f]};
        ChainView bc;
        bc.frags = fr.data(); bc.chain_len = chain_len[i]; bc.aux = nullptr; bc.spec = nullptr;
        ret[i] = cm_circ_exact_row(v, r1[i], r2[i], pm[i], dir[i], qspos[i], (const uint8_t *) seqs + (size_t) i * stride, rlen[i], whole_len[i], bc);
    }
    return 0;
}

// the single-split loop (cm_circ_orch.h) over the host build of the three pieces
struct HsCircEngine {
    HsCtx *c;
    const uint32_t *circ_pos, *circ_off;
};
static int hs_eng_chain(void *self, int contig, uint32_t n, const char *seqs, uint32_t stride, const uint32_t *qspos, const uint32_t *qepos,
                        const uint32_t *gs, const uint32_t *ge, int ws, int step, int F, int32_t *count, float *score, uint32_t *clen, uint32_t *rpos,
                        int32_t *qpos) {
    HsCircEngine *e = (HsCircEngine *) self;
    return hs_circ_chain(e->c, contig, e->circ_pos, e->circ_off, n, seqs, stride, qspos, qepos, gs, ge, ws, step, F, count, nullptr, score, clen, rpos, qpos);
}
static int hs_eng_exact(void *self, int contig, uint32_t n, const char *seqs, uint32_t stride, cm_circ_mate *r1, cm_circ_mate *r2, const int32_t *dir,
                        const uint32_t *qspos, const int32_t *rlen, const int32_t *wl, uint32_t F, const uint32_t *clen, const uint32_t *rpos,
                        const int32_t *qpos, int ws, cm_circ_mate *pm, int32_t *ret) {
    return hs_circ_exact_mates(((HsCircEngine *) self)->c, contig, n, seqs, stride, r1, r2, dir, qspos, rlen, wl, F, clen, rpos, qpos, ws, pm, ret);
}
static int hs_eng_check(void *self, int contig, uint32_t n, const char *rem, const char *ful, uint32_t stride, const int32_t *rl, const int32_t *fl,
                        const uint8_t *r1p, cm_circ_mate *r1, cm_circ_mate *r2, cm_circ_mate *pm, int ws, int32_t *ret, cm_circ_res *cr, uint8_t *sw) {
    HsCtx *c = ((HsCircEngine *) self)->c;
    DevView v = make_view(c, contig);
    v.p.kmer = ws;
    for (uint32_t i = 0; i < n; i++)
        ret[i] = cm_circ_check_row(v, (const uint8_t *) rem + (size_t) i * stride, rl[i], (const uint8_t *) ful + (size_t) i * stride, fl[i], r1p[i] != 0,
                                   r1[i], r2[i], pm[i], cr[i], sw + i);
    return 0;
}
static int hs_eng_dbl(void *self, int contig, uint32_t n, const char *r1seq, const char *r2seq, uint32_t stride, cm_circ_mate *r1, cm_circ_mate *r2,
                      const uint32_t *qs1, const uint32_t *qe1, const uint32_t *qs2, const uint32_t *qe2, uint32_t F, const uint32_t *clen1,
                      const uint32_t *rpos1, const int32_t *qpos1, const uint32_t *clen2, const uint32_t *rpos2, const int32_t *qpos2, int ws,
                      cm_circ_mate *p1, cm_circ_mate *p2, int32_t *code) {
    HsCtx *c = ((HsCircEngine *) self)->c;
    DevView v = make_view(c, contig);
    v.p.kmer = ws;
    std::vector<FragDev> f1(F), f2(F);
    for (uint32_t i = 0; i < n; i++) {
        for (uint32_t f = 0; f < F; f++) {
            f1[f] = FragDev{rpos1[(size_t) i * F + f], qpos1[(size_t) i * F + f]};
            f2[f] = FragDev{rpos2[(size_t) i * F + f], qpos2[(size_t) i * F + f]};
        }
        ChainView c1, c2;
        c1.frags = f1.data(); c1.chain_len = clen1[i]; c1.aux = nullptr; c1.spec = nullptr;
        c2.frags = f2.data(); c2.chain_len = clen2[i]; c2.aux = nullptr; c2.spec = nullptr;
        code[i] = cm_circ_double_row(v, r1[i], r2[i], qs1[i], qe1[i], qs2[i], qe2[i], (const uint8_t *) r1seq + (size_t) i * stride,
                                     (const uint8_t *) r2seq + (size_t) i * stride, c1, c2, p1[i], p2[i]);
    }
    return 0;
}
static int hs_eng_check4(void *self, int contig, uint32_t n, const char *r1seq, const char *r2seq, uint32_t stride, const int32_t *l1, const int32_t *l2,
                         const cm_circ_mate *r1, const cm_circ_mate *r2, const cm_circ_mate *p1, const cm_circ_mate *p2, int ws, int32_t *ret,
                         cm_circ_res *cr) {
    HsCtx *c = ((HsCircEngine *) self)->c;
    DevView v = make_view(c, contig);
    v.p.kmer = ws;
    for (uint32_t i = 0; i < n; i++)
        ret[i] = cm_circ_check4_row(v, (const uint8_t *) r1seq + (size_t) i * stride, l1[i], (const uint8_t *) r2seq + (size_t) i * stride, l2[i], nullptr, 0,
                                    r1[i], r2[i], p1[i], p2[i], cr[i]);
    return 0;
}
// check_split_map on four mates as a function of its arguments; fullmap_sel: 0 none, 1 = R1, 2 = R2 (see cmref_circ_check4)
int hs_circ_check4(HsCtx *c, int contig_id, uint32_t n, const char *r1seq, const char *r2seq, uint32_t stride, const int32_t *l1, const int32_t *l2,
                   const uint8_t *fullmap_sel, const cm_circ_mate *r1_1, const cm_circ_mate *r2_1, const cm_circ_mate *r1_2, const cm_circ_mate *r2_2,
                   int ws, int32_t *ret, cm_circ_res *cr) {
    DevView v = make_view(c, contig_id);
    v.p.kmer = ws;
    for (uint32_t i = 0; i < n; i++) {
        const uint8_t *a = (const uint8_t *) r1seq + (size_t) i * stride, *b = (const uint8_t *) r2seq + (size_t) i * stride;
        const uint8_t *f = fullmap_sel[i] == 1 ? a : fullmap_sel[i] == 2 ? b : nullptr;
        const int fl = fullmap_sel[i] == 1 ? l1[i] : fullmap_sel[i] == 2 ? l2[i] : 0;
        ret[i] = cm_circ_check4_row(v, a, l1[i], b, l2[i], f, fl, r1_1[i], r2_1[i], r1_2[i], r2_2[i], cr[i]);
    }
    return 0;
}
// call_circ_double_split for a batch: -> number of events (<= cap written), or < 0
int hs_circ_double(HsCtx *c, int contig_id, const uint32_t *circ_pos, const uint32_t *circ_off, uint32_t n, const char *seqs1, const char *seqs2,
                   uint32_t stride, const uint16_t *len1, const uint16_t *len2, const cm_mread *mr, const uint32_t *gene_off, const uint32_t *gene_start,
                   const uint32_t *gene_end, int ws, int step, uint32_t cap, uint32_t *ev_pair, uint8_t *ev_kind, cm_circ_mate *ev_r1, cm_circ_mate *ev_r2,
                   cm_circ_mate *ev_p1, cm_circ_mate *ev_p2, int32_t *ev_ret, cm_circ_res *ev_cr, int32_t *push_type, cm_circ_res *push_cr) {
    HsCircEngine he{c, circ_pos, circ_off};
    CircEngine E{&he, contig_id, hs_eng_chain, hs_eng_exact, hs_eng_check, hs_eng_dbl, hs_eng_check4};
    CircSingleIn in{n, seqs1, seqs2, stride, len1, len2, mr, gene_off, gene_start, gene_end, ws, step, c->p.max_read_len, c->p.max_chain_len, 4};
    std::vector<CircEvent> ev;
    int r = circ_orch_double(E, in, ev, push_type, push_cr);
    if (r != 0) return r < 0 ? r : -r;
    for (size_t k = 0; k < ev.size() && k < cap; k++) {
        ev_pair[k] = ev[k].pair; ev_kind[k] = ev[k].kind; ev_r1[k] = ev[k].r1; ev_r2[k] = ev[k].r2; ev_p1[k] = ev[k].p1; ev_p2[k] = ev[k].p2;
        ev_ret[k] = ev[k].ret; ev_cr[k] = ev[k].cr;
    }
    return (int) ev.size();
}

// -> number of attempts (<= cap written), or < 0
int hs_circ_single(HsCtx *c, int contig_id, const uint32_t *circ_pos, const uint32_t *circ_off, uint32_t n, const char *seqs1, const char *seqs2,
                   uint32_t stride, const uint16_t *len1, const uint16_t *len2, const cm_mread *mr, const uint32_t *gene_off, const uint32_t *gene_start,
                   const uint32_t *gene_end, int ws, int step, uint32_t cap, uint32_t *att_pair, uint8_t *att_r1_partial, cm_circ_mate *in_r1,
                   cm_circ_mate *in_r2, cm_circ_mate *in_pm, int32_t *ret, cm_circ_res *cr, cm_circ_mate *out_r1, cm_circ_mate *out_r2,
                   cm_circ_mate *out_pm, int32_t *best_type, cm_circ_res *best_cr) {
    HsCircEngine he{c, circ_pos, circ_off};
    CircEngine E{&he, contig_id, hs_eng_chain, hs_eng_exact, hs_eng_check};
    CircSingleIn in{n, seqs1, seqs2, stride, len1, len2, mr, gene_off, gene_start, gene_end, ws, step, c->p.max_read_len, c->p.max_chain_len, 4};
    std::vector<CircAttempt> att;
    int r = circ_orch_single(E, in, att, best_type, best_cr);
    if (r != 0) return r < 0 ? r : -r;
    for (size_t k = 0; k < att.size() && k < cap; k++) {
        att_pair[k] = att[k].pair; att_r1_partial[k] = att[k].r1_partial;
        in_r1[k] = att[k].in_r1; in_r2[k] = att[k].in_r2; in_pm[k] = att[k].in_pm;
        out_r1[k] = att[k].out_r1; out_r2[k] = att[k].out_r2; out_pm[k] = att[k].out_pm;
        ret[k] = att[k].ret; cr[k] = att[k].cr;
    }
    return (int) att.size();
}

// which paths of the stage-2 classification the rows so far reached (see CM_CIRC_HIT in cm_circ_split.h); resets the counters
void hs_circ_trace(long *out16) {
    for (int k = 0; k < 16; k++) { out16[k] = g_circ_trace[k]; g_circ_trace[k] = 0; }
}

int hs_process_batch_pe(HsCtx *c, int contig_id, uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride,
                        const uint16_t *len1, const uint16_t *len2, const cm_mread *in, cm_mread *out, int32_t *state,
                        cm_counters *ctr) {
    DevView v = make_view(c, contig_id);
    HsBatch B;
    const char *srcs[2] = {seqs1, seqs2};
    const uint16_t *lens[2] = {len1, len2};
    uint64_t links = 0;
    build_batch(c, v, B, n, srcs, lens, 2, stride, &links);
    AlignStats tot{0, 0, 0, 0};
    for (uint32_t i = 0; i < n; i++) {
        cm_mread mr = in[i];
        AlignStats st;
        state[i] = cm_pair_process(v, B.b, i, mr, st);
        out[i] = mr;
        tot.xdrop_cells += st.xdrop_cells; tot.edit_cells += st.edit_cells;
        tot.xdrop_calls += st.xdrop_calls; tot.edit_calls += st.edit_calls;
    }
    if (ctr) {
        memset(ctr, 0, sizeof *ctr);
        ctr->chain_links = links; ctr->xdrop_cells = tot.xdrop_cells; ctr->edit_cells = tot.edit_cells;
        ctr->xdrop_calls = tot.xdrop_calls; ctr->edit_calls = tot.edit_calls; ctr->pairs = n;
    }
    return 0;
}

// cm_gpu_map_rounds_pe on the host: the same skip rule / header round trip (cm_rounds.h), survivors as an index list
int hs_map_rounds_pe(HsCtx *c, int first, int n_contigs, uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride,
                     const uint16_t *len1, const uint16_t *len2, const cm_mread *in, cm_mread *out, int32_t *state,
                     int32_t *final_round, cm_mread *round_out, int32_t *round_state) {
    std::vector<uint32_t> ids(n);
    for (uint32_t i = 0; i < n; i++) ids[i] = i;
    std::vector<cm_mread> cur(in, in + n);
    if (round_state) for (size_t k = 0; k < (size_t) n_contigs * n; k++) round_state[k] = -1;
    for (int r = 0; r < n_contigs && !ids.empty(); r++) {
        const uint32_t m = (uint32_t) ids.size();
        std::vector<char> s1((size_t) m * stride), s2((size_t) m * stride);
        std::vector<uint16_t> l1(m), l2(m);
        std::vector<cm_mread> o(m);
        std::vector<int32_t> st(m);
        for (uint32_t k = 0; k < m; k++) {
            memcpy(&s1[(size_t) k * stride], seqs1 + (size_t) ids[k] * stride, stride);
            memcpy(&s2[(size_t) k * stride], seqs2 + (size_t) ids[k] * stride, stride);
            l1[k] = len1[ids[k]];
            l2[k] = len2[ids[k]];
        }
        hs_process_batch_pe(c, first + r, m, s1.data(), s2.data(), stride, l1.data(), l2.data(), cur.data(), o.data(), st.data(), nullptr);
        const bool last = r == n_contigs - 1;
        std::vector<uint32_t> nids;
        std::vector<cm_mread> ncur;
        for (uint32_t k = 0; k < m; k++) {
            const uint32_t id = ids[k];
            if (round_out) round_out[(size_t) r * n + id] = o[k];
            if (round_state) round_state[(size_t) r * n + id] = st[k];
            if (cm_skip_rule(c->p, st[k], o[k], (int) l1[k] + (int) l2[k]) || last) {
                out[id] = o[k];
                state[id] = st[k];
                if (final_round) final_round[id] = first + r;
            } else {
                nids.push_back(id);
                ncur.push_back(cm_header_round_trip(o[k], c->p));
            }
        }
        ids.swap(nids);
        cur.swap(ncur);
    }
    return 0;
}

int hs_process_batch_se(HsCtx *c, int contig_id, uint32_t n, const char *seqs, uint32_t stride, const uint16_t *len, int32_t *state) {
    DevView v = make_view(c, contig_id);
    HsBatch B;
    const char *srcs[1] = {seqs};
    const uint16_t *lens[1] = {len};
    build_batch(c, v, B, n, srcs, lens, 1, stride, nullptr);
    for (uint32_t i = 0; i < n; i++) {
        AlignStats st;
        state[i] = cm_read_process_se(v, B.b, i, st);
    }
    return 0;
}

int hs_align_batch(HsCtx *c, int kind, uint32_t n, const char *s, const uint16_t *n_s, const char *t, const uint16_t *n_t,
                   uint32_t stride, const int32_t *aux, int32_t *out_ed, int32_t *out_sclen, int32_t *out_indel, int32_t *out_score) {
    for (uint32_t i = 0; i < n; i++) {
        const uint8_t *si = (const uint8_t *) s + (size_t) i * stride, *ti = (const uint8_t *) t + (size_t) i * stride;
        int ed = 0, sc = 0, indel = 0, score = 0;
        if (kind == 0 || kind == 1) ed = cm_drop_alignment_sc(si, n_s[i], ti, n_t[i], kind == 1, c->p, sc, indel, score, nullptr);
        else if (kind == 2 || kind == 3) ed = cm_local_alignment(si, n_s[i], ti, n_t[i], kind == 3, c->p, indel, score, nullptr);
        else if (kind == 4) ed = cm_one_side_banded(si, n_s[i], ti, n_t[i], aux[i], nullptr);
        else if (kind == 5 || kind == 6) ed = cm_edit_alignment_sc(si, n_s[i], ti, n_t[i], kind == 6, c->p, sc, indel, score, nullptr);
        out_ed[i] = ed; out_sclen[i] = sc; out_indel[i] = indel; out_score[i] = score;
    }
    return 0;
}

}  // extern "C"
