// The reference-side binding of libcm_gpu.so: a batched replacement of map_reads (src/circminer.cpp:354-406).
// Compiled against the reference's own headers and linked with its unmodified objects (oracle/Makefile target
// `gpu`): the reference keeps reading FASTQ, keeping the per-round MatchedRead state in the read headers and
// writing mapping.pam / *_remain_R?.fastq; only FilterRead::process_read is replaced, N pairs at a time, by
// cm_gpu_process_batch_pe.  This file is what INTEGRATION.md describes, in working form.
#define private public
#define protected public
#include "common.h"
#include "filter.h"
#include "fastq_parser.h"
#include "gene_annotation.h"
#include "match_read.h"
#undef private
#undef protected

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../include/cm_gpu.h"

extern "C" {
#include "mrsfast/Common.h"
#include "mrsfast/HashTable.h"
extern IHashTable *_ih_hashTable;
extern GeneralIndex *_ih_hashTableMem;
extern CompressedSeq *_ih_crefGen;
extern int _ih_refGenLen;
extern int _ih_maxHashTableSize;
}
extern pthread_mutex_t buffer_lock;

namespace {

cm_gpu_ctx *g_ctx = nullptr;
cm_params g_params;
const bool g_verbose = getenv("CM_GPU_VERBOSE") != nullptr;   // phase times on stderr
double now_s() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}
FilterArgs *g_fa = nullptr;   // mapping thread 0's arguments (the reference's per-thread scratch)
std::vector<std::string> g_chr;   // chromosome names <-> cm_mread::chr_r1 / chr_r2 indices
int g_loaded_contig = -1;

[[noreturn]] void die(const char *what) {
    fprintf(stderr, "circminer_gpu: %s: %s\n", what, cm_gpu_last_error(g_ctx));
    exit(1);
}

// Called by the parsing threads of host_feed.h: g_chr never reallocates (reserved once), readers scan the
// published prefix, a new name is appended under the mutex.
std::atomic<size_t> g_nchr(0);
std::mutex g_chr_mu;
constexpr size_t CHR_CAP = 1 << 16;

int chr_index(const std::string &name) {
    if (name == "-") return -1;
    const size_t n0 = g_nchr.load(std::memory_order_acquire);
    for (size_t j = 0; j < n0; j++)
        if (g_chr[j] == name) return (int) j;
    std::lock_guard<std::mutex> lk(g_chr_mu);
    if (g_chr.capacity() < CHR_CAP) g_chr.reserve(CHR_CAP);
    for (size_t j = n0; j < g_chr.size(); j++)
        if (g_chr[j] == name) return (int) j;
    if (g_chr.size() >= CHR_CAP) { fprintf(stderr, "circminer_gpu: more than %zu chromosome names\n", CHR_CAP); exit(1); }
    g_chr.push_back(name);
    g_nchr.store(g_chr.size(), std::memory_order_release);
    return (int) g_chr.size() - 1;
}

void to_pod(const MatchedRead *r, cm_mread *m) {
    memset(m, 0, sizeof *m);
    m->spos_r1 = r->spos_r1; m->spos_r2 = r->spos_r2; m->epos_r1 = r->epos_r1; m->epos_r2 = r->epos_r2;
    m->qspos_r1 = r->qspos_r1; m->qspos_r2 = r->qspos_r2; m->qepos_r1 = r->qepos_r1; m->qepos_r2 = r->qepos_r2;
    m->mlen_r1 = r->mlen_r1; m->mlen_r2 = r->mlen_r2;
    m->r1_forward = r->r1_forward; m->r2_forward = r->r2_forward;
    m->ed_r1 = r->ed_r1; m->ed_r2 = r->ed_r2; m->type = r->type; m->tlen = r->tlen;
    m->junc_num = r->junc_num; m->gm_compatible = r->gm_compatible; m->contig_num = r->contig_num;
    m->chr_r1 = chr_index(r->chr_r1);
    m->chr_r2 = chr_index(r->chr_r2);
}

void from_pod(const cm_mread *m, MatchedRead *r) {
    r->spos_r1 = m->spos_r1; r->spos_r2 = m->spos_r2; r->epos_r1 = m->epos_r1; r->epos_r2 = m->epos_r2;
    r->qspos_r1 = m->qspos_r1; r->qspos_r2 = m->qspos_r2; r->qepos_r1 = m->qepos_r1; r->qepos_r2 = m->qepos_r2;
    r->mlen_r1 = m->mlen_r1; r->mlen_r2 = m->mlen_r2;
    r->r1_forward = m->r1_forward; r->r2_forward = m->r2_forward;
    r->ed_r1 = m->ed_r1; r->ed_r2 = m->ed_r2; r->type = m->type; r->tlen = m->tlen;
    r->junc_num = m->junc_num; r->gm_compatible = m->gm_compatible; r->contig_num = m->contig_num;
    r->chr_r1 = m->chr_r1 >= 0 ? g_chr[m->chr_r1] : std::string("-");
    r->chr_r2 = m->chr_r2 >= 0 ? g_chr[m->chr_r2] : std::string("-");
}

}   // namespace
#include "host_feed.h"
namespace {

// GTFParser tables of the current contig as the flat arrays of cm_annot_flat
void upload_annotation(int c) {
    FlatIntervalTree<UniqSeg> &T = gtf_parser.exons_int_map[c];
    const size_t n_iv = T.disjoint_intervals.size();
    std::vector<uint32_t> iv_spos(n_iv), iv_epos(n_iv), iv_max_end(n_iv), iv_min_end(n_iv), iv_next(n_iv), iv_seg_off(n_iv + 1);
    std::vector<uint32_t> seg_start, seg_end, seg_next, seg_gene, seg_tid_off, tid;
    for (size_t i = 0; i < n_iv; i++) {
        auto &iv = T.disjoint_intervals[i];
        iv_spos[i] = iv.spos; iv_epos[i] = iv.epos; iv_max_end[i] = iv.max_end; iv_min_end[i] = iv.min_end;
        iv_next[i] = iv.max_next_exon;
        iv_seg_off[i] = (uint32_t) seg_start.size();
        for (auto &s : iv.seg_list) {
            seg_start.push_back(s.start); seg_end.push_back(s.end); seg_next.push_back(s.next_exon_beg);
            seg_gene.push_back(s.gene_id);
            seg_tid_off.push_back((uint32_t) tid.size());
            for (auto t : s.trans_id) tid.push_back(t);
        }
    }
    iv_seg_off[n_iv] = (uint32_t) seg_start.size();
    seg_tid_off.push_back((uint32_t) tid.size());
    const size_t ntr = (size_t) c < gtf_parser.trans2seg.size() ? gtf_parser.trans2seg[c].size() : 0;
    std::vector<int32_t> trans_start(ntr);
    std::vector<uint32_t> t2s_off(ntr + 1);
    std::vector<uint8_t> t2s;
    for (size_t i = 0; i < ntr; i++) {
        trans_start[i] = gtf_parser.trans_start_ind[c][i];
        t2s_off[i] = (uint32_t) t2s.size();
        for (auto v : gtf_parser.trans2seg[c][i]) t2s.push_back((uint8_t) v);
    }
    t2s_off[ntr] = (uint32_t) t2s.size();
    std::vector<uint32_t> gene_start, gene_end;
    if ((size_t) c < gtf_parser.gid2ginfo.size())
        for (auto &g : gtf_parser.gid2ginfo[c]) { gene_start.push_back(g.start); gene_end.push_back(g.end); }
    std::vector<uint32_t> chr_shift;
    std::vector<int32_t> chr_id;
    for (auto &cc : gtf_parser.con2chr[c]) { chr_shift.push_back(cc.shift); chr_id.push_back(chr_index(cc.contig)); }
    const uint64_t n_bits = (uint64_t) _ih_refGenLen + 2;
    std::vector<uint64_t> nb((n_bits + 63) / 64, 0), in((n_bits + 63) / 64, 0);
    {   // std::bitset has no word access: the bits are read one by one, whole words per thread
        const int T = std::max(1, threadCount);
        const size_t nw = nb.size();
        const bool has_nb = (size_t) c < near_border_bs.size(), has_in = (size_t) c < intronic_bs.size();
        feed::parallel_for(T, [&](int t) {
            for (size_t w = nw * t / T; w < nw * (t + 1) / T; w++) {
                uint64_t a = 0, b = 0;
                const uint64_t k0 = (uint64_t) w << 6, k1 = std::min<uint64_t>(k0 + 64, n_bits);
                for (uint64_t k = k0; k < k1; k++) {
                    if (has_nb && near_border_bs[c][k]) a |= 1ull << (k & 63);
                    if (has_in && intronic_bs[c][k]) b |= 1ull << (k & 63);
                }
                nb[w] = a; in[w] = b;
            }
        });
    }
    auto pad = [](auto &v) { if (v.empty()) v.resize(1); };
    pad(seg_start); pad(seg_end); pad(seg_next); pad(seg_gene); pad(tid); pad(trans_start); pad(t2s); pad(gene_start); pad(gene_end);
    pad(chr_shift); pad(chr_id);
    cm_annot_flat A;
    memset(&A, 0, sizeof A);
    A.n_iv = (uint32_t) n_iv;
    A.iv_spos = iv_spos.data(); A.iv_epos = iv_epos.data(); A.iv_max_end = iv_max_end.data(); A.iv_min_end = iv_min_end.data();
    A.iv_max_next_exon = iv_next.data(); A.iv_seg_off = iv_seg_off.data();
    A.n_seg = iv_seg_off[n_iv];
    A.seg_start = seg_start.data(); A.seg_end = seg_end.data(); A.seg_next_exon_beg = seg_next.data(); A.seg_gene_id = seg_gene.data();
    A.seg_tid_off = seg_tid_off.data();
    A.n_tid = seg_tid_off.back();
    A.tid = tid.data();
    A.n_trans = (uint32_t) ntr;
    A.trans_start_ind = trans_start.data(); A.t2s_off = t2s_off.data(); A.t2s = t2s.data();
    A.n_genes = (size_t) c < gtf_parser.gid2ginfo.size() ? (uint32_t) gtf_parser.gid2ginfo[c].size() : 0;
    A.gene_start = gene_start.data(); A.gene_end = gene_end.data();
    A.n_chr = (uint32_t) gtf_parser.con2chr[c].size();
    A.chr_shift = chr_shift.data(); A.chr_id = chr_id.data();
    A.n_bits = n_bits;
    A.near_border = nb.data(); A.intronic = in.data();
    if (cm_gpu_load_annotation(g_ctx, c, &A) != CM_OK) die("cm_gpu_load_annotation");
}

int gpu_device() {
    const char *dev = getenv("CM_GPU_DEVICE");
    return dev ? atoi(dev) : 0;
}

// CUDA context creation overlaps the reference's own start-up (GTF, index meta, first loadHashTable)
struct Warmup {
    std::thread th;
    Warmup() : th([] { cm_gpu_warmup(gpu_device()); }) {}
    void wait() { if (th.joinable()) th.join(); }
    ~Warmup() { wait(); }
} g_warmup;

void ensure_round(int c) {
    const double t_in = now_s();
    g_warmup.wait();
    if (!g_ctx) {
        g_chr.reserve(CHR_CAP);
        cm_params &p = g_params;
        cm_params_default(&p);
        p.kmer = kmer; p.max_read_len = maxReadLength; p.max_ed = maxEd; p.max_sc = maxSc; p.band_width = bandWidth;
        p.seed_lim = seedLim; p.max_tlen = maxTlen; p.max_intron_len = maxIntronLen; p.max_chain_len = maxChainLen;
        p.scan_level = scanLevel;
        if (cm_gpu_create(gpu_device(), &p, &g_ctx) != CM_OK) die("cm_gpu_create");
    }
    if (g_loaded_contig == c) return;
    const double t_ctx = now_s();
    // Index of this contig on the device.  Default: built there from the 3-bit packed contig the reference has just
    // loaded (38 MB over PCIe per 100 Mbp).  CM_GPU_INDEX=upload converts and uploads the reference's own bucket table
    // instead (the two are compared entry by entry in tests/test_gpu_seed.py).
    double t_scan = now_s();
    const char *how = getenv("CM_GPU_INDEX");
    if (how && !strcmp(how, "upload")) {
        // entries in use = end of the furthest bucket (_ih_hashTableMemSize is only maintained by the index builder)
        uint64_t n = 0;
        const int T = std::max(1, threadCount);
        std::vector<uint64_t> part(T, 0);
        feed::parallel_for(T, [&](int t) {
            uint64_t m = 0;
            for (int64_t hv = (int64_t) _ih_maxHashTableSize * t / T; hv < (int64_t) _ih_maxHashTableSize * (t + 1) / T; hv++)
                if (_ih_hashTable[hv].list) {
                    uint64_t end = (uint64_t) (_ih_hashTable[hv].list - _ih_hashTableMem) + _ih_hashTable[hv].list[0].info + 1;
                    if (end > m) m = end;
                }
            part[t] = m;
        });
        for (uint64_t m : part) n = std::max(n, m);
        t_scan = now_s();
        if (cm_gpu_load_contig(g_ctx, c, (const void *const *) _ih_hashTable, _ih_hashTableMem, n, (const uint64_t *) _ih_crefGen,
                               (uint32_t) _ih_refGenLen) != CM_OK)
            die("cm_gpu_load_contig");
    } else if (cm_gpu_build_contig_packed(g_ctx, c, (const uint64_t *) _ih_crefGen, (uint32_t) _ih_refGenLen) != CM_OK)
        die("cm_gpu_build_contig_packed");
    const double t_idx = now_s();
    upload_annotation(c);
    g_loaded_contig = c;
    if (g_verbose)
        fprintf(stderr, "circminer_gpu: contig %d: context %.3f s, table scan %.3f s, index on device %.3f s, annotation %.3f s\n", c,
                t_ctx - t_in, t_scan - t_ctx, t_idx - t_scan, now_s() - t_idx);
}

// a Record that outlives FASTQParser's per-thread buffer
struct Held {
    Record rec;
    std::vector<char> buf;
    void take(const Record *src) {
        const size_t ln = strlen(src->rname) + 1, ls = src->seq_len + 1, lc = strlen(src->comment) + 1, lq = strlen(src->qual) + 1;
        buf.resize(ln + 2 * ls + lc + 2 * lq);
        char *p = buf.data();
        rec.rname = p; memcpy(p, src->rname, ln); p += ln;
        rec.seq = p; memcpy(p, src->seq, ls); p += ls;
        rec.rcseq = p; memcpy(p, src->rcseq, ls); p += ls;
        rec.comment = p; memcpy(p, src->comment, lc); p += lc;
        rec.qual = p; memcpy(p, src->qual, lq); p += lq;
        rec.rqual = p; memcpy(p, src->rqual, lq);
        rec.seq_len = src->seq_len;
        *rec.mr = *src->mr;
    }
};

template <class T> T *pinned(size_t n) {
    void *p = nullptr;
    if (cm_gpu_host_alloc(n * sizeof(T), &p) != CM_OK) die("cm_gpu_host_alloc");
    return (T *) p;
}

// Batched rounds with the parallel host feed of host_feed.h (PAM output or none).
void gpu_map_reads_parallel(size_t B) {
    const int is_last = filter_read.get_last_round();
    const int T = std::max(1, threadCount);
    const uint32_t stride = (uint32_t) maxReadLength;
    const size_t cap = std::min<size_t>(std::max<size_t>(B * 512, 1u << 20), 1u << 30);
    // block buffers, pinned batch arrays and text buffers live across rounds (one allocation per run)
    static feed::RawFile raw[2][2];
    static size_t B_alloc = 0;
    static char *s1, *s2;
    static uint16_t *l1, *l2;
    static cm_mread *in, *out;
    static int32_t *state;
    static std::vector<feed::RecMeta> m1, m2;
    static std::vector<feed::TextOut> text;
    const double t_setup = now_s();
    for (int s = 0; s < 2; s++) {
        raw[s][0].init(fq_parser1.gzinput, cap);
        raw[s][1].init(fq_parser2.gzinput, cap);
    }
    if (B_alloc < B) {
        if (B_alloc) {
            cm_gpu_host_free(s1); cm_gpu_host_free(s2); cm_gpu_host_free(l1); cm_gpu_host_free(l2);
            cm_gpu_host_free(in); cm_gpu_host_free(out); cm_gpu_host_free(state);
        }
        s1 = pinned<char>(B * stride); s2 = pinned<char>(B * stride);
        l1 = pinned<uint16_t>(B); l2 = pinned<uint16_t>(B);
        in = pinned<cm_mread>(B); out = pinned<cm_mread>(B);
        state = pinned<int32_t>(B);
        m1.resize(B); m2.resize(B);
        B_alloc = B;
    }
    text.resize(T);
    cm_mread fresh;
    cm_mread_init(&fresh, &g_params);
    int cur = 0;
    const double t_begin = now_s();
    if (g_verbose) fprintf(stderr, "circminer_gpu: host buffers %.3f s\n", t_begin - t_setup);
    feed::parallel_for(2, [&](int k) { raw[cur][k].fill(); });
    feed::LineIndex ix1, ix2;
    double ph[6] = {0, 0, 0, 0, 0, 0};   // first read, parse, map, format, write, wait for the next block
    ph[0] = now_s() - t_begin;
    size_t n_batches = 0, n_total = 0;
    while (true) {
        double t0 = now_s();
        feed::count_lines(raw[cur][0], T, ix1);
        feed::count_lines(raw[cur][1], T, ix2);
        const size_t n = (size_t) std::min<uint64_t>(B, std::min(ix1.lines, ix2.lines) / 4);
        if (!n) break;
        size_t e[2] = {raw[cur][0].len, raw[cur][1].len};
        feed::parse_records(raw[cur][0], ix1, T, n, stride, m1.data(), s1, l1, in, fresh, e[0]);
        feed::parse_records(raw[cur][1], ix2, T, n, stride, m2.data(), s2, l2, nullptr, fresh, e[1]);
        // the next block of both files is read while this batch is mapped and written
        std::thread pre[2];
        for (int k = 0; k < 2; k++)
            pre[k] = std::thread([&, k] {
                feed::RawFile &a = raw[cur][k], &b = raw[cur ^ 1][k];
                b.len = a.len - e[k];
                memcpy(b.buf.data(), a.buf.data() + e[k], b.len);
                b.eof = a.eof;
                b.fill();
            });
        double t1 = now_s();
        ph[1] += t1 - t0;
        if (cm_gpu_process_batch_pe(g_ctx, contigNum, (uint32_t) n, s1, s2, stride, l1, l2, in, out, state) != CM_OK)
            die("cm_gpu_process_batch_pe");
        t0 = now_s();
        ph[2] += t0 - t1;
        const char *b1 = raw[cur][0].buf.data(), *b2 = raw[cur][1].buf.data();
        feed::parallel_for(T, [&](int t) {
            feed::TextOut &o = text[t];
            o.pam.clear(); o.r1.clear(); o.r2.clear();
            feed::format_range(n * t / T, n * (t + 1) / T, b1, b2, m1.data(), m2.data(), out, state, is_last, o);
        });
        t1 = now_s();
        ph[3] += t1 - t0;
        for (int t = 0; t < T; t++) {   // input order == the reference's -t 1 order
            if (!text[t].pam.empty()) fwrite(text[t].pam.data(), 1, text[t].pam.size(), filter_read.sam_output.outsam);
            if (!text[t].r1.empty()) fwrite(text[t].r1.data(), 1, text[t].r1.size(), filter_read.temp_fq_r1);
            if (!text[t].r2.empty()) fwrite(text[t].r2.data(), 1, text[t].r2.size(), filter_read.temp_fq_r2);
        }
        t0 = now_s();
        ph[4] += t0 - t1;
        for (auto &p : pre) p.join();
        ph[5] += now_s() - t0;
        cur ^= 1;
        n_batches++;
        n_total += n;
    }
    if (g_verbose)
        fprintf(stderr, "circminer_gpu: round: %zu pairs in %zu batches, %d host threads: first read %.3f s, parse %.3f s, map %.3f s, "
                "format %.3f s, write %.3f s, wait for next block %.3f s\n", n_total, n_batches, T, ph[0], ph[1], ph[2], ph[3], ph[4], ph[5]);
}

void *gpu_map_reads(void *args) {
    FilterArgs *fa = (FilterArgs *) args;
    g_fa = fa;
    ensure_round(contigNum);
    const int is_last = filter_read.get_last_round();
    const char *bs = getenv("CM_GPU_BATCH");
    const size_t B = bs ? (size_t) atol(bs) : 262144;
    if (reportMapping != SAMFORMAT && !getenv("CM_GPU_SERIAL_HOST")) {
        gpu_map_reads_parallel(B);
        return NULL;
    }
    // SAM output (needs the reference's Record-based writer) or CM_GPU_SERIAL_HOST=1: the reference's own parser and
    // writers, one pair at a time, around the batched call
    const uint32_t stride = (uint32_t) maxReadLength;
    std::vector<Held> h1(B), h2(B);
    std::vector<char> s1(B * stride), s2(B * stride);
    std::vector<uint16_t> l1(B), l2(B);
    std::vector<cm_mread> in(B), out(B);
    std::vector<int32_t> state(B);
    bool more = true;
    while (more) {
        size_t n = 0;
        for (; n < B; n++) {   // the reference's serial section, unchanged (src/circminer.cpp:373-379)
            Record *r1 = fq_parser1.get_next_read(fa->id);
            Record *r2 = r1 ? fq_parser2.get_next_read(fa->id) : nullptr;
            if (!r1) { more = false; break; }
            h1[n].take(r1);
            h2[n].take(r2);
            memcpy(&s1[n * stride], r1->seq, r1->seq_len);
            memcpy(&s2[n * stride], r2->seq, r2->seq_len);
            l1[n] = (uint16_t) r1->seq_len;
            l2[n] = (uint16_t) r2->seq_len;
            to_pod(r1->mr, &in[n]);
        }
        if (!n) break;
        if (cm_gpu_process_batch_pe(g_ctx, contigNum, (uint32_t) n, s1.data(), s2.data(), stride, l1.data(), l2.data(), in.data(),
                                    out.data(), state.data()) != CM_OK)
            die("cm_gpu_process_batch_pe");
        for (size_t i = 0; i < n; i++) {   // the reference's policy, in input order (src/circminer.cpp:386-397)
            Record *r1 = &h1[i].rec, *r2 = &h2[i].rec;
            from_pod(&out[i], r1->mr);
            const int st = state[i];
            bool skip = (scanLevel == 0 and st == CONCRD) or
                        (scanLevel == 1 and st == CONCRD and r1->mr->gm_compatible and (r1->mr->ed_r1 + r1->mr->ed_r2 == 0) and
                         (r1->mr->mlen_r1 + r1->mr->mlen_r2 == r1->seq_len + r2->seq_len));
            if (skip or is_last) {
                if (reportMapping == SAMFORMAT) filter_read.print_sam(r1, r2);
                else if (reportMapping == PAMFORMAT) filter_read.print_pam(r1, r2);
            }
            if ((!is_last and !skip) or (is_last and (r1->mr->type == CHIBSJ or r1->mr->type == CHI2BSJ)))
                filter_read.write_read_category(r1, r2, *(r1->mr));
        }
    }
    return NULL;
}

void *idle(void *) { return NULL; }

}   // namespace

int cm_shim_pthread_create(pthread_t *t, const pthread_attr_t *a, void *(*)(void *), void *arg) {
    FilterArgs *fa = (FilterArgs *) arg;
    if (!pairedEnd) {
        fprintf(stderr, "circminer_gpu: single-end mode writes no output in the reference (SURVEY.md 0.5b); use the C-ABI directly\n");
        exit(1);
    }
    return (::pthread_create)(t, a, fa->id == 0 ? gpu_map_reads : idle, arg);
}
