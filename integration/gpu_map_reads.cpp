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
#include <sys/stat.h>
#include <unistd.h>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <set>
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

cm_gpu_ctx *g_ctx = nullptr;            // first root context (serial / SAM path, error text)
std::set<int> g_resident;               // contigs built on the devices so far
std::vector<cm_gpu_ctx *> g_roots;      // one per device in use: owns that GPU's resident contigs and annotation
std::vector<cm_gpu_ctx *> g_workers;    // contexts batches are dealt to: every root plus its shared contexts
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

// Validity stamp of an annotation cache file: the GTF and the packed reference's layout it was flattened from.
uint64_t annot_tag() {
    uint64_t h = 1469598103934665603ull;
    auto mix = [&](uint64_t v) { h = (h ^ v) * 1099511628211ull; };
    struct stat st;
    if (stat(gtfFilename, &st) == 0) { mix((uint64_t) st.st_size); mix((uint64_t) st.st_mtime); }
    for (auto &per_contig : gtf_parser.con2chr)
        for (auto &cc : per_contig) {
            for (char ch : cc.contig) mix((unsigned char) ch);
            mix(cc.shift);
        }
    mix((uint64_t) kmer);
    return h;
}
std::string annot_cache_path(int c) {
    const char *dir = getenv("CM_GPU_ANNOT_CACHE");   // "0": no cache; a directory: files go there; unset: next to the GTF
    if (dir && !strcmp(dir, "0")) return "";
    std::string base = gtfFilename;
    if (dir && *dir) {
        const size_t slash = base.find_last_of('/');
        base = std::string(dir) + "/" + (slash == std::string::npos ? base : base.substr(slash + 1));
    }
    return base + ".cmannot." + std::to_string(c + 1) + ".l" + std::to_string(maxReadLength);
}

// GTFParser tables of the current contig as the flat arrays of cm_annot_flat.  The flattening (two 1.3-Gbit masks read bit by
// bit out of std::bitset) is done once per (GTF, packed reference, -l): the result is kept as a file beside the GTF
// (cm_annot_file_write) and later runs upload that file as it is.
// The words of a std::bitset.  The standard gives no word access; libstdc++ keeps the bits as an array of unsigned long with bit k
// in word k / 64, and that is checked on the object itself (size, and 64 words spread over the range against operator[]) before
// the words are copied -- false means "read it bit by bit".  Bits from n_bits on are cleared, as the bit-by-bit copy leaves them.
bool g_annot_by_words = false;   // for the CM_GPU_VERBOSE line
template <size_t N>
bool bitset_words(const std::bitset<N> &b, uint64_t *dst, size_t nw, uint64_t n_bits) {
    if (sizeof(unsigned long) != 8 || sizeof(b) != ((N + 63) / 64) * 8 || nw == 0 || nw * 64 > ((N + 63) / 64) * 64 || n_bits > N) return false;
    const unsigned char *raw = reinterpret_cast<const unsigned char *>(&b);
    for (int s = 0; s < 64; s++) {
        const size_t i = (nw - 1) * (size_t) s / 63;
        uint64_t a = 0, w;
        for (uint64_t k = (uint64_t) i << 6; k < ((uint64_t) i << 6) + 64 && k < N; k++)
            if (b[k]) a |= 1ull << (k & 63);
        memcpy(&w, raw + i * 8, 8);
        if (a != w) return false;
    }
    memcpy(dst, raw, nw * 8);
    if (n_bits < nw * 64) {
        const size_t last = (size_t) (n_bits >> 6);
        if (n_bits & 63) dst[last] &= (1ull << (n_bits & 63)) - 1; else dst[last] = 0;
        for (size_t i = last + 1; i < nw; i++) dst[i] = 0;
    }
    return true;
}

void upload_annotation(int c) {
    std::vector<int32_t> chr_now;   // this run's chromosome-table indices of the contig's chromosomes (registers the names)
    for (auto &cc : gtf_parser.con2chr[c]) chr_now.push_back(chr_index(cc.contig));
    const std::string cache = annot_cache_path(c);
    const uint64_t tag = annot_tag();
    if (!cache.empty()) {
        cm_annot_flat F;
        uint32_t mrl = 0;
        uint64_t ftag = 0;
        void *st = nullptr;
        if (cm_annot_file_read(cache.c_str(), &F, &mrl, &ftag, &st) == CM_OK) {
            bool ok = (int) mrl == maxReadLength && ftag == tag && F.n_chr == chr_now.size() && F.n_bits == (uint64_t) _ih_refGenLen + 2;
            for (size_t i = 0; ok && i < chr_now.size(); i++) ok = F.chr_id[i] == chr_now[i];
            if (ok) {
                for (cm_gpu_ctx *root : g_roots)
                    if (cm_gpu_load_annotation(root, c, &F) != CM_OK) { g_ctx = root; die("cm_gpu_load_annotation"); }
                cm_annot_file_free(st);
                if (g_verbose) fprintf(stderr, "circminer_gpu: annotation of contig %d from %s\n", c, cache.c_str());
                return;
            }
            cm_annot_file_free(st);
        }
    }
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
    const bool has_nb = (size_t) c < near_border_bs.size(), has_in = (size_t) c < intronic_bs.size();
    const bool by_words = (!has_nb || bitset_words(near_border_bs[c], nb.data(), nb.size(), n_bits)) &&
                          (!has_in || bitset_words(intronic_bs[c], in.data(), in.size(), n_bits));
    g_annot_by_words = by_words;
    if (!by_words) {   // the bits read one by one, whole words per thread (0.3 s per Gbp and bitset on 16 threads)
        const int T = std::max(1, threadCount);
        const size_t nw = nb.size();
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
    for (cm_gpu_ctx *root : g_roots)
        if (cm_gpu_load_annotation(root, c, &A) != CM_OK) { g_ctx = root; die("cm_gpu_load_annotation"); }
    if (!cache.empty()) {   // written beside the mapping (to a temporary name first: a reader never sees half a file)
        const std::string tmp = cache + ".tmp" + std::to_string((long) getpid());
        const uint32_t mrl = (uint32_t) maxReadLength;
        const bool ok = cm_annot_file_write(tmp.c_str(), &A, mrl, tag) == CM_OK && rename(tmp.c_str(), cache.c_str()) == 0;
        if (!ok) remove(tmp.c_str());
        if (g_verbose) fprintf(stderr, "circminer_gpu: annotation of contig %d %s %s\n", c, ok ? "written to" : "could not be written to", cache.c_str());
    }
}

// Devices the mapping round uses: CM_GPU_DEVICES = "all" or a comma list ("0,1,2,3"; a device may be named twice, which gives
// it two independent root contexts); else CM_GPU_DEVICE (one device); else every visible GPU.  The reference's counterpart
// is `-t threadCount` mapping threads over one shared index (src/circminer.cpp:290-297): here every GPU holds the index
// and batches of pairs are dealt to them round-robin.
std::vector<int> gpu_devices() {
    std::vector<int> d;
    const char *list = getenv("CM_GPU_DEVICES"), *one = getenv("CM_GPU_DEVICE");
    if (list && strcmp(list, "all")) {
        for (const char *p = list; *p;) {
            d.push_back(atoi(p));
            while (*p && *p != ',') p++;
            if (*p == ',') p++;
        }
    } else if (!list && one) {
        d.push_back(atoi(one));
    } else {
        const int n = cm_gpu_device_count();
        for (int i = 0; i < n; i++) d.push_back(i);
    }
    if (d.empty()) d.push_back(0);
    return d;
}
int gpu_device() { return gpu_devices()[0]; }

void warmup_wait();

void ensure_round(int c) {
    const double t_in = now_s();
    warmup_wait();
    if (g_roots.empty()) {
        g_chr.reserve(CHR_CAP);
        cm_params &p = g_params;
        cm_params_default(&p);
        p.kmer = kmer; p.max_read_len = maxReadLength; p.max_ed = maxEd; p.max_sc = maxSc; p.band_width = bandWidth;
        p.seed_lim = seedLim; p.max_tlen = maxTlen; p.max_intron_len = maxIntronLen; p.max_chain_len = maxChainLen;
        p.scan_level = scanLevel;
        for (int dev : gpu_devices()) {
            cm_gpu_ctx *root = nullptr;
            if (cm_gpu_create(dev, &p, &root) != CM_OK) die("cm_gpu_create");
            g_roots.push_back(root);
        }
        g_ctx = g_roots[0];
    }
    if (g_loaded_contig == c) return;
    const double t_ctx = now_s();
    // Index of this contig on every device.  Default: built there from the 3-bit packed contig the reference has just
    // loaded (38 MB over PCIe per 100 Mbp), all devices at once.  CM_GPU_INDEX=upload converts and uploads the reference's own
    // bucket table instead (the two are compared entry by entry in tests/test_gpu_seed.py).
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
        for (cm_gpu_ctx *root : g_roots)
            if (cm_gpu_load_contig(root, c, (const void *const *) _ih_hashTable, _ih_hashTableMem, n, (const uint64_t *) _ih_crefGen,
                                   (uint32_t) _ih_refGenLen) != CM_OK) { g_ctx = root; die("cm_gpu_load_contig"); }
    } else {
        std::atomic<int> bad(-1);
        feed::parallel_for((int) g_roots.size(), [&](int k) {
            if (cm_gpu_build_contig_packed(g_roots[k], c, (const uint64_t *) _ih_crefGen, (uint32_t) _ih_refGenLen) != CM_OK) bad = k;
        });
        if (bad.load() >= 0) { g_ctx = g_roots[bad.load()]; die("cm_gpu_build_contig_packed"); }
    }
    const double t_idx = now_s();
    upload_annotation(c);
    g_loaded_contig = c;
    g_resident.insert(c);
    if (g_workers.empty()) {
        // CM_GPU_CTX_PER_DEVICE (default 2): contexts per device; the extra ones share their root's index and let one batch's
        // copies run beside another batch's kernels
        const char *e = getenv("CM_GPU_CTX_PER_DEVICE");
        const int per = std::max(1, e ? atoi(e) : 2);
        for (int k = 0; k < per; k++)
            for (cm_gpu_ctx *root : g_roots) {
                cm_gpu_ctx *w = root;
                if (k && cm_gpu_create_shared(root, &w) != CM_OK) { g_ctx = root; die("cm_gpu_create_shared"); }
                g_workers.push_back(w);
            }
    }
    if (g_verbose)
        fprintf(stderr, "circminer_gpu: contig %d on %zu device(s), %zu contexts: context %.3f s, table scan %.3f s, index on device %.3f s, annotation %.3f s (bit vectors %s)\n",
                c, g_roots.size(), g_workers.size(), t_ctx - t_in, t_scan - t_ctx, t_idx - t_scan, now_s() - t_idx,
                g_annot_by_words ? "copied by words" : "read bit by bit or from the cache file");
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

// One batch in flight: its block of both FASTQ files, the pinned batch arrays, and where it stands.
struct Slot {
    feed::RawFile raw[2];
    size_t B = 0;
    uint32_t stride_ = 0;
    char *s1 = nullptr, *s2 = nullptr;
    uint16_t *l1 = nullptr, *l2 = nullptr;
    cm_mread *in = nullptr, *out = nullptr;
    int32_t *state = nullptr;
    std::vector<feed::RecMeta> m1, m2;
    size_t n = 0;
    enum { EMPTY, READY, DONE } st = EMPTY;
    bool quit = false;
    std::mutex mu;
    std::condition_variable cv;
    std::thread prep;   // allocates this slot's buffers beside the first block read (page-locking 0.6 GB takes 0.15 s)
    void reserve_async(size_t b, uint32_t stride, size_t cap) {
        if (prep.joinable()) prep.join();
        if (B >= b && stride_ >= stride && raw[0].buf.size() >= cap + 1) return;
        prep = std::thread([this, b, stride, cap] { reserve(b, stride, cap); });
    }
    void ready() { if (prep.joinable()) prep.join(); }
    ~Slot() { ready(); }
    void reserve(size_t b, uint32_t stride, size_t cap) {
        raw[0].buf.reserve_bytes(cap + 1);
        raw[1].buf.reserve_bytes(cap + 1);
        if (B >= b && stride_ >= stride) return;
        if (B) {
            cm_gpu_host_free(s1); cm_gpu_host_free(s2); cm_gpu_host_free(l1); cm_gpu_host_free(l2);
            cm_gpu_host_free(in); cm_gpu_host_free(out); cm_gpu_host_free(state);
        }
        s1 = pinned<char>(b * stride); s2 = pinned<char>(b * stride);
        l1 = pinned<uint16_t>(b); l2 = pinned<uint16_t>(b);
        in = pinned<cm_mread>(b); out = pinned<cm_mread>(b);
        state = pinned<int32_t>(b);
        m1.resize(b); m2.resize(b);
        B = b;
        stride_ = stride;
    }
};

std::vector<std::unique_ptr<Slot>> g_slots;

// CUDA context creation (and kernel loading) on every device in use, and the page-locking of the batch buffers, overlap
// the reference's own start-up (GTF, index meta, first loadHashTable): threads started from a static initialiser.  The
// buffers are sized for the default maximum read length; a run with a larger -l re-allocates them in its first round.
size_t batch_pairs() {
    const char *bs = getenv("CM_GPU_BATCH");
    return bs ? (size_t) atol(bs) : 524288;
}
size_t block_bytes(size_t B) { return std::min<size_t>(std::max<size_t>(B * 512, 1u << 20), 1u << 30); }
struct Warmup {
    std::vector<std::thread> th;
    Warmup() {
        if (getenv("CM_GPU_NO_WARMUP")) return;
        const std::vector<int> devs = gpu_devices();
        for (int dev : devs) th.emplace_back([dev] { cm_gpu_warmup(dev); });
        const char *e = getenv("CM_GPU_CTX_PER_DEVICE");
        const size_t K = devs.size() * (size_t) std::max(1, e ? atoi(e) : 2), B = batch_pairs();
        while (g_slots.size() < K) g_slots.emplace_back(new Slot);
        for (auto &sl : g_slots) sl->reserve_async(B, 300, block_bytes(B));
    }
    void wait() { for (auto &t : th) if (t.joinable()) t.join(); }
    ~Warmup() { wait(); }
} g_warmup;
void warmup_wait() { g_warmup.wait(); }

// Batched rounds with the parallel host feed of host_feed.h (PAM output or none).  Batches are dealt round-robin to the
// worker contexts (every GPU in use, CM_GPU_CTX_PER_DEVICE contexts each): batch b goes to slot / context b mod K, the
// host reads and parses batch b + 1 while earlier batches are on the GPUs, and results are formatted and written in batch
// order -- the reference's `-t 1` order -- so the files do not depend on how many GPUs took part.
void gpu_map_reads_parallel(size_t B) {
    const int is_last = filter_read.get_last_round();
    const int T = std::max(1, threadCount);
    const uint32_t stride = (uint32_t) maxReadLength;
    const size_t cap = block_bytes(B);
    const size_t K = g_workers.size();
    // slots (block buffers, pinned batch arrays) and text buffers live across rounds
    auto &slots = g_slots;
    static std::vector<feed::TextOut> text;
    while (slots.size() < K) slots.emplace_back(new Slot);
    for (size_t k = 0; k < K; k++) slots[k]->reserve_async(B, stride, cap);
    text.resize(T);
    cm_mread fresh;
    cm_mread_init(&fresh, &g_params);
    const int contig = contigNum;
    double ph[6] = {0, 0, 0, 0, 0, 0};   // read, parse, wait for GPUs, format, write, host buffers
    std::vector<double> gpu_busy(K, 0.0);
    std::vector<std::vector<float>> call_ms(K);
    std::atomic<int> failed(-1);

    std::vector<std::thread> workers;
    for (size_t k = 0; k < K; k++) {
        slots[k]->st = Slot::EMPTY;
        slots[k]->quit = false;
        workers.emplace_back([&, k] {
            Slot &sl = *slots[k];
            for (;;) {
                std::unique_lock<std::mutex> lk(sl.mu);
                sl.cv.wait(lk, [&] { return sl.st == Slot::READY || sl.quit; });
                if (sl.st != Slot::READY) return;
                lk.unlock();
                const double t0 = now_s();
                if (cm_gpu_process_batch_pe(g_workers[k], contig, (uint32_t) sl.n, sl.s1, sl.s2, stride, sl.l1, sl.l2, sl.in, sl.out, sl.state) != CM_OK)
                    failed = (int) k;
                gpu_busy[k] += now_s() - t0;
                call_ms[k].push_back((float) ((now_s() - t0) * 1e3));
                lk.lock();
                sl.st = Slot::DONE;
                sl.cv.notify_all();
            }
        });
    }
    auto stop_workers = [&] {
        for (size_t k = 0; k < K; k++) {
            { std::lock_guard<std::mutex> lk(slots[k]->mu); slots[k]->quit = true; }
            slots[k]->cv.notify_all();
        }
        for (auto &w : workers) w.join();
    };
    // batch `b` (in slot b mod K): wait for its GPU call, format, write, free the slot
    auto retire = [&](size_t b) {
        Slot &sl = *slots[b % K];
        double t0 = now_s();
        {
            std::unique_lock<std::mutex> lk(sl.mu);
            sl.cv.wait(lk, [&] { return sl.st == Slot::DONE; });
        }
        if (failed.load() >= 0) { g_ctx = g_workers[failed.load()]; die("cm_gpu_process_batch_pe"); }
        double t1 = now_s();
        ph[2] += t1 - t0;
        const char *b1 = sl.raw[0].buf.data(), *b2 = sl.raw[1].buf.data();
        const size_t n = sl.n;
        feed::parallel_for(T, [&](int t) {
            feed::TextOut &o = text[t];
            o.pam.clear(); o.r1.clear(); o.r2.clear();
            feed::format_range(n * t / T, n * (t + 1) / T, b1, b2, sl.m1.data(), sl.m2.data(), sl.out, sl.state, is_last, o);
        });
        t0 = now_s();
        ph[3] += t0 - t1;
        for (int t = 0; t < T; t++) {   // input order == the reference's -t 1 order
            if (!text[t].pam.empty()) fwrite(text[t].pam.data(), 1, text[t].pam.size(), filter_read.sam_output.outsam);
            if (!text[t].r1.empty()) fwrite(text[t].r1.data(), 1, text[t].r1.size(), filter_read.temp_fq_r1);
            if (!text[t].r2.empty()) fwrite(text[t].r2.data(), 1, text[t].r2.size(), filter_read.temp_fq_r2);
        }
        ph[4] += now_s() - t0;
        {
            std::lock_guard<std::mutex> lk(sl.mu);
            sl.st = Slot::EMPTY;
        }
        sl.cv.notify_all();
    };
    // batches are retired in order by a thread of their own, so that formatting and writing batch b run beside the reading
    // and parsing of batch b + K
    std::mutex flow_mu;
    std::condition_variable flow_cv;
    size_t dispatched = 0;
    bool input_done = false;
    std::thread retirer([&] {
        for (size_t b = 0;; b++) {
            {
                std::unique_lock<std::mutex> lk(flow_mu);
                flow_cv.wait(lk, [&] { return dispatched > b || input_done; });
                if (dispatched <= b) return;
            }
            retire(b);
        }
    });

    const double t_begin = now_s();
    feed::LineIndex ix1, ix2;
    size_t n_batches = 0, n_total = 0;
    size_t carry_len[2] = {0, 0};
    const char *carry_src[2] = {nullptr, nullptr};
    bool eof[2] = {false, false};
    for (size_t b = 0;; b++) {
        Slot &sl = *slots[b % K];
        double t0 = now_s();
        sl.ready();
        {   // the slot batch b needs is free again once batch b - K has been written
            std::unique_lock<std::mutex> lk(sl.mu);
            sl.cv.wait(lk, [&] { return sl.st == Slot::EMPTY; });
        }
        double t1 = now_s();
        ph[5] += t1 - t0;
        gzFile gz[2] = {fq_parser1.gzinput, fq_parser2.gzinput};
        feed::parallel_for(2, [&](int k) {
            feed::RawFile &r = sl.raw[k];
            r.gz = gz[k];
            if (carry_len[k]) memmove(r.buf.data(), carry_src[k], carry_len[k]);   // the unread tail of the previous block
            r.len = carry_len[k];
            r.eof = eof[k];
            r.fill();
            eof[k] = r.eof;
        });
        t0 = now_s();
        ph[0] += t0 - t1;
        feed::count_lines(sl.raw[0], T, ix1);
        feed::count_lines(sl.raw[1], T, ix2);
        const size_t n = (size_t) std::min<uint64_t>(B, std::min(ix1.lines, ix2.lines) / 4);
        if (!n) break;
        size_t e[2] = {sl.raw[0].len, sl.raw[1].len};
        feed::parse_records(sl.raw[0], ix1, T, n, stride, sl.m1.data(), sl.s1, sl.l1, sl.in, fresh, e[0]);
        feed::parse_records(sl.raw[1], ix2, T, n, stride, sl.m2.data(), sl.s2, sl.l2, nullptr, fresh, e[1]);
        for (int k = 0; k < 2; k++) {
            carry_src[k] = sl.raw[k].buf.data() + e[k];
            carry_len[k] = sl.raw[k].len - e[k];
        }
        sl.n = n;
        ph[1] += now_s() - t0;
        {
            std::lock_guard<std::mutex> lk(sl.mu);
            sl.st = Slot::READY;
        }
        sl.cv.notify_all();
        n_batches++;
        n_total += n;
        {
            std::lock_guard<std::mutex> lk(flow_mu);
            dispatched = n_batches;
        }
        flow_cv.notify_all();
    }
    {
        std::lock_guard<std::mutex> lk(flow_mu);
        input_done = true;
    }
    flow_cv.notify_all();
    retirer.join();
    stop_workers();
    if (g_verbose) {
        double busy = 0;
        for (double x : gpu_busy) busy += x;
        fprintf(stderr, "circminer_gpu: round: %zu pairs in %zu batches over %zu contexts on %zu device(s), %d host threads, %.3f s: feeding thread: wait for a free slot %.3f s, "
                "read %.3f s, parse %.3f s | retiring thread: wait for GPUs %.3f s, format %.3f s, write %.3f s | GPU calls %.3f s in total\n", n_total, n_batches, K,
                g_roots.size(), T, now_s() - t_begin, ph[5], ph[0], ph[1], ph[2], ph[3], ph[4], busy);
        for (size_t k = 0; k < K; k++) {
            fprintf(stderr, "circminer_gpu:   context %zu: GPU call ms", k);
            for (size_t j = 0; j < call_ms[k].size() && j < 12; j++) fprintf(stderr, " %.1f", call_ms[k][j]);
            fprintf(stderr, "\n");
        }
    }
}

void *gpu_map_reads(void *args) {
    FilterArgs *fa = (FilterArgs *) args;
    g_fa = fa;
    ensure_round(contigNum);
    const int is_last = filter_read.get_last_round();
    const size_t B = batch_pairs();
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

// stage 2 (gpu_circ_detect.cpp): contig c with its annotation on the first device; the contigs of the mapping rounds are still
// resident there and are not rebuilt (shared contexts hold them)
cm_gpu_ctx *cm_shim_stage2_contig(int c) {
    if (!g_resident.count(c)) ensure_round(c);
    return g_roots[0];
}

// mapping()'s loadHashTable(&tmpTime) (src/circminer.cpp:232), renamed by gpu_shim.h.  The drop-in builds its index on the device
// from the 3-bit packed contig (ensure_round), so of an index-file record it needs the contig alone: the reference's own
// loadCompressedRefGenome (src/mrsfast/HashTable.c, what stage 2's load_genome calls, src/process_circ.cpp:1668) reads the same
// record, leaves the same globals (_ih_refGenName / Off / Len, _ih_crefGen) and the file at the next record, and seeks over the
// k-mer table instead of reading 8 bytes per base into memory (or, with a compact index, re-deriving it with all threads).
// CM_GPU_INDEX=upload, which converts the reference's own table, and CM_GPU_REF_TABLE=1 keep the full load.
extern "C" int cm_shim_loadHashTable(double *load_time) {
#ifdef CM_SHIM_KEEP_REF_TABLE
    return loadHashTable(load_time);
#else
    const char *how = getenv("CM_GPU_INDEX"), *keep = getenv("CM_GPU_REF_TABLE");
    if ((how && !strcmp(how, "upload")) || (keep && atoi(keep))) return loadHashTable(load_time);
    return loadCompressedRefGenome(load_time);
#endif
}

// mapping()'s genome_seeder.pac2char_whole_contig() (src/circminer.cpp:244), renamed by gpu_shim.h (the member itself is in
// gpu_shim_members.cpp): the contig as characters is what the reference's extension reads; the drop-in's rounds read the packed
// contig on the device, so the unpacking is skipped unless the reference's table is kept as well.
bool cm_shim_unpack_contig() {
#ifdef CM_SHIM_KEEP_REF_TABLE
    return genome_seeder.pac2char_whole_contig();
#else
    const char *how = getenv("CM_GPU_INDEX"), *keep = getenv("CM_GPU_REF_TABLE");
    if ((how && !strcmp(how, "upload")) || (keep && atoi(keep))) return genome_seeder.pac2char_whole_contig();
    return true;
#endif
}

int cm_shim_pthread_create(pthread_t *t, const pthread_attr_t *a, void *(*)(void *), void *arg) {
    FilterArgs *fa = (FilterArgs *) arg;
    if (!pairedEnd) {
        fprintf(stderr, "circminer_gpu: single-end mode writes no output in the reference (SURVEY.md 0.5b); use the C-ABI directly\n");
        exit(1);
    }
    return (::pthread_create)(t, a, fa->id == 0 ? gpu_map_reads : idle, arg);
}
