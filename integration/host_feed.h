// Parallel host feed for the drop-in binary: what the reference does per pair inside its serial
// sections -- FASTQParser::get_next_read (src/fastq_parser.cpp:86-178, header tokens by extract_map_info / fill_map_info
// :181-269), SAMOutput::write_pam_rec_pe (src/output.cpp:280-299) and FilterRead::write_read_category
// (src/filter.cpp:413-455) -- restated over whole batches:
//   raw FASTQ bytes are read in large blocks (gzread on the parser's own handle, so .gz inputs keep working) by one
//   thread per file while the previous batch is on the GPU; the block is cut into lines by all host threads (count
//   newlines per slice, prefix sum, then every thread handles the lines that end in its slice), each record's fields go
//   straight into the batch arrays of cm_gpu_process_batch_pe; after the GPU call the text of mapping.pam and of the two
//   remain FASTQs is formatted by all threads into per-thread buffers (the reference's own printf formats) and written
//   in input order.  Output is byte-identical to the reference at -t 1 (tests/test_gpu_dropin.py).
// Included by gpu_map_reads.cpp only (needs its chr_index / g_chr / die).
#pragma once
#include <atomic>
#include <cinttypes>
#include <functional>
#include <thread>
#include <zlib.h>

namespace feed {

template <class F> void parallel_for(int T, F &&fn) {
    std::vector<std::thread> th;
    th.reserve(T > 0 ? T - 1 : 0);
    for (int t = 1; t < T; t++) th.emplace_back([&fn, t] { fn(t); });
    fn(0);
    for (auto &x : th) x.join();
}

struct RawFile {
    gzFile gz = nullptr;
    struct Bytes {   // uninitialised storage: a std::vector would zero-fill (and page in) the whole block first
        char *p = nullptr;
        size_t n = 0;
        char *data() { return p; }
        const char *data() const { return p; }
        size_t size() const { return n; }
        char &operator[](size_t i) { return p[i]; }
        void reserve_bytes(size_t m) { if (m > n) { free(p); p = (char *) malloc(m); n = p ? m : 0; if (!p) { fprintf(stderr, "circminer_gpu: out of host memory\n"); exit(1); } } }
        ~Bytes() { free(p); }
    } buf;
    size_t len = 0;
    bool eof = false;
    void init(gzFile g, size_t cap) { gz = g; buf.reserve_bytes(cap + 1); len = 0; eof = false; }
    void fill() {
        const size_t cap = buf.size() - 1;
        while (!eof && len < cap) {
            const size_t want = std::min<size_t>(cap - len, 1u << 30);
            const int r = gzread(gz, buf.data() + len, (unsigned) want);
            if (r < 0) { int err; fprintf(stderr, "gzread error: %s\n", gzerror(gz, &err)); exit(1); }
            if (r == 0) { eof = true; break; }
            len += (size_t) r;
        }
        if (eof && len && buf[len - 1] != '\n') buf[len++] = '\n';
    }
};

struct RecMeta { uint32_t name_off, name_len, seq_off, seq_len, qual_off, qual_len; };

struct LineIndex {
    std::vector<size_t> cut;      // T + 1 slice borders
    std::vector<uint64_t> before; // newlines before each slice
    uint64_t lines = 0;
};

inline void count_lines(const RawFile &f, int T, LineIndex &ix) {
    ix.cut.assign(T + 1, 0);
    ix.before.assign(T + 1, 0);
    for (int t = 0; t <= T; t++) ix.cut[t] = f.len * (size_t) t / (size_t) T;
    std::vector<uint64_t> cnt(T, 0);
    parallel_for(T, [&](int t) {
        const char *p = f.buf.data() + ix.cut[t], *e = f.buf.data() + ix.cut[t + 1];
        uint64_t c = 0;
        while (p < e && (p = (const char *) memchr(p, '\n', (size_t) (e - p)))) { c++; p++; }
        cnt[t] = c;
    });
    for (int t = 0; t < T; t++) ix.before[t + 1] = ix.before[t] + cnt[t];
    ix.lines = ix.before[T];
}

// header line -> read name and the carried MatchedRead (extract_map_info + fill_map_info)
inline void parse_header(const char *h, size_t hl, uint32_t base_off, RecMeta &m, cm_mread *mr, const cm_mread &fresh) {
    const char *tok[24];
    uint32_t tl[24];
    int cnt = 0;
    size_t i = 0;
    while (i < hl) {
        while (i < hl && h[i] == ' ') i++;
        if (i >= hl) break;
        const size_t s = i;
        while (i < hl && h[i] != ' ') i++;
        if (cnt < 24) { tok[cnt] = h + s; tl[cnt] = (uint32_t) (i - s); }
        cnt++;
    }
    uint32_t nl = cnt ? tl[0] : 0;
    m.name_off = base_off + (cnt ? (uint32_t) (tok[0] - h) : 0);
    if (nl >= 2 && tok[0][nl - 2] == '/') nl -= 2;   // "/1", "/2" suffix (src/fastq_parser.cpp:196-197)
    m.name_len = nl;
    if (!mr) return;
    *mr = fresh;
    if (cnt != FQCOMMENTCNT) return;   // type NOPROC_NOMATCH, tlen INF, junc 0, not gm-compatible
    const int type = atoi(tok[2]);
    mr->type = type;
    if (type == CONCRD or type == DISCRD or type == CHIORF or type == CHIBSJ or type == CHI2BSJ or type == CONGNM or type == CONGEN) {
        mr->chr_r1 = chr_index(std::string(tok[3], tl[3]));
        mr->spos_r1 = (uint32_t) strtoul(tok[4], nullptr, 10);
        mr->epos_r1 = (uint32_t) strtoul(tok[5], nullptr, 10);
        mr->mlen_r1 = (uint32_t) atoi(tok[6]);
        mr->qspos_r1 = (uint32_t) strtoul(tok[7], nullptr, 10);
        mr->qepos_r1 = (uint32_t) strtoul(tok[8], nullptr, 10);
        mr->r1_forward = tok[9][0] == '+';
        mr->ed_r1 = atoi(tok[10]);
        mr->chr_r2 = chr_index(std::string(tok[11], tl[11]));
        mr->spos_r2 = (uint32_t) strtoul(tok[12], nullptr, 10);
        mr->epos_r2 = (uint32_t) strtoul(tok[13], nullptr, 10);
        mr->mlen_r2 = (uint32_t) atoi(tok[14]);
        mr->qspos_r2 = (uint32_t) strtoul(tok[15], nullptr, 10);
        mr->qepos_r2 = (uint32_t) strtoul(tok[16], nullptr, 10);
        mr->r2_forward = tok[17][0] == '+';
        mr->ed_r2 = atoi(tok[18]);
        mr->tlen = atoi(tok[19]);
        mr->junc_num = (uint16_t) strtoul(tok[20], nullptr, 10);
        mr->gm_compatible = tok[21][0] == '1';
        mr->contig_num = atoi(tok[22]);
    }
    // any other type: the fresh state ("-", zeros, forward, ed maxEd + 1, tlen INF)
}

// every thread handles the lines that END in its slice; records >= n are left for the next batch
inline void parse_records(const RawFile &f, const LineIndex &ix, int T, size_t n, uint32_t stride, RecMeta *meta, char *seqs,
                          uint16_t *lens, cm_mread *mr, const cm_mread &fresh, size_t &carry_off) {
    const char *b = f.buf.data();
    std::atomic<int> bad(0);
    parallel_for(T, [&](int t) {
        const size_t c0 = ix.cut[t], c1 = ix.cut[t + 1];
        if (c0 >= c1) return;
        uint64_t line = ix.before[t];
        if (line >= 4 * (uint64_t) n) return;
        size_t start;   // start of the line that ends first in this slice
        if (c0 == 0) start = 0;
        else {
            const void *q = memrchr(b, '\n', c0);
            start = q ? (size_t) ((const char *) q - b) + 1 : 0;
        }
        const char *p = b + c0, *e = b + c1;
        while (p < e && (p = (const char *) memchr(p, '\n', (size_t) (e - p)))) {
            const size_t end = (size_t) (p - b);
            const size_t r = (size_t) (line >> 2);
            if (r >= n) break;
            RecMeta &m = meta[r];
            const size_t ll = end - start;
            switch (line & 3) {
            case 0:
                if (!ll || b[start] != '@') { bad = 1; break; }
                parse_header(b + start + 1, ll - 1, (uint32_t) start + 1, m, mr ? mr + r : nullptr, fresh);
                break;
            case 1:
                m.seq_off = (uint32_t) start; m.seq_len = (uint32_t) ll;
                if (ll > stride) { bad = 2; break; }
                memcpy(seqs + r * stride, b + start, ll);
                lens[r] = (uint16_t) ll;
                break;
            case 2:
                if (!ll || b[start] != '+') bad = 3;
                break;
            default:
                m.qual_off = (uint32_t) start; m.qual_len = (uint32_t) ll;
                if (line + 1 == 4 * (uint64_t) n) carry_off = end + 1;
                break;
            }
            line++;
            start = end + 1;
            p++;
        }
    });
    if (bad.load()) {
        fprintf(stderr, "circminer_gpu: malformed FASTQ record (%s)\n",
                bad.load() == 1 ? "header does not start with '@'" : bad.load() == 2 ? "read longer than the maximum read length" : "third line does not start with '+'");
        exit(1);
    }
    for (size_t r = 0; r < n; r++)
        if (meta[r].seq_len != meta[r].qual_len) {   // src/fastq_parser.cpp:161-165
            fprintf(stderr, "ERROR: read: %.*s, length of sequence (%d) does not match with quality (%d)!\nAborting\n",
                    (int) meta[r].name_len, b + meta[r].name_off, (int) meta[r].seq_len, (int) meta[r].qual_len);
            exit(1);
        }
}

inline bool mapped_type(int type) {
    return type == CONCRD or type == DISCRD or type == CHIORF or type == CHIBSJ or type == CHI2BSJ or type == CONGNM or type == CONGEN;
}

struct TextOut { std::string pam, r1, r2; };

inline void append_remain(std::string &o, const char *b, const RecMeta &m, const char *info, int info_len) {
    o.push_back('@');
    o.append(b + m.name_off, m.name_len);
    o.append(info, (size_t) info_len);
    o.push_back('\n');
    o.append(b + m.seq_off, m.seq_len);
    o.append("\n+\n", 3);
    o.append(b + m.qual_off, m.qual_len);
    o.push_back('\n');
}

// the reference's per-pair policy (src/circminer.cpp:386-397) and its writers, for pairs [lo, hi)
inline void format_range(size_t lo, size_t hi, const char *b1, const char *b2, const RecMeta *m1, const RecMeta *m2,
                         const cm_mread *out, const int32_t *state, int is_last, TextOut &o) {
    char line[2048];
    for (size_t i = lo; i < hi; i++) {
        const cm_mread &m = out[i];
        const int st = state[i];
        const bool skip = (scanLevel == 0 and st == CONCRD) or
                          (scanLevel == 1 and st == CONCRD and m.gm_compatible and (m.ed_r1 + m.ed_r2 == 0) and
                           ((int) (m.mlen_r1 + m.mlen_r2) == (int) (m1[i].seq_len + m2[i].seq_len)));
        const bool mapped = mapped_type(m.type);
        const char r1_dir = m.r1_forward ? '+' : '-', r2_dir = m.r2_forward ? '+' : '-';
        const char *c1 = "-", *c2 = "-";
        if (mapped) {
            c1 = m.chr_r1 >= 0 ? g_chr[m.chr_r1].c_str() : "-";
            c2 = m.chr_r2 >= 0 ? g_chr[m.chr_r2].c_str() : "-";
        }
        if ((skip or is_last) and reportMapping == PAMFORMAT) {
            o.pam.append(b1 + m1[i].name_off, m1[i].name_len);
            int k;
            if (mapped)
                k = snprintf(line, sizeof line, "\t%s\t%u\t%u\t%d\t%u\t%u\t%c\t%d\t%s\t%u\t%u\t%d\t%u\t%u\t%c\t%d\t%d\t%d\t%d\t%d\n",
                             c1, m.spos_r1, m.epos_r1, (int) m.mlen_r1, m.qspos_r1, m.qepos_r1, r1_dir, m.ed_r1,
                             c2, m.spos_r2, m.epos_r2, (int) m.mlen_r2, m.qspos_r2, m.qepos_r2, r2_dir, m.ed_r2,
                             m.tlen, (int) m.junc_num, (int) m.gm_compatible, m.type);
            else
                k = snprintf(line, sizeof line, "\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t*\t%d\n", m.type);
            o.pam.append(line, (size_t) k);
        }
        if ((!is_last and !skip) or (is_last and (m.type == CHIBSJ or m.type == CHI2BSJ))) {
            int k;
            if (mapped) {
                uint32_t con_spos = m.spos_r1;
                auto it = gtf_parser.chr2con.find(c1);
                if (it != gtf_parser.chr2con.end()) con_spos += it->second.shift;
                const uint64_t gspos = uint64_t(m.contig_num) * DEF_CONTIG_SIZE + con_spos;
                k = snprintf(line, sizeof line, " %" PRId64 " %d %s %u %u %d %u %u %c %d %s %u %u %d %u %u %c %d %d %d %d %d",
                             gspos, m.type, c1, m.spos_r1, m.epos_r1, (int) m.mlen_r1, m.qspos_r1, m.qepos_r1, r1_dir, m.ed_r1,
                             c2, m.spos_r2, m.epos_r2, (int) m.mlen_r2, m.qspos_r2, m.qepos_r2, r2_dir, m.ed_r2,
                             m.tlen, (int) m.junc_num, (int) m.gm_compatible, m.contig_num);
            } else
                k = snprintf(line, sizeof line, " * %d * * * * * * * * * * * * * * * * * * * *", m.type);
            append_remain(o.r1, b1, m1[i], line, k);
            append_remain(o.r2, b2, m2[i], line, k);
        }
    }
}

}   // namespace feed
