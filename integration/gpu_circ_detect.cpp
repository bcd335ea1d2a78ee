// Stage 2 of the drop-in binary: ProcessCirc::do_process (src/process_circ.cpp:195-332) with the single-split pairs of every
// contig going through libcm_gpu.so in one batch (cm_gpu_circ_single_batch) instead of call_circ_single_split one pair at a time.
//
// How it is bound: integration/gpu_shim.h renames the one call `process_circ.do_process()` of circ_detect (src/circminer.cpp:349-352)
// to cm_shim_do_process, a member this file defines; no reference source is edited.  Everything around the per-pair work is
// the reference's own code called in do_process's order: index / GTF set-up, the sorted remain FASTQ through FASTQParser,
// load_genome per contig, report_events; the double-split pairs go through cm_gpu_circ_double_batch the same way (CM_GPU_STAGE2_DOUBLE=0: through the
// reference's call_circ).  For a CHIBSJ
// pair the batch returns every check_split_map attempt with its three mates and type -- the line print_split_mapping +
// fprintf(candid_file, "%d\n", type) write -- and the CircRes call_circ_single_split pushes; they are written / pushed here
// when the pair's turn comes, so candidates.pam and circ_report are byte-identical.
// CM_GPU_STAGE2=0 runs the reference's do_process instead.
#define private public
#define protected public
#define do_process cm_shim_do_process
#include "common.h"
#include "fastq_parser.h"
#include "gene_annotation.h"
#include "genome.h"
#include "match_read.h"
#include "process_circ.h"
#include "utils.h"
#undef do_process
#undef private
#undef protected

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../include/cm_gpu.h"

extern "C" {
#include "mrsfast/Common.h"
#include "mrsfast/HashTable.h"
}

cm_gpu_ctx *cm_shim_stage2_contig(int c);       // gpu_map_reads.cpp: index + annotation of contig c on the device, the root context
void cm_ref_do_process(ProcessCirc *p);         // ref_do_process.cpp: the reference's own do_process

namespace {

struct Pair {                 // one pair of the sorted remain files
    std::string rname, seq1, seq2, rc1, rc2;
    MatchedRead mr;
};

double g_t_batch = 0, g_t_ref_pairs = 0, g_t_write = 0, g_t_load = 0;
long g_n_redone = 0;          // pairs redone by the reference's call_circ because a device list overflowed
bool g_text_ready = false;    // genome_seeder holds the current contig as characters
void die2(cm_gpu_ctx *ctx, const char *what) {
    fprintf(stderr, "circminer_gpu: %s failed: %s\n", what, cm_gpu_last_error(ctx));
    exit(1);
}

// inputs of the batch entries for a set of pairs: reads as read, MatchedRead in contig coordinates, genes under mr.spos_r1
struct Inputs {
    uint32_t n = 0, stride = 16;
    std::vector<char> s1, s2;
    std::vector<uint16_t> l1, l2;
    std::vector<cm_mread> mr;
    std::vector<uint32_t> goff, gs, ge;
    void build(const std::vector<Pair> &P, const std::vector<size_t> &who) {
        n = (uint32_t) who.size();
        for (size_t k : who) stride = std::max<uint32_t>(stride, (uint32_t) std::max(P[k].seq1.size(), P[k].seq2.size()));
        stride = (stride + 15) & ~15u;
        s1.assign((size_t) n * stride, 0); s2.assign((size_t) n * stride, 0);
        l1.resize(n); l2.resize(n); mr.resize(n); goff.assign(n + 1, 0);
        for (uint32_t k = 0; k < n; k++) {
            const Pair &p = P[who[k]];
            memcpy(&s1[(size_t) k * stride], p.seq1.data(), p.seq1.size());
            memcpy(&s2[(size_t) k * stride], p.seq2.data(), p.seq2.size());
            l1[k] = (uint16_t) p.seq1.size(); l2[k] = (uint16_t) p.seq2.size();
            MatchedRead m = p.mr;
            gtf_parser.chrloc2conloc(m.chr_r1, m.spos_r1, m.epos_r1);      // call_circ_single_split :372-373, call_circ_double_split :518-519
            gtf_parser.chrloc2conloc(m.chr_r2, m.spos_r2, m.epos_r2);
            cm_mread &o = mr[k];
            memset(&o, 0, sizeof(o));
            o.spos_r1 = m.spos_r1; o.epos_r1 = m.epos_r1; o.spos_r2 = m.spos_r2; o.epos_r2 = m.epos_r2;
            o.qspos_r1 = m.qspos_r1; o.qepos_r1 = m.qepos_r1; o.qspos_r2 = m.qspos_r2; o.qepos_r2 = m.qepos_r2;
            o.mlen_r1 = m.mlen_r1; o.mlen_r2 = m.mlen_r2; o.ed_r1 = m.ed_r1; o.ed_r2 = m.ed_r2;
            o.type = m.type; o.tlen = m.tlen; o.contig_num = m.contig_num;
            o.r1_forward = m.r1_forward; o.r2_forward = m.r2_forward;
            const IntervalInfo<GeneInfo> *gi = gtf_parser.get_gene_overlap(m.spos_r1, false);   // :404, :521
            if (gi)
                for (const GeneInfo &g : gi->seg_list) { gs.push_back(g.start); ge.push_back(g.end); }
            goff[k + 1] = (uint32_t) gs.size();
        }
        if (gs.empty()) { gs.push_back(0); ge.push_back(0); }
    }
};

// call_circ_single_split's answer for every pair of a set
struct SingleRes {
    std::vector<uint32_t> first;                    // attempts of pair k: [first[k], first[k + 1])
    std::vector<cm_circ_mate> in1, r1, r2, pm;
    std::vector<int32_t> ret, best;
    std::vector<cm_circ_res> best_cr;
    uint32_t n_att = 0;
    void run(ProcessCirc &pc, cm_gpu_ctx *ctx, Inputs &in) {
        const uint32_t n = in.n;
        best.assign(n, -1); best_cr.resize(n); first.assign(n + 1, 0);
        if (!n) return;
        std::vector<uint32_t> att_pair;
        uint32_t cap = 4 * n + 1024;
        for (int pass = 0; pass < 2; pass++) {      // a second pass only when more attempts came back than the first guess holds
            att_pair.assign(cap, 0); r1.resize(cap); r2.resize(cap); pm.resize(cap); in1.resize(cap); ret.assign(cap, 0);
            if (cm_gpu_circ_single_batch(ctx, contigNum, n, in.s1.data(), in.s2.data(), in.stride, in.l1.data(), in.l2.data(), in.mr.data(),
                                         in.goff.data(), in.gs.data(), in.ge.data(), pc.window_size, pc.step, cap, &n_att, att_pair.data(), nullptr,
                                         in1.data(), nullptr, nullptr, ret.data(), nullptr, r1.data(), r2.data(), pm.data(), best.data(),
                                         best_cr.data()) != CM_OK)
                die2(ctx, "cm_gpu_circ_single_batch");
            if (n_att <= cap) break;
            cap = n_att;
        }
        uint32_t a = 0;                             // attempts are in (pair, gene, chain) order
        for (uint32_t k = 0; k < n; k++) {
            first[k] = a;
            while (a < n_att && att_pair[a] == k) a++;
        }
        first[n] = a;
    }
    // what one call_circ_single_split(record1, record2) writes and pushes for pair k
    void emit(ProcessCirc &pc, const Pair &p, uint32_t k) const {
        for (uint32_t t = first[k]; t < first[k + 1]; t++) {
            // print_split_mapping (:1685-1696) + the type; con_shift = get_shift(contigNum, mm_r1.spos) as it was before the call
            ConShift cs = gtf_parser.get_shift(contigNum, in1[t].m.spos);
            const cm_mate &q = pm[t].m, &m1 = r1[t].m, &m2 = r2[t].m;
            fprintf(pc.candid_file, "%s\t%s\t%u\t%u\t%d\t%d\t%d\t%u\t%u\t%d\t%d\t%d\t%u\t%u\t%d\t%d\t%d\t", p.rname.c_str(), cs.contig.c_str(),
                    q.spos - cs.shift, q.epos - cs.shift, q.qspos, q.matched_len, q.dir, m1.spos - cs.shift, m1.epos - cs.shift, m1.qspos,
                    m1.matched_len, m1.dir, m2.spos - cs.shift, m2.epos - cs.shift, m2.qspos, m2.matched_len, m2.dir);
            fprintf(pc.candid_file, "%d\n", ret[t]);
        }
        if (best[k] >= CR && best[k] <= MCR) {
            uint32_t t_best = first[k];             // the attempt that set best_cr: the first one of the final type (types only improve)
            for (uint32_t t = first[k]; t < first[k + 1]; t++)
                if (ret[t] == best[k]) { t_best = t; break; }
            push(pc, p, best[k], best_cr[k], in1[t_best].m.spos);
        }
    }
    static void push(ProcessCirc &pc, const Pair &p, int type, const cm_circ_res &c, uint32_t r1_spos) {
        ConShift cs = gtf_parser.get_shift(contigNum, r1_spos);
        CircRes cr;
        cr.chr = cs.contig;
        cr.rname = p.rname;
        cr.spos = c.spos - cs.shift;
        cr.epos = c.epos - cs.shift;
        cr.type = type;
        cr.start_signal = std::string((const char *) c.start_signal, c.start_signal_n);
        cr.end_signal = std::string((const char *) c.end_signal, c.end_signal_n);
        cr.start_bp_ref = std::string((const char *) c.start_bp_ref, c.ref_n);
        cr.end_bp_ref = std::string((const char *) c.end_bp_ref, c.ref_n);
        pc.circ_res.push_back(cr);
    }
};

// the pairs of one contig: the CHIBSJ and CHI2BSJ ones through the batch entries, then everything in file order
void process_contig(ProcessCirc &pc, std::vector<Pair> &P) {
    const size_t n_all = P.size();
    const char *sw2 = getenv("CM_GPU_STAGE2_DOUBLE");
    const bool gpu_double = !(sw2 && !strcmp(sw2, "0"));       // CM_GPU_STAGE2_DOUBLE=0: double-split pairs through the reference's call_circ
    std::vector<size_t> who1, who2;
    for (size_t i = 0; i < n_all; i++) {
        if (P[i].mr.type == CHIBSJ) who1.push_back(i);
        else if (P[i].mr.type == CHI2BSJ && gpu_double) who2.push_back(i);
    }
    const double t_b0 = get_real_time();
    Inputs in1, in2;
    SingleRes S1, S2;
    std::vector<uint32_t> ev_first(who2.size() + 1, 0), ev_pair;
    std::vector<uint8_t> ev_kind;
    std::vector<cm_circ_mate> e_r1, e_r2, e_p1, e_p2;
    std::vector<int32_t> e_ret, push_type(who2.size(), -1);
    std::vector<cm_circ_res> push_cr(who2.size());
    uint32_t n_ev = 0;
    if (!who1.empty() || !who2.empty()) {
        double tp[6];
        tp[0] = get_real_time();
        cm_gpu_ctx *ctx = cm_shim_stage2_contig(contigNum);
        tp[1] = get_real_time();
        in1.build(P, who1);
        in2.build(P, who2);
        tp[2] = get_real_time();
        S1.run(pc, ctx, in1);
        tp[3] = get_real_time();
        S2.run(pc, ctx, in2);                                   // what call_circ_single_split gives on the double-split pairs
        tp[4] = get_real_time();
        if (in2.n) {
            uint32_t cap = 8 * in2.n + 1024;
            for (int pass = 0; pass < 2; pass++) {
                ev_pair.assign(cap, 0); ev_kind.assign(cap, 0); e_r1.resize(cap); e_r2.resize(cap); e_p1.resize(cap); e_p2.resize(cap); e_ret.assign(cap, 0);
                if (cm_gpu_circ_double_batch(ctx, contigNum, in2.n, in2.s1.data(), in2.s2.data(), in2.stride, in2.l1.data(), in2.l2.data(), in2.mr.data(),
                                             in2.goff.data(), in2.gs.data(), in2.ge.data(), pc.window_size, pc.step, cap, &n_ev, ev_pair.data(),
                                             ev_kind.data(), e_r1.data(), e_r2.data(), e_p1.data(), e_p2.data(), e_ret.data(), nullptr, push_type.data(),
                                             push_cr.data()) != CM_OK)
                    die2(ctx, "cm_gpu_circ_double_batch");
                if (n_ev <= cap) break;
                cap = n_ev;
            }
            uint32_t a = 0;
            for (uint32_t k = 0; k < in2.n; k++) {
                ev_first[k] = a;
                while (a < n_ev && ev_pair[a] == k) a++;
            }
            ev_first[in2.n] = a;
        }
        tp[5] = get_real_time();
        if (getenv("CM_GPU_VERBOSE"))
            fprintf(stderr, "circminer_gpu: stage 2, contig %d: %u single-split pairs of %zu in one batch, %u check_split_map attempts; %u double-split "
                    "pairs, %u events; contig on the device %.3f s, gene lists %.3f s, single-split batch %.3f s, the same on the double-split pairs "
                    "%.3f s, double-split batch %.3f s\n", contigNum, in1.n, n_all, S1.n_att, in2.n, n_ev, tp[1] - tp[0], tp[2] - tp[1], tp[3] - tp[2],
                    tp[4] - tp[3], tp[5] - tp[4]);
    }
    g_t_batch += get_real_time() - t_b0;
    // file order
    Record r1, r2;
    r1.init();
    r2.init();
    uint32_t k1 = 0, k2 = 0;
    for (size_t i = 0; i < n_all; i++) {
        Pair &p = P[i];
        const double t_p0 = get_real_time();
        // a pair whose classification overflowed a fixed-size list of the device code (type -2) goes through the reference's own
        // call_circ below, like the pairs that were never sent to the device
        bool on_device = false;
        uint32_t k = 0;
        if (p.mr.type == CHIBSJ) {
            k = k1++;
            on_device = S1.best[k] != -2;
        } else if (p.mr.type == CHI2BSJ && gpu_double) {
            k = k2++;
            bool single_used = false;
            for (uint32_t t = ev_first[k]; t < ev_first[k + 1]; t++) single_used |= ev_kind[t] == 0;
            on_device = push_type[k] != -2 && !(single_used && S2.best[k] == -2);
        }
        if (!on_device && (p.mr.type == CHIBSJ || (p.mr.type == CHI2BSJ && gpu_double))) g_n_redone++;
        if (on_device && p.mr.type == CHIBSJ) {
            S1.emit(pc, p, k);
            g_t_write += get_real_time() - t_p0;
        } else if (on_device) {
            for (uint32_t t = ev_first[k]; t < ev_first[k + 1]; t++) {
                if (ev_kind[t] == 0) { S2.emit(pc, p, k); continue; }
                // print_split_mapping, four mates (:1698-1711) + the type
                ConShift cs = gtf_parser.get_shift(contigNum, in2.mr[k].spos_r1);
                const cm_mate &a = e_p1[t].m, &b = e_p2[t].m, &m1 = e_r1[t].m, &m2 = e_r2[t].m;
                fprintf(pc.candid_file, "%s\t%s\t%u\t%u\t%d\t%d\t%d\t%u\t%u\t%d\t%d\t%d\t%u\t%u\t%d\t%d\t%d\t%u\t%u\t%d\t%d\t%d\t", p.rname.c_str(),
                        cs.contig.c_str(), a.spos - cs.shift, a.epos - cs.shift, a.qspos, a.matched_len, a.dir, b.spos - cs.shift, b.epos - cs.shift,
                        b.qspos, b.matched_len, b.dir, m1.spos - cs.shift, m1.epos - cs.shift, m1.qspos, m1.matched_len, m1.dir, m2.spos - cs.shift,
                        m2.epos - cs.shift, m2.qspos, m2.matched_len, m2.dir);
                fprintf(pc.candid_file, "%d\n", e_ret[t]);
            }
            if (push_type[k] >= CR && push_type[k] <= MCR) SingleRes::push(pc, p, push_type[k], push_cr[k], in2.mr[k].spos_r1);
            g_t_write += get_real_time() - t_p0;
        } else {
            strcpy(r1.rname, p.rname.c_str()); strcpy(r2.rname, p.rname.c_str());
            strcpy(r1.seq, p.seq1.c_str()); strcpy(r1.rcseq, p.rc1.c_str());
            strcpy(r2.seq, p.seq2.c_str()); strcpy(r2.rcseq, p.rc2.c_str());
            r1.seq_len = (uint32_t) p.seq1.size(); r2.seq_len = (uint32_t) p.seq2.size();
            *r1.mr = p.mr;
            *r2.mr = p.mr;
            if (!g_text_ready && (p.mr.type == CHIBSJ || p.mr.type == CHI2BSJ)) {   // call_circ does nothing for other types
                genome_seeder.pac2char_whole_contig();
                g_text_ready = true;
            }
            pc.call_circ(&r1, &r2);
            g_t_ref_pairs += get_real_time() - t_p0;
        }
    }
    r1.finalize();
    r2.finalize();
}

}  // namespace

void ProcessCirc::cm_shim_do_process(void) {
    const char *sw = getenv("CM_GPU_STAGE2");
    if ((sw && !strcmp(sw, "0")) || !pairedEnd) {
        cm_ref_do_process(this);
        return;
    }
    // ---- set-up, as do_process :200-262 -----------------------------------------------------------------------------------
    double cputime_start = get_cpu_time(), realtime_start = get_real_time();
    std::vector<ContigLen> orig_contig_len;
    genome_packer.load_index_info(orig_contig_len);
    char index_file[FILE_NAME_MAX_LEN];
    strcpy(index_file, genome_packer.get_index_fname().c_str());
    initCommon();
    THREAD_COUNT = threadCount;
    for (int i = 0; i < 255; i++) THREAD_ID[i] = i;
    if (!checkHashTable(index_file)) return;
    if (!initLoadingCompressedGenomeMeta(index_file)) return;
    kmer = WINDOW_SIZE + checkSumLength;
    genome_seeder.init(LOADCONTIGSTRINMEM);
    if (stage == 1) {
        Logger::instance().info("Loading GTF file...\n");
        gtf_parser.init(gtfFilename, orig_contig_len);
        if (!gtf_parser.load_gtf()) {
            fprintf(stdout, "Error in reading GTF file.\n");
            exit(1);
        }
        const double c = get_cpu_time(), r = get_real_time();
        Logger::instance().info("Completed! (CPU time: %.2lfs; Real time: %.2lfs)\n", c - cputime_start, r - realtime_start);
        cputime_start = c;
        realtime_start = r;
    }
    const double fq_cputime_start = cputime_start, fq_realtime_start = realtime_start;
    // ---- the sorted remain files, contig by contig --------------------------------------------------------------------------
    FASTQParser fq_parser1, fq_parser2;
    fq_parser1.init();
    fq_parser1.reset(fq_file1);
    fq_parser2.init();
    fq_parser2.reset(fq_file2);
    open_candid_file();
    std::vector<Pair> group;
    int group_contig = -1;
    auto flush = [&]() {
        if (group.empty()) return;
        const double t_l0 = get_real_time();
        while (pre_contig != group_contig) {      // :315-318
            // load_genome (:1659-1683) without its last step: the contig's text as characters (pac2char_whole_contig, about a
            // second per Gbp) is only needed by the reference's own per-pair code, and is made when a pair goes there
            double tmp;
            loadCompressedRefGenome(&tmp);
            pre_contig = atoi(getRefGenomeName()) - 1;
            contigNum = pre_contig;
            g_text_ready = false;
            refresh_hash_table_list();
        }
        g_t_load += get_real_time() - t_l0;
        process_contig(*this, group);
        group.clear();
    };
    while (true) {
        Record *c1 = fq_parser1.get_next_read(0);
        if (c1 == NULL) break;
        Record *c2 = fq_parser2.get_next_read(0);
        if ((int) c1->mr->contig_num != group_contig) {
            flush();
            group_contig = c1->mr->contig_num;
        }
        group.emplace_back();
        Pair &p = group.back();
        p.rname = c1->rname;
        p.seq1.assign(c1->seq, c1->seq_len);
        p.rc1.assign(c1->rcseq, c1->seq_len);
        if (c2) {
            p.seq2.assign(c2->seq, c2->seq_len);
            p.rc2.assign(c2->rcseq, c2->seq_len);
        }
        p.mr = *c1->mr;
    }
    flush();
    refresh_hash_table_list();
    const double t_r0 = get_real_time();
    report_events();
    if (getenv("CM_GPU_VERBOSE"))
        fprintf(stderr, "circminer_gpu: stage 2: load_genome %.2f s, single-split batches (device + index) %.2f s, their lines / CircRes %.2f s, "
                "other pairs through the reference's call_circ %.2f s (%ld of them redone after a device list overflowed), report_events %.2f s\n",
                g_t_load, g_t_batch, g_t_write, g_t_ref_pairs, g_n_redone, get_real_time() - t_r0);
    Logger::instance().info.set_prefix("[INFO] ");
    Logger::instance().info("CircRNA detection Completed! (CPU time: %.2lfs; Real time: %.2lfs)\n", get_cpu_time() - fq_cputime_start,
                            get_real_time() - fq_realtime_start);
    finalizeLoadingCompressedGenome();
    genome_seeder.finalize();
}
