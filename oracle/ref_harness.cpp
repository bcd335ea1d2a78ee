// TEST INFRASTRUCTURE ONLY -- never linked, imported or executed by the product path.
//
// Thin extern "C" harness compiled TOGETHER WITH the unmodified reference objects
// (see Makefile) into oracle/_ref/libcm_ref.so.  It drives the reference's own
// globals/singletons (gtf_parser, genome_seeder, filter_read, mrsFAST _ih_*) the way
// mapping() does (/root/reference/src/circminer.cpp:98-345) and exposes stage-level
// results as POD arrays so that tests can compare the CUDA path with the reference
// itself: seed lists, chains, alignments, process_read, plus the in-memory index and
// annotation a maintainer's shim would hand to cm_gpu_load_contig/_annotation.
#define private public
#define protected public
#include "common.h"
#include "align.h"
#include "chain.h"
#include "extend.h"
#include "filter.h"
#include "genome.h"
#include "gene_annotation.h"
#include "match_read.h"
#include "utils.h"
#include "fastq_parser.h"
#include "hash_table.h"
#include "process_circ.h"
#undef private
#undef protected

#include <cmath>
#include <string>
#include <vector>

#include "../include/cm_gpu.h"

extern "C" {
#include "mrsfast/Common.h"
#include "mrsfast/HashTable.h"
extern IHashTable *_ih_hashTable;
extern GeneralIndex *_ih_hashTableMem;
extern unsigned int _ih_hashTableMemSize;
extern CompressedSeq *_ih_crefGen;
extern int _ih_refGenLen;
extern int _ih_maxHashTableSize;
}
extern unsigned int CONTIG_SIZE;
extern unsigned int CONTIG_MAX_SIZE;
extern int maxCheckSumLen;
int cm_ref_main(int argc, char **argv);

static std::vector<ContigLen> g_contig_len;
static std::vector<GIMatchedKmer> g_fl, g_bl;
static chain_list g_cl[4];
static int g_max_seg_cnt = 0;
static bool g_filter_inited = false;
static std::string g_tmp_prefix;
static Record *g_rec1 = NULL, *g_rec2 = NULL;
static char g_comp[128];

static void alloc_chain_list(chain_list &cl) {
    cl.chains = (chain_t *) malloc(maxChainLen * sizeof(chain_t));
    for (int i = 0; i < maxChainLen; i++)
        cl.chains[i].frags = (fragment_t *) malloc(g_max_seg_cnt * sizeof(fragment_t));
    cl.best_chain_count = 0;
}

static void revcomp(const char *s, int len, char *rc) {
    // FASTQParser::set_comp / set_reverse_comp (fastq_parser.cpp:141-164)
    for (int i = 0; i < len; i++) rc[len - i - 1] = g_comp[(unsigned char) s[i] & 127];
    rc[len] = '\0';
}

extern "C" {

// run the unmodified main() (index build, whole pipeline)
int cmref_main(int argc, char **argv) {
    optind = 1;
    return cm_ref_main(argc, argv);
}

// Globals that parse_command would set (commandline_parser.cpp:7-30,247-262)
int cmref_set_params(const cm_params *p, int threads) {
    kmer = p->kmer;
    maxReadLength = p->max_read_len;
    maxEd = p->max_ed;
    maxSc = p->max_sc;
    bandWidth = p->band_width;
    seedLim = p->seed_lim;
    maxTlen = p->max_tlen;
    maxIntronLen = p->max_intron_len;
    maxChainLen = p->max_chain_len;
    scanLevel = p->scan_level;
    threadCount = threads;
    THREAD_COUNT = threads;
    for (int i = 0; i < 255; i++) THREAD_ID[i] = i;
    checkSumLength = (WINDOW_SIZE > kmer) ? 0 : kmer - WINDOW_SIZE;
    pairedEnd = true;
    initCommon();
    score_mat.init(p->mat_sc, p->mis_sc, p->ind_sc, p->xdrop);
    edit_mat.init(0, 1, 1, 10000);
    static bool fq_inited = false;
    if (!fq_inited) {   // main() does this (circminer.cpp:59-60); keeps the global destructors well-defined
        fq_parser1.init();
        fq_parser2.init();
        fq_inited = true;
    }
    memset(g_comp, 0, sizeof(g_comp));
    g_comp['A'] = 'T'; g_comp['C'] = 'G'; g_comp['G'] = 'C'; g_comp['T'] = 'A'; g_comp['N'] = 'N';
    g_comp['a'] = 'T'; g_comp['c'] = 'G'; g_comp['g'] = 'C'; g_comp['t'] = 'A'; g_comp['n'] = 'N';
    return 0;
}

// GenomePacker::build_index with a caller-chosen CONTIG_SIZE so that small synthetic
// genomes exercise the multi-contig "rounds" (genome.cpp:96-145). contig_size 0 = default.
int cmref_build_index(const char *ref_fa, unsigned contig_size, int compact) {
    CONTIG_SIZE = contig_size ? contig_size : DEF_CONTIG_SIZE;
    CONTIG_MAX_SIZE = DEF_CONTIG_MAX_SIZE;
    indexMode = true;
    genome_packer.init((char *) ref_fa);
    int r = genome_packer.build_index(compact != 0);
    indexMode = false;
    return r ? 0 : -1;
}

// mapping() prologue (circminer.cpp:125-209): index meta, seeder, scratch, GTF.
int cmref_open(const char *ref_fa, const char *gtf, const char *tmp_prefix) {
    genome_packer.init((char *) ref_fa);
    genome_packer.load_index_info(g_contig_len);
    char index_file[FILE_NAME_MAX_LEN];
    strcpy(index_file, genome_packer.get_index_fname().c_str());
    if (!checkHashTable(index_file)) return -1;
    if (!initLoadingHashTableMeta(index_file)) return -2;
    kmer = WINDOW_SIZE + checkSumLength;
    genome_seeder.init(LOADCONTIGSTRINMEM);
    g_max_seg_cnt = 2 * (ceil(1.0 * maxReadLength / kmer)) - 1;
    g_fl.resize(g_max_seg_cnt);
    g_bl.resize(g_max_seg_cnt);
    for (int i = 0; i < 4; i++) alloc_chain_list(g_cl[i]);
    strcpy(gtfFilename, gtf);
    gtf_parser.init((char *) gtf, g_contig_len);
    if (!gtf_parser.load_gtf()) return -3;
    g_tmp_prefix = tmp_prefix;
    strcpy(outputFilename, tmp_prefix);
    if (!g_rec1) {
        g_rec1 = new Record; g_rec1->init();
        g_rec2 = new Record; g_rec2->init();
    }
    return 0;
}

int cmref_kmer(void) { return kmer; }
int cmref_max_seg_cnt(void) { return g_max_seg_cnt; }
int cmref_n_chr(void) { return (int) g_contig_len.size(); }
const char *cmref_chr_name(int i) { return g_contig_len[i].name.c_str(); }
int cmref_chr_info(int i, uint32_t *contig_id, uint32_t *start, uint32_t *end) {
    *contig_id = g_contig_len[i].contig_id; *start = g_contig_len[i].start_pos; *end = g_contig_len[i].end_pos;
    return 0;
}

// one "round" (circminer.cpp:229-286): returns 1 if more contigs follow, 0 if last, <0 error
int cmref_next_contig(void) {
    double t;
    int flag = loadHashTable(&t);
    if (!genome_seeder.pac2char_whole_contig()) return -1;
    contigName = getRefGenomeName();
    contigNum = atoi(contigName) - 1;
    if (g_filter_inited) filter_read.finalize();
    char fq1[FILE_NAME_MAX_LEN], fq2[FILE_NAME_MAX_LEN];
    filter_read.init((char *) g_tmp_prefix.c_str(), true, contigNum + 1, !g_filter_inited, !flag, fq1, fq2,
                     g_contig_len);
    g_filter_inited = true;
    return flag ? 1 : 0;
}
int cmref_contig_num(void) { return contigNum; }

// what a maintainer's shim passes to cm_gpu_load_contig
const void *cmref_hash_table(void) { return _ih_hashTable; }
const void *cmref_hash_mem(void) { return _ih_hashTableMem; }
uint64_t cmref_hash_mem_size(void) {
    // _ih_hashTableMemSize is only maintained by the builder; recompute from the directory
    uint64_t n = 0;
    for (int hv = 0; hv < _ih_maxHashTableSize; hv++)
        if (_ih_hashTable[hv].list) {
            uint64_t end = (uint64_t) (_ih_hashTable[hv].list - _ih_hashTableMem) + _ih_hashTable[hv].list[0].info + 1;
            if (end > n) n = end;
        }
    return n;
}
const uint64_t *cmref_cref_gen(void) { return (const uint64_t *) _ih_crefGen; }
uint32_t cmref_ref_gen_len(void) { return (uint32_t) _ih_refGenLen; }
int cmref_genome_chars(uint32_t start, int len, char *out) { return genome_seeder.pac2char(start, len, out) ? 0 : -1; }

// ---- annotation dump (current contigNum) -------------------------------------------
// call with NULL pointers to get sizes first
int cmref_annot_sizes(uint32_t *n_iv, uint32_t *n_seg, uint32_t *n_tid, uint32_t *n_trans, uint32_t *n_t2s,
                      uint32_t *n_genes, uint32_t *n_chr) {
    FlatIntervalTree<UniqSeg> &T = gtf_parser.exons_int_map[contigNum];
    *n_iv = T.disjoint_intervals.size();
    uint32_t ns = 0, nt = 0;
    for (auto &iv : T.disjoint_intervals) {
        ns += iv.seg_list.size();
        for (auto &s : iv.seg_list) nt += s.trans_id.size();
    }
    *n_seg = ns; *n_tid = nt;
    uint32_t ntr = (uint32_t) contigNum < gtf_parser.trans2seg.size() ? gtf_parser.trans2seg[contigNum].size() : 0;
    *n_trans = ntr;
    uint32_t n2 = 0;
    for (uint32_t i = 0; i < ntr; i++) n2 += gtf_parser.trans2seg[contigNum][i].size();
    *n_t2s = n2;
    *n_genes = (uint32_t) contigNum < gtf_parser.gid2ginfo.size() ? gtf_parser.gid2ginfo[contigNum].size() : 0;
    *n_chr = gtf_parser.con2chr[contigNum].size();
    return 0;
}

int cmref_annot_fill(uint32_t *iv_spos, uint32_t *iv_epos, uint32_t *iv_max_end, uint32_t *iv_min_end,
                     uint32_t *iv_max_next_exon, uint32_t *iv_seg_off, uint32_t *seg_start, uint32_t *seg_end,
                     uint32_t *seg_next, uint32_t *seg_gene, uint32_t *seg_tid_off, uint32_t *tid,
                     int32_t *trans_start_ind, uint32_t *t2s_off, uint8_t *t2s, uint32_t *gene_start,
                     uint32_t *gene_end, uint32_t *chr_shift, int32_t *chr_id) {
    FlatIntervalTree<UniqSeg> &T = gtf_parser.exons_int_map[contigNum];
    uint32_t si = 0, ti = 0;
    for (size_t i = 0; i < T.disjoint_intervals.size(); i++) {
        auto &iv = T.disjoint_intervals[i];
        iv_spos[i] = iv.spos; iv_epos[i] = iv.epos;
        iv_max_end[i] = iv.max_end; iv_min_end[i] = iv.min_end; iv_max_next_exon[i] = iv.max_next_exon;
        iv_seg_off[i] = si;
        for (auto &s : iv.seg_list) {
            seg_start[si] = s.start; seg_end[si] = s.end; seg_next[si] = s.next_exon_beg; seg_gene[si] = s.gene_id;
            seg_tid_off[si] = ti;
            for (auto t : s.trans_id) tid[ti++] = t;
            si++;
        }
    }
    iv_seg_off[T.disjoint_intervals.size()] = si;
    seg_tid_off[si] = ti;
    uint32_t ntr = (uint32_t) contigNum < gtf_parser.trans2seg.size() ? gtf_parser.trans2seg[contigNum].size() : 0;
    uint32_t o = 0;
    for (uint32_t i = 0; i < ntr; i++) {
        trans_start_ind[i] = gtf_parser.trans_start_ind[contigNum][i];
        t2s_off[i] = o;
        for (auto v : gtf_parser.trans2seg[contigNum][i]) t2s[o++] = v;
    }
    t2s_off[ntr] = o;
    if ((uint32_t) contigNum < gtf_parser.gid2ginfo.size())
        for (size_t i = 0; i < gtf_parser.gid2ginfo[contigNum].size(); i++) {
            gene_start[i] = gtf_parser.gid2ginfo[contigNum][i].start;
            gene_end[i] = gtf_parser.gid2ginfo[contigNum][i].end;
        }
    for (size_t i = 0; i < gtf_parser.con2chr[contigNum].size(); i++) {
        chr_shift[i] = gtf_parser.con2chr[contigNum][i].shift;
        chr_id[i] = -1;
        for (size_t j = 0; j < g_contig_len.size(); j++)
            if (g_contig_len[j].name == gtf_parser.con2chr[contigNum][i].contig) chr_id[i] = (int32_t) j;
    }
    return 0;
}

// copy the first n_bits of the two std::bitset masks (common.h:117-118) into 64-bit words
int cmref_annot_masks(uint64_t n_bits, uint64_t *near_border, uint64_t *intronic) {
    uint64_t nw = (n_bits + 63) / 64;
    memset(near_border, 0, nw * 8);
    memset(intronic, 0, nw * 8);
    if ((size_t) contigNum < near_border_bs.size())
        for (uint64_t k = 0; k < n_bits; k++)
            if (near_border_bs[contigNum][k]) near_border[k >> 6] |= 1ull << (k & 63);
    if ((size_t) contigNum < intronic_bs.size())
        for (uint64_t k = 0; k < n_bits; k++)
            if (intronic_bs[contigNum][k]) intronic[k >> 6] |= 1ull << (k & 63);
    return 0;
}

// the look-ups of cm_gpu_annot_lookup_batch by the reference's own GTFParser (current contigNum)
int cmref_annot_lookup(uint32_t n, const uint32_t *pos, uint32_t mlen, uint32_t rlen, uint32_t *ub, uint32_t *max_end, int32_t *ol_iv,
                       int32_t *loc_iv, int32_t *loc_ind, int32_t *shift_idx, uint8_t *intronic) {
    FlatIntervalTree<UniqSeg> &T = gtf_parser.exons_int_map[contigNum];
    const IntervalInfo<UniqSeg> *base = T.disjoint_intervals.data();
    for (uint32_t i = 0; i < n; i++) {
        const IntervalInfo<UniqSeg> *ol = NULL;
        uint32_t me = 0;
        ub[i] = gtf_parser.get_upper_bound(pos[i], mlen, rlen, me, ol);
        max_end[i] = me;
        ol_iv[i] = ol ? (int32_t) (ol - base) : -1;
        int ind = -1;
        const IntervalInfo<UniqSeg> *lo = gtf_parser.get_location_overlap_ind(pos[i], false, ind);
        loc_iv[i] = lo ? (int32_t) (lo - base) : -1;
        loc_ind[i] = ind;
        ConShift cs = gtf_parser.get_shift(contigNum, pos[i]);
        shift_idx[i] = -1;
        for (size_t k = 0; k < gtf_parser.con2chr[contigNum].size(); k++)
            if (gtf_parser.con2chr[contigNum][k].shift == cs.shift && gtf_parser.con2chr[contigNum][k].contig == cs.contig) shift_idx[i] = (int32_t) k;
        intronic[i] = intronic_bs[contigNum][pos[i]] ? 1 : 0;
    }
    return 0;
}

// cm_gpu_extend_batch by the reference's own TransExtension (thread 0's, created by FilterRead::init)
static void mate_export(const MatchedMate &m, cm_mate *o) {
    o->spos = m.spos; o->epos = m.epos; o->qspos = m.qspos; o->qepos = m.qepos; o->matched_len = m.matched_len;
    o->left_ed = m.left_ed; o->middle_ed = m.middle_ed; o->right_ed = m.right_ed; o->sclen_left = m.sclen_left; o->sclen_right = m.sclen_right;
    o->dir = m.dir; o->type = m.type;
    o->is_concord = m.is_concord; o->left_ok = m.left_ok; o->right_ok = m.right_ok; o->pad_ = 0;
}
static void mate_fresh(MatchedMate &m) {   // MatchedMate() leaves these unset; the device starts them at 0
    m.spos = m.epos = m.qspos = m.qepos = 0;
    m.matched_len = 0;
    m.dir = 0;
    m.is_concord = false;
}
int cmref_extend(int mode, uint32_t n, const char *lseq, const char *rseq, uint32_t stride, const uint16_t *llen, const uint16_t *rlen,
                 uint32_t F, const uint32_t *lcl, const uint32_t *lrp, const int32_t *lqp, const uint32_t *rcl, const uint32_t *rrp,
                 const int32_t *rqp, const int32_t *iv_a, const int32_t *iv_b, const int32_t *dir, cm_mate *lmm, cm_mate *rmm, int32_t *ret) {
    std::vector<fragment_t> lf(F), rf(F);
    std::vector<char> ls(stride + 8), rs(stride + 8);
    FlatIntervalTree<UniqSeg> &T = gtf_parser.exons_int_map[contigNum];
    for (uint32_t i = 0; i < n; i++) {
        chain_t lc, rc;
        lc.frags = lf.data(); lc.chain_len = lcl[i]; lc.score = 0;
        for (uint32_t f = 0; f < lcl[i]; f++) { lf[f].rpos = lrp[(size_t) i * F + f]; lf[f].qpos = lqp[(size_t) i * F + f]; lf[f].len = kmer; }
        memcpy(ls.data(), lseq + (size_t) i * stride, llen[i]); ls[llen[i]] = 0;
        MatchedMate a, b;
        mate_fresh(a);
        mate_fresh(b);
        if (mode == 0) {
            rc.frags = rf.data(); rc.chain_len = rcl[i]; rc.score = 0;
            for (uint32_t f = 0; f < rcl[i]; f++) { rf[f].rpos = rrp[(size_t) i * F + f]; rf[f].qpos = rqp[(size_t) i * F + f]; rf[f].len = kmer; }
            memcpy(rs.data(), rseq + (size_t) i * stride, rlen[i]); rs[rlen[i]] = 0;
            std::vector<uint32_t> common_tid;
            if (iv_a[i] >= 0 && iv_b[i] >= 0) same_transcript(&T.disjoint_intervals[iv_a[i]], &T.disjoint_intervals[iv_b[i]], common_tid);
            ret[i] = filter_read.extension[0].extend_both_mates(lc, rc, common_tid, ls.data(), rs.data(), 1, 1, llen[i], rlen[i], a, b) ? 1 : 0;
            mate_export(b, &rmm[i]);
        } else {
            ret[i] = filter_read.extension[0].extend_chain_both_sides(lc, ls.data(), llen[i], a, dir[i]);
        }
        mate_export(a, &lmm[i]);
    }
    return 0;
}

// ---- stage dumps ---------------------------------------------------------------------
// GenomeSeeder::split_match_hash on one strand; arrays of max_seg_cnt
int cmref_seed(const char *seq, int len, int32_t *qpos, uint32_t *frag_count, uint8_t *has_frags,
               uint32_t *first_info, uint32_t *last_info, uint64_t *sum_info) {
    std::vector<char> s(seq, seq + len);
    s.push_back(0);
    genome_seeder.split_match_hash(s.data(), len, kmer, g_fl.data());
    for (int i = 0; i < g_max_seg_cnt; i++) {
        GIMatchedKmer &m = g_fl[i];
        qpos[i] = (i % 2 == 0 && i / 2 * kmer + kmer <= len) ? m.qpos : 0;   // odd/unused slots keep stale qpos
        frag_count[i] = m.frag_count;
        has_frags[i] = m.frags != NULL;
        first_info[i] = last_info[i] = 0;
        sum_info[i] = 0;
        if (m.frags && m.frag_count) {
            first_info[i] = m.frags[0].info;
            last_info[i] = m.frags[m.frag_count - 1].info;
            for (uint32_t j = 0; j < m.frag_count; j++) sum_info[i] += m.frags[j].info;
        }
    }
    return 0;
}

// copy the hit list of one slot after cmref_seed (positions), up to cap
int cmref_seed_slot(int slot, uint32_t cap, uint32_t *info) {
    GIMatchedKmer &m = g_fl[slot];
    uint32_t n = m.frag_count < cap ? m.frag_count : cap;
    for (uint32_t j = 0; j < n; j++) info[j] = m.frags[j].info;
    return (int) n;
}

// FilterRead::get_best_chains on one strand (filter.cpp:469-482)
int cmref_chain(const char *seq, int len, int32_t *chain_count, int32_t *high_hits, float *score,
                uint32_t *chain_len, uint32_t *frag_rpos, int32_t *frag_qpos) {
    std::vector<char> s(seq, seq + len);
    s.push_back(0);
    int hh = 0;
    filter_read.get_best_chains(s.data(), len, kmer, g_cl[0], g_fl.data(), hh);
    int F = (g_max_seg_cnt + 1) / 2;
    *chain_count = g_cl[0].best_chain_count;
    *high_hits = hh;
    for (int c = 0; c < maxChainLen; c++) {
        score[c] = 0; chain_len[c] = 0;
        for (int f = 0; f < F; f++) { frag_rpos[c * F + f] = 0; frag_qpos[c * F + f] = 0; }
    }
    for (int c = 0; c < g_cl[0].best_chain_count; c++) {
        score[c] = g_cl[0].chains[c].score;
        chain_len[c] = g_cl[0].chains[c].chain_len;
        for (uint32_t f = 0; f < g_cl[0].chains[c].chain_len && (int) f < F; f++) {
            frag_rpos[c * F + f] = g_cl[0].chains[c].frags[f].rpos;
            frag_qpos[c * F + f] = g_cl[0].chains[c].frags[f].qpos;
        }
    }
    return 0;
}

static int chr_index(const std::string &name) {
    for (size_t j = 0; j < g_contig_len.size(); j++)
        if (g_contig_len[j].name == name) return (int) j;
    return -1;
}

static void mread_to_ref(const cm_mread *m, MatchedRead *r) {
    r->spos_r1 = m->spos_r1; r->spos_r2 = m->spos_r2; r->epos_r1 = m->epos_r1; r->epos_r2 = m->epos_r2;
    r->qspos_r1 = m->qspos_r1; r->qspos_r2 = m->qspos_r2; r->qepos_r1 = m->qepos_r1; r->qepos_r2 = m->qepos_r2;
    r->mlen_r1 = m->mlen_r1; r->mlen_r2 = m->mlen_r2;
    r->r1_forward = m->r1_forward; r->r2_forward = m->r2_forward;
    r->ed_r1 = m->ed_r1; r->ed_r2 = m->ed_r2; r->type = m->type; r->tlen = m->tlen;
    r->junc_num = m->junc_num; r->gm_compatible = m->gm_compatible; r->contig_num = m->contig_num;
    r->chr_r1 = m->chr_r1 >= 0 ? g_contig_len[m->chr_r1].name : std::string("-");
    r->chr_r2 = m->chr_r2 >= 0 ? g_contig_len[m->chr_r2].name : std::string("-");
}

static void ref_to_mread(const MatchedRead *r, cm_mread *m) {
    memset(m, 0, sizeof(*m));
    m->spos_r1 = r->spos_r1; m->spos_r2 = r->spos_r2; m->epos_r1 = r->epos_r1; m->epos_r2 = r->epos_r2;
    m->qspos_r1 = r->qspos_r1; m->qspos_r2 = r->qspos_r2; m->qepos_r1 = r->qepos_r1; m->qepos_r2 = r->qepos_r2;
    m->mlen_r1 = r->mlen_r1; m->mlen_r2 = r->mlen_r2;
    m->r1_forward = r->r1_forward; m->r2_forward = r->r2_forward;
    m->ed_r1 = r->ed_r1; m->ed_r2 = r->ed_r2; m->type = r->type; m->tlen = r->tlen;
    m->junc_num = r->junc_num; m->gm_compatible = r->gm_compatible; m->contig_num = r->contig_num;
    m->chr_r1 = chr_index(r->chr_r1);
    m->chr_r2 = chr_index(r->chr_r2);
}

// FilterRead::process_read, PE (filter.cpp:124-241), n pairs, rows of `stride` bytes
int cmref_process_pe(uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride, const uint16_t *len1,
                     const uint16_t *len2, const cm_mread *in, cm_mread *out, int32_t *state) {
    for (uint32_t i = 0; i < n; i++) {
        int l1 = len1[i], l2 = len2[i];
        memcpy(g_rec1->seq, seqs1 + (size_t) i * stride, l1); g_rec1->seq[l1] = 0;
        memcpy(g_rec2->seq, seqs2 + (size_t) i * stride, l2); g_rec2->seq[l2] = 0;
        revcomp(g_rec1->seq, l1, g_rec1->rcseq);
        revcomp(g_rec2->seq, l2, g_rec2->rcseq);
        g_rec1->seq_len = l1; g_rec2->seq_len = l2;
        strcpy(g_rec1->rname, "r"); strcpy(g_rec2->rname, "r");
        mread_to_ref(&in[i], g_rec1->mr);
        state[i] = filter_read.process_read(0, g_rec1, g_rec2, kmer, g_fl.data(), g_bl.data(), g_cl[0], g_cl[1],
                                            g_cl[2], g_cl[3]);
        ref_to_mread(g_rec1->mr, &out[i]);
    }
    return 0;
}

// SE overload (filter.cpp:86-121)
int cmref_process_se(uint32_t n, const char *seqs, uint32_t stride, const uint16_t *len, int32_t *state) {
    for (uint32_t i = 0; i < n; i++) {
        int l1 = len[i];
        memcpy(g_rec1->seq, seqs + (size_t) i * stride, l1); g_rec1->seq[l1] = 0;
        revcomp(g_rec1->seq, l1, g_rec1->rcseq);
        g_rec1->seq_len = l1;
        strcpy(g_rec1->rname, "r");
        state[i] = filter_read.process_read(0, g_rec1, kmer, g_fl.data(), g_bl.data(), g_cl[0], g_cl[1]);
    }
    return 0;
}

// alignment kinds as in cm_gpu_align_batch
int cmref_align(int kind, uint32_t n, const char *s, const uint16_t *n_s, const char *t, const uint16_t *n_t,
                uint32_t stride, const int32_t *aux, int32_t *out_ed, int32_t *out_sclen, int32_t *out_indel,
                int32_t *out_score) {
    static DropAlignment *al = NULL;
    static EditDistAlignment *eal = NULL;
    if (!al) { al = new DropAlignment(); eal = new EditDistAlignment(); }
    al->init();
    eal->init();
    std::vector<char> sb(stride + 8), tb(stride + 8);
    for (uint32_t i = 0; i < n; i++) {
        int ns = n_s[i], nt = n_t[i];
        memcpy(sb.data(), s + (size_t) i * stride, ns); sb[ns] = 0;
        memcpy(tb.data(), t + (size_t) i * stride, nt); tb[nt] = 0;
        int sc = 0, indel = 0, ed = 0;
        if (kind == 0) ed = al->local_alignment_right_sc(sb.data(), ns, tb.data(), nt, sc, indel);
        else if (kind == 1) ed = al->local_alignment_left_sc(sb.data(), ns, tb.data(), nt, sc, indel);
        else if (kind == 2) ed = al->local_alignment_right(sb.data(), ns, tb.data(), nt, indel);
        else if (kind == 3) ed = al->local_alignment_left(sb.data(), ns, tb.data(), nt, indel);
        else if (kind == 4) ed = al->global_one_side_banded_alignment(sb.data(), ns, tb.data(), nt, aux[i]);
        else if (kind == 5) ed = eal->local_alignment_right_sc(sb.data(), ns, tb.data(), nt, sc, indel);
        else if (kind == 6) ed = eal->local_alignment_left_sc(sb.data(), ns, tb.data(), nt, sc, indel);
        out_ed[i] = ed; out_sclen[i] = sc; out_indel[i] = indel;
        out_score[i] = (kind == 4) ? 0 : (kind >= 5 ? eal->get_score() : al->get_score());
    }
    return 0;
}


// Stage 2: ProcessCirc::chaining (process_circ.cpp:678-737) on one RegionalHashTable per gene, built the way
// get_hash_table does (:814-830: pac2char_otf of the gene, create_table(seq, 0, len)).  One table object is reused:
// create_table resets it (hash_table.cpp:58-59).
static ProcessCirc *g_pc = NULL;
static RegionalHashTable *g_ht = NULL;
static uint32_t g_ht_s = 0, g_ht_e = 0;
static int circ_pc(int ws) {
    if (!g_pc) {
        for (int m = 1; m <= 2; m++) {   // the constructor sorts the last round's remain files
            char fn[FILE_NAME_MAX_LEN];
            sprintf(fn, "%s_1_remain_R%d.fastq", outputFilename, m);
            FILE *f = fopen(fn, "w");
            if (f) fclose(f);
        }
        finalCleaning = false;
        g_pc = new ProcessCirc(1, ws);
    }
    return g_pc->window_size == ws ? 0 : -1;
}

// ProcessCirc::find_exact_coord (process_circ.cpp:739-789) per row: the two mates of the pair are given by their contig
// coordinates (all find_exact_coord reads of them), the chain as fragments of length ws
int cmref_circ_exact(uint32_t n, const char *seqs, uint32_t stride, const uint32_t *r1_spos, const uint32_t *r1_epos, const uint32_t *r2_spos,
                     const uint32_t *r2_epos, const int32_t *dir, const uint32_t *qspos, const int32_t *rlen, const int32_t *whole_len,
                     uint32_t F, const uint32_t *cl, const uint32_t *rp, const int32_t *qp, int ws, cm_mate *pmm, int32_t *ret) {
    if (circ_pc(ws) != 0) return -1;
    std::vector<fragment_t> fr(F);
    std::vector<char> row(stride + 8);
    for (uint32_t i = 0; i < n; i++) {
        chain_t bc;
        bc.frags = fr.data(); bc.chain_len = cl[i]; bc.score = 0;
        for (uint32_t f = 0; f < cl[i]; f++) { fr[f].rpos = rp[(size_t) i * F + f]; fr[f].qpos = qp[(size_t) i * F + f]; fr[f].len = ws; }
        memcpy(row.data(), seqs + (size_t) i * stride, stride);
        row[stride] = 0;
        MatchedMate a, b, pm;
        mate_fresh(a); mate_fresh(b); mate_fresh(pm);
        a.spos = r1_spos[i]; a.epos = r1_epos[i];
        b.spos = r2_spos[i]; b.epos = r2_epos[i];
        ret[i] = g_pc->find_exact_coord(a, b, pm, dir[i], qspos[i], row.data(), rlen[i], whole_len[i], bc) ? 1 : 0;
        mate_export(pm, &pmm[i]);
    }
    return 0;
}

int cmref_circ_chain(uint32_t n, const char *seqs, uint32_t stride, const uint32_t *qspos, const uint32_t *qepos,
                     const uint32_t *gene_start, const uint32_t *gene_end, int ws, int step, int F, int32_t *chain_count,
                     float *score, uint32_t *chain_len, uint32_t *frag_rpos, int32_t *frag_qpos) {
    if (circ_pc(ws) != 0) return -1;
    g_pc->step = step;
    std::vector<char> row(stride + 1);
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t gs = gene_start[i], ge = gene_end[i];
        const int gene_len = (int) (ge - gs + 1);
        if (!g_ht || gs != g_ht_s || ge != g_ht_e) {
            std::vector<char> gene_seq(gene_len + 1);
            gene_seq[gene_len] = 0;
            if (!genome_seeder.pac2char_otf(gs, gene_len, gene_seq.data())) return -2;
            if (!g_ht) g_ht = new RegionalHashTable(ws, gs, ge);
            g_ht->gene_spos = gs;
            g_ht->gene_epos = ge;
            g_ht->create_table(gene_seq.data(), 0, gene_len);
            g_ht_s = gs;
            g_ht_e = ge;
        }
        memcpy(row.data(), seqs + (size_t) i * stride, stride);
        row[stride] = 0;
        chain_list &bc = g_pc->bc1;
        bc.best_chain_count = 0;
        g_pc->chaining(qspos[i], qepos[i], g_ht, row.data(), gene_len, gs, bc);
        chain_count[i] = bc.best_chain_count;
        const size_t o = (size_t) i * maxChainLen;
        for (int c = 0; c < maxChainLen; c++) {
            score[o + c] = 0;
            chain_len[o + c] = 0;
            for (int f = 0; f < F; f++) { frag_rpos[(o + c) * F + f] = 0; frag_qpos[(o + c) * F + f] = 0; }
        }
        for (int c = 0; c < bc.best_chain_count; c++) {
            score[o + c] = bc.chains[c].score;
            chain_len[o + c] = bc.chains[c].chain_len;
            for (uint32_t f = 0; f < bc.chains[c].chain_len && (int) f < F; f++) {
                frag_rpos[(o + c) * F + f] = bc.chains[c].frags[f].rpos;
                frag_qpos[(o + c) * F + f] = bc.chains[c].frags[f].qpos;
                if (bc.chains[c].frags[f].len != ws) return -3;
            }
        }
    }
    return 0;
}

// ---- stage 2, the single-split path with every check_split_map call recorded -------------------------------------------
static void circ_mate_export(const MatchedMate &m, cm_circ_mate *o) {
    FlatIntervalTree<UniqSeg> &T = gtf_parser.exons_int_map[contigNum];
    const IntervalInfo<UniqSeg> *base = T.disjoint_intervals.data();
    mate_export(m, &o->m);
    o->exons_spos = m.exons_spos ? (int32_t) (m.exons_spos - base) : -1;
    o->exons_epos = m.exons_epos ? (int32_t) (m.exons_epos - base) : -1;
    o->exon_ind_spos = m.exon_ind_spos;
    o->exon_ind_epos = m.exon_ind_epos;
    o->looked_up_spos = m.looked_up_spos;
    o->looked_up_epos = m.looked_up_epos;
    o->pad_[0] = o->pad_[1] = 0;
}
static void circ_res_export(const CircRes &cr, bool was_set, cm_circ_res *o) {
    memset(o, 0, sizeof(*o));
    if (!was_set) return;
    o->spos = cr.spos; o->epos = cr.epos;
    o->start_signal_n = (uint8_t) cr.start_signal.size(); o->end_signal_n = (uint8_t) cr.end_signal.size();
    for (size_t k = 0; k < 2; k++) {
        if (k < cr.start_signal.size()) o->start_signal[k] = cr.start_signal[k];
        if (k < cr.end_signal.size()) o->end_signal[k] = cr.end_signal[k];
        if (k < cr.start_bp_ref.size()) o->start_bp_ref[k] = cr.start_bp_ref[k];
        if (k < cr.end_bp_ref.size()) o->end_bp_ref[k] = cr.end_bp_ref[k];
    }
    o->ref_n = (uint8_t) (cr.start_bp_ref.size() == 2 && cr.end_bp_ref.size() == 2 ? 2 : 0);
    o->was_set = 1;
}

// The control flow of ProcessCirc::call_circ_single_split (process_circ.cpp:360-483) over the reference's own member functions
// (get_gene_overlap, RegionalHashTable, chaining, find_exact_coord, check_split_map), for every pair whose MatchedRead is a
// CHIBSJ of the current contig (positions of `mr` are chromosome coordinates as process_read leaves them).  Every
// check_split_map call is recorded: the three mates as they were handed in, r1_partial, the two oriented reads, what it
// returned, the CircRes it filled and the mates as it left them.  best_type[i] = the type call_circ_single_split ends with
// (NF = 40 when nothing was found), 0xFF.. when the pair was not a candidate.  Returns the number of records (<= cap).
int cmref_circ_single(uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride, const uint16_t *len1, const uint16_t *len2,
                      const cm_mread *mr_in, int ws, uint32_t cap, uint32_t *call_pair, uint8_t *r1_partial_out, char *remain_rows,
                      char *fullmap_rows, int32_t *remain_len, int32_t *fullmap_len, cm_circ_mate *in_r1, cm_circ_mate *in_r2,
                      cm_circ_mate *in_pm, int32_t *ret, cm_circ_res *cr_out, cm_circ_mate *out_r1, cm_circ_mate *out_r2,
                      cm_circ_mate *out_pm, int32_t *best_type, cm_mread *mr_con, uint32_t gene_cap, uint32_t *gene_cnt, uint32_t *gene_se) {
    // mr_con[i]: the MatchedRead with contig coordinates (type -1 for pairs that are no candidates here); gene_cnt[i] /
    // gene_se[i][gene_cap][2]: start / end of the genes get_gene_overlap(mm_r1.spos) lists, in its order
    if (circ_pc(ws) != 0) return -1;
    ProcessCirc &pc = *g_pc;
    uint32_t n_rec = 0;
    std::vector<char> s1(stride + 8), s2(stride + 8), rc1(stride + 8), rc2(stride + 8);
    for (uint32_t i = 0; i < n; i++) {
        best_type[i] = -1;
        gene_cnt[i] = 0;
        mr_con[i] = mr_in[i];
        mr_con[i].type = -1;
        MatchedRead mr;
        mread_to_ref(&mr_in[i], &mr);
        if (mr.type == CHI2BSJ && (int) mr.contig_num == contigNum) {
            // inputs of the double-split path for the batch entry: contig coordinates and the genes under mr.spos_r1 (:521-524)
            gtf_parser.chrloc2conloc(mr.chr_r1, mr.spos_r1, mr.epos_r1);
            gtf_parser.chrloc2conloc(mr.chr_r2, mr.spos_r2, mr.epos_r2);
            mr_con[i].type = mr.type;
            mr_con[i].spos_r1 = mr.spos_r1; mr_con[i].epos_r1 = mr.epos_r1; mr_con[i].spos_r2 = mr.spos_r2; mr_con[i].epos_r2 = mr.epos_r2;
            const IntervalInfo<GeneInfo> *gi = gtf_parser.get_gene_overlap(mr.spos_r1, false);
            if (gi) {
                if (gi->seg_list.size() > gene_cap) return -3;
                gene_cnt[i] = (uint32_t) gi->seg_list.size();
                for (unsigned int g = 0; g < gi->seg_list.size(); ++g) {
                    gene_se[((size_t) i * gene_cap + g) * 2] = gi->seg_list[g].start;
                    gene_se[((size_t) i * gene_cap + g) * 2 + 1] = gi->seg_list[g].end;
                }
            }
            continue;
        }
        if (mr.type != CHIBSJ || (int) mr.contig_num != contigNum) continue;
        const int l1 = len1[i], l2 = len2[i];
        memcpy(s1.data(), seqs1 + (size_t) i * stride, l1); s1[l1] = 0;
        memcpy(s2.data(), seqs2 + (size_t) i * stride, l2); s2[l2] = 0;
        revcomp(s1.data(), l1, rc1.data()); rc1[l1] = 0;
        revcomp(s2.data(), l2, rc2.data()); rc2[l2] = 0;
        // call_circ (:334-358)
        pc.fullmap_seq = NULL; pc.remain_seq = NULL; pc.r1_seq = NULL; pc.r2_seq = NULL;
        pc.fullmap_seq_len = 0; pc.remain_seq_len = 0; pc.r1_seq_len = 0; pc.r2_seq_len = 0;
        const bool r1_partial = mr.mlen_r1 < mr.mlen_r2;
        pc.remain_seq = r1_partial ? (mr.r1_forward ? s1.data() : rc1.data()) : (mr.r2_forward ? s2.data() : rc2.data());
        pc.fullmap_seq = !r1_partial ? (mr.r1_forward ? s1.data() : rc1.data()) : (mr.r2_forward ? s2.data() : rc2.data());
        pc.remain_seq_len = r1_partial ? l1 : l2;
        pc.fullmap_seq_len = !r1_partial ? l1 : l2;
        gtf_parser.chrloc2conloc(mr.chr_r1, mr.spos_r1, mr.epos_r1);
        gtf_parser.chrloc2conloc(mr.chr_r2, mr.spos_r2, mr.epos_r2);
        mr_con[i].type = mr.type;
        mr_con[i].spos_r1 = mr.spos_r1; mr_con[i].epos_r1 = mr.epos_r1; mr_con[i].spos_r2 = mr.spos_r2; mr_con[i].epos_r2 = mr.epos_r2;
        MatchedMate mm_r1(mr, 1, l1, r1_partial);
        MatchedMate mm_r2(mr, 2, l2, !r1_partial);
        // this constructor leaves is_concord / left_ok / right_ok unset (src/common.cpp:192-235); nothing on the path reads them
        mm_r1.is_concord = mm_r1.left_ok = mm_r1.right_ok = false;
        mm_r2.is_concord = mm_r2.left_ok = mm_r2.right_ok = false;
        const MatchedMate &pmate = r1_partial ? mm_r1 : mm_r2;
        const uint32_t plen = r1_partial ? l1 : l2;
        const bool right_side = (pmate.qspos - 1) > (plen - pmate.qepos);
        uint32_t qspos = right_side ? 1 : pmate.qepos + 1;
        uint32_t qepos = right_side ? pmate.qspos - 1 : plen;
        const int whole_seq_len = plen;
        const int remain_len_i = qepos - qspos + 1;
        best_type[i] = NF;
        if (qepos < qspos || remain_len_i < ws) continue;
        const IntervalInfo<GeneInfo> *gene_info = gtf_parser.get_gene_overlap(mm_r1.spos, false);
        if (gene_info == NULL) continue;
        if (gene_info->seg_list.size() > gene_cap) return -3;
        gene_cnt[i] = (uint32_t) gene_info->seg_list.size();
        for (unsigned int g = 0; g < gene_info->seg_list.size(); ++g) {
            gene_se[((size_t) i * gene_cap + g) * 2] = gene_info->seg_list[g].start;
            gene_se[((size_t) i * gene_cap + g) * 2 + 1] = gene_info->seg_list[g].end;
        }
        int best = NF;
        bool done = false;
        for (unsigned int g = 0; g < gene_info->seg_list.size() && !done; ++g) {
            const GeneInfo &gi = gene_info->seg_list[g];
            const uint32_t gene_len = gi.end - gi.start + 1;
            if (!g_ht || gi.start != g_ht_s || gi.end != g_ht_e) {
                std::vector<char> gene_seq(gene_len + 1);
                gene_seq[gene_len] = 0;
                if (!genome_seeder.pac2char_otf(gi.start, gene_len, gene_seq.data())) return -2;
                if (!g_ht) g_ht = new RegionalHashTable(ws, gi.start, gi.end);
                g_ht->gene_spos = gi.start; g_ht->gene_epos = gi.end;
                g_ht->create_table(gene_seq.data(), 0, gene_len);
                g_ht_s = gi.start; g_ht_e = gi.end;
            }
            chain_list &bc = pc.bc1;
            pc.chaining(qspos, qepos, g_ht, pc.remain_seq, gene_len, gi.start, bc);
            if (bc.best_chain_count <= 0) continue;
            const bool forward = r1_partial ? mr.r1_forward : mr.r2_forward;
            const int dir = forward ? 1 : -1;
            for (int j = 0; j < (bc.best_chain_count < 10 ? bc.best_chain_count : 10) && !done; ++j) {   // TOPCHAIN
                MatchedMate partial_mm;
                mate_fresh(partial_mm);
                pc.find_exact_coord(mm_r1, mm_r2, partial_mm, dir, qspos, pc.remain_seq, remain_len_i, whole_seq_len, bc.chains[j]);
                if (partial_mm.type != CONCRD) continue;
                CircRes cr;
                if (n_rec < cap) {
                    call_pair[n_rec] = i;
                    r1_partial_out[n_rec] = r1_partial;
                    memset(remain_rows + (size_t) n_rec * stride, 0, stride);
                    memset(fullmap_rows + (size_t) n_rec * stride, 0, stride);
                    memcpy(remain_rows + (size_t) n_rec * stride, pc.remain_seq, pc.remain_seq_len);
                    memcpy(fullmap_rows + (size_t) n_rec * stride, pc.fullmap_seq, pc.fullmap_seq_len);
                    remain_len[n_rec] = pc.remain_seq_len; fullmap_len[n_rec] = pc.fullmap_seq_len;
                    circ_mate_export(mm_r1, &in_r1[n_rec]); circ_mate_export(mm_r2, &in_r2[n_rec]); circ_mate_export(partial_mm, &in_pm[n_rec]);
                }
                char *keep_remain = pc.remain_seq, *keep_full = pc.fullmap_seq;   // the four-mate path repoints remain_seq (:973, :983)
                const uint32_t keep_rl = pc.remain_seq_len;
                const int type = pc.check_split_map(mm_r1, mm_r2, partial_mm, r1_partial, cr);
                if (n_rec < cap) {
                    ret[n_rec] = type;
                    circ_res_export(cr, !cr.start_bp_ref.empty(), &cr_out[n_rec]);
                    circ_mate_export(mm_r1, &out_r1[n_rec]); circ_mate_export(mm_r2, &out_r2[n_rec]); circ_mate_export(partial_mm, &out_pm[n_rec]);
                }
                n_rec++;
                (void) keep_remain; (void) keep_full; (void) keep_rl;
                if (type < CR) { best = type; done = true; break; }
                if (type >= CR && type <= MCR && type < best) {
                    best = type;
                    if (type == CR) { done = true; break; }
                }
            }
        }
        best_type[i] = best;
    }
    return (int) (n_rec < cap ? n_rec : cap);
}

static void circ_mate_import(const cm_circ_mate *i, MatchedMate &m) {
    FlatIntervalTree<UniqSeg> &T = gtf_parser.exons_int_map[contigNum];
    const IntervalInfo<UniqSeg> *base = T.disjoint_intervals.data();
    m.spos = i->m.spos; m.epos = i->m.epos; m.qspos = i->m.qspos; m.qepos = i->m.qepos; m.matched_len = i->m.matched_len;
    m.left_ed = i->m.left_ed; m.middle_ed = i->m.middle_ed; m.right_ed = i->m.right_ed; m.sclen_left = i->m.sclen_left; m.sclen_right = i->m.sclen_right;
    m.dir = i->m.dir; m.type = i->m.type;
    m.is_concord = i->m.is_concord; m.left_ok = i->m.left_ok; m.right_ok = i->m.right_ok;
    m.looked_up_spos = i->looked_up_spos; m.looked_up_epos = i->looked_up_epos;
    m.exons_spos = i->exons_spos >= 0 ? base + i->exons_spos : NULL;
    m.exons_epos = i->exons_epos >= 0 ? base + i->exons_epos : NULL;
    m.exon_ind_spos = i->exon_ind_spos; m.exon_ind_epos = i->exon_ind_epos;
}

// ProcessCirc::check_split_map(mm_r1, mm_r2, partial_mm, r1_partial, cr) on given inputs (the arguments of cm_gpu_circ_check_batch)
int cmref_circ_check(uint32_t n, const char *remain_seq, const char *fullmap_seq, uint32_t stride, const int32_t *remain_len,
                     const int32_t *fullmap_len, const uint8_t *r1_partial, cm_circ_mate *mm_r1, cm_circ_mate *mm_r2, cm_circ_mate *partial_mm,
                     int ws, int32_t *ret, cm_circ_res *cr_out) {
    if (circ_pc(ws) != 0) return -1;
    ProcessCirc &pc = *g_pc;
    std::vector<char> rs(stride + 8), fs(stride + 8);
    for (uint32_t i = 0; i < n; i++) {
        memcpy(rs.data(), remain_seq + (size_t) i * stride, stride); rs[remain_len[i]] = 0;
        memcpy(fs.data(), fullmap_seq + (size_t) i * stride, stride); fs[fullmap_len[i]] = 0;
        pc.remain_seq = rs.data(); pc.fullmap_seq = fs.data(); pc.r1_seq = NULL; pc.r2_seq = NULL;
        pc.remain_seq_len = remain_len[i]; pc.fullmap_seq_len = fullmap_len[i]; pc.r1_seq_len = 0; pc.r2_seq_len = 0;
        MatchedMate a, b, p;
        circ_mate_import(&mm_r1[i], a); circ_mate_import(&mm_r2[i], b); circ_mate_import(&partial_mm[i], p);
        CircRes cr;
        ret[i] = pc.check_split_map(a, b, p, r1_partial[i] != 0, cr);
        circ_res_export(cr, !cr.start_bp_ref.empty(), &cr_out[i]);
        circ_mate_export(a, &mm_r1[i]); circ_mate_export(b, &mm_r2[i]); circ_mate_export(p, &partial_mm[i]);
    }
    return 0;
}

// ProcessCirc::check_split_map(mm_r1_1, mm_r2_1, mm_r1_2, mm_r2_2, cr) (process_circ.cpp:922-1133) on given inputs: the two oriented
// reads and the four mates; fullmap_sel picks what ProcessCirc::fullmap_seq holds when the call is made (0 none, 1 = R1, 2 = R2) --
// in the reference it is whatever an earlier call left, here it is an argument so that the call is a function of its inputs.
int cmref_circ_check4(uint32_t n, const char *r1seq, const char *r2seq, uint32_t stride, const int32_t *l1, const int32_t *l2,
                      const uint8_t *fullmap_sel, const cm_circ_mate *r1_1, const cm_circ_mate *r2_1, const cm_circ_mate *r1_2,
                      const cm_circ_mate *r2_2, int ws, int32_t *ret, cm_circ_res *cr_out) {
    if (circ_pc(ws) != 0) return -1;
    ProcessCirc &pc = *g_pc;
    std::vector<char> s1(stride + 8), s2(stride + 8);
    for (uint32_t i = 0; i < n; i++) {
        memcpy(s1.data(), r1seq + (size_t) i * stride, stride); s1[l1[i]] = 0;
        memcpy(s2.data(), r2seq + (size_t) i * stride, stride); s2[l2[i]] = 0;
        pc.r1_seq = s1.data(); pc.r2_seq = s2.data(); pc.r1_seq_len = l1[i]; pc.r2_seq_len = l2[i];
        pc.remain_seq = NULL; pc.remain_seq_len = 0;
        pc.fullmap_seq = fullmap_sel[i] == 1 ? s1.data() : fullmap_sel[i] == 2 ? s2.data() : NULL;
        pc.fullmap_seq_len = fullmap_sel[i] == 1 ? l1[i] : fullmap_sel[i] == 2 ? l2[i] : 0;
        MatchedMate a, b, c, d;
        circ_mate_import(&r1_1[i], a); circ_mate_import(&r2_1[i], b); circ_mate_import(&r1_2[i], c); circ_mate_import(&r2_2[i], d);
        CircRes cr;
        ret[i] = pc.check_split_map(a, b, c, d, cr);
        circ_res_export(cr, !cr.start_bp_ref.empty(), &cr_out[i]);
    }
    return 0;
}

// The reference's own ProcessCirc::call_circ_single_split (process_circ.cpp:360-483), unmodified, on the same pairs: what it
// writes to candidates.pam per check_split_map call (the type at the end of every line) and what it pushes to circ_res.  Pins
// the control flow restated in cmref_circ_single.  types[i][type_cap], n_types[i]; pushed[i] = 1 with the pushed CircRes's
// type / spos / epos (chromosome coordinates, as the reference stores them).
int cmref_circ_single_real(uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride, const uint16_t *len1, const uint16_t *len2,
                           const cm_mread *mr_in, int ws, uint32_t type_cap, uint32_t *n_types, int32_t *types, uint8_t *pushed, int32_t *push_type,
                           uint32_t *push_spos, uint32_t *push_epos) {
    if (circ_pc(ws) != 0) return -1;
    ProcessCirc &pc = *g_pc;
    pc.refresh_hash_table_list();   // do_process drops the per-gene tables when the contig changes (gene ids are per contig)
    for (uint32_t i = 0; i < n; i++) {
        n_types[i] = 0; pushed[i] = 0; push_type[i] = -1; push_spos[i] = push_epos[i] = 0;
        MatchedRead mr;
        mread_to_ref(&mr_in[i], &mr);
        if (mr.type != CHIBSJ || (int) mr.contig_num != contigNum) continue;
        const int l1 = len1[i], l2 = len2[i];
        memcpy(g_rec1->seq, seqs1 + (size_t) i * stride, l1); g_rec1->seq[l1] = 0;
        memcpy(g_rec2->seq, seqs2 + (size_t) i * stride, l2); g_rec2->seq[l2] = 0;
        revcomp(g_rec1->seq, l1, g_rec1->rcseq); g_rec1->rcseq[l1] = 0;
        revcomp(g_rec2->seq, l2, g_rec2->rcseq); g_rec2->rcseq[l2] = 0;
        g_rec1->seq_len = l1; g_rec2->seq_len = l2;
        strcpy(g_rec1->rname, "r"); strcpy(g_rec2->rname, "r");
        *(g_rec1->mr) = mr;
        *(g_rec2->mr) = mr;
        char *buf = NULL;
        size_t blen = 0;
        FILE *mem = open_memstream(&buf, &blen);
        FILE *keep = pc.candid_file;
        pc.candid_file = mem;
        const size_t before = pc.circ_res.size();
        pc.call_circ(g_rec1, g_rec2);
        fflush(mem);
        pc.candid_file = keep;
        fclose(mem);
        // one line per attempt: ... \t<type>\n
        for (char *ln = buf; ln && *ln;) {
            char *nl = strchr(ln, '\n');
            if (!nl) break;
            char *tab = nl;
            while (tab > ln && *(tab - 1) != '\t') tab--;
            if (n_types[i] < type_cap) types[(size_t) i * type_cap + n_types[i]] = atoi(tab);
            n_types[i]++;
            ln = nl + 1;
        }
        free(buf);
        if (pc.circ_res.size() > before) {
            const CircRes &cr = pc.circ_res.back();
            pushed[i] = 1; push_type[i] = cr.type; push_spos[i] = cr.spos; push_epos[i] = cr.epos;
            pc.circ_res.pop_back();
        }
    }
    return 0;
}

// The unmodified ProcessCirc::call_circ on every pair of type want_type (CHIBSJ 3 / CHI2BSJ 4) of the current contig: the text it
// writes to candidates.pam (text[text_off[i] .. text_off[i + 1])) and everything it pushes to circ_res (up to push_cap per pair:
// type, spos, epos in chromosome coordinates, the two signals and reference bases as 8 bytes).
int cmref_circ_call_real(uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride, const uint16_t *len1, const uint16_t *len2,
                         const cm_mread *mr_in, int ws, int want_type, uint64_t text_cap, uint64_t *text_off, char *text, uint32_t push_cap,
                         uint32_t *n_push, int32_t *push_type, uint32_t *push_spos, uint32_t *push_epos, uint8_t *push_sig) {
    if (circ_pc(ws) != 0) return -1;
    ProcessCirc &pc = *g_pc;
    pc.refresh_hash_table_list();
    uint64_t at = 0;
    for (uint32_t i = 0; i < n; i++) {
        text_off[i] = at;
        n_push[i] = 0;
        MatchedRead mr;
        mread_to_ref(&mr_in[i], &mr);
        if ((int) mr.type != want_type || (int) mr.contig_num != contigNum) continue;
        const int l1 = len1[i], l2 = len2[i];
        memcpy(g_rec1->seq, seqs1 + (size_t) i * stride, l1); g_rec1->seq[l1] = 0;
        memcpy(g_rec2->seq, seqs2 + (size_t) i * stride, l2); g_rec2->seq[l2] = 0;
        revcomp(g_rec1->seq, l1, g_rec1->rcseq); g_rec1->rcseq[l1] = 0;
        revcomp(g_rec2->seq, l2, g_rec2->rcseq); g_rec2->rcseq[l2] = 0;
        g_rec1->seq_len = l1; g_rec2->seq_len = l2;
        strcpy(g_rec1->rname, "r"); strcpy(g_rec2->rname, "r");
        *(g_rec1->mr) = mr;
        *(g_rec2->mr) = mr;
        char *buf = NULL;
        size_t blen = 0;
        FILE *mem = open_memstream(&buf, &blen);
        FILE *keep = pc.candid_file;
        pc.candid_file = mem;
        const size_t before = pc.circ_res.size();
        pc.call_circ(g_rec1, g_rec2);
        fflush(mem);
        pc.candid_file = keep;
        fclose(mem);
        if (at + blen > text_cap) { free(buf); return -4; }
        memcpy(text + at, buf, blen);
        at += blen;
        free(buf);
        for (size_t q = before; q < pc.circ_res.size(); q++) {
            const CircRes &cr = pc.circ_res[q];
            const uint32_t u = n_push[i];
            if (u < push_cap) {
                const size_t o = (size_t) i * push_cap + u;
                push_type[o] = cr.type; push_spos[o] = cr.spos; push_epos[o] = cr.epos;
                uint8_t *sg = push_sig + o * 8;
                memset(sg, 0, 8);
                for (size_t c = 0; c < 2; c++) {
                    if (c < cr.start_signal.size()) sg[c] = cr.start_signal[c];
                    if (c < cr.end_signal.size()) sg[2 + c] = cr.end_signal[c];
                    if (c < cr.start_bp_ref.size()) sg[4 + c] = cr.start_bp_ref[c];
                    if (c < cr.end_bp_ref.size()) sg[6 + c] = cr.end_bp_ref[c];
                }
            }
            n_push[i]++;
        }
        pc.circ_res.resize(before);
    }
    text_off[n] = at;
    return 0;
}
}  // extern "C"
