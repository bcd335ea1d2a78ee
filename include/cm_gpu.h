/*
 * cm_gpu.h -- C-ABI of the B200 mapping hot path (seed lookup -> chaining ->
 * extension/alignment -> per-pair classification) of CircMiner.
 *
 * The reference has no plugin/FFI layer; this boundary sits exactly on the
 * member-function seam its thread body calls once per pair
 * (/root/reference/src/circminer.cpp:354-406 map_reads ->
 *  FilterRead::process_read, src/filter.cpp:86 and :124).  Each export names
 * the reference interface it replaces.  Plain pointers and sizes only; every
 * function returns 0 on success and a negative cm_status on failure (the
 * reference exit(1)s / asserts instead); cm_gpu_last_error() gives the text.
 * All host buffers are caller-owned; the context owns all device memory; no
 * pointer handed back outlives the call.  One context per GPU; calls on one
 * context must be serialised by the caller, different contexts are concurrent.
 */
#ifndef CM_GPU_H
#define CM_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum cm_status {
    CM_OK = 0,
    CM_ERR_ARG = -1,       /* bad argument / precondition violated               */
    CM_ERR_CUDA = -2,      /* CUDA runtime error (no device, launch failure ...)  */
    CM_ERR_ALLOC = -3,     /* host or device allocation failed                   */
    CM_ERR_STATE = -4      /* call order violated (contig / annotation missing)  */
} cm_status;

/* output categories, order matters (src/common.h:59-72) */
enum {
    CM_CONCRD = 0, CM_DISCRD = 1, CM_CHIORF = 2, CM_CHIBSJ = 3, CM_CHI2BSJ = 4, CM_CONGEN = 5, CM_CHIFUS = 6,
    CM_CONGNM = 7, CM_OEA2 = 8, CM_CANDID = 9, CM_OEANCH = 10, CM_ORPHAN = 11, CM_NOPROC_MANYHIT = 12,
    CM_NOPROC_NOMATCH = 13
};

/* The reference's mutable globals (src/common.h:92-114, defaults in
 * src/commandline_parser.cpp:7-30) and score matrix (src/circminer.cpp:74),
 * frozen at context creation. */
typedef struct cm_params {
    int32_t kmer;            /* 14..22 ; WINDOW_SIZE 14 + checkSumLength            */
    int32_t max_read_len;    /* maxReadLength (-l), default 300                     */
    int32_t max_ed;          /* maxEd (-e), 4                                       */
    int32_t max_sc;          /* maxSc (-c), 7                                       */
    int32_t band_width;      /* bandWidth (-w), 3                                   */
    int32_t seed_lim;        /* seedLim (-S), 500                                   */
    int32_t max_tlen;        /* maxTlen (-T), 500                                   */
    int32_t max_intron_len;  /* maxIntronLen (-I), 2000000                          */
    int32_t max_chain_len;   /* maxChainLen (-C), 30                                */
    int32_t scan_level;      /* scanLevel (-a), 0                                   */
    int32_t mat_sc, mis_sc, ind_sc, xdrop;   /* score_mat.init(1,-3,-3,8)           */
} cm_params;

void cm_params_default(cm_params *p);

/* POD mirror of MatchedRead (src/common.h:311-352).  chr_r1/chr_r2 are indices
 * into the host's chromosome table (row order of <ref>.packed.fa.index.info,
 * src/genome.cpp:147-167); -1 stands for "-".  genome_spos is derived by the
 * host when it writes the remain-FASTQ header (src/filter.cpp:425-430). */
typedef struct cm_mread {
    uint32_t spos_r1, spos_r2, epos_r1, epos_r2;
    uint32_t qspos_r1, qspos_r2, qepos_r1, qepos_r2;
    uint32_t mlen_r1, mlen_r2;
    int32_t ed_r1, ed_r2;
    int32_t type;
    int32_t tlen;
    int32_t contig_num;
    int32_t chr_r1, chr_r2;
    uint16_t junc_num;
    uint8_t r1_forward, r2_forward, gm_compatible;
    uint8_t pad_[3];
} cm_mread;

/* MatchedRead::MatchedRead(void) (src/common.cpp:268-271) */
void cm_mread_init(cm_mread *m, const cm_params *p);

/* Flattened, read-only copy of what GTFParser holds for ONE packed contig
 * (src/gene_annotation.h:55-72): FlatIntervalTree<UniqSeg>::disjoint_intervals
 * as CSR (src/interval_tree.h:16-18, src/interval_info.h:9-18), trans2seg,
 * trans_start_ind, gid2ginfo, con2chr and the two bitsets
 * (src/common.h:117-118).  All positions contig-relative, 1-based. */
typedef struct cm_annot_flat {
    uint32_t n_iv;                 /* disjoint intervals (including the dummy)   */
    const uint32_t *iv_spos, *iv_epos, *iv_max_end, *iv_min_end, *iv_max_next_exon;
    const uint32_t *iv_seg_off;    /* n_iv + 1                                   */
    uint32_t n_seg;
    const uint32_t *seg_start, *seg_end, *seg_next_exon_beg, *seg_gene_id;
    const uint32_t *seg_tid_off;   /* n_seg + 1                                  */
    uint32_t n_tid;
    const uint32_t *tid;           /* UniqSeg::trans_id, concatenated            */
    uint32_t n_trans;
    const int32_t *trans_start_ind;/* n_trans                                    */
    const uint32_t *t2s_off;       /* n_trans + 1                                */
    const uint8_t *t2s;            /* trans2seg states 0/1/2/3                   */
    uint32_t n_genes;
    const uint32_t *gene_start, *gene_end;  /* gid2ginfo                         */
    uint32_t n_chr;                /* con2chr[contig]                            */
    const uint32_t *chr_shift;
    const int32_t *chr_id;         /* global chromosome index of each entry      */
    uint64_t n_bits;               /* bits available in the two masks            */
    const uint64_t *near_border;   /* bit k of word k/64 = near_border_bs[c][k]  */
    const uint64_t *intronic;
} cm_annot_flat;

typedef struct cm_counters {
    uint64_t seeds;          /* seeds looked up                                    */
    uint64_t seed_sectors;   /* algorithmic 32-B sectors, SURVEY.md 8(d)-1         */
    uint64_t seed_hits;      /* kept hit-list entries                              */
    uint64_t chain_links;    /* (i,j) candidates reaching src/chain.cpp:173        */
    uint64_t xdrop_cells;    /* src/align.cpp:296-325 cell visits                  */
    uint64_t edit_cells;     /* banded / one-side / full edit DP cells             */
    uint64_t xdrop_calls, edit_calls;
    uint64_t pairs;
    uint64_t kernel_launches;
} cm_counters;

typedef struct cm_gpu_ctx cm_gpu_ctx;

/* ---- life cycle ------------------------------------------------------- */
/* replaces the per-thread scratch set-up in mapping() (src/circminer.cpp:163-191)
 * and FilterRead::init's TransExtension array (src/filter.cpp:41-44) */
int cm_gpu_create(int device, const cm_params *params, cm_gpu_ctx **out);
/* A context on the same device that shares `parent`'s resident contigs and annotation
 * (read-only on this path) but owns its stream and batch buffers, so that two host threads
 * alternating batches overlap one batch's copies with the other's kernels -- the role the
 * reference's threadCount map_reads threads over shared globals play (src/circminer.cpp:290).
 * Loading entry points are refused on it; destroy it before its parent. */
/* Creates the CUDA context of `device` (a second or two on a large part) and nothing else.  Optional: a host
 * program calls it from a side thread at start-up so that the cost overlaps its own index / GTF loading
 * (src/circminer.cpp:125-260) instead of landing in the first cm_gpu_create. */
int cm_gpu_warmup(int device);
/* number of CUDA devices visible to the process (0 when there is none): how many contexts a multi-GPU host creates, the
 * role of `-t threadCount` for the reference's mapping threads (src/circminer.cpp:290-297) */
int cm_gpu_device_count(void);
int cm_gpu_create_shared(cm_gpu_ctx *parent, cm_gpu_ctx **out);
int cm_gpu_destroy(cm_gpu_ctx *ctx);
const char *cm_gpu_last_error(const cm_gpu_ctx *ctx);   /* ctx may be NULL: last create error */
/* page-locked host memory for the batch buffers of the callers below (the reference allocates its per-thread
 * scratch with malloc, src/circminer.cpp:163-191; pinned buffers let the H2D / D2H copies of a batch run at
 * full PCIe speed and overlap kernels).  Any host pointer is accepted by the batch calls; these are optional. */
int cm_gpu_host_alloc(size_t bytes, void **out);
int cm_gpu_host_free(void *p);
int cm_gpu_get_counters(cm_gpu_ctx *ctx, cm_counters *out, int reset);
/* device time (ms, CUDA events on the context's stream) of the stages of the last
 * batch call: [0] prep, [1] seed, [2] chain, [3] extend+classify, [4] total. */
int cm_gpu_last_stage_ms(cm_gpu_ctx *ctx, float out[5]);
/* device time (ms) of the dense alignment kernels of the last chain stage: [0] k_chain_spec (middle edit
 * distance + X-drop of both read ends on the genome text, src/align.cpp:166-723), [1] k_chain_walk
 * (the same DPs along transcripts, src/extend.cpp:37-920); with cm_counters.xdrop_cells + edit_cells
 * this gives the GCUPS of the alignment stage. */
int cm_gpu_last_align_ms(cm_gpu_ctx *ctx, float out[2]);
/* INT32 roofline denominator measured on this device: lane-operations per second of an add + max chain kernel
 * ("calibrate with an IADD3/IMNMX micro-benchmark on the box and use the lower of the two"). */
int cm_gpu_int32_peak(cm_gpu_ctx *ctx, double *lane_ops_per_s);

/* ---- index / genome --------------------------------------------------- */
/* Upload one packed contig straight from the reference's in-memory index right
 * after loadHashTable() (src/mrsfast/HashTable.c:971-1057):
 *   bucket_lists = _ih_hashTable   (IHashTable[4^14], 8-byte pointers or NULL)
 *   table_mem    = _ih_hashTableMem (6-byte GeneralIndex; list[0].info = count)
 *   cref_gen     = _ih_crefGen     (3 bits / base, 21 per word, Common.c:174-196)
 * Replaces getCandidates (HashTable.c:1093) and GenomeSeeder::pac2char
 * (src/match_read.cpp:288-299) as data sources. contig_id is 0-based (contigNum). */
int cm_gpu_load_contig(cm_gpu_ctx *ctx, int contig_id, const void *const *bucket_lists, const void *table_mem,
                       uint64_t table_mem_entries, const uint64_t *cref_gen, uint32_t ref_gen_len);
/* Same index built on the device from the packed contig's bases ('A','C','G','T',
 * anything else = N): count -> scan -> scatter -> sort by (checksum, position),
 * the order src/mrsfast/Sort.c:110-176 produces (HashTable.c:769-839). */
int cm_gpu_build_contig(cm_gpu_ctx *ctx, int contig_id, const char *contig_seq, uint32_t ref_gen_len);
/* The same device-side build from the reference's 3-bit packed contig (_ih_crefGen / _ih_refGenLen,
 * lib/mrsfast/RefGenome.c; 21 bases per 64-bit word, Common.c:174-196): what a host that has just run
 * loadHashTable() holds, without converting and uploading its bucket table. */
int cm_gpu_build_contig_packed(cm_gpu_ctx *ctx, int contig_id, const uint64_t *cref_gen, uint32_t ref_gen_len);
/* The reference's on-disk index, <ref>.packed.fa.index, without the reference's builder and loader:
 *   cm_gpu_write_index      writes the file `circminer --index` would write (generateHashTableOnDisk / generateHashTable,
 *                           src/mrsfast/HashTable.c:257-481; format :106-254) from resident contigs 0 .. n_contigs - 1
 *                           (names "1", "2", ... as GenomePacker::pack_genome numbers them, src/genome.cpp:96-145);
 *                           compact != 0: counts only (magic 2, `-m`).  Every packed contig must fit one piece
 *                           (non-N bases <= CONTIG_SIZE), as GenomePacker guarantees for chromosomes <= 1.1 Gbp.
 *   cm_gpu_load_index_file  reads the header and every contig's packed text from such a file (the bucket counts and
 *                           the six-byte table are skipped, not read) and builds the device index of each, ids
 *                           first_id, first_id + 1, ...: what loadHashTable (HashTable.c:971-1057) + cm_gpu_build_contig_packed
 *                           do, at 0.4 bytes of file I/O per base instead of ~8. */
int cm_gpu_write_index(cm_gpu_ctx *ctx, const char *path, int n_contigs, int compact, uint32_t contig_max_size);
int cm_gpu_load_index_file(cm_gpu_ctx *ctx, const char *path, int first_id, int *n_loaded);
int cm_gpu_contig_stats(cm_gpu_ctx *ctx, int contig_id, uint64_t *n_entries, uint32_t *ref_gen_len);
/* debugging / parity: copy the positions of bucket hv (all checksums) back */
int cm_gpu_read_bucket(cm_gpu_ctx *ctx, int contig_id, uint32_t hv, uint32_t cap, uint16_t *checksum,
                       uint32_t *pos, uint32_t *n);
/* after GTFParser::load_gtf (src/gene_annotation.cpp:191) */
int cm_gpu_load_annotation(cm_gpu_ctx *ctx, int contig_id, const cm_annot_flat *annot);

/* The flattened annotation as a build-once artefact: GTFParser::load_gtf (src/gene_annotation.cpp:191-399)
 * rebuilds its tables and sets the 1.3-Gbit masks bit by bit in every run; a host that has flattened them once writes one
 * file per packed contig and later runs upload it as it is.  `tag` is the caller's validity stamp (e.g. GTF size ^ mtime):
 * cm_gpu_load_annotation_file refuses a file with another tag or another max_read_len (near_border_bs depends on -l,
 * src/gene_annotation.cpp:273-278).  cm_annot_file_* are host-only (no device needed). */
int cm_annot_file_write(const char *path, const cm_annot_flat *annot, uint32_t max_read_len, uint64_t tag);
int cm_annot_file_read(const char *path, cm_annot_flat *out, uint32_t *max_read_len, uint64_t *tag, void **storage);
void cm_annot_file_free(void *storage);
int cm_gpu_load_annotation_file(cm_gpu_ctx *ctx, int contig_id, const char *path, uint64_t expect_tag);

/* ---- stage-level batch entries (parity tests) -------------------------- */
/* Reads are ASCII, n strands, fixed row stride `stride` bytes, length len[i].
 * The device derives the reverse complement exactly as
 * FASTQParser::set_reverse_comp does (src/fastq_parser.cpp:141-153). */

/* GenomeSeeder::split_match_hash (src/match_read.cpp:270; get_exact_locs_hash :54).
 * Outputs are [n][max_seg_cnt] with max_seg_cnt = 2*ceil(max_read_len/kmer)-1:
 *   qpos, frag_count (0 when > seed_lim), has_frags (frags != NULL),
 *   first/last/sum of frags[].info over the kept run. */
int cm_gpu_seed_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride,
                      const uint16_t *len, int32_t *qpos, uint32_t *frag_count, uint8_t *has_frags,
                      uint32_t *first_info, uint32_t *last_info, uint64_t *sum_info);

/* get_best_chains = split_match_hash + chain_seeds_sorted_kbest (src/filter.cpp:469,
 * src/chain.cpp:73).  Outputs: chain_count[n], high_hits[n], and per strand
 * max_chain_len slots: score, chain_len, frags (rpos, qpos) [n][C][F],
 * F = (max_seg_cnt+1)/2; fragment len is always kmer. */
int cm_gpu_chain_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride,
                       const uint16_t *len, int32_t *chain_count, int32_t *high_hits, float *score,
                       uint32_t *chain_len, uint32_t *frag_rpos, int32_t *frag_qpos);

/* The annotation look-ups of the hot loops, one position per row (contig-relative, 1-based), for parity with GTFParser:
 *   ub, max_end, ol_iv   GTFParser::get_upper_bound(pos, mlen, rlen, max_end, ol_exons) (src/gene_annotation.h:121-131 ->
 *                        get_upper_bound_lookup src/gene_annotation.cpp:464-533); ol_iv = index of ol_exons in
 *                        FlatIntervalTree::disjoint_intervals, -1 for NULL
 *   loc_iv, loc_ind      get_location_overlap_ind(pos, false, ind) (src/gene_annotation.cpp:555-568)
 *   shift_idx            index into con2chr[contig] of get_shift(contig, pos) (:451-457)
 *   intronic             intronic_bs[contig][pos] (src/common.h:118) */
int cm_gpu_annot_lookup_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const uint32_t *pos, uint32_t mlen, uint32_t rlen,
                              uint32_t *ub, uint32_t *max_end, int32_t *ol_iv, int32_t *loc_iv, int32_t *loc_ind,
                              int32_t *shift_idx, uint8_t *intronic);

/* POD image of the MatchedMate fields the extension writes (src/common.h:260-307) */
typedef struct cm_mate {
    uint32_t spos, epos, qspos, qepos, matched_len;
    int32_t left_ed, middle_ed, right_ed, sclen_left, sclen_right, dir, type;
    uint8_t is_concord, left_ok, right_ok, pad_;
} cm_mate;

/* The extension stage, one problem per row (src/extend.h:66-70):
 *   mode 0: bool TransExtension::extend_both_mates(lch, rch, common_tid, lseq, rseq, 1, 1, lseq_len, rseq_len, lmm, rmm)
 *           (src/extend.cpp:37-125); common_tid = same_transcript(interval iv_a, interval iv_b) (src/utils.cpp:335-354), empty when
 *           iv_a < 0; ret = the return value
 *   mode 1: int extend_chain_both_sides(lch, lseq, lseq_len, lmm, dir) (:131-213); ret = the returned type; the r* arguments,
 *           iv_a, iv_b and rmm are unused (may be NULL)
 * Chains are given as chain_len + fragments (rpos, qpos) [n][frag_slots] (fragment length is always kmer); the mates start
 * from MatchedMate() with the never-initialised fields at 0. */
int cm_gpu_extend_batch(cm_gpu_ctx *ctx, int contig_id, int mode, uint32_t n, const char *lseq, const char *rseq, uint32_t stride,
                        const uint16_t *llen, const uint16_t *rlen, uint32_t frag_slots, const uint32_t *l_chain_len,
                        const uint32_t *l_rpos, const int32_t *l_qpos, const uint32_t *r_chain_len, const uint32_t *r_rpos,
                        const int32_t *r_qpos, const int32_t *iv_a, const int32_t *iv_b, const int32_t *dir, cm_mate *lmm,
                        cm_mate *rmm, int32_t *ret);

/* Stage 2, first piece: the split-read re-seeding and chaining of ProcessCirc::chaining
 * (src/process_circ.cpp:678-737) -- RegionalHashTable look-ups of every `step`-th window_size-mer of seq[qspos-1 .. qepos-1]
 * inside one gene (src/hash_table.cpp:58-112: positions relative to the gene start, a k-mer with more than 1000 hits in
 * the gene has none; more than seedLim hits: none), chain_seeds_sorted_kbest2 (src/chain.cpp:310-539) and the cut at the
 * first chain that misses more k-mers than its predecessor.  One problem per row: the read as ProcessCirc passes it
 * (remain_seq, already oriented), the 1-based inclusive range [qspos, qepos] of its unmapped part, and the gene
 * [gene_start, gene_end] in contig coordinates (GeneInfo::start / end of gtf_parser.get_gene_overlap).
 * Outputs as cm_gpu_chain_batch: chain_count[n] (bc.best_chain_count after the cut), kmer_cnt[n] (rows without N; may be
 * NULL), score / chain_len [n][max_chain_len], fragments (rpos in contig coordinates = shift + info, qpos) [n][max_chain_len][F];
 * F >= (max_read_len - window_size) / step + 1, at most 128; fragment len is always window_size.
 * cm_gpu_build_circ_index builds the per-contig window_size-mer index the look-ups run on (4 bytes per base, once per
 * contig, on the root context; cm_gpu_circ_chain_batch builds it on first use when called on a root context). */
int cm_gpu_build_circ_index(cm_gpu_ctx *ctx, int contig_id, int window_size);
int cm_gpu_circ_chain_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride, const uint32_t *qspos,
                            const uint32_t *qepos, const uint32_t *gene_start, const uint32_t *gene_end, int window_size, int step,
                            int F, int32_t *chain_count, int32_t *kmer_cnt, float *score, uint32_t *chain_len,
                            uint32_t *frag_rpos, int32_t *frag_qpos);

/* Stage 2, second piece: ProcessCirc::find_exact_coord (src/process_circ.cpp:739-789) per row -- set_mm, the
 * three-mate same_transcript (src/utils.cpp:447-494), calc_middle_ed and extend_chain_left / _right of stage 2's
 * TransExtension(0, EDIT_ALIGNMENT), whose soft-clipping aligner is EditDistAlignment (src/align.cpp:602-667), and
 * update_match_mate_info.  Inputs: the read as ProcessCirc passes it (remain_seq), the two mates of the pair by their contig
 * coordinates (all the function reads of them), dir, qspos (1-based), rlen = length of the unmapped part, whole_len, and one
 * chain of cm_gpu_circ_chain_batch (chain_len + fragments [n][frag_slots], fragment length window_size).
 * Outputs: partial_mm (from MatchedMate() with the never-initialised fields at 0) and the return value. */
int cm_gpu_circ_exact_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride, const uint32_t *r1_spos,
                            const uint32_t *r1_epos, const uint32_t *r2_spos, const uint32_t *r2_epos, const int32_t *dir,
                            const uint32_t *qspos, const int32_t *rlen, const int32_t *whole_len, uint32_t frag_slots,
                            const uint32_t *chain_len, const uint32_t *frag_rpos, const int32_t *frag_qpos, int window_size,
                            cm_mate *pmm, int32_t *ret);

/* MatchedMate as stage 2's classification reads and writes it: the cm_mate fields plus the cached exon look-ups
 * (looked_up_spos / _epos, exons_spos / _epos as indices into FlatIntervalTree::disjoint_intervals or -1, exon_ind_spos / _epos),
 * which stay as they are when final_check trims the positions (src/process_circ.cpp:1158-1177). */
typedef struct cm_circ_mate {
    cm_mate m;
    int32_t exons_spos, exons_epos, exon_ind_spos, exon_ind_epos;
    uint8_t looked_up_spos, looked_up_epos, pad_[2];
} cm_circ_mate;

/* CircRes::set_bp (src/common.cpp:468-477): breakpoints, the two signals (0 or 2 characters) and the reference bases */
typedef struct cm_circ_res {
    uint32_t spos, epos;
    uint8_t start_signal[2], end_signal[2], start_bp_ref[2], end_bp_ref[2];
    uint8_t start_signal_n, end_signal_n, ref_n, was_set;
} cm_circ_res;

/* Stage 2, third piece: ProcessCirc::check_split_map(mm_r1, mm_r2, partial_mm, r1_partial, cr) (src/process_circ.cpp:892-919)
 * with everything under it -- final_check (:1135-1341), both split_realignment overloads (:1343-1489),
 * rescue_overlapping_bsj (:1491-1552), the four-mate check_split_map (:922-1133), get_junctions (src/utils.cpp:697-757).
 * One problem per row: the three mates (in / out: final_check trims the full mate and fills look-up caches), r1_partial,
 * remain_seq and fullmap_seq (rows of `stride` bytes) with their lengths.  Outputs: the return value (FR 0, RF 1, CR 20,
 * NCR 21, MCR 22, UD 30; -1 when a fixed-size list of the device code overflowed and the row must be redone by the caller)
 * and cr (was_set = 0 when the reference would not have called set_bp). */
int cm_gpu_circ_check_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *remain_seq, const char *fullmap_seq, uint32_t stride,
                            const int32_t *remain_len, const int32_t *fullmap_len, const uint8_t *r1_partial, cm_circ_mate *mm_r1,
                            cm_circ_mate *mm_r2, cm_circ_mate *partial_mm, int window_size, int32_t *ret, cm_circ_res *cr);

/* The four-mate ProcessCirc::check_split_map(mm_r1_1, mm_r2_1, mm_r1_2, mm_r2_2, cr) (src/process_circ.cpp:922-1133) on its own, one
 * problem per row: the two oriented reads (ProcessCirc::r1_seq / r2_seq) and the four mates (read only).  Its two single-BSJ arms
 * end in final_check, which reads ProcessCirc::fullmap_seq -- in the reference whatever an earlier call left there; here
 * fullmap_sel[i] says what it holds (0 nothing, 1 = R1, 2 = R2; NULL = 0 for every row, as call_circ_double_split's own calls).
 * ret / cr as cm_gpu_circ_check_batch. */
int cm_gpu_circ_check4_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *r1_seq, const char *r2_seq, uint32_t stride,
                             const int32_t *r1_len, const int32_t *r2_len, const uint8_t *fullmap_sel, const cm_circ_mate *mm_r1_1,
                             const cm_circ_mate *mm_r2_1, const cm_circ_mate *mm_r1_2, const cm_circ_mate *mm_r2_2, int window_size,
                             int32_t *ret, cm_circ_res *cr);

/* find_exact_coord as cm_gpu_circ_exact_batch, with the two mates of the pair going in and out as ProcessCirc keeps them from one
 * chain of a read to the next (their cached look-ups survive a trim by final_check) and partial_mm coming back with the look-ups
 * find_exact_coord leaves on it -- what check_split_map is then called with. */
int cm_gpu_circ_exact_mates_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride, cm_circ_mate *mm_r1,
                                  cm_circ_mate *mm_r2, const int32_t *dir, const uint32_t *qspos, const int32_t *rlen, const int32_t *whole_len,
                                  uint32_t frag_slots, const uint32_t *chain_len, const uint32_t *frag_rpos, const int32_t *frag_qpos,
                                  int window_size, cm_circ_mate *partial_mm, int32_t *ret);

/* ProcessCirc::call_circ_single_split (src/process_circ.cpp:360-483) for a batch of pairs: for every pair whose MatchedRead is a
 * CHIBSJ, per gene of gtf_parser.get_gene_overlap(mm_r1.spos, false)->seg_list (given by the caller in that order: gene_off[n + 1],
 * gene_start / gene_end), chaining, then chain by chain (TOPCHAIN = 10) find_exact_coord and, when it is CONCRD, check_split_map,
 * until FR / RF or CR.  mr: stage 1's MatchedRead with its positions in CONTIG coordinates (after GTFParser::chrloc2conloc).
 * Outputs: every check_split_map call in the reference's order (pair, gene, chain) -- the pair, the mates before and after,
 * the type it returned and its CircRes: what print_split_mapping and the type column of candidates.pam are written from --
 * n_attempts (the first attempt_cap are stored; the att_* arrays may be NULL), and per pair best_type (what the function ends
 * with: 40 = nothing, 0 / 1 = FR / RF, 20..22 = the CircRes pushed to circ_res, in best_cr, contig coordinates; -1 = not a
 * candidate; -2 = a check_split_map of this pair needed more than a fixed-size list of the device code holds (CM_CIRC_TID_CAP
 * transcripts at one breakpoint, CM_CIRC_JUNC_CAP junctions): its attempts are not to be used and the caller redoes the pair with
 * ProcessCirc::call_circ_single_split itself). */
int cm_gpu_circ_single_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride,
                             const uint16_t *len1, const uint16_t *len2, const cm_mread *mr, const uint32_t *gene_off, const uint32_t *gene_start,
                             const uint32_t *gene_end, int window_size, int step, uint32_t attempt_cap, uint32_t *n_attempts, uint32_t *att_pair,
                             uint8_t *att_r1_partial, cm_circ_mate *att_in_r1, cm_circ_mate *att_in_r2, cm_circ_mate *att_in_pm, int32_t *att_ret,
                             cm_circ_res *att_cr, cm_circ_mate *att_out_r1, cm_circ_mate *att_out_r2, cm_circ_mate *att_out_pm, int32_t *best_type,
                             cm_circ_res *best_cr);

/* ProcessCirc::call_circ_double_split (src/process_circ.cpp:485-645) for a batch of CHI2BSJ pairs (inputs as cm_gpu_circ_single_batch).
 * Per pair, in the function's order, the events that produce output: kind 0 = "call_circ_single_split(record1, record2) is called
 * here" (before the gene loop when one unmapped part is shorter than window_size, for a gene where only one part chains, at the
 * end when nothing was pushed; what that call writes is cm_gpu_circ_single_batch's answer for the pair -- it accepts CHI2BSJ
 * pairs too -- and is the same every time), kind 1 = a four-mate check_split_map call with mm_r1, mm_r2, r1_partial_mm,
 * r2_partial_mm (what print_split_mapping prints), its type and CircRes.  push_type[i] = 20..22 when the function pushes its own
 * best_cr (push_cr[i], contig coordinates; con_shift from mm_r1.spos = mr.spos_r1), -2 when a four-mate check_split_map the
 * function reaches overflowed a fixed-size list (as best_type -2 above: the pair's events are not to be used), else -1. */
int cm_gpu_circ_double_batch(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride,
                             const uint16_t *len1, const uint16_t *len2, const cm_mread *mr, const uint32_t *gene_off, const uint32_t *gene_start,
                             const uint32_t *gene_end, int window_size, int step, uint32_t event_cap, uint32_t *n_events, uint32_t *ev_pair,
                             uint8_t *ev_kind, cm_circ_mate *ev_r1, cm_circ_mate *ev_r2, cm_circ_mate *ev_p1, cm_circ_mate *ev_p2, int32_t *ev_ret,
                             cm_circ_res *ev_cr, int32_t *push_type, cm_circ_res *push_cr);

/* Alignment kernels, one problem per row (src/align.h:165-176):
 *   kind 0: DropAlignment::local_alignment_right_sc  (src/align.cpp:669)
 *   kind 1: DropAlignment::local_alignment_left_sc   (:694)
 *   kind 2: Alignment::local_alignment_right         (:556)
 *   kind 3: Alignment::local_alignment_left          (:578)
 *   kind 4: Alignment::global_one_side_banded_alignment (:219), w in aux[i]
 * s/t are ASCII rows (stride bytes); out_ed, out_sclen, out_indel, out_score as the
 * reference returns them (unused ones are 0). */
int cm_gpu_align_batch(cm_gpu_ctx *ctx, int kind, uint32_t n, const char *s, const uint16_t *n_s, const char *t,
                       const uint16_t *n_t, uint32_t stride, const int32_t *aux, int32_t *out_ed,
                       int32_t *out_sclen, int32_t *out_indel, int32_t *out_score);

/* cm_gpu_align_batch for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments
 * and results as cm_gpu_align_batch; inputs and results are staged in one fixed device array of
 * cm_gpu_noalloc_scratch_bytes() = 4 MiB that every GPU loading the library holds until exit.  The call makes no
 * cudaMalloc / cudaFree; a batch that needs more than the array (about 2 * n * stride + 24 * n bytes) is refused with
 * CM_ERR_ARG.  All calls on a device share the array: calls that overlap on the GPU race, the caller serialises them. */
int cm_gpu_align_batch_noalloc(cm_gpu_ctx *ctx, int kind, uint32_t n, const char *s, const uint16_t *n_s, const char *t,
                               const uint16_t *n_t, uint32_t stride, const int32_t *aux, int32_t *out_ed,
                               int32_t *out_sclen, int32_t *out_indel, int32_t *out_score);
size_t cm_gpu_noalloc_scratch_bytes(void);

/* ---- the fused hot path ------------------------------------------------ */
/* FilterRead::process_read, PE overload (src/filter.cpp:124-241) for n pairs.
 * seqs1/seqs2: mate rows; in[i] = the pair's carried MatchedRead (what
 * FASTQParser::fill_map_info reconstructed, src/fastq_parser.cpp:203-269);
 * out[i] = updated MatchedRead; state[i] = return value.  in may alias out. */
int cm_gpu_process_batch_pe(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs1, const char *seqs2,
                            uint32_t stride, const uint16_t *len1, const uint16_t *len2, const cm_mread *in,
                            cm_mread *out, int32_t *state);
/* SE overload (src/filter.cpp:86-121): only the return value is observable. */
int cm_gpu_process_batch_se(cm_gpu_ctx *ctx, int contig_id, uint32_t n, const char *seqs, uint32_t stride,
                            const uint16_t *len, int32_t *state);

/* Device-resident variant used for kernel-only timing: stage the batch once,
 * then run the pipeline any number of times without host copies. */
int cm_gpu_stage_batch_pe(cm_gpu_ctx *ctx, uint32_t n, const char *seqs1, const char *seqs2, uint32_t stride,
                          const uint16_t *len1, const uint16_t *len2, const cm_mread *in);
int cm_gpu_run_staged_pe(cm_gpu_ctx *ctx, int contig_id, int upto_stage /* 1 seed, 2 chain, 3 all */);
int cm_gpu_fetch_staged_pe(cm_gpu_ctx *ctx, cm_mread *out, int32_t *state);

/* ---- rounds --------------------------------------------------------------- */
/* The reference's round loop for one batch with every packed contig resident (mapping(), src/circminer.cpp:229-308):
 * contigs first_contig .. first_contig + n_contigs - 1 in order.  After each contig the skip rule of map_reads
 * (src/circminer.cpp:386-390: scan-lev 0 -> a CONCRD pair is final; scan-lev 1 -> CONCRD, gm-compatible, no edits,
 * fully matched; scan-lev 2 -> never) picks the pairs the reference prints in that round; every other pair goes on to
 * the next contig with the MatchedRead the reference would read back from the remain-FASTQ header
 * (FilterRead::write_read_category src/filter.cpp:413-455 -> FASTQParser::fill_map_info src/fastq_parser.cpp:203-269:
 * unmapped types keep only `type`).  Survivors are compacted on the device; nothing returns to the host between contigs.
 *   out[i], state[i]   MatchedRead / process_read return of pair i in the round it was printed (skip, or last contig)
 *   final_round[i]     that contig (nullable)
 *   round_out, round_state   optional [n_contigs][n] snapshots after every contig, state -1 where the pair was already
 *                      final (what the per-round remain files of a `-z` run hold); for parity tests. */
int cm_gpu_map_rounds_pe(cm_gpu_ctx *ctx, int first_contig, int n_contigs, uint32_t n, const char *seqs1, const char *seqs2,
                         uint32_t stride, const uint16_t *len1, const uint16_t *len2, const cm_mread *in, cm_mread *out,
                         int32_t *state, int32_t *final_round, cm_mread *round_out, int32_t *round_state);
/* the same on the batch staged by cm_gpu_stage_batch_pe (kernel-only timing; the staged batch is left intact) */
int cm_gpu_run_staged_rounds_pe(cm_gpu_ctx *ctx, int first_contig, int n_contigs);
int cm_gpu_fetch_rounds_pe(cm_gpu_ctx *ctx, cm_mread *out, int32_t *state, int32_t *final_round);
/* device ms of the last rounds call: [0] unused, [1] seed, [2] chain, [3] extend+classify, [4] sum of the contigs' stage
 * totals, [5] the whole call on the stream (commit / compaction included), [6] k_chain_spec, [7] k_chain_walk;
 * pairs_per_round[r] = pairs that visited the r-th contig (cap entries).  Returns the number of contigs visited. */
int cm_gpu_last_rounds_ms(cm_gpu_ctx *ctx, float out[8], uint32_t *pairs_per_round, int cap);

#ifdef __cplusplus
}
#endif
#endif /* CM_GPU_H */
