// Shared POD types of the device hot path.  Everything in cm_*.h is written as plain
// C++ over raw pointers and marked CM_HD so that (a) nvcc compiles it for sm_100a and
// (b) tests/hostsim can compile the very same functions with g++ to debug the per-read
// logic against the reference in the GPU-less build container.  The product library
// (cm_gpu.cu) only ever runs them on the device.
#pragma once
#include <stdint.h>
#include "../../include/cm_gpu.h"

#if defined(__CUDACC__)
#define CM_HD __host__ __device__ __forceinline__
#define CM_HDN __host__ __device__ __noinline__
#else
#define CM_HD inline
#define CM_HDN
#endif

#define CM_WINDOW 14                       // WINDOW_SIZE, src/common.cpp:7
#define CM_NBUCKETS (1u << (2 * CM_WINDOW))  // 4^14 directory entries
#define CM_INF 1000000000                  // INF 1e9, src/common.h:34
#define CM_MAXUB 4294967295u               // src/common.h:37
#define CM_MAXDISCRDTLEN 20000             // src/common.h:40
#define CM_LARIAT2BEGTH 1000               // src/common.h:53
#define CM_DPTINF 10000000                 // src/align.cpp:12
#define CM_MAX_CHAINS 30                   // BESTCHAINLIM, src/common.h:51 (max_chain_len <= this)
#define CM_MAX_SEEDS 22                    // seeds per strand we support: ceil(300/14)
#define CM_MAX_BAND 15                     // largest band_width the register DP supports

// Fused directory entry: bucket offset and size plus the bucket's first CM_DIRX_INLINE checksums in one 16-byte (or
// 32-byte) read, so a look-up in a small bucket (or whose key sorts before the last inline checksum) never touches the
// checksum array -- one random DRAM access per seed instead of the reference's two dependent ones (directory -> bucket).
// A bucket of >= ContigDev::big_th entries (repeat families: up to 10^5 entries) keeps, instead of inline checksums, the
// number of its row in ContigDev::big_tab: the first entry of every value of the checksum's top big_bits bits, which
// narrows the bisection from 17 dependent DRAM probes to a run of cnt / 2^big_bits entries.
#ifndef CM_DIRX_INLINE
#define CM_DIRX_INLINE 4    // 4: 16-byte entry, 12: 32-byte entry.  Measured on C3 (28 M seeds, contig 1): 1.70 ms against 2.11 ms --
                            // the smaller table wins although 13 % of the GRCh38 buckets overflow four inline checksums
#endif
struct DirX {
    uint32_t off, cnt;
    uint16_t c[CM_DIRX_INLINE];
};
static_assert(sizeof(DirX) == 16 || sizeof(DirX) == 32, "the directory entry is one or two 16-byte loads");
#define CM_BIG_TH_DEFAULT 96

// One packed contig resident in HBM.
struct ContigDev {
    const uint32_t *dir;     // [4^14 + 1] exclusive prefix of bucket sizes
    const DirX *dirx;        // [4^14] fused directory entry (nullptr: look-ups go through dir + cks)
    const uint32_t *big_tab; // [n_big][2^big_bits + 1] coarse checksum index of the large buckets (bucket-relative)
    uint32_t big_th;         // buckets with cnt >= big_th are "large"
    int big_bits, big_shift; // table index of a checksum = cks >> big_shift
    const uint16_t *cks;     // [n_entries] checksum of each hit, sorted by (bucket, checksum, pos)
    const uint32_t *pos;     // [n_entries] 1-based contig position of the k-mer
    const uint8_t *genome;   // [ref_len + 2] ASCII, genome[p] = base at 1-based p; genome[0] unused
    uint64_t n_entries;
    uint32_t ref_len;
};

// Flat annotation of one contig (device or host pointers), mirrors cm_annot_flat.
struct AnnotDev {
    uint32_t n_iv;
    const uint32_t *iv_spos, *iv_epos, *iv_max_end, *iv_min_end, *iv_max_next_exon, *iv_seg_off;
    // optional radix index over iv_spos (device builds): iv_lut[b] = number of intervals with spos < (b << iv_lut_shift),
    // b in [0, iv_lut_n]; narrows FlatIntervalTree::search from ~17 dependent probes to the few intervals of one bucket
    const uint32_t *iv_lut;
    uint32_t iv_lut_n, iv_lut_shift;
    uint32_t n_seg;
    const uint32_t *seg_start, *seg_end, *seg_next_exon_beg, *seg_gene_id, *seg_tid_off;
    const uint32_t *tid;
    uint32_t n_trans;
    const int32_t *trans_start_ind;
    const uint32_t *t2s_off;
    const uint8_t *t2s;
    uint32_t n_genes;
    const uint32_t *gene_start, *gene_end;
    uint32_t n_chr;
    const uint32_t *chr_shift;
    const int32_t *chr_id;
    uint64_t n_bits;
    const uint64_t *near_border, *intronic;
};

struct DevView {
    cm_params p;
    ContigDev c;
    AnnotDev a;
    int contig_id;
    int max_seg_cnt;   // 2*ceil(max_read_len/kmer)-1
    int n_slots;       // (max_seg_cnt+1)/2 : even slots only are ever filled (ll_step 2)
};

// One chained fragment: reference position + read offset (len is always kmer).
struct FragDev {
    uint32_t rpos;
    int32_t qpos;
};

// chain_t (src/common.h:150-154) with the fragments inline
struct ChainDev {
    float score;
    uint32_t chain_len;
    FragDev frags[CM_MAX_SEEDS];
};

CM_HD int cm_max_seg_cnt(const cm_params &p) {
    // 2 * ceil(1.0 * maxReadLength / kmer) - 1   (src/circminer.cpp:161)
    return 2 * ((p.max_read_len + p.kmer - 1) / p.kmer) - 1;
}
