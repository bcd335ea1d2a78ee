// Seed lookup: hashVal / checkSumVal (lib/mrsfast/Common.c:199-225), getCandidates
// (src/mrsfast/HashTable.c:1093-1098) and GenomeSeeder::get_exact_locs_hash
// (src/match_read.cpp:54-110) over the SoA index of ContigDev.
#pragma once
#include "cm_types.h"

// nHVal (Common.c:241-245): upper-case ACGT only, everything else -1
CM_HD int cm_nhval(uint8_t c) {
    switch (c) {
        case 'A': return 0;
        case 'C': return 1;
        case 'G': return 2;
        case 'T': return 3;
        default: return -1;
    }
}

// FASTQParser::set_comp (src/fastq_parser.cpp:141-153); unknown bytes map to 0
CM_HD uint8_t cm_comp(uint8_t c) {
    switch (c) {
        case 'A': case 'a': return 'T';
        case 'C': case 'c': return 'G';
        case 'G': case 'g': return 'C';
        case 'T': case 't': return 'A';
        case 'N': case 'n': return 'N';
        default: return 0;
    }
}

struct SeedHit {
    uint32_t off;      // first matching entry in ContigDev::pos / cks
    uint32_t cnt;      // raw size of the equal range (0: frags == NULL)
};

// hv = first 14 bases, cv = next kmer-14 bases, -1 on any non-ACGT
CM_HD void cm_hash_kmer(const uint8_t *seq, int kmer, int &hv, int &cv) {
    hv = 0;
    cv = 0;
    bool bad_h = false, bad_c = false;
    for (int i = 0; i < CM_WINDOW; i++) {
        int v = cm_nhval(seq[i]);
        if (v < 0) bad_h = true;
        hv = (hv << 2) | (v & 3);
    }
    for (int i = CM_WINDOW; i < kmer; i++) {
        int v = cm_nhval(seq[i]);
        if (v < 0) bad_c = true;
        cv = (cv << 2) | (v & 3);
    }
    if (bad_h) hv = -1;
    if (bad_c) cv = -1;
}

// get_exact_locs_hash.  `probes` counts checksum words touched (information bytes).
CM_HD SeedHit cm_seed_lookup(const ContigDev &c, int hv, int cv, uint32_t *sectors) {
    SeedHit h;
    h.off = 0;
    h.cnt = 0;
    if (sectors) *sectors = 0;
    if (hv < 0 || cv < 0) return h;
    uint32_t lo = c.dir[hv], hi = c.dir[hv + 1];
    if (sectors) *sectors = 1;                 // directory sector
    if (hi == lo) return h;                    // getCandidates == NULL
    if (sectors) *sectors = 2;                 // bucket head
    // `int16_t target = cv` (match_read.cpp:77): checksums >= 32768 never compare equal
    if (cv >= 32768) return h;
    uint16_t target = (uint16_t) cv;
    uint32_t a = lo, b = hi;                   // lower bound in [lo, hi)
    while (a < b) {
        uint32_t mid = (a + b) >> 1;
        if (c.cks[mid] < target) a = mid + 1; else b = mid;
    }
    if (a >= hi || c.cks[a] != target) return h;
    uint32_t LB = a;
    b = hi;                                    // upper bound in [LB, hi)
    while (a < b) {
        uint32_t mid = (a + b) >> 1;
        if (c.cks[mid] <= target) a = mid + 1; else b = mid;
    }
    h.off = LB;
    h.cnt = a - LB;
    return h;
}

// The same look-up through the fused directory (ContigDev::dirx).  Equal range of `target` in the sorted bucket:
// positions < 4 are answered from the inline checksums, the rest by bisection of cks[off + 4, off + cnt).
CM_HD SeedHit cm_seed_lookup_x(const ContigDev &c, int hv, int cv, uint32_t *sectors) {
    SeedHit h;
    h.off = 0;
    h.cnt = 0;
    if (sectors) *sectors = 0;
    if (hv < 0 || cv < 0) return h;
#if defined(__CUDA_ARCH__)
    const uint4 *dp = reinterpret_cast<const uint4 *>(c.dirx) + hv;
    uint4 raw;
    switch ((CM_DBG_FLAGS >> 8) & 7) {   // load-path A/B (CM_GPU_DBG = 256 * variant); measured equal in DRAM bytes, 0 is 4 % faster
        default: asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(raw.x), "=r"(raw.y), "=r"(raw.z), "=r"(raw.w) : "l"(dp)); break;
        case 1: raw = *dp; break;
        case 2: raw = __ldcg(dp); break;
        case 3: raw = __ldg(dp); break;
    }
    const uint32_t off = raw.x, cnt = raw.y;
    const uint16_t c0 = (uint16_t) (raw.z & 0xFFFFu), c1 = (uint16_t) (raw.z >> 16), c2 = (uint16_t) (raw.w & 0xFFFFu), c3 = (uint16_t) (raw.w >> 16);
#else
    const DirX e = c.dirx[hv];
    const uint32_t off = e.off, cnt = e.cnt;
    const uint16_t c0 = e.c[0], c1 = e.c[1], c2 = e.c[2], c3 = e.c[3];
#endif
    if (sectors) *sectors = 1;                 // directory sector
    if (cnt == 0) return h;                    // getCandidates == NULL
    if (sectors) *sectors = 2;                 // bucket head
    if (cv >= 32768) return h;                 // `int16_t target = cv` (match_read.cpp:77)
    const uint16_t target = (uint16_t) cv;
    const uint32_t ni = cnt < CM_DIRX_INLINE ? cnt : CM_DIRX_INLINE;
    // inline part: entries 0..ni-1 (unused slots are never counted)
    const uint32_t lt = (uint32_t) (c0 < target) + (uint32_t) (ni > 1 && c1 < target) + (uint32_t) (ni > 2 && c2 < target) +
                        (uint32_t) (ni > 3 && c3 < target);
    const uint32_t le = (uint32_t) (c0 <= target) + (uint32_t) (ni > 1 && c1 <= target) + (uint32_t) (ni > 2 && c2 <= target) +
                        (uint32_t) (ni > 3 && c3 <= target);
    uint32_t LB = lt, UB = le;                 // lower / upper bound as indices into the bucket
    if (cnt > CM_DIRX_INLINE && le == CM_DIRX_INLINE) {
        // the range may continue (or start) past the inline entries
        const uint32_t lo0 = off + CM_DIRX_INLINE, hi0 = off + cnt;
        uint32_t a = lo0, b = hi0;
        if (lt == CM_DIRX_INLINE) {            // all four inline entries are smaller: lower bound is in memory too
            while (a < b) {
                uint32_t mid = (a + b) >> 1;
                if (c.cks[mid] < target) a = mid + 1; else b = mid;
            }
            LB = a - off;
            b = hi0;
        }
        while (a < b) {
            uint32_t mid = (a + b) >> 1;
            if (c.cks[mid] <= target) a = mid + 1; else b = mid;
        }
        UB = a - off;
    }
    if (UB <= LB) return h;
    h.off = off + LB;
    h.cnt = UB - LB;
    return h;
}

// number of seeds kmer_match_skip_hash visits (match_read.cpp:201-212): q = 0, k, 2k, ... while len - q >= k
CM_HD int cm_seed_count(int len, int kmer) {
    return len < kmer ? 0 : (len - kmer) / kmer + 1;
}
