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
// positions < CM_DIRX_INLINE are answered from the inline checksums, the rest by bisection of the bucket's tail; a large
// bucket goes through its coarse checksum table first.
// the directory entry of bucket hv, as words (off, cnt, checksums two per word); all zero for hv < 0
#define CM_DIRX_WORDS (sizeof(DirX) / 4)
CM_HD void cm_dirx_load(const ContigDev &c, int hv, uint32_t *w) {
    constexpr unsigned NW = CM_DIRX_WORDS;
    if (hv < 0) {
        for (unsigned k = 0; k < NW; k++) w[k] = 0;
        return;
    }
#if defined(__CUDA_ARCH__)
    const uint4 *dp = reinterpret_cast<const uint4 *>(c.dirx + hv);
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "l"(dp));
    if (NW == 8)
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[4 % NW]), "=r"(w[5 % NW]), "=r"(w[6 % NW]), "=r"(w[7 % NW]) : "l"(dp + 1));
#else
    const DirX &e = c.dirx[hv];
    w[0] = e.off; w[1] = e.cnt;
    for (int k = 0; k < CM_DIRX_INLINE / 2; k++) w[2 + k] = (uint32_t) e.c[2 * k] | ((uint32_t) e.c[2 * k + 1] << 16);
#endif
}

CM_HD SeedHit cm_seed_resolve_x(const ContigDev &c, const uint32_t *w, int hv, int cv, uint32_t *sectors);

CM_HD SeedHit cm_seed_lookup_x(const ContigDev &c, int hv, int cv, uint32_t *sectors) {
    uint32_t w[CM_DIRX_WORDS];
    cm_dirx_load(c, (hv < 0 || cv < 0) ? -1 : hv, w);
    return cm_seed_resolve_x(c, w, hv, cv, sectors);
}

// the look-up given the loaded directory entry (split off so that a thread can have several entries in flight)
CM_HD SeedHit cm_seed_resolve_x(const ContigDev &c, const uint32_t *w, int hv, int cv, uint32_t *sectors) {
    SeedHit h;
    h.off = 0;
    h.cnt = 0;
    if (sectors) *sectors = 0;
    if (hv < 0 || cv < 0) return h;
    const uint32_t off = w[0], cnt = w[1];
    if (sectors) *sectors = 1;                 // directory sector
    if (cnt == 0) return h;                    // getCandidates == NULL
    if (sectors) *sectors = 2;                 // bucket head
    if (cv >= 32768) return h;                 // `int16_t target = cv` (match_read.cpp:77)
    const uint32_t target = (uint32_t) cv;
    uint32_t LB, UB;                           // lower / upper bound as indices into the bucket
    if (cnt >= c.big_th && c.big_tab) {
        // large bucket: w[2] = its row of the coarse table; the equal range lies inside one table cell
        const uint32_t *row = c.big_tab + (size_t) w[2] * ((1u << c.big_bits) + 1u);
        const uint32_t t = target >> c.big_shift;
        uint32_t a = off + row[t], b = off + row[t + 1];
        const uint32_t hi0 = b;
        while (a < b) {
            uint32_t mid = (a + b) >> 1;
            if (c.cks[mid] < target) a = mid + 1; else b = mid;
        }
        LB = a - off;
        b = hi0;
        while (a < b) {
            uint32_t mid = (a + b) >> 1;
            if (c.cks[mid] <= target) a = mid + 1; else b = mid;
        }
        UB = a - off;
    } else {
        const uint32_t ni = cnt < CM_DIRX_INLINE ? cnt : CM_DIRX_INLINE;
        uint32_t lt = 0, le = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (uint32_t k = 0; k < CM_DIRX_INLINE; k++) {   // unused slots are never counted
            const uint32_t ck = (w[2 + (k >> 1)] >> (16 * (k & 1))) & 0xFFFFu;
            lt += (uint32_t) (k < ni && ck < target);
            le += (uint32_t) (k < ni && ck <= target);
        }
        LB = lt;
        UB = le;
        if (cnt > CM_DIRX_INLINE && le == CM_DIRX_INLINE) {
            // the range may continue (or start) past the inline entries
            const uint32_t lo0 = off + CM_DIRX_INLINE, hi0 = off + cnt;
            uint32_t a = lo0, b = hi0;
            if (lt == CM_DIRX_INLINE) {        // every inline entry is smaller: the lower bound is in memory too
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
