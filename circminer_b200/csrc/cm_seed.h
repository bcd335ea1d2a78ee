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

// number of seeds kmer_match_skip_hash visits (match_read.cpp:201-212): q = 0, k, 2k, ... while len - q >= k
CM_HD int cm_seed_count(int len, int kmer) {
    return len < kmer ? 0 : (len - kmer) / kmer + 1;
}
