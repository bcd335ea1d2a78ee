// Annotation look-ups over the flat per-contig tables (AnnotDev):
// FlatIntervalTree::search/find/find_ind/get_node (src/interval_tree_impl.h:135-182),
// GTFParser::get_upper_bound (src/gene_annotation.h:121-131) and get_upper_bound_lookup
// (src/gene_annotation.cpp:464-533), get_shift (:451), check_junction (src/chain.cpp:28-64).
// An IntervalInfo<UniqSeg>* of the reference is an interval index here (-1 == NULL).
#pragma once
#include "cm_types.h"

CM_HD bool cm_bit(const uint64_t *w, uint64_t n_bits, uint64_t k) {
    if (!w || k >= n_bits) return false;
    return (w[k >> 6] >> (k & 63)) & 1ull;
}

// FlatIntervalTree::search: returns i such that target in [spos[i-1], spos[i])
CM_HD int cm_iv_search(const AnnotDev &a, uint32_t target) {
    if (a.iv_lut && a.n_iv > 1) {
        // the loop below returns max(1, #{spos <= target}); count inside the bucket of `target` only
        uint32_t b = target >> a.iv_lut_shift;
        if (b >= a.iv_lut_n) b = a.iv_lut_n - 1;
        uint32_t lo = a.iv_lut[b], hi = a.iv_lut[b + 1];   // spos[k] <= target for k < lo; spos[k] > target for k >= hi
        if (b == a.iv_lut_n - 1) hi = a.n_iv;
        while (lo < hi) {
            uint32_t mid = (lo + hi) >> 1;
            if (a.iv_spos[mid] <= target) lo = mid + 1; else hi = mid;
        }
        return lo < 1 ? 1 : (int) lo;
    }
    int beg = 0, end = (int) a.n_iv;
    while (end - beg > 1) {
        int mid = (beg + end) / 2;
        if (target < a.iv_spos[mid]) end = mid; else beg = mid;
    }
    return end;
}

// FlatIntervalTree::find_ind; returns the interval or -1, `ind` as the reference leaves it
CM_HD int cm_find_ind(const AnnotDev &a, uint32_t pos, int &ind) {
    ind = -1;
    if (pos < a.iv_spos[0]) return -1;
    ind = cm_iv_search(a, pos) - 1;
    if (ind < 0 || a.iv_epos[ind] < pos) return -1;
    return ind;
}

// GTFParser::get_location_overlap (gene_annotation.cpp:538)
CM_HD int cm_location_overlap(const AnnotDev &a, uint32_t pos) {
    int ind;
    return cm_find_ind(a, pos, ind);
}

// GTFParser::get_interval: valid index or -1
CM_HD int cm_get_interval(const AnnotDev &a, int ind) { return (ind < 0 || ind >= (int) a.n_iv) ? -1 : ind; }

// GTFParser::get_upper_bound_lookup
CM_HD uint32_t cm_upper_bound_lookup(const AnnotDev &a, const cm_params &p, uint32_t spos, uint32_t mlen, uint32_t rlen,
                                     uint32_t &max_end, int &ol_iv) {
    max_end = 0;
    int it_ind = -1;
    int ov = cm_find_ind(a, spos, it_ind);
    uint32_t epos = spos + mlen - 1;
    if (ov < 0) {   // not found => intronic
        ol_iv = -1;
        int nx = cm_get_interval(a, it_ind + 1);
        // the reference dereferences get_interval(it_ind+1) unconditionally; past the last interval it would crash
        max_end = (nx >= 0 ? a.iv_spos[nx] : 0u) - 1;
        if (max_end < epos) return 0;
        uint32_t x = spos + rlen + (uint32_t) p.max_ed, y = max_end - mlen + 1;
        return x < y ? x : y;
    }
    ol_iv = -1;
    uint32_t min_end = 1000000000u, max_next_exon = 0;
    if (epos > a.iv_epos[ov]) {
        for (uint32_t s = a.iv_seg_off[ov]; s < a.iv_seg_off[ov + 1]; s++) {
            if (a.seg_end[s] >= epos) {
                if (a.seg_end[s] > max_end) max_end = a.seg_end[s];
                if (a.seg_end[s] < min_end) min_end = a.seg_end[s];
                if (a.seg_next_exon_beg[s] > max_next_exon) max_next_exon = a.seg_next_exon_beg[s];
            }
        }
    } else {
        max_end = a.iv_max_end[ov];
        min_end = a.iv_min_end[ov];
        max_next_exon = a.iv_max_next_exon[ov];
    }
    if (max_end > 0 && max_end >= epos) {   // exonic
        ol_iv = ov;
        if (min_end < rlen + epos && max_next_exon != 0) return max_next_exon + mlen - 1;
        return max_end - mlen + 1;
    }
    max_end = 0;   // on exon boundary
    ol_iv = -1;
    return 0;
}

// GTFParser::get_upper_bound
CM_HD uint32_t cm_upper_bound(const AnnotDev &a, const cm_params &p, uint32_t spos, uint32_t mlen, uint32_t rlen,
                              uint32_t &max_end, int &ol_iv) {
    if (cm_bit(a.near_border, a.n_bits, spos)) return cm_upper_bound_lookup(a, p, spos, mlen, rlen, max_end, ol_iv);
    max_end = 0;
    ol_iv = -1;
    return spos + rlen + (uint32_t) p.max_ed;
}

// check_junction (chain.cpp:28-64)
CM_HD bool cm_check_junction(const AnnotDev &a, const cm_params &p, uint32_t s1, uint32_t s2, int ol_iv, int kmer,
                             int read_dist, int &trans_dist) {
    trans_dist = CM_INF;
    if (ol_iv < 0) return false;
    uint32_t e1 = s1 + kmer - 1;
    if (s2 <= e1) return false;
    int trans_dist2intron = -1;
    for (uint32_t s = a.iv_seg_off[ol_iv]; s < a.iv_seg_off[ol_iv + 1]; s++) {
        int e12end = (int) (a.seg_end[s] - e1);
        int beg2s2 = (int) (s2 - a.seg_next_exon_beg[s]);
        if (e12end >= 0 && e12end < read_dist && beg2s2 + kmer < 0) trans_dist2intron = (int) (s2 - e1 - 1);
        if (e12end < 0 || beg2s2 < 0) continue;
        trans_dist = e12end + beg2s2;
        int d = trans_dist - read_dist;
        if ((d < 0 ? -d : d) <= p.max_ed) return true;
    }
    if (trans_dist2intron != -1) {
        trans_dist = trans_dist2intron;
        return true;
    }
    trans_dist = CM_INF;
    return false;
}

// GTFParser::get_shift (gene_annotation.cpp:451-457): entry of con2chr[contig] covering loc
CM_HD int cm_get_shift(const AnnotDev &a, uint32_t loc) {
    uint32_t i;
    for (i = 1; i < a.n_chr; i++)
        if (loc < a.chr_shift[i]) return (int) i - 1;
    return (int) i - 1;
}
