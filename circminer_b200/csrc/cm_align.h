// Alignment arithmetic of the hot path, restated for one thread with the live band in
// registers / local memory instead of the reference's 600x600 tables:
//   ScoreMatrix::init                          src/align.cpp:725-760
//   Alignment::global_alignment[_reverse]      src/align.cpp:166-212
//   global_one_side_banded_alignment           src/align.cpp:219-252
//   global_banded_alignment_drop (X-drop)      src/align.cpp:254-390
//   global_banded_alignment[_reverse]          src/align.cpp:395-509
//   local_alignment_right / _left              src/align.cpp:556-600
//   DropAlignment::local_alignment_*_sc        src/align.cpp:669-723
//   AlignCandid                                src/align.h:123-153
// The reference's tables persist between calls, but every cell these functions read is written
// in the same call (|i-j| = w+1 guard lines, row/column 0), so a rolling band is exact.
#pragma once
#include "cm_types.h"

#define CM_BANDSZ (2 * CM_MAX_BAND + 3)
#define CM_FULLDP_MAX 64   // largest s-length the full-DP fallback accepts (reference needs <= 2w)

// score_ch / diff_ch: equal iff both in "ACGTacgt" and equal ignoring case; N never matches
CM_HD bool cm_base_eq(uint8_t a, uint8_t b) {
    uint8_t ua = a & 0xDF, ub = b & 0xDF;
    if (ua != ub) return false;
    return ua == 'A' || ua == 'C' || ua == 'G' || ua == 'T';
}

// logical character x of a sequence of length len, optionally read back to front
CM_HD uint8_t cm_at(const uint8_t *p, int len, int x, bool rev) { return rev ? p[len - 1 - x] : p[x]; }

struct AlignCand {   // AlignCandid
    int ed, sclen, indel, score;
};
CM_HD AlignCand cm_cand(int e, int s, int i) { AlignCand c; c.ed = e; c.sclen = s; c.indel = i; c.score = -s - 2 * e; return c; }
CM_HD AlignCand cm_cand4(int e, int s, int i, int scr) { AlignCand c; c.ed = e; c.sclen = s; c.indel = i; c.score = scr; return c; }
CM_HD bool cm_cand_less(const AlignCand &a, const AlignCand &r) {
    if (a.score != r.score) return a.score > r.score;
    if (a.ed != r.ed) return a.ed < r.ed;
    int x = a.indel < 0 ? -a.indel : a.indel, y = r.indel < 0 ? -r.indel : r.indel;
    return x < y;
}
CM_HD void cm_cand_update(AlignCand &b, const AlignCand &r) { if (cm_cand_less(r, b)) b = r; }

struct AlignStats {      // roofline numerators
    uint64_t xdrop_cells, edit_cells, xdrop_calls, edit_calls;
};

// ---- banded edit DP with the band in registers ---------------------------------------------
// Same recurrence, same arithmetic and the same DPTINF guard values as the banded branch of
// cm_edit_last_column below, for a compile-time band W: one array of 2W+3 slots updated in place
// (slot d of column j reads its own old value = dp[i-1][j-1], the old slot d+1 = dp[i][j-1] and the new
// slot d-1 = dp[i-1][j]), and the 2W+1 characters of s under the band in a sliding window of codes
// (0..3 = ACGT ignoring case, 4 = anything else in s, 5 = anything else in t: never equal).
CM_HD uint32_t cm_base_code(uint8_t c, uint32_t other) {
    // 'A' 0x41, 'C' 0x43, 'G' 0x47, 'T' 0x54: bits 1..2 tell them apart (A 0, C 1, T 2, G 3; only equality is ever tested);
    // membership is one bit test in the mask of the letters A, C, G, T
    const uint32_t u = c & 0xDFu, idx = u - 'A';
    return (idx < 26u && ((0x80045u >> idx) & 1u)) ? ((u >> 1) & 3u) : other;
}
template <int W, bool CHECK>
CM_HD void cm_edit_col_reg(uint32_t (&a)[2 * W + 3], const uint32_t (&win)[2 * W + 1], uint32_t tc, int j, int n) {
    // rows of this column: i = j - W - 1 + d for slot d in [1, 2W+1]
#pragma unroll
    for (int d = 1; d <= 2 * W + 1; d++) {
        uint32_t v = a[d] + (win[d - 1] == tc ? 0u : 1u);
        const uint32_t xy = (a[d - 1] < a[d + 1] ? a[d - 1] : a[d + 1]) + 1;   // min(x + 1, y + 1); values <= DPTINF + m
        v = v < xy ? v : xy;
        if (CHECK) {
            const int i = j - W - 1 + d;
            if (i < 1 || i > n) v = CM_DPTINF;
            if (i == 0) v = (uint32_t) j;   // row 0: dp[0][j] = j (only while j <= W)
        }
        a[d] = v;
    }
}
template <int W>
CM_HD void cm_edit_banded_reg(const uint8_t *s, int n, const uint8_t *t, int m, bool rev, int lo, int hi, uint32_t *out, AlignStats *st) {
    uint32_t a[2 * W + 3], win[2 * W + 1];
#pragma unroll
    for (int d = 0; d < 2 * W + 3; d++) a[d] = CM_DPTINF;
#pragma unroll
    for (int i = 0; i <= W; i++) a[i + W + 1] = (uint32_t) i;
    // window of column 1: slot d holds s[i-1] with i = d - W, so characters -W .. W of s
#pragma unroll
    for (int d = 1; d <= 2 * W + 1; d++) {
        const int x = d - W - 1;
        win[d - 1] = (x >= 0 && x < n) ? cm_base_code(cm_at(s, n, x, rev), 4u) : 4u;
    }
    const int sstep = rev ? -1 : 1, tstep = rev ? -1 : 1;
    const uint8_t *sp = rev ? s + (n - 1 - (W + 1)) : s + (W + 1);   // character W+1 of s: enters the window at column 2
    const uint8_t *tp = rev ? t + (m - 1) : t;
    uint64_t cells = 0;
    for (int j = 1; j <= m; j++) {
        const uint32_t tc = cm_base_code(*tp, 5u);
        tp += tstep;
        if (j > W && j + W <= n) cm_edit_col_reg<W, false>(a, win, tc, j, n);
        else cm_edit_col_reg<W, true>(a, win, tc, j, n);
        const int i0 = j - W < 1 ? 1 : j - W, i1 = j + W > n ? n : j + W;
        cells += (uint64_t) (i1 >= i0 ? i1 - i0 + 1 : 0);
        // slide: column j+1 sees characters j-W .. j+W
#pragma unroll
        for (int d = 0; d < 2 * W; d++) win[d] = win[d + 1];
        win[2 * W] = (j + W < n) ? cm_base_code(*sp, 4u) : 4u;
        sp += sstep;
    }
    if (st) st->edit_cells += cells;
    for (int i = lo; i <= hi; i++) {
        const int d = i - m + W + 1;
        if (i >= 0 && i <= n && d >= 1 && d <= 2 * W + 1) {
            uint32_t v = 0;
#pragma unroll
            for (int q = 1; q <= 2 * W + 1; q++) v = q == d ? a[q] : v;
            out[i - lo] = v;
        }
    }
}

// ---- unit-cost edit DP, last column ------------------------------------------------------
// Fills col[i] = dp[i][m] for the rows the callers read, i in [lo, hi]; entries outside are DPTINF.
// Banded when the reference bands (n > 2w and m > w), full otherwise.  rev: the *_reverse variant.
CM_HDN void cm_edit_last_column_generic_banded(const uint8_t *s, int n, const uint8_t *t, int m, bool rev, int w, int lo, int hi,
                                               uint32_t *out, AlignStats *st);
CM_HDN void cm_edit_last_column(const uint8_t *s, int n, const uint8_t *t, int m, bool rev, int w, int lo, int hi,
                                uint32_t *out /* [hi-lo+1] */, AlignStats *st) {
    if (st) st->edit_calls++;
    for (int i = lo; i <= hi; i++) out[i - lo] = CM_DPTINF;
    if (w < 0 || n <= 2 * w || m <= w) {
        // global_alignment: full table, rolling column over j
        if (n > CM_FULLDP_MAX) return;
        uint32_t col[CM_FULLDP_MAX + 1];
        for (int i = 0; i <= n; i++) col[i] = (uint32_t) i;
        for (int j = 1; j <= m; j++) {
            uint32_t diag = col[0];
            col[0] = (uint32_t) j;
            uint8_t tc = cm_at(t, m, j - 1, rev);
            for (int i = 1; i <= n; i++) {
                uint32_t up = col[i];   // dp[i][j-1]
                uint32_t v = diag + (cm_base_eq(cm_at(s, n, i - 1, rev), tc) ? 0u : 1u);
                uint32_t a = col[i - 1] + 1, b = up + 1;
                v = v < a ? v : a;
                v = v < b ? v : b;
                diag = up;
                col[i] = v;
            }
        }
        if (st) st->edit_cells += (uint64_t) n * (uint64_t) (m > 0 ? m : 0);
        for (int i = lo; i <= hi; i++)
            if (i >= 0 && i <= n) out[i - lo] = col[i];
        return;
    }
#if !defined(CM_EDIT_GENERIC_ONLY)
    if (w == 3) {
#ifdef CM_EDIT_CHECK
        uint32_t chk[2 * CM_MAX_BAND + 1];
        for (int i = lo; i <= hi; i++) chk[i - lo] = CM_DPTINF;
        AlignStats st2 = {0, 0, 0, 0};
        cm_edit_banded_reg<3>(s, n, t, m, rev, lo, hi, chk, &st2);
        uint32_t ref_out[2 * CM_MAX_BAND + 1];
        AlignStats st3 = {0, 0, 0, 0};
        cm_edit_last_column_generic_banded(s, n, t, m, rev, w, lo, hi, ref_out, &st3);
        for (int i = lo; i <= hi; i++)
            if (chk[i - lo] != ref_out[i - lo] || st2.edit_cells != st3.edit_cells) __builtin_trap();
#endif
        cm_edit_banded_reg<3>(s, n, t, m, rev, lo, hi, out, st);
        return;
    }
#endif
    cm_edit_last_column_generic_banded(s, n, t, m, rev, w, lo, hi, out, st);
}

// the banded branch for any band width (arrays indexed at run time)
CM_HDN void cm_edit_last_column_generic_banded(const uint8_t *s, int n, const uint8_t *t, int m, bool rev, int w, int lo, int hi,
                                               uint32_t *out, AlignStats *st) {
    for (int i = lo; i <= hi; i++) out[i - lo] = CM_DPTINF;
    // band: column j holds rows i in [j-w, j+w]; slot d = i - j + w + 1, slots 0 and 2w+2 are the guard lines
    uint32_t prev[CM_BANDSZ], cur[CM_BANDSZ];
    const int B = 2 * w + 3;
    for (int d = 0; d < B; d++) prev[d] = CM_DPTINF;
    for (int i = 0; i <= w; i++) prev[i + w + 1] = (uint32_t) i;   // column 0: dp[i][0] = i, dp[w+1][0] = INF
    uint64_t cells = 0;
    for (int j = 1; j <= m; j++) {
        for (int d = 0; d < B; d++) cur[d] = CM_DPTINF;
        if (j <= w) cur[-j + w + 1] = (uint32_t) j;                // row 0: dp[0][j] = j
        uint8_t tc = cm_at(t, m, j - 1, rev);
        int i0 = j - w < 1 ? 1 : j - w, i1 = j + w > n ? n : j + w;
        for (int i = i0; i <= i1; i++) {
            int d = i - j + w + 1;
            uint32_t v = prev[d] + (cm_base_eq(cm_at(s, n, i - 1, rev), tc) ? 0u : 1u);   // dp[i-1][j-1]
            uint32_t a = cur[d - 1] + 1;                                                  // dp[i-1][j]
            uint32_t b = prev[d + 1] + 1;                                                 // dp[i][j-1]
            v = v < a ? v : a;
            v = v < b ? v : b;
            cur[d] = v;
        }
        cells += (uint64_t) (i1 >= i0 ? i1 - i0 + 1 : 0);
        for (int d = 0; d < B; d++) prev[d] = cur[d];
    }
    if (st) st->edit_cells += cells;
    for (int i = lo; i <= hi; i++) {
        int d = i - m + w + 1;
        if (i >= 0 && i <= n && d >= 1 && d <= 2 * w + 1) out[i - lo] = prev[d];
    }
}

// Alignment::local_alignment_right (rev = false) / _left (rev = true): s = read piece, t = exon text
CM_HDN int cm_local_alignment(const uint8_t *s, int n, const uint8_t *t, int m, bool rev, const cm_params &p, int &indel,
                              int &score, AlignStats *st) {
    const int w = p.band_width;
    int lo = m - w < 0 ? 0 : m - w, hi = m + w > n ? n : m + w;
    uint32_t colv[2 * CM_MAX_BAND + 1];
    if (hi >= lo) cm_edit_last_column(s, n, t, m, rev, w, lo, hi, colv, st);
    AlignCand best = cm_cand(p.max_ed + 1, p.max_sc + 1, w + 1);
    for (int i = lo; i <= hi; i++)
        if (colv[i - lo] <= (uint32_t) p.max_ed) cm_cand_update(best, cm_cand((int) colv[i - lo], 0, m - i));
    score = -best.ed;
    indel = best.indel;
    return best.ed;
}

// EditDistAlignment::local_alignment_right_sc (rev = false, align.cpp:602-633) / _left_sc (rev = true, :635-667): the
// aligner of stage 2's TransExtension (ProcessCirc: extension(0, EDIT_ALIGNMENT)).  s = reference text, t = read text:
// unit-cost banded (or, for short texts, full) global DP, then the best cell over the last min(maxSc, m) + 1 columns
// inside the band.  The reference scans columns m down to m - max_sclen and rows upward; candidates of different columns
// never tie (same score and ed => same soft clip), so scanning the columns as they are produced picks the same one.
CM_HDN int cm_edit_alignment_sc(const uint8_t *s, int n, const uint8_t *t, int m, bool rev, const cm_params &p, int &sc_len, int &indel,
                                int &score, AlignStats *st) {
    const int w = p.band_width, max_sclen = p.max_sc < m ? p.max_sc : m;
    AlignCand best = cm_cand(p.max_ed + 1, p.max_sc + 1, w + 1);
    if (st) st->edit_calls++;
    uint64_t cells = 0;
    if (w < 0 || n <= 2 * w || m <= w) {            // global_alignment: full table, rolling column over j
        // only rows <= j + w are read below and a row depends on the rows above it: rows past m + w are not computed
        const int nn = (w >= 0 && m + w < n) ? m + w : n;
        if (nn <= CM_FULLDP_MAX) {
            uint32_t col[CM_FULLDP_MAX + 1];
            for (int i = 0; i <= nn; i++) col[i] = (uint32_t) i;
            for (int j = 0; j <= m; j++) {
                if (j > 0) {
                    uint32_t diag = col[0];
                    col[0] = (uint32_t) j;
                    const uint8_t tc = cm_at(t, m, j - 1, rev);
                    for (int i = 1; i <= nn; i++) {
                        const uint32_t up = col[i];
                        uint32_t v = diag + (cm_base_eq(cm_at(s, n, i - 1, rev), tc) ? 0u : 1u);
                        const uint32_t a = col[i - 1] + 1, b = up + 1;
                        v = v < a ? v : a;
                        v = v < b ? v : b;
                        diag = up;
                        col[i] = v;
                    }
                    cells += (uint64_t) nn;
                }
                if (j >= m - max_sclen) {
                    const int i0 = j - w > 0 ? j - w : 0, i1 = j + w < nn ? j + w : nn;
                    for (int i = i0; i <= i1; i++)
                        if (col[i] <= (uint32_t) p.max_ed) cm_cand_update(best, cm_cand((int) col[i], m - j, j - i));
                }
            }
        }
    } else {
        // band: column j holds rows i in [j-w, j+w]; slot d = i - j + w + 1, slots 0 and 2w+2 are the guard lines
        uint32_t prev[CM_BANDSZ], cur[CM_BANDSZ];
        const int B = 2 * w + 3;
        for (int d = 0; d < B; d++) prev[d] = CM_DPTINF;
        for (int i = 0; i <= w; i++) prev[i + w + 1] = (uint32_t) i;   // column 0
        for (int j = 0; j <= m; j++) {
            if (j > 0) {
                for (int d = 0; d < B; d++) cur[d] = CM_DPTINF;
                if (j <= w) cur[-j + w + 1] = (uint32_t) j;                // row 0
                const uint8_t tc = cm_at(t, m, j - 1, rev);
                const int i0 = j - w < 1 ? 1 : j - w, i1 = j + w > n ? n : j + w;
                for (int i = i0; i <= i1; i++) {
                    const int d = i - j + w + 1;
                    uint32_t v = prev[d] + (cm_base_eq(cm_at(s, n, i - 1, rev), tc) ? 0u : 1u);
                    const uint32_t a = cur[d - 1] + 1, b = prev[d + 1] + 1;
                    v = v < a ? v : a;
                    v = v < b ? v : b;
                    cur[d] = v;
                }
                cells += (uint64_t) (i1 >= i0 ? i1 - i0 + 1 : 0);
                for (int d = 0; d < B; d++) prev[d] = cur[d];
            }
            if (j >= m - max_sclen) {
                const int i0 = j - w > 0 ? j - w : 0, i1 = j + w < n ? j + w : n;
                for (int i = i0; i <= i1; i++) {
                    const uint32_t v = prev[i - j + w + 1];
                    if (v <= (uint32_t) p.max_ed) cm_cand_update(best, cm_cand((int) v, m - j, j - i));
                }
            }
        }
    }
    if (st) st->edit_cells += cells;
    if (m <= p.max_ed) cm_cand_update(best, cm_cand(m, 0, 0));   // very small text that cannot pass the threshold
    score = m - best.sclen - 2 * best.ed;
    sc_len = best.sclen;
    indel = best.indel;
    return best.ed;
}

// One-side band (columns j in [i, i+W] of row i) in registers: slot d = j - i + 1, updated in place (its own old
// value = dp[i-1][j-1], old slot d+1 = dp[i-1][j], new slot d-1 = dp[i][j-1]); the W+1 characters of t under the band
// slide by one per row.  Same arithmetic and guards as the run-time-indexed loop in cm_one_side_banded.
template <int W>
CM_HD int cm_one_side_banded_reg(const uint8_t *s, int n, const uint8_t *t, int m) {
    uint32_t a[W + 3], tw[W + 1];
    a[0] = CM_DPTINF;
    a[W + 2] = CM_DPTINF;
#pragma unroll
    for (int j = 0; j <= W; j++) {
        a[j + 1] = (uint32_t) j;
        tw[j] = j < m ? cm_base_code(t[j], 5u) : 5u;
    }
    for (int i = 1; i <= n; i++) {
        const uint32_t sc = cm_base_code(s[i - 1], 4u);
#pragma unroll
        for (int d = 1; d <= W + 1; d++) {
            uint32_t x = a[d] + (sc == tw[d - 1] ? 0u : 1u);
            const uint32_t y = a[d + 1] + 1, z = a[d - 1] + 1;
            x = x < y ? x : y;
            x = x < z ? x : z;
            a[d] = x;
        }
#pragma unroll
        for (int k = 0; k < W; k++) tw[k] = tw[k + 1];
        tw[W] = i + W < m ? cm_base_code(t[i + W], 5u) : 5u;
    }
    const int d = m - n + 1;
    if (d < 1 || d > W + 1) return CM_DPTINF;
    uint32_t v = 0;
#pragma unroll
    for (int q = 1; q <= W + 1; q++) v = q == d ? a[q] : v;
    return (int) v;
}

// Alignment::global_one_side_banded_alignment: dp[n][m], rows i, columns j in [i, i+w]
CM_HDN int cm_one_side_banded_generic(const uint8_t *s, int n, const uint8_t *t, int m, int w);
CM_HDN int cm_one_side_banded(const uint8_t *s, int n, const uint8_t *t, int m, int w, AlignStats *st) {
    if (st) st->edit_calls++;
    if (w < 0 || n <= w) {
        uint32_t v = CM_DPTINF;
        if (n < 0) n = 0;
        if (n > CM_FULLDP_MAX) return (int) v;
        uint32_t col[CM_FULLDP_MAX + 1];
        for (int i = 0; i <= n; i++) col[i] = (uint32_t) i;
        for (int j = 1; j <= m; j++) {
            uint32_t diag = col[0];
            col[0] = (uint32_t) j;
            for (int i = 1; i <= n; i++) {
                uint32_t up = col[i];
                uint32_t x = diag + (cm_base_eq(s[i - 1], t[j - 1]) ? 0u : 1u);
                uint32_t a = col[i - 1] + 1, b = up + 1;
                x = x < a ? x : a;
                x = x < b ? x : b;
                diag = up;
                col[i] = x;
            }
        }
        if (st) st->edit_cells += (uint64_t) n * (uint64_t) (m > 0 ? m : 0);
        return (int) col[n];
    }
    if (w > CM_MAX_BAND) return CM_DPTINF;
#if !defined(CM_EDIT_GENERIC_ONLY)
    if (w <= 3) {
        const int r = w == 0 ? cm_one_side_banded_reg<0>(s, n, t, m) : w == 1 ? cm_one_side_banded_reg<1>(s, n, t, m)
                    : w == 2 ? cm_one_side_banded_reg<2>(s, n, t, m) : cm_one_side_banded_reg<3>(s, n, t, m);
#ifdef CM_EDIT_CHECK
        if (r != cm_one_side_banded_generic(s, n, t, m, w)) __builtin_trap();
#endif
        if (st) st->edit_cells += (uint64_t) n * (uint64_t) (w + 1);
        return r;
    }
#endif
    if (st) st->edit_cells += (uint64_t) n * (uint64_t) (w + 1);
    return cm_one_side_banded_generic(s, n, t, m, w);
}

// the banded branch for any band width (arrays indexed at run time)
CM_HDN int cm_one_side_banded_generic(const uint8_t *s, int n, const uint8_t *t, int m, int w) {
    // slot d = j - i + 1 in [1, w+1]; slots 0 and w+2 are dp[i][i-1] and dp[i-1][i+w] (= DPTINF)
    uint32_t prev[CM_MAX_BAND + 3], cur[CM_MAX_BAND + 3];
    for (int d = 0; d <= w + 2; d++) prev[d] = CM_DPTINF;
    for (int j = 0; j <= w; j++) prev[j + 1] = (uint32_t) j;
    for (int i = 1; i <= n; i++) {
        cur[0] = CM_DPTINF;
        cur[w + 2] = CM_DPTINF;
        uint8_t sc = s[i - 1];
        for (int d = 1; d <= w + 1; d++) {
            int j = i + d - 1;
            uint32_t x = prev[d] + ((j <= m && cm_base_eq(sc, t[j - 1])) ? 0u : 1u);   // dp[i-1][j-1]; columns > m never reach dp[n][m]
            uint32_t a = prev[d + 1] + 1;                                  // dp[i-1][j]
            uint32_t b = cur[d - 1] + 1;                                   // dp[i][j-1]
            x = x < a ? x : a;
            x = x < b ? x : b;
            cur[d] = x;
        }
        for (int d = 0; d <= w + 2; d++) prev[d] = cur[d];
    }
    int d = m - n + 1;
    if (d < 1 || d > w + 1) return CM_DPTINF;
    return (int) prev[d];
}

// ---- X-drop banded score DP --------------------------------------------------------------
// Alignment::global_banded_alignment_drop: s = reference text (n), t = read piece (m).  Anti-diagonal
// k = i + j; slot d = i - j + w + 1 in [1, 2w+1]; slots 0 and 2w+2 are the -DPTINF guard lines
// dpx[x][x+w+1] / dpx[x+w+1][x].  Returns the last cell (in (k, i) visit order) holding the running
// maximum, and that maximum (== dpx[on_s][on_t], which is never pruned).
CM_HDN void cm_xdrop_generic(const uint8_t *s, int n, const uint8_t *t, int m, bool rev, const cm_params &p, int &on_s, int &on_t,
                             int &score, AlignStats *st) {
    const int w = p.band_width, B = 2 * w + 3;
    const int ind = p.ind_sc, NEG = -CM_DPTINF;
    if (st) st->xdrop_calls++;
    on_s = 0;
    on_t = 0;
    score = 0;
    if (m <= 0 || n <= 0) return;
    int buf[3][CM_BANDSZ];
    int *a2 = buf[0], *a1 = buf[1], *a0 = buf[2];
    for (int d = 0; d < B; d++) { a2[d] = NEG; a1[d] = NEG; }
    a2[w + 1] = 0;                                  // dpx[0][0]
    if (w >= 1) { a1[w + 2] = ind; a1[w] = ind; }   // dpx[1][0], dpx[0][1]
    int pre_optimum = 0, cur_optimum = 0;
    int lb = 1, ub = 1;
    int new_ub, pre_ub = 0, best_i = 0, best_j = 0;
    uint64_t cells = 0;
    for (int k = 2; k <= m + n; ++k) {
        for (int d = 0; d < B; d++) a0[d] = NEG;
        if (k <= w) { a0[k + w + 1] = k * ind; a0[-k + w + 1] = k * ind; }
        new_ub = -1;
        for (int i = lb; i <= ub; ++i) {
            int j = k - i;
            int d = i - j + w + 1;
            if (d < 1 || d > 2 * w + 1) continue;   // cannot happen with the reference's lb/ub evolution
            int sc = cm_base_eq(cm_at(s, n, i - 1, rev), cm_at(t, m, j - 1, rev)) ? p.mat_sc : p.mis_sc;
            int v = a2[d] + sc;
            int x = a1[d - 1] + ind;   // dpx[i-1][j]
            int y = a1[d + 1] + ind;   // dpx[i][j-1]
            v = v > x ? v : x;
            v = v > y ? v : y;
            cells++;
            if (v > cur_optimum) cur_optimum = v;
            if (v >= cur_optimum) {
                cur_optimum = v;
                best_i = i;
                best_j = j;
            }
            if (v + p.xdrop < pre_optimum) v = NEG;
            if (v > NEG) new_ub = i;
            a0[d] = v;
        }
        int lb_t = k - lb;
        if (lb_t == m || (k > w && ((k - w) % 2 == 0))) ++lb;
        if (ub < n && (k <= w || (k > w && ((k - w) % 2 == 1)))) ++ub;
        if ((pre_ub == -1 && new_ub == -1) || lb > ub) break;
        pre_ub = new_ub;
        if (cur_optimum > pre_optimum) pre_optimum = cur_optimum;
        int *tmp = a2;
        a2 = a1;
        a1 = a0;
        a0 = tmp;
    }
    if (st) st->xdrop_cells += cells;
    on_s = best_i;
    on_t = best_j;
    score = cur_optimum;
}

// Register-resident form for a compile-time band W.  Cell (i, j) lives in slot e = i - j + W + 1 of ONE
// array: slots whose parity matches anti-diagonal k hold diagonal k, the others diagonal k-1, so the
// update X[e] = max(X[e] + sc, X[e-1] + ind, X[e+1] + ind) is in place (X[e] still holds diagonal k-2)
// and every index is a compile-time constant after unrolling.  Same visit order, same lb/ub evolution,
// same pruning as cm_xdrop_generic, hence the same (on_s, on_t, score).
//
// The two strings are not read cell by cell: anti-diagonals k = 2q and 2q + 1 only touch characters
// q + LO .. q + HI of s and of t (LO = -(W/2 + 1), HI = (W + 1)/2 - 1), so a window of W + 1 character
// codes per string sits in registers and slides by one character per q (one byte load per string per
// two anti-diagonals, issued ahead of its use).  Codes: 'A','C','G','T' (case folded) or a value that
// never compares equal (N never matches N, align.cpp:745-759).
CM_HD int cm_code(uint8_t c, int bad) {
    const uint32_t u = c & 0xDFu, idx = u - 'A';
    return (idx < 26u && ((0x80045u >> idx) & 1u)) ? (int) u : bad;
}
CM_HD int cm_code_at(const uint8_t *p0, int step, int len, int c, int bad) { return (c >= 0 && c < len) ? cm_code(p0[c * step], bad) : bad; }

// State of one X-drop run (registers after inlining).
template <int W>
struct XdropState {
    int X[2 * W + 3];
    int SW[W + 1], TW[W + 1];
    int pre, cur, lb, ub, pre_ub, best_i, best_k;
    uint32_t cells;
};

// One anti-diagonal k = 2q + KP.  EARLY: k may be <= W (boundary cells dpx[k][0], dpx[0][k] still enter the band).
// Row indices are handled relative to q (h = i - q is a compile-time constant per slot); the cell count of a diagonal
// is ub - lb + 1 (every i in [lb, ub] has a slot in the band); the best cell is kept as (i, k).
#ifndef CM_XDROP_FAST
#define CM_XDROP_FAST 1
#endif
// FULL: every slot of this parity is inside [lb, ub] (the usual state once the band has opened), so the per-slot range
// tests and the boundary values drop out.
template <int W, int KP, bool EARLY, bool FULL>
CM_HD void cm_xdrop_step_impl(XdropState<W> &S, int k, int q, int mat, int mis, int ind, int xd, int &new_ub) {
    const int NEG = -CM_DPTINF;
    constexpr int LO = -(W / 2 + 1);
    const int lq = S.lb - q, uq = S.ub - q, lim = S.pre - xd;   // lim >= -xd > NEG: a cell that survives the X-drop is live
    int nub_h = -1000;
    // best cell of this diagonal (last one among equals, as the running "v >= cur" of the reference picks), merged into
    // the running optimum once per diagonal: the same (cur, i, k) as updating per cell
    int dmax = -2 * CM_DPTINF, dh = 0;
    S.cells += (uint32_t) (S.ub - S.lb + 1);
#pragma unroll
    for (int e = 1; e <= 2 * W + 1; e++) {
        const int d = e - W - 1;
        if (((d + KP) & 1) != 0) continue;          // slot belongs to the other parity
        const int h = (KP + d) / 2;                 // exact: KP + d is even
        int v;
        if (FULL || (h >= lq && h <= uq)) {
            int sc = S.SW[h - 1 - LO] == S.TW[(KP - d) / 2 - 1 - LO] ? mat : mis;
            v = S.X[e] + sc;
            int ab = S.X[e - 1] > S.X[e + 1] ? S.X[e - 1] : S.X[e + 1];
            ab += ind;                              // max(a + ind, b + ind); |values| <= DPTINF + a few hundred, no overflow
            v = v > ab ? v : ab;
            if (v >= dmax) { dmax = v; dh = h; }
            const bool live = v >= lim;             // otherwise v + xdrop < pre_optimum
            v = live ? v : NEG;
            if (live) nub_h = h;                    // h ascends with e: the last live cell wins
        } else {
            v = (EARLY && k <= W && (d == k || d == -k)) ? k * ind : NEG;
        }
        S.X[e] = v;
    }
    if (dmax >= S.cur) { S.cur = dmax; S.best_i = q + dh; S.best_k = k; }
    new_ub = nub_h > -1000 ? q + nub_h : -1;
}

template <int W, int KP, bool EARLY>
CM_HD void cm_xdrop_step(XdropState<W> &S, int k, int q, int mat, int mis, int ind, int xd, int &new_ub) {
#if CM_XDROP_FAST
    if (!EARLY) {
        // rows of the slots of parity KP, relative to q: h = (KP + d) / 2 over d = -W .. W with (d + KP) even
        constexpr int DMIN = ((W + KP) & 1) ? -W + 1 : -W, DMAX = ((W + KP) & 1) ? W - 1 : W;
        constexpr int HMIN = (KP + DMIN) / 2, HMAX = (KP + DMAX) / 2;
        static_assert((KP + DMIN) % 2 == 0 && (KP + DMAX) % 2 == 0, "slot parity");
        if (S.lb - q <= HMIN && S.ub - q >= HMAX) {
            cm_xdrop_step_impl<W, KP, false, true>(S, k, q, mat, mis, ind, xd, new_ub);
            return;
        }
    }
#endif
    cm_xdrop_step_impl<W, KP, EARLY, false>(S, k, q, mat, mis, ind, xd, new_ub);
}

// anti-diagonals 2q and 2q + 1; true = the run is over
template <int W, bool EARLY>
CM_HD bool cm_xdrop_pair(XdropState<W> &S, int q, int K, int n, int m, int mat, int mis, int ind, int xd) {
#pragma unroll
    for (int kp = 0; kp < 2; kp++) {
        const int k = 2 * q + kp;
        if (k > K) return true;
        int new_ub;
        if (kp) cm_xdrop_step<W, 1, EARLY>(S, k, q, mat, mis, ind, xd, new_ub);
        else cm_xdrop_step<W, 0, EARLY>(S, k, q, mat, mis, ind, xd, new_ub);
        const bool even_kw = ((kp ^ W) & 1) == 0;   // (k - W) % 2 == 0
        const bool past = !EARLY || k > W;
        if (k - S.lb == m || (past && even_kw)) ++S.lb;
        if (S.ub < n && (!past || !even_kw)) ++S.ub;
        if ((S.pre_ub == -1 && new_ub == -1) || S.lb > S.ub) return true;
        S.pre_ub = new_ub;
        if (S.cur > S.pre) S.pre = S.cur;
    }
    return false;
}

template <int W>
CM_HD void cm_xdrop_reg(const uint8_t *s, int n, const uint8_t *t, int m, bool rev, const cm_params &p, int &on_s, int &on_t,
                        int &score, AlignStats *st) {
    const int ind = p.ind_sc, NEG = -CM_DPTINF, mat = p.mat_sc, mis = p.mis_sc, xd = p.xdrop;
    if (st) st->xdrop_calls++;
    on_s = 0;
    on_t = 0;
    score = 0;
    if (m <= 0 || n <= 0) return;
    XdropState<W> S;
#pragma unroll
    for (int e = 0; e < 2 * W + 3; e++) S.X[e] = NEG;
    S.X[W + 1] = 0;                                    // dpx[0][0]  (anti-diagonal 0)
    if (W >= 1) { S.X[W + 2] = ind; S.X[W] = ind; }    // dpx[1][0], dpx[0][1]  (anti-diagonal 1)
    constexpr int LO = -(W / 2 + 1), HI = (W + 1) / 2 - 1;
    const uint8_t *s0 = rev ? s + n - 1 : s, *t0 = rev ? t + m - 1 : t;
    S.pre = 0; S.cur = 0; S.lb = 1; S.ub = 1; S.pre_ub = 0; S.best_i = 0; S.best_k = 0; S.cells = 0;
    const int K = m + n;
    const int step = rev ? -1 : 1;
#pragma unroll
    for (int x = 0; x <= W; x++) {                     // window of q = 1
        S.SW[x] = cm_code_at(s0, step, n, 1 + LO + x, 1);
        S.TW[x] = cm_code_at(t0, step, m, 1 + LO + x, 2);
    }
    // one q = anti-diagonals 2q and 2q + 1.  The raw bytes of the characters that enter the windows at q + 1 are loaded
    // first and used after the two anti-diagonals.  (A queue of aligned 32-bit words, one load per four characters, was
    // measured on C3: k_chain_spec 101 -> 114 ms per step -- the kernel is bound by issued instructions, not by these loads.)
#define CM_XDROP_Q(EARLY)                                                                                             \
    {                                                                                                                 \
        const int cn = q + 1 + HI;                                                                                    \
        const bool in_s = cn < n, in_t = cn < m;                                                                      \
        const uint8_t rs = in_s ? s0[cn * step] : (uint8_t) 0, rt = in_t ? t0[cn * step] : (uint8_t) 0;               \
        if (cm_xdrop_pair<W, EARLY>(S, q, K, n, m, mat, mis, ind, xd)) { stop = true; break; }                        \
        _Pragma("unroll") for (int x = 0; x < W; x++) { S.SW[x] = S.SW[x + 1]; S.TW[x] = S.TW[x + 1]; }               \
        S.SW[W] = in_s ? cm_code(rs, 1) : 1;                                                                          \
        S.TW[W] = in_t ? cm_code(rt, 2) : 2;                                                                          \
    }
    // the anti-diagonals that still see the boundary cells (k <= W) are peeled off: the steady-state loop carries one
    // variant of the step only
    int q = 1;
    bool stop = false;
    for (; 2 * q <= K && 2 * q <= W; ++q) CM_XDROP_Q(true)
    if (!stop)
        for (; 2 * q <= K; ++q) CM_XDROP_Q(false)
#undef CM_XDROP_Q
    if (st) st->xdrop_cells += S.cells;
    on_s = S.best_i;
    on_t = S.best_k - S.best_i;
    score = S.cur;
}

CM_HDN void cm_xdrop(const uint8_t *s, int n, const uint8_t *t, int m, bool rev, const cm_params &p, int &on_s, int &on_t, int &score,
                     AlignStats *st) {
#ifdef CM_XDROP_CHECK
    if (p.band_width == 3) {
        int a, b, c;
        cm_xdrop_generic(s, n, t, m, rev, p, a, b, c, nullptr);
        cm_xdrop_reg<3>(s, n, t, m, rev, p, on_s, on_t, score, st);
        if (a != on_s || b != on_t || c != score) __builtin_trap();
        return;
    }
#endif
    if (p.band_width == 3) cm_xdrop_reg<3>(s, n, t, m, rev, p, on_s, on_t, score, st);
    else cm_xdrop_generic(s, n, t, m, rev, p, on_s, on_t, score, st);
}

// A result computed ahead of time by the dense alignment kernel for exactly this problem (same text
// pointers, lengths and direction): a pure memo, consulted before running the DP inline.
struct DropHint {
    const uint8_t *s, *t;
    int n, m, rev, valid;
    int ed, sclen, indel, score;
};

// DropAlignment::local_alignment_right_sc (rev = false) / _left_sc (rev = true: both strings reversed)
CM_HDN int cm_drop_alignment_sc(const uint8_t *s, int n, const uint8_t *t, int m, bool rev, const cm_params &p, int &sc_len,
                                int &indel, int &score, AlignStats *st, const DropHint *h = nullptr) {
    if (h && h->valid && h->s == s && h->t == t && h->n == n && h->m == m && h->rev == (int) rev) {
        sc_len = h->sclen;
        indel = h->indel;
        score = h->score;
        return h->ed;
    }
    int on_s, on_t;
    cm_xdrop(s, n, t, m, rev, p, on_s, on_t, score, st);
    int mx = on_s > on_t ? on_s : on_t;
    uint32_t ed = (uint32_t) ((p.mat_sc * mx - score) / (p.mat_sc - p.mis_sc));
    if (ed <= (uint32_t) p.max_ed) {
        // best.update / best.set always take the candidate: its score is >= 0 == best.score and ed <= max_ed
        sc_len = m - on_t;
        indel = on_t - on_s;
        return (int) ed;
    }
    sc_len = (p.max_sc > m ? p.max_sc : m) + 1;
    indel = p.band_width + 1;
    return p.max_ed + 1;
}
