// Stage 2, third piece: the classification of a refined split read --
//   ProcessCirc::check_split_map  (src/process_circ.cpp:892-919 three mates, :922-1133 four mates)
//   ProcessCirc::final_check      (:1135-1341)
//   ProcessCirc::split_realignment (:1343-1392 around a breakpoint of one transcript, :1394-1489 of the full mate)
//   ProcessCirc::rescue_overlapping_bsj (:1491-1552), get_junctions (src/utils.cpp:697-757),
//   MatchedMate::merge_to_right (src/common.cpp:163-189)
// as device functions.  The reference's functions call each other in a cycle (final_check -> rescue_overlapping_bsj ->
// split_realignment -> check_split_map -> final_check ...); here the cycle is a loop over an explicit stack of rescue frames
// (cm_circ_final_check), so that no kernel needs a device-wide stack limit; the depth is capped at CM_CIRC_MAX_DEPTH.
//
// Where the reference would dereference a NULL interval (get_interval(-1) after a mate end that lies in no exon) these
// functions stop the walk instead; such inputs do not come out of find_exact_coord on the reference's own path.
#pragma once
#include "cm_circ.h"
#include "../../include/cm_gpu.h"

#define CM_FR 0       // src/process_circ.h:14-20
#define CM_RF 1
#define CM_CR 20
#define CM_NCR 21
#define CM_MCR 22
#define CM_UD 30
#define CM_NF 40
#define CM_BPRES 5    // src/common.h:42
#define CM_INDELTH 3  // src/common.h:45
// Fixed sizes of per-thread lists.  A row that needs more comes back as -1, the pair as type -2 from the batch entries, and the caller
// redoes that pair with the reference's code (integration/gpu_circ_detect.cpp does).  The host build of the tests overrides them to
// make that path run.
#ifndef CM_CIRC_TID_CAP
#define CM_CIRC_TID_CAP 192    // (transcript, offset) pairs collected per side of a breakpoint
#endif
#ifndef CM_CIRC_JUNC_CAP
#define CM_CIRC_JUNC_CAP 64
#endif
#define CM_CIRC_MAX_DEPTH 6

// branch counters of the host build (tests report which paths their inputs reached); nothing on the device
#if defined(CM_CIRC_TRACE) && !defined(__CUDA_ARCH__)
static long g_circ_trace[16];
#define CM_CIRC_HIT(k) (g_circ_trace[k]++)
#else
#define CM_CIRC_HIT(k) ((void) 0)
#endif

struct CircResD {              // CircRes::set_bp (src/common.cpp:468-477); chr / rname / type are the caller's
    uint32_t spos, epos;
    uint8_t ss[2], es[2];      // start_signal, end_signal
    uint8_t sref[2], eref[2];  // start_bp_ref, end_bp_ref
    uint8_t ss_n, es_n, ref_n, was_set;
};

struct CircSplitCtx {          // the ProcessCirc members these functions read and write
    ExtCtx *x;
    const uint8_t *remain_seq, *fullmap_seq, *r1_seq, *r2_seq;
    int remain_len, fullmap_len, r1_len, r2_len;
    int depth;
    int overflow;              // a fixed-size list was too small: the result is not to be used
};

struct CircSig {               // a string of 0 or 2 characters, as scalars (no byte array to index)
    uint32_t c0, c1, n;
};
CM_HD CircSig cm_sig2(const uint8_t *seq, int len, int i) {   // string() + seq[i] + seq[i + 1]
    CircSig s;
    s.n = 2;
    s.c0 = (i >= 0 && i < len) ? (uint32_t) seq[i] : 0u;
    s.c1 = (i + 1 >= 0 && i + 1 < len) ? (uint32_t) seq[i + 1] : 0u;
    return s;
}
CM_HD CircSig cm_sig_empty() { CircSig s; s.n = 0; s.c0 = s.c1 = 0; return s; }
// get_consensus(s1, s2) (src/utils.cpp:759-769)
CM_HD CircSig cm_sig_consensus(const CircSig &a, const CircSig &b) {
    CircSig r = cm_sig_empty();
    if (a.n != b.n) return r;
    r.n = a.n;
    if (a.n >= 1) r.c0 = a.c0 == b.c0 ? a.c0 : (uint32_t) 'N';
    if (a.n >= 2) r.c1 = a.c1 == b.c1 ? a.c1 : (uint32_t) 'N';
    return r;
}
CM_HD void cm_circ_set_bp(const DevView &v, CircResD &cr, uint32_t beg_bp, uint32_t end_bp, const CircSig &ss, const CircSig &es) {
    cr.spos = beg_bp;
    cr.epos = end_bp;
    cr.ss_n = (uint8_t) ss.n; cr.ss[0] = (uint8_t) ss.c0; cr.ss[1] = (uint8_t) ss.c1;
    cr.es_n = (uint8_t) es.n; cr.es[0] = (uint8_t) es.c0; cr.es[1] = (uint8_t) es.c1;
    // pac2char_otf(beg_bp, 2) / pac2char_otf(end_bp - 1, 2)
    const uint8_t *a = cm_pac2char(v, beg_bp, 2), *b = cm_pac2char(v, end_bp - 1, 2);
    cr.ref_n = (a && b) ? 2 : 0;
    cr.sref[0] = a ? a[0] : 0; cr.sref[1] = a ? a[1] : 0;
    cr.eref[0] = b ? b[0] : 0; cr.eref[1] = b ? b[1] : 0;
    cr.was_set = 1;
}

// MatchedMate::merge_to_right: the interval pointer of the end is NOT taken over (only the look-up flag and the index)
CM_HD bool cm_mate_merge_to_right(MateD &m, const MateD &r, const cm_params &p) {
    if (m.dir != r.dir) return false;
    m.epos = r.epos;
    m.qepos = r.qepos;
    m.middle_ed += m.right_ed + r.left_ed;
    m.right_ed = r.right_ed;
    m.matched_len += r.matched_len + (uint32_t) m.sclen_right + (uint32_t) r.sclen_left;
    m.middle_ed += m.sclen_right + r.sclen_left;
    m.sclen_right = r.sclen_right;
    m.right_ok = r.right_ok;
    m.looked_up_epos = r.looked_up_epos;
    m.exon_ind_epos = r.exon_ind_epos;
    return m.left_ed + m.middle_ed + m.right_ed <= p.max_ed;
}

// one base of the contig as pac2char_otf(pos, 1) gives it (0 when out of range)
CM_HD uint8_t cm_circ_base(const DevView &v, uint32_t pos) {
    const uint8_t *q = cm_pac2char(v, pos, 1);
    return q ? q[0] : 0;
}

// split_realignment around a breakpoint [beg_bp, end_bp] of one transcript (process_circ.cpp:1343-1392)
CM_HDN int cm_circ_split_realign7(CircSplitCtx &c, uint32_t qcutpos, uint32_t beg_bp, uint32_t end_bp, const uint8_t *seq, uint32_t seq_len,
                                  uint32_t tid) {
    ExtCtx &x = *c.x;
    const cm_params &p = x.v->p;
    if (qcutpos <= 0 || qcutpos >= seq_len) return p.max_ed + 1;
    const int last_bp_err = seq[qcutpos - 1] == cm_circ_base(*x.v, end_bp) ? 0 : 1;
    const int first_bp_err = seq[qcutpos] == cm_circ_base(*x.v, beg_bp) ? 0 : 1;
    uint32_t lm_pos = end_bp, rm_pos = beg_bp;
    const uint32_t lb = beg_bp, ub = end_bp;
    AlignResD bl = ar_init(lb), br = ar_init(ub);
    const TidSet ts = cm_tid_single(tid);
    const bool lok = cm_extend_left<true>(x, ts, seq, lm_pos, (int) qcutpos - 1, p.max_ed - last_bp_err, lb, bl);
    const bool rok = cm_extend_right<true>(x, ts, seq + qcutpos + 1, rm_pos, (int) (seq_len - qcutpos - 1), p.max_ed - first_bp_err, ub, br);
    bl.ed += last_bp_err;
    br.ed += first_bp_err;
    if (lok && rok && bl.ed + br.ed <= p.max_ed) return bl.ed + br.ed;
    return p.max_ed + 1;
}

// ---- get_junctions (src/utils.cpp:697-757) --------------------------------------------------------------------------------
struct CircJuncs {
    uint32_t count;
    uint32_t beg[CM_CIRC_JUNC_CAP], end[CM_CIRC_JUNC_CAP], bp_matched[CM_CIRC_JUNC_CAP];
};
CM_HD void cm_juncs_push(CircJuncs &j, uint32_t b, uint32_t e, uint32_t m, int &overflow) {
    if (b >= e) return;                        // JunctionInfoList::push_back (src/common.h:206-213)
    if (j.count >= CM_CIRC_JUNC_CAP) { overflow = 1; return; }
    j.beg[j.count] = b; j.end[j.count] = e; j.bp_matched[j.count] = m;
    j.count++;
}
CM_HDN void cm_circ_get_junctions(const AnnotDev &a, MateD &mm, CircJuncs &ji, int &overflow) {
    cm_overlap_to_spos(a, mm);
    cm_overlap_to_epos(a, mm);
    ji.count = 0;
    if (mm.exons_spos < 0 || mm.exons_epos < 0) return;
    const int ivs = mm.exons_spos, ive = mm.exons_epos;
    for (uint32_t sg = a.iv_seg_off[ivs]; sg < a.iv_seg_off[ivs + 1]; sg++)
        for (uint32_t k = a.seg_tid_off[sg]; k < a.seg_tid_off[sg + 1]; k++) {
            const uint32_t tid = a.tid[k];
            const int start_ind = a.trans_start_ind[tid];
            const uint32_t start_table_ind = (uint32_t) (mm.exon_ind_spos - start_ind);   // unsigned in the reference: never < 0
            const uint32_t end_table_ind = (uint32_t) (mm.exon_ind_epos - start_ind);
            const uint32_t tlen = a.t2s_off[tid + 1] - a.t2s_off[tid];
            const uint8_t *t2s = a.t2s + a.t2s_off[tid];
            if (mm.exon_ind_epos < start_ind || end_table_ind >= tlen || t2s[end_table_ind] == 0) continue;
            if (start_table_ind == end_table_ind) return;                                  // no junction
            uint32_t junc_start = a.iv_epos[ivs];
            uint32_t covered = a.iv_epos[ivs] - mm.spos + 1;
            int this_it_ind = mm.exon_ind_spos;
            for (uint32_t q = start_table_ind + 1; q < end_table_ind; q++) {
                this_it_ind++;
                // trans2seg[tid][q] for q past the table is an out-of-range read in the reference; taken as 0 here
                if (q < tlen && t2s[q] != 0) {
                    const int r = cm_get_interval(a, this_it_ind);
                    if (r < 0) break;
                    cm_juncs_push(ji, junc_start, a.iv_spos[r], covered, overflow);
                    covered += a.iv_epos[r] - a.iv_spos[r] + 1;
                    junc_start = a.iv_epos[r];
                }
            }
            cm_juncs_push(ji, junc_start, a.iv_spos[ive], covered, overflow);
            covered += mm.epos - a.iv_spos[ive] + 1;
            int d = (int) (covered - mm.matched_len);
            if ((d < 0 ? -d : d) <= CM_INDELTH) return;
            ji.count = 0;
        }
}

// (transcript, offset) pairs of the exon ends / starts within BPRES of a mate's end / start, walking over the intervals the mate
// covers (process_circ.cpp:1196-1242 and :996-1031)
struct CircTidList {
    uint32_t n;
    uint32_t tid[CM_CIRC_TID_CAP];
    int32_t diff[CM_CIRC_TID_CAP];
};
// ends: intervals ind, ind - 1, ... while walk_from < interval.epos; diff = edge - seg.end
CM_HD void cm_circ_end_tids(const AnnotDev &a, int ind, uint32_t walk_from, uint32_t edge, CircTidList &L, int &overflow) {
    L.n = 0;
    int r = cm_get_interval(a, ind);
    while (r >= 0 && walk_from < a.iv_epos[r]) {
        for (uint32_t sg = a.iv_seg_off[r]; sg < a.iv_seg_off[r + 1]; sg++) {
            const int diff = (int) (edge - a.seg_end[sg]);
            if ((diff < 0 ? -diff : diff) <= CM_BPRES)
                for (uint32_t k = a.seg_tid_off[sg]; k < a.seg_tid_off[sg + 1]; k++) {
                    if (L.n >= CM_CIRC_TID_CAP) { overflow = 1; break; }
                    L.tid[L.n] = a.tid[k]; L.diff[L.n] = diff; L.n++;
                }
        }
        r = cm_get_interval(a, --ind);
    }
}
// starts: intervals ind, ind + 1, ... while walk_to > interval.spos; diff = edge - seg.start
CM_HD void cm_circ_start_tids(const AnnotDev &a, int ind, uint32_t walk_to, uint32_t edge, CircTidList &L, int &overflow) {
    L.n = 0;
    int r = cm_get_interval(a, ind);
    while (r >= 0 && walk_to > a.iv_spos[r]) {
        for (uint32_t sg = a.iv_seg_off[r]; sg < a.iv_seg_off[r + 1]; sg++) {
            const int diff = (int) (edge - a.seg_start[sg]);
            if ((diff < 0 ? -diff : diff) <= CM_BPRES)
                for (uint32_t k = a.seg_tid_off[sg]; k < a.seg_tid_off[sg + 1]; k++) {
                    if (L.n >= CM_CIRC_TID_CAP) { overflow = 1; break; }
                    L.tid[L.n] = a.tid[k]; L.diff[L.n] = diff; L.n++;
                }
        }
        r = cm_get_interval(a, ++ind);
    }
}

// A call of final_check that one of the functions below ends in (its value is theirs): handed back to the driver loop
struct CircNested {
    bool set;
    MateD full, left, right;
};

// ---- final_check (process_circ.cpp:1135-1341) without its last line: full = the mate that is not split, left / right = the
// two sides of the split read; rescue = true when the reference goes on to rescue_overlapping_bsj
CM_HDN int cm_circ_final_check_body(CircSplitCtx &c, MateD &full, MateD &left, MateD &right, CircResD &cr, bool &rescue) {
    rescue = false;
    ExtCtx &x = *c.x;
    const cm_params &p = x.v->p;
    const AnnotDev &a = x.v->a;
    if (left.epos < right.spos) {
        CM_CIRC_HIT(0);
        if (full.dir == 1) {
            if (full.spos <= left.spos) return CM_FR;
            else if (full.epos >= right.epos) return CM_RF;
        }
        if (full.dir == -1) {
            if (full.epos >= right.epos) return CM_FR;
            else if (full.spos <= left.spos) return CM_RF;
        }
    } else if (right.spos <= left.spos && left.epos >= right.epos) {
        if (full.spos < right.spos) {
            const int off = (int) (right.spos - full.spos), sc_remained = p.max_sc - full.sclen_left;
            if (off <= sc_remained) {
                full.spos = right.spos;
                full.sclen_left += off;
                full.qspos += (uint32_t) off;
                full.matched_len -= (uint32_t) off;
            }
        }
        if (full.epos > left.epos) {
            const int off = (int) (full.epos - left.epos), sc_remained = p.max_sc - full.sclen_right;
            if (off <= sc_remained) {
                full.epos = left.epos;
                full.sclen_right += off;
                full.qepos -= (uint32_t) off;
                full.matched_len -= (uint32_t) off;
            }
        }
        if (full.spos >= right.spos && full.epos <= left.epos) {
            CM_CIRC_HIT(1);
            cm_overlap_to_spos(a, full);  cm_overlap_to_epos(a, full);
            cm_overlap_to_spos(a, right); cm_overlap_to_epos(a, right);
            cm_overlap_to_spos(a, left);  cm_overlap_to_epos(a, left);
            CircTidList end_tids, start_tids;
            cm_circ_end_tids(a, left.exon_ind_epos, left.spos, left.epos + (uint32_t) left.sclen_right, end_tids, c.overflow);
            cm_circ_start_tids(a, right.exon_ind_spos, right.epos, right.spos - (uint32_t) right.sclen_left, start_tids, c.overflow);
            int best_ed = p.max_ed + 1;
            for (uint32_t i = 0; i < start_tids.n; i++)
                for (uint32_t j = 0; j < end_tids.n; j++) {
                    const int sdiff = start_tids.diff[i], ediff = end_tids.diff[j];
                    if (start_tids.tid[i] != end_tids.tid[j] || sdiff != ediff) continue;
                    const uint32_t tid = start_tids.tid[i];
                    const uint32_t qcutpos = left.qepos + (uint32_t) left.sclen_right - (uint32_t) ediff;
                    const uint32_t beg_bp = right.spos - (uint32_t) right.sclen_left - (uint32_t) sdiff;
                    const uint32_t end_bp = left.epos + (uint32_t) left.sclen_right - (uint32_t) ediff;
                    if (full.sclen_right > 0) {
                        if (full.epos + (uint32_t) full.sclen_right > end_bp) {   // the other mate crosses end_bp: split realignment
                            const uint32_t fm_qcutpos = full.qepos + (end_bp - full.epos);
                            if (cm_circ_split_realign7(c, fm_qcutpos, beg_bp, end_bp, c.fullmap_seq, (uint32_t) c.fullmap_len, tid) > p.max_ed) continue;
                        } else if (full.sclen_right > p.max_sc)
                            continue;
                    }
                    if (full.sclen_left > 0) {
                        if (full.spos - (uint32_t) full.sclen_left < beg_bp) {
                            const uint32_t fm_qcutpos = (uint32_t) full.sclen_left + (full.spos - beg_bp);
                            if (cm_circ_split_realign7(c, fm_qcutpos, beg_bp, end_bp, c.fullmap_seq, (uint32_t) c.fullmap_len, tid) > p.max_ed) continue;
                        } else if (full.sclen_left > p.max_sc)
                            continue;
                    }
                    const int ed = cm_circ_split_realign7(c, qcutpos, beg_bp, end_bp, c.remain_seq, (uint32_t) c.remain_len, tid);
                    if (ed < best_ed) {
                        const CircSig es = cm_sig2(c.remain_seq, c.remain_len, (int) qcutpos - 2), ss = cm_sig2(c.remain_seq, c.remain_len, (int) qcutpos);
                        cm_circ_set_bp(*x.v, cr, beg_bp, end_bp, ss, es);
                        if (ed == 0) return CM_CR;
                        best_ed = ed;
                    }
                }
            if (best_ed <= p.max_ed) return CM_CR;
            const uint32_t qcutpos = left.qepos + (uint32_t) left.sclen_right;
            const uint32_t beg_bp = right.spos - (uint32_t) right.sclen_left;
            const uint32_t end_bp = left.epos + (uint32_t) left.sclen_right;
            if (qcutpos < 2 || qcutpos > (uint32_t) (c.remain_len - 2)) return CM_MCR;
            // the names are swapped here in the reference (ssignal takes the two bases before the cut)
            const CircSig ss = cm_sig2(c.remain_seq, c.remain_len, (int) qcutpos - 2), es = cm_sig2(c.remain_seq, c.remain_len, (int) qcutpos);
            cm_circ_set_bp(*x.v, cr, beg_bp, end_bp, ss, es);
            if (start_tids.n > 0 && end_tids.n > 0) return CM_NCR;
            return CM_MCR;
        }
    }
    CM_CIRC_HIT(2);
    rescue = true;
    return CM_UD;
}

// same_transcript(segments, 2, common_tid) (src/utils.cpp:421-445): start/start, start/end, end/start, end/end
CM_HD bool cm_circ_same_transcript2(const AnnotDev &a, MateD &m0, MateD &m1, TidSet &ts) {
    cm_overlap_to_spos(a, m0); cm_overlap_to_spos(a, m1);
    cm_overlap_to_epos(a, m0); cm_overlap_to_epos(a, m1);
    const int c0[4] = {m0.exons_spos, m0.exons_spos, m0.exons_epos, m0.exons_epos};
    const int c1[4] = {m1.exons_spos, m1.exons_epos, m1.exons_spos, m1.exons_epos};
    for (int k = 0; k < 4; k++)
        if (cm_same_transcript(a, c0[k], c1[k])) { ts.iv_a = c0[k]; ts.iv_b = c1[k]; return true; }
    ts.iv_a = ts.iv_b = -1;
    return false;
}

// split_realignment of the full mate at qcutpos (process_circ.cpp:1394-1489)
CM_HDN int cm_circ_check_split_map4(CircSplitCtx &c, MateD &r1_1, MateD &r2_1, MateD &r1_2, MateD &r2_2, CircResD &cr, CircNested &nest);
CM_HDN int cm_circ_split_realign5(CircSplitCtx &c, uint32_t qcutpos, MateD &full, MateD &left, MateD &right, CircResD &cr, CircNested &nest) {
    ExtCtx &x = *c.x;
    const cm_params &p = x.v->p;
    const AnnotDev &a = x.v->a;
    const uint32_t flen = (uint32_t) c.fullmap_len;
    if (qcutpos <= 0 || qcutpos >= flen) return CM_UD;
    qcutpos += full.qspos - 1;
    if (qcutpos <= 0 || qcutpos >= flen) return CM_UD;
    CM_CIRC_HIT(3);
    TidSet ts;
    if (!cm_circ_same_transcript2(a, left, right, ts)) return CM_UD;
    const int last_bp_err = c.fullmap_seq[qcutpos - 1] == cm_circ_base(*x.v, left.epos) ? 0 : 1;
    const int first_bp_err = c.fullmap_seq[qcutpos] == cm_circ_base(*x.v, right.spos) ? 0 : 1;
    uint32_t lm_pos = left.epos, rm_pos = right.spos;
    const uint32_t lb = right.spos, ub = left.epos;
    AlignResD bl = ar_init(lb), br = ar_init(ub);
    const bool lok = cm_extend_left<true>(x, ts, c.fullmap_seq, lm_pos, (int) qcutpos - 1, p.max_ed - last_bp_err, lb, bl);
    const bool rok = cm_extend_right<true>(x, ts, c.fullmap_seq + qcutpos + 1, rm_pos, (int) (flen - qcutpos - 1), p.max_ed - first_bp_err, ub, br);
    bl.ed += last_bp_err;
    br.ed += first_bp_err;
    CM_CIRC_HIT(4);
    if (!lok || !rok || bl.ed + br.ed > p.max_ed) return CM_UD;
    CM_CIRC_HIT(5);
    MateD nl, nr;                 // MatchedMate(): what is not assigned below keeps the constructor's value
    mate_init(nl, p);
    mate_init(nr, p);
    nl.spos = lm_pos; nl.epos = left.epos;
    nl.qspos = (uint32_t) bl.sclen; nl.qepos = qcutpos;
    nl.dir = full.dir;
    nl.matched_len = qcutpos - (uint32_t) bl.sclen;
    nl.sclen_left = bl.sclen; nl.sclen_right = 0;
    nl.left_ed = bl.ed; nl.right_ed = 0; nl.middle_ed = 0;
    nl.left_ok = true; nl.right_ok = true;
    nr.spos = right.spos; nr.epos = rm_pos;
    nr.qspos = qcutpos + 1; nr.qepos = flen - (uint32_t) br.sclen;
    nr.dir = full.dir;
    nr.matched_len = flen - qcutpos - (uint32_t) br.sclen;
    nr.sclen_left = 0; nr.sclen_right = br.sclen;
    nr.left_ed = 0; nr.right_ed = br.ed; nr.middle_ed = 0;
    nr.left_ok = true; nr.right_ok = true;
    c.r1_seq = c.remain_seq;  c.r1_len = c.remain_len;
    c.r2_seq = c.fullmap_seq; c.r2_len = c.fullmap_len;
    return cm_circ_check_split_map4(c, right, nr, left, nl, cr, nest);
}

// check_split_map for overlapping split mates (process_circ.cpp:922-1133)
CM_HDN int cm_circ_check_split_map4(CircSplitCtx &c, MateD &r1_1, MateD &r2_1, MateD &r1_2, MateD &r2_2, CircResD &cr, CircNested &nest) {
    ExtCtx &x = *c.x;
    const cm_params &p = x.v->p;
    const AnnotDev &a = x.v->a;
    const int r1_ed = r1_1.right_ed + r1_1.left_ed + r1_1.middle_ed + r1_2.right_ed + r1_2.left_ed + r1_2.middle_ed;
    const int r2_ed = r2_1.right_ed + r2_1.left_ed + r2_1.middle_ed + r2_2.right_ed + r2_2.left_ed + r2_2.middle_ed;
    if (r1_ed > p.max_ed || r2_ed > p.max_ed) return CM_UD;
    MateD r1_l = (r1_1.spos <= r1_2.spos) ? r1_1 : r1_2, r1_r = (r1_1.spos <= r1_2.spos) ? r1_2 : r1_1;
    MateD r2_l = (r2_1.spos <= r2_2.spos) ? r2_1 : r2_2, r2_r = (r2_1.spos <= r2_2.spos) ? r2_2 : r2_1;
    const bool r1_regular = r1_l.qspos < r1_r.qspos, r2_regular = r2_l.qspos < r2_r.qspos;
    CM_CIRC_HIT(6);
    if (r1_regular && r2_regular) {                       // RF or FR
        CM_CIRC_HIT(7);
        if (r1_l.dir == 1) {
            if (r1_r.spos <= r2_l.spos) return CM_FR;
            else if (r1_l.epos >= r2_r.epos) return CM_RF;
        }
        if (r1_l.dir == -1) {
            if (r2_r.spos <= r1_l.spos) return CM_FR;
            else if (r2_l.epos >= r1_r.epos) return CM_RF;
        }
    } else if (r1_regular && !r2_regular) {               // single BSJ on R2
        MateD full = r1_l;
        if (!cm_mate_merge_to_right(full, r1_r, p)) return CM_UD;
        CM_CIRC_HIT(8);
        c.remain_seq = c.r2_seq;
        c.remain_len = c.r2_len;
        nest.set = true; nest.full = full; nest.left = r2_l; nest.right = r2_r;   // return final_check(full, r2_l, r2_r, cr)
        return CM_UD;
    } else if (!r1_regular && r2_regular) {               // single BSJ on R1
        MateD full = r2_l;
        if (!cm_mate_merge_to_right(full, r2_r, p)) return CM_UD;
        CM_CIRC_HIT(9);
        c.remain_seq = c.r1_seq;
        c.remain_len = c.r1_len;
        nest.set = true; nest.full = full; nest.left = r1_l; nest.right = r1_r;   // return final_check(full, r1_l, r1_r, cr)
        return CM_UD;
    } else {                                              // BSJ on the overlap
        CM_CIRC_HIT(10);
        if (r1_l.spos == r2_l.spos && r1_r.epos == r2_r.epos) {
            CM_CIRC_HIT(11);
            cm_overlap_to_spos(a, r1_l);
            cm_overlap_to_epos(a, r1_r);
            CircTidList end_tids, start_tids;
            cm_circ_end_tids(a, r1_r.exon_ind_epos, r1_r.spos, r1_r.epos + (uint32_t) r1_r.sclen_right, end_tids, c.overflow);
            cm_circ_start_tids(a, r1_l.exon_ind_spos, r1_l.epos, r1_l.spos - (uint32_t) r1_l.sclen_left, start_tids, c.overflow);
            int best_ed1 = p.max_ed + 1, best_ed2 = p.max_ed + 1;
            CircSig ss1 = cm_sig_empty(), es1 = cm_sig_empty(), ss2 = cm_sig_empty(), es2 = cm_sig_empty();
            for (uint32_t i = 0; i < start_tids.n; i++)
                for (uint32_t j = 0; j < end_tids.n; j++) {
                    const int sdiff = start_tids.diff[i], ediff = end_tids.diff[j];
                    if (start_tids.tid[i] != end_tids.tid[j] || sdiff != ediff) continue;
                    const uint32_t tid = start_tids.tid[i];
                    const uint32_t beg_bp = r1_l.spos - (uint32_t) r1_l.sclen_left - (uint32_t) sdiff;
                    const uint32_t end_bp = r1_r.epos + (uint32_t) r1_r.sclen_right - (uint32_t) ediff;
                    uint32_t qcutpos = r1_r.qepos + (uint32_t) r1_r.sclen_right - (uint32_t) ediff;
                    const int ed1 = cm_circ_split_realign7(c, qcutpos, beg_bp, end_bp, c.r1_seq, (uint32_t) c.r1_len, tid);
                    if (qcutpos < 2 || qcutpos + 2 > (uint32_t) c.r1_len) { es1 = cm_sig_empty(); ss1 = cm_sig_empty(); }
                    else { es1 = cm_sig2(c.r1_seq, c.r1_len, (int) qcutpos - 2); ss1 = cm_sig2(c.r1_seq, c.r1_len, (int) qcutpos); }
                    qcutpos = r2_r.qepos + (uint32_t) r2_r.sclen_right - (uint32_t) ediff;
                    const int ed2 = cm_circ_split_realign7(c, qcutpos, beg_bp, end_bp, c.r2_seq, (uint32_t) c.r2_len, tid);
                    if (qcutpos < 2 || qcutpos + 2 > (uint32_t) c.r2_len) { ss2 = cm_sig_empty(); es2 = cm_sig_empty(); }
                    else { es2 = cm_sig2(c.r2_seq, c.r2_len, (int) qcutpos - 2); ss2 = cm_sig2(c.r2_seq, c.r2_len, (int) qcutpos); }
                    if (ed1 < best_ed1 && ed2 < best_ed2) {
                        if (ss1.n == 0) cm_circ_set_bp(*x.v, cr, beg_bp, end_bp, ss2, es2);
                        else if (ss2.n == 0) cm_circ_set_bp(*x.v, cr, beg_bp, end_bp, ss1, es1);
                        else cm_circ_set_bp(*x.v, cr, beg_bp, end_bp, cm_sig_consensus(ss1, ss2), cm_sig_consensus(es1, es2));
                        best_ed1 = ed1;
                        best_ed2 = ed2;
                    }
                }
            if (best_ed1 <= p.max_ed && best_ed2 <= p.max_ed) return CM_CR;
            const uint32_t qcutpos = r1_r.qepos + (uint32_t) r1_r.sclen_right;
            const uint32_t beg_bp = r1_l.spos - (uint32_t) r1_l.sclen_left;
            const uint32_t end_bp = r1_r.epos + (uint32_t) r1_r.sclen_right;
            if (qcutpos < 2 || qcutpos > (uint32_t) (c.r1_len - 2) || qcutpos > (uint32_t) (c.r2_len - 2)) return CM_MCR;
            es1 = cm_sig2(c.r1_seq, c.r1_len, (int) qcutpos - 2); ss1 = cm_sig2(c.r1_seq, c.r1_len, (int) qcutpos);
            es2 = cm_sig2(c.r2_seq, c.r2_len, (int) qcutpos - 2); ss2 = cm_sig2(c.r2_seq, c.r2_len, (int) qcutpos);
            cm_circ_set_bp(*x.v, cr, beg_bp, end_bp, cm_sig_consensus(ss1, ss2), cm_sig_consensus(es1, es2));
            if (start_tids.n > 0 && end_tids.n > 0) return CM_NCR;
            return CM_MCR;
        }
    }
    return CM_UD;
}

// final_check with rescue_overlapping_bsj (process_circ.cpp:1491-1552) as a loop: level 0 works on the caller's mates, a nested
// final_check (the tail of the four-mate check_split_map) on copies kept in its frame.  blk[k] = the next block of the rescue
// of level k to try (0: start of the breakpoint inside the full mate, 1: end of it, 2: none left).
CM_HDN int cm_circ_final_check(CircSplitCtx &c, MateD &full0, MateD &left0, MateD &right0, CircResD &cr) {
    const AnnotDev &a = c.x->v->a;
    MateD F[CM_CIRC_MAX_DEPTH], L[CM_CIRC_MAX_DEPTH], R[CM_CIRC_MAX_DEPTH];
    int blk[CM_CIRC_MAX_DEPTH];
    int lvl = 0;
    F[0] = full0; L[0] = left0; R[0] = right0;
    int value = CM_UD;
    bool have_value = false;       // value = what final_check of level lvl returned
    bool enter = true;             // run final_check's body on level lvl
    CircJuncs ji;
    for (;;) {
        if (enter) {
            bool rescue;
            value = cm_circ_final_check_body(c, F[lvl], L[lvl], R[lvl], cr, rescue);
            enter = false;
            if (rescue) { blk[lvl] = 0; have_value = false; }
            else have_value = true;
        }
        if (!have_value) {
            // the rescue of level lvl: next block
            MateD &full = F[lvl], &left = L[lvl], &right = R[lvl];
            int res = CM_UD;
            bool ran = false;
            CircNested nest;
            nest.set = false;
            while (blk[lvl] < 2 && !ran) {
                const int b = blk[lvl]++;
                if (b == 0 && right.spos <= full.epos && right.spos > full.spos) {
                    cm_circ_get_junctions(a, full, ji, c.overflow);
                    uint32_t qcutpos = 0;
                    for (uint32_t i = 0; i < ji.count; i++)
                        if (ji.end[i] == right.spos) qcutpos = ji.bp_matched[i];
                    if (qcutpos == 0) qcutpos = right.spos - full.spos;            // intron retention
                    res = cm_circ_split_realign5(c, qcutpos, full, left, right, cr, nest);
                    ran = true;
                } else if (b == 1 && left.epos >= full.spos && left.epos < full.epos) {
                    cm_circ_get_junctions(a, full, ji, c.overflow);
                    uint32_t qcutpos = 0;
                    for (uint32_t i = 0; i < ji.count; i++)
                        if (ji.beg[i] == left.epos) qcutpos = ji.bp_matched[i];
                    if (qcutpos == 0) qcutpos = full.matched_len - (full.epos - left.epos);
                    res = cm_circ_split_realign5(c, qcutpos, full, left, right, cr, nest);
                    ran = true;
                }
            }
            if (!ran) { value = CM_UD; have_value = true; }                        // rescue returns UD
            else if (nest.set) {
                if (lvl + 1 >= CM_CIRC_MAX_DEPTH) { c.overflow = 1; value = CM_UD; have_value = true; lvl = 0; break; }
                lvl++;
                F[lvl] = nest.full; L[lvl] = nest.left; R[lvl] = nest.right;
                enter = true;
                continue;
            } else if (res == CM_CR) { value = CM_CR; have_value = true; }         // rescue returns CR
            else continue;                                                         // next block of the same rescue
        }
        // final_check of level lvl returned `value`
        if (lvl == 0) break;
        lvl--;                      // it was the value of a block of the rescue of the level above
        if (value == CM_CR) have_value = true;          // that rescue returns CR, and so does its final_check
        else have_value = false;                        // go on with that rescue's next block
    }
    full0 = F[0]; left0 = L[0]; right0 = R[0];
    return value;
}

// check_split_map for non-overlapping split mates (process_circ.cpp:892-919); r1 / r2 may be trimmed by final_check
CM_HDN int cm_circ_check_split_map3(CircSplitCtx &c, MateD &r1, MateD &r2, MateD &pm, bool r1_partial, CircResD &cr) {
    const cm_params &p = c.x->v->p;
    int valid, split_read_ed;
    if (r1_partial) {
        split_read_ed = r1.right_ed + r1.left_ed + r1.middle_ed + pm.right_ed + pm.left_ed + pm.middle_ed;
        valid = (r1.qspos < pm.qspos) ? cm_circ_final_check(c, r2, r1, pm, cr) : cm_circ_final_check(c, r2, pm, r1, cr);
    } else {
        split_read_ed = r2.right_ed + r2.left_ed + r2.middle_ed + pm.right_ed + pm.left_ed + pm.middle_ed;
        valid = (r2.qspos < pm.qspos) ? cm_circ_final_check(c, r1, r2, pm, cr) : cm_circ_final_check(c, r1, pm, r2, cr);
    }
    if (split_read_ed > p.max_ed) valid = CM_UD;
    return valid;
}

// ---- one row of cm_gpu_circ_check_batch -----------------------------------------------------------------------------------
CM_HD void cm_circ_mate_in(const cm_circ_mate &i, MateD &m) {
    m.spos = i.m.spos; m.epos = i.m.epos; m.qspos = i.m.qspos; m.qepos = i.m.qepos; m.matched_len = i.m.matched_len;
    m.left_ed = i.m.left_ed; m.middle_ed = i.m.middle_ed; m.right_ed = i.m.right_ed; m.sclen_left = i.m.sclen_left; m.sclen_right = i.m.sclen_right;
    m.dir = i.m.dir; m.type = i.m.type;
    m.is_concord = i.m.is_concord != 0; m.left_ok = i.m.left_ok != 0; m.right_ok = i.m.right_ok != 0;
    m.looked_up_spos = i.looked_up_spos != 0; m.looked_up_epos = i.looked_up_epos != 0;
    m.exons_spos = i.exons_spos; m.exons_epos = i.exons_epos; m.exon_ind_spos = i.exon_ind_spos; m.exon_ind_epos = i.exon_ind_epos;
}
CM_HD void cm_circ_mate_out(const MateD &m, cm_circ_mate &o) {
    o.m.spos = m.spos; o.m.epos = m.epos; o.m.qspos = m.qspos; o.m.qepos = m.qepos; o.m.matched_len = m.matched_len;
    o.m.left_ed = m.left_ed; o.m.middle_ed = m.middle_ed; o.m.right_ed = m.right_ed; o.m.sclen_left = m.sclen_left; o.m.sclen_right = m.sclen_right;
    o.m.dir = m.dir; o.m.type = m.type;
    o.m.is_concord = m.is_concord; o.m.left_ok = m.left_ok; o.m.right_ok = m.right_ok; o.m.pad_ = 0;
    o.looked_up_spos = m.looked_up_spos; o.looked_up_epos = m.looked_up_epos; o.pad_[0] = o.pad_[1] = 0;
    o.exons_spos = m.exons_spos; o.exons_epos = m.exons_epos; o.exon_ind_spos = m.exon_ind_spos; o.exon_ind_epos = m.exon_ind_epos;
}
// v.p.kmer must be window_size; fullmap / remain = the two reads as call_circ_single_split orients them
// remain_is_fullmap (nullable): the four-mate path may leave ProcessCirc::remain_seq pointing at the other read
// (process_circ.cpp:973, :983), which the next chain of the same read then sees
CM_HDN int cm_circ_check_row(const DevView &v, const uint8_t *remain_seq, int remain_len, const uint8_t *fullmap_seq, int fullmap_len,
                             bool r1_partial, cm_circ_mate &r1, cm_circ_mate &r2, cm_circ_mate &pm, cm_circ_res &out,
                             uint8_t *remain_is_fullmap = nullptr) {
    ExtCtx x;
    ext_ctx_init(x, &v);
    CircSplitCtx c;
    c.x = &x;
    c.remain_seq = remain_seq; c.remain_len = remain_len;
    c.fullmap_seq = fullmap_seq; c.fullmap_len = fullmap_len;
    c.r1_seq = c.r2_seq = nullptr; c.r1_len = c.r2_len = 0;     // call_circ sets them to NULL / 0 (process_circ.cpp:335-343)
    c.depth = 0;
    c.overflow = 0;
    MateD a, b, p;
    cm_circ_mate_in(r1, a);
    cm_circ_mate_in(r2, b);
    cm_circ_mate_in(pm, p);
    CircResD cr;
    cr.spos = cr.epos = 0;
    cr.ss_n = cr.es_n = cr.ref_n = cr.was_set = 0;
    cr.ss[0] = cr.ss[1] = cr.es[0] = cr.es[1] = cr.sref[0] = cr.sref[1] = cr.eref[0] = cr.eref[1] = 0;
    const int ret = cm_circ_check_split_map3(c, a, b, p, r1_partial, cr);
    cm_circ_mate_out(a, r1);
    cm_circ_mate_out(b, r2);
    cm_circ_mate_out(p, pm);
    if (remain_is_fullmap) *remain_is_fullmap = (c.remain_seq == fullmap_seq && fullmap_seq != remain_seq) ? 1 : 0;
    out.spos = cr.spos; out.epos = cr.epos;
    for (int k = 0; k < 2; k++) {
        out.start_signal[k] = cr.ss[k]; out.end_signal[k] = cr.es[k]; out.start_bp_ref[k] = cr.sref[k]; out.end_bp_ref[k] = cr.eref[k];
    }
    out.start_signal_n = cr.ss_n; out.end_signal_n = cr.es_n; out.ref_n = cr.ref_n; out.was_set = cr.was_set;
    return c.overflow ? -1 : ret;
}

// find_exact_coord with the pair's mates as ProcessCirc keeps them from chain to chain (their cached start look-ups survive a
// trim by final_check); pm comes back with the caches find_exact_coord leaves on it
CM_HDN int cm_circ_exact_row(const DevView &v, cm_circ_mate &r1, cm_circ_mate &r2, cm_circ_mate &pm_out, int dir, uint32_t qspos, const uint8_t *rseq,
                             int rlen, int whole_len, const ChainView &bc) {
    ExtCtx x;
    ext_ctx_init(x, &v);
    MateD a, b, pm;
    cm_circ_mate_in(r1, a);
    cm_circ_mate_in(r2, b);
    mate_init(pm, v.p);
    const int ret = cm_circ_exact(x, a, b, pm, dir, qspos, rseq, rlen, whole_len, bc) ? 1 : 0;
    cm_circ_mate_out(a, r1);
    cm_circ_mate_out(b, r2);
    cm_circ_mate_out(pm, pm_out);
    return ret;
}

// ---- call_circ_double_split (process_circ.cpp:485-645): one (chain j of R1's unmapped part, chain k of R2's) -----------------
// same_transcript(segments, 4, common_tid) (src/utils.cpp:496-604): the sixteen start / end combinations in the reference's
// order (the fourth mate's end only in the second half); common_tid = transcripts of (seg0 & seg1) that are also in (seg2 & seg3).
// -> TidSet{i0, i1} with the two further intervals in x.ts_c / x.ts_d.
CM_HD bool cm_circ_same_transcript4(ExtCtx &x, MateD &m0, MateD &m1, MateD &m2, MateD &m3, TidSet &ts) {
    const AnnotDev &a = x.v->a;
    cm_overlap_to_spos(a, m0); cm_overlap_to_spos(a, m1); cm_overlap_to_spos(a, m2); cm_overlap_to_spos(a, m3);
    MateD c0 = m0, c1 = m1, c2 = m2, c3 = m3;     // the end look-ups are made on same_transcript's copies
    cm_overlap_to_epos(a, c0); cm_overlap_to_epos(a, c1); cm_overlap_to_epos(a, c2); cm_overlap_to_epos(a, c3);
    for (int e3 = 0; e3 < 2; e3++)
        for (int e0 = 0; e0 < 2; e0++)
            for (int e1 = 0; e1 < 2; e1++)
                for (int e2 = 0; e2 < 2; e2++) {
                    const int i0 = e0 ? c0.exons_epos : c0.exons_spos, i1 = e1 ? c1.exons_epos : c1.exons_spos,
                              i2 = e2 ? c2.exons_epos : c2.exons_spos, i3 = e3 ? c3.exons_epos : c3.exons_spos;
                    if (i0 < 0 || i1 < 0 || i2 < 0 || i3 < 0) continue;
                    const uint32_t b = a.seg_tid_off[a.iv_seg_off[i0]], e = a.seg_tid_off[a.iv_seg_off[i0 + 1]];
                    for (uint32_t k = b; k < e; k++) {
                        const uint32_t t = a.tid[k];
                        if (cm_iv_has_tid(a, i1, t) && cm_iv_has_tid(a, i2, t) && cm_iv_has_tid(a, i3, t)) {
                            ts.iv_a = i0; ts.iv_b = i1; x.ts_c = i2; x.ts_d = i3;
                            return true;
                        }
                    }
                }
    ts.iv_a = ts.iv_b = -1;
    x.ts_c = x.ts_d = -1;
    return false;
}

CM_HD void cm_circ_set_mm(const ChainView &ch, uint32_t qspos, int rlen, int dir, int ws, MateD &mm) {   // set_mm, :1713-1752
    mm.spos = ch.frags[0].rpos;
    mm.epos = ch.frags[ch.chain_len - 1].rpos + (uint32_t) ws - 1;
    mm.qspos = qspos;
    mm.qepos = qspos + (uint32_t) rlen - 1;
    mm.matched_len = (mm.qepos + 1 >= mm.qspos) ? (mm.qepos - mm.qspos + 1) : 0;
    mm.dir = dir;
}

// returns 0: no common transcript (continue), 1: extend_both_mates failed (continue), 2: extended (the caller goes on when both
// partial mates are CONCRD).  r1 / r2 keep the start look-ups made on them.
CM_HDN int cm_circ_double_row(const DevView &v, cm_circ_mate &r1io, cm_circ_mate &r2io, uint32_t r1_qspos, uint32_t r1_qepos, uint32_t r2_qspos,
                              uint32_t r2_qepos, const uint8_t *r1_seq, const uint8_t *r2_seq, const ChainView &c1, const ChainView &c2,
                              cm_circ_mate &p1_out, cm_circ_mate &p2_out) {
    ExtCtx x;
    ext_ctx_init(x, &v);
    const int ws = v.p.kmer;
    MateD r1, r2, p1, p2;
    cm_circ_mate_in(r1io, r1);
    cm_circ_mate_in(r2io, r2);
    mate_init(p1, v.p);
    mate_init(p2, v.p);
    cm_circ_set_mm(c1, r1_qspos, (int) (r1_qepos - r1_qspos + 1), r1.dir, ws, p1);
    cm_circ_set_mm(c2, r2_qspos, (int) (r2_qepos - r2_qspos + 1), r2.dir, ws, p2);
    TidSet ts;
    const bool same = cm_circ_same_transcript4(x, r1, r2, p1, p2, ts);
    cm_circ_mate_out(r1, r1io);
    cm_circ_mate_out(r2, r2io);
    int ret = 0;
    if (same) {
        bool ok;
        if (c1.frags[0].rpos <= c2.frags[0].rpos)
            ok = cm_extend_both_mates<true>(x, c1, c2, ts, r1_seq, r2_seq, (int) r1_qspos, (int) r2_qspos, (int) r1_qepos, (int) r2_qepos, p1, p2);
        else
            ok = cm_extend_both_mates<true>(x, c2, c1, ts, r2_seq, r1_seq, (int) r2_qspos, (int) r1_qspos, (int) r2_qepos, (int) r1_qepos, p2, p1);
        ret = ok ? 2 : 1;
    }
    cm_circ_mate_out(p1, p1_out);
    cm_circ_mate_out(p2, p2_out);
    return ret;
}

// check_split_map(mm_r1, mm_r2, r1_partial_mm, r2_partial_mm, cr) as call_circ_double_split calls it (:602): r1_seq / r2_seq are the
// two oriented reads; fullmap_seq is whatever an earlier call_circ_single_split of the same pair left (NULL / 0 when none)
CM_HDN int cm_circ_check4_row(const DevView &v, const uint8_t *r1_seq, int r1_len, const uint8_t *r2_seq, int r2_len, const uint8_t *fullmap_seq,
                              int fullmap_len, const cm_circ_mate &r1, const cm_circ_mate &r2, const cm_circ_mate &p1, const cm_circ_mate &p2,
                              cm_circ_res &out) {
    ExtCtx x;
    ext_ctx_init(x, &v);
    CircSplitCtx c;
    c.x = &x;
    c.remain_seq = nullptr; c.remain_len = 0;
    c.fullmap_seq = fullmap_seq; c.fullmap_len = fullmap_len;
    c.r1_seq = r1_seq; c.r1_len = r1_len;
    c.r2_seq = r2_seq; c.r2_len = r2_len;
    c.depth = 0;
    c.overflow = 0;
    MateD a, b, q1, q2;
    cm_circ_mate_in(r1, a); cm_circ_mate_in(r2, b); cm_circ_mate_in(p1, q1); cm_circ_mate_in(p2, q2);
    CircResD cr;
    cr.spos = cr.epos = 0;
    cr.ss_n = cr.es_n = cr.ref_n = cr.was_set = 0;
    cr.ss[0] = cr.ss[1] = cr.es[0] = cr.es[1] = cr.sref[0] = cr.sref[1] = cr.eref[0] = cr.eref[1] = 0;
    CircNested nest;
    nest.set = false;
    int ret = cm_circ_check_split_map4(c, a, b, q1, q2, cr, nest);
    if (nest.set) ret = cm_circ_final_check(c, nest.full, nest.left, nest.right, cr);   // the tail call of its two single-BSJ arms
    out.spos = cr.spos; out.epos = cr.epos;
    for (int k = 0; k < 2; k++) {
        out.start_signal[k] = cr.ss[k]; out.end_signal[k] = cr.es[k]; out.start_bp_ref[k] = cr.sref[k]; out.end_bp_ref[k] = cr.eref[k];
    }
    out.start_signal_n = cr.ss_n; out.end_signal_n = cr.es_n; out.ref_n = cr.ref_n; out.was_set = cr.was_set;
    return c.overflow ? -1 : ret;
}
