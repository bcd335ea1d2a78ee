// Transcript-aware extension of chains (TransExtension, src/extend.cpp:37-920) restated for one
// thread.  Genome text is read in place from ContigDev::genome (GenomeSeeder::pac2char only copies,
// src/match_read.cpp:288-299); an IntervalInfo<UniqSeg>* of the reference is an interval index
// (-1 == NULL); common_tid vectors are never materialised: they are the tids of one interval that
// also occur in another, enumerated in the reference's order (same_transcript, src/utils.cpp:335-354).
#pragma once
#include "cm_types.h"
#include "cm_annot.h"
#include "cm_align.h"

#define CM_MINLB 0u
#define CM_NEGINF (-1000000000)   // int(-INF), AlignRes(uint32_t p) (src/align.h:24)

struct AlignResD {   // AlignRes, src/align.h:12-121
    uint32_t pos;
    int ed, sclen, indel, qcovlen, rcovlen, score;
};

CM_HD void ar_set(AlignResD &a, uint32_t p, int e, int s, int i, int qc, int scr) {
    a.pos = p; a.ed = e; a.sclen = s; a.indel = i; a.qcovlen = qc; a.rcovlen = qc - i; a.score = scr;
}
CM_HD AlignResD ar_init(uint32_t p) {
    AlignResD a;
    a.pos = p; a.ed = 0; a.sclen = 0; a.indel = 0; a.qcovlen = 0; a.rcovlen = 0; a.score = CM_NEGINF;
    return a;
}
CM_HD void ar_update(AlignResD &a, int edit_dist, int sclength, uint32_t newpos, int indel, int qcovlen, int scr) {
    a.pos = newpos; a.ed += edit_dist; a.sclen = sclength; a.indel += indel; a.qcovlen += qcovlen;
    a.rcovlen += qcovlen - indel; a.score = scr;
}
CM_HD bool ar_update_by_score(AlignResD &a, const AlignResD &r, bool right) {
    if (a.score < r.score || (a.score == r.score && (right ? r.pos < a.pos : r.pos > a.pos))) {
        ar_set(a, r.pos, r.ed, r.sclen, r.indel, r.qcovlen, r.score);
        return true;
    }
    return false;
}
// AlignRes::update_right / update_left (src/align.h:73-115)
CM_HD void ar_update_side(AlignResD &a, const AlignResD &r, bool right, const cm_params &p) {
    if (r.qcovlen > a.qcovlen) {
        if (r.ed <= p.max_ed && r.sclen <= p.max_sc && 2 * (r.ed - a.ed) < (r.qcovlen - a.qcovlen))
            ar_set(a, r.pos, r.ed, r.sclen, r.indel, r.qcovlen, r.score);
    } else if (r.qcovlen < a.qcovlen) {
        if (r.ed <= p.max_ed && r.sclen <= p.max_sc && 2 * (a.ed - r.ed) >= (a.qcovlen - r.qcovlen))
            ar_set(a, r.pos, r.ed, r.sclen, r.indel, r.qcovlen, r.score);
    } else {
        if ((r.ed < a.ed) || (r.ed == a.ed && r.sclen < a.sclen) ||
            (r.ed == a.ed && r.sclen == a.sclen && (right ? r.pos < a.pos : r.pos > a.pos)))
            ar_set(a, r.pos, r.ed, r.sclen, r.indel, r.qcovlen, r.score);
    }
}

struct MateD {   // MatchedMate, src/common.h:260-307 (junc_num is always 0 on this path)
    uint32_t spos, epos, qspos, qepos;
    int right_ed, left_ed, middle_ed;
    int sclen_right, sclen_left;
    uint32_t matched_len;
    int dir, type;
    bool is_concord, left_ok, right_ok, looked_up_spos, looked_up_epos;
    int exon_ind_spos, exon_ind_epos;
    int exons_spos, exons_epos;   // interval index or -1
};

CM_HD void mate_init(MateD &m, const cm_params &p) {   // MatchedMate::MatchedMate(void), src/common.cpp:147-152
    m.spos = m.epos = m.qspos = m.qepos = 0;
    m.right_ed = m.left_ed = m.middle_ed = p.max_ed + 1;
    m.sclen_right = m.sclen_left = 0;
    m.matched_len = 0;
    m.dir = 0;
    m.type = CM_ORPHAN;
    m.is_concord = false;
    m.left_ok = m.right_ok = m.looked_up_spos = m.looked_up_epos = false;
    m.exon_ind_spos = m.exon_ind_epos = -1;
    m.exons_spos = m.exons_epos = -1;
}

// Per-chain results of the dense pre-alignment kernel (k_chain_spec): the middle edit distance of
// calc_middle_ed(ch, maxEd) and the genome-text X-drop alignment of the two read ends.
struct EndRes {
    int32_t ed, sclen, indel, score;
};
struct ChainAux {
    int32_t mid_ed;
    uint8_t has_mid, has_l, has_r, pad_;
    EndRes l, r;
};

// DP results of the transcript walk of one chain end (extend_right / extend_left over every transcript of
// the end's exon interval), recorded by the dense pre-alignment kernel.  An entry is the value of one DP as
// a pure function of its inputs (texts, lengths, direction, kind); the per-pair kernel uses it in place of
// running that DP and otherwise proceeds exactly as if it had computed it.
struct SpecEntry {
    uint32_t s_off;          // reference text: offset into ContigDev::genome
    uint16_t t_off, n, m;    // read text: offset into the strand row; lengths of the two texts
    uint8_t kind;            // 0 X-drop, 1 X-drop reversed, 2 banded edit, 3 banded edit reversed
    int8_t ed;
    int16_t sclen, indel;
    int32_t score;
};
#define CM_SPEC_CAP 4
struct EndSpec {
    uint32_t n;
    SpecEntry e[CM_SPEC_CAP];
};

// chain view used by the extension: fragment len is always kmer
struct ChainView {
    const FragDev *frags;
    uint32_t chain_len;
    const ChainAux *aux;   // nullable
    const EndSpec *spec;   // nullable: [0] left end, [1] right end; valid where aux->has_l / has_r
};

// common_tid of a MatePair: tids of interval iv_a that also occur in interval iv_b (iv_a == -1: empty).
// iv_a <= -2: exactly one transcript, tid = -2 - iv_a (stage 2's split_realignment, src/process_circ.cpp:1254-1256)
struct TidSet {
    int iv_a, iv_b;
};
CM_HD TidSet cm_tid_single(uint32_t tid) { TidSet t; t.iv_a = -2 - (int) tid; t.iv_b = -1; return t; }

struct AlignMemoEntry {   // map<AllCoord, AlignRes> align_res (src/extend.cpp:299)
    uint32_t rspos, rlen, qspos, qlen;
    AlignResD res;
};
#define CM_MEMO_CAP 12
struct AlignMemo {
    int n;
    AlignMemoEntry e[CM_MEMO_CAP];
};
CM_HD int memo_find(const AlignMemo &m, uint32_t rs, uint32_t rl, uint32_t qs, uint32_t ql) {
    for (int i = 0; i < m.n; i++)
        if (m.e[i].rspos == rs && m.e[i].rlen == rl && m.e[i].qspos == qs && m.e[i].qlen == ql) return i;
    return -1;
}
CM_HD void memo_insert(AlignMemo &m, uint32_t rs, uint32_t rl, uint32_t qs, uint32_t ql, const AlignResD &r) {
    if (m.n >= CM_MEMO_CAP) return;   // beyond the cap the memo degrades to recomputation (it is a cache)
    AlignMemoEntry &e = m.e[m.n++];
    e.rspos = rs; e.rlen = rl; e.qspos = qs; e.qlen = ql; e.res = r;
}

// GenomeSeeder::pac2char: text of [start, start+len-1] (1-based) or NULL when out of range
CM_HD const uint8_t *cm_pac2char(const DevView &v, uint32_t start, int len) {
    if ((int) start < 0 || (int) start + len - 1 > (int) v.c.ref_len) return nullptr;
    return v.c.genome + start;
}

struct ExtCtx {
    const DevView *v;
    AlignStats st;
    DropHint hint;          // valid only while the end of a chain with ChainAux is being extended
    const EndSpec *seed;    // pre-computed DPs of that end's transcript walk (nullable)
    EndSpec *capture;       // pre-alignment kernel only: where the walk's DPs are recorded (nullable)
    const uint8_t *tbase;   // strand row the t_off of seed / capture entries refer to
    int ts_c, ts_d;         // stage 2 (EDIT instantiations only): two more intervals a transcript of the TidSet must occur in
                            // (the four-interval same_transcript, src/utils.cpp:378-398); -1 = none
};
CM_HD void ext_ctx_init(ExtCtx &x, const DevView *v) {
    x.v = v;
    x.st.xdrop_cells = x.st.edit_cells = x.st.xdrop_calls = x.st.edit_calls = 0;
    x.hint.valid = 0;
    x.seed = nullptr;
    x.capture = nullptr;
    x.tbase = nullptr;
    x.ts_c = x.ts_d = -1;
}

// The genome-text alignment problem of one chain end as extend_left / extend_right pose it in their
// "intron retention" step (src/extend.cpp:336, :411) -- and, when the exon reaches far enough, in the
// transcript walk too.  false when there is nothing to extend or the text is out of range.
CM_HD bool cm_end_problem(const DevView &v, const ChainView &ch, const uint8_t *seq, int seq_len, bool right, const uint8_t *&s,
                          int &n, const uint8_t *&t, int &m) {
    const int k = v.p.kmer, w = v.p.band_width;
    if (ch.chain_len == 0) return false;
    if (right) {
        const FragDev &fl = ch.frags[ch.chain_len - 1];
        m = seq_len - (fl.qpos + k);
        if (m <= 0) return false;
        n = m + w;
        t = seq + seq_len - m;
        s = cm_pac2char(v, fl.rpos + (uint32_t) k, n);
    } else {
        m = ch.frags[0].qpos;
        if (m <= 0) return false;
        n = m + w;
        t = seq;
        s = cm_pac2char(v, ch.frags[0].rpos - (uint32_t) n, n);
    }
    return s != nullptr;
}

CM_HD void cm_set_hint(ExtCtx &x, const ChainView &ch, const uint8_t *seq, int seq_len, bool right) {
    x.hint.valid = 0;
    if (!ch.aux || !(right ? ch.aux->has_r : ch.aux->has_l)) return;
    if (!cm_end_problem(*x.v, ch, seq, seq_len, right, x.hint.s, x.hint.n, x.hint.t, x.hint.m)) return;
    const EndRes &r = right ? ch.aux->r : ch.aux->l;
    x.hint.rev = right ? 0 : 1;
    x.hint.ed = r.ed; x.hint.sclen = r.sclen; x.hint.indel = r.indel; x.hint.score = r.score;
    x.hint.valid = 1;
}

// get_location_overlap_ind: find_ind + "seg_list.size() == 0 -> NULL" (src/gene_annotation.cpp:555-568)
CM_HD int cm_location_overlap_ind(const AnnotDev &a, uint32_t pos, int &ind) {
    int r = cm_find_ind(a, pos, ind);
    if (r >= 0 && a.iv_seg_off[r + 1] == a.iv_seg_off[r]) return -1;
    return r;
}

// estimate_middle_error (src/utils.cpp:35-49)
CM_HD int cm_estimate_middle_error(const ChainView &ch, const cm_params &p) {
    int mid_err = 0;
    for (uint32_t i = 0; i + 1 < ch.chain_len; i++) {
        if (ch.frags[i + 1].qpos > ch.frags[i].qpos + p.kmer) {
            int diff = (int) (ch.frags[i + 1].rpos - ch.frags[i].rpos) - (ch.frags[i + 1].qpos - ch.frags[i].qpos);
            if (diff == 0) mid_err++;
            else if (diff > 0 && diff <= p.band_width) mid_err += diff;
            else if (diff < 0 && diff >= -p.band_width) mid_err -= diff;
        }
    }
    return mid_err;
}

// is_concord (concord2 = false, src/utils.cpp:116) / is_concord2 (:134)
CM_HD bool cm_is_concord(const ChainView &a, uint32_t seq_len, MateD &mr, const cm_params &p, bool concord2) {
    const FragDev &f0 = a.frags[0], &fl = a.frags[a.chain_len - 1];
    if (a.chain_len < 2) {
        mr.is_concord = false;
    } else if ((uint32_t) (fl.qpos + p.kmer - f0.qpos) >= seq_len) {
        mr.is_concord = true;
        mr.type = CM_CONCRD;
        mr.spos = f0.rpos;
        mr.epos = fl.rpos + p.kmer - 1;
        mr.matched_len = (uint32_t) (fl.qpos + p.kmer - f0.qpos);
        mr.qspos = (uint32_t) f0.qpos;
        mr.qepos = (uint32_t) (fl.qpos + p.kmer - 1);
    } else {
        mr.is_concord = false;
        if (concord2 && (f0.qpos == 0 || (uint32_t) (fl.qpos + p.kmer) == seq_len)) mr.type = CM_CANDID;
    }
    return mr.is_concord;
}

// update_match_mate_info (src/utils.cpp:22-33)
CM_HD void cm_update_match_mate_info(bool lok, bool rok, int err, MateD &mm, const cm_params &p) {
    mm.left_ok = lok && (mm.sclen_left <= p.max_sc);
    mm.right_ok = rok && (mm.sclen_right <= p.max_sc);
    if (lok && rok && (err <= p.max_ed) && (mm.sclen_right <= p.max_sc) && (mm.sclen_left <= p.max_sc)) {
        mm.is_concord = true;
        mm.type = CM_CONCRD;
    } else if (lok || rok) {
        mm.type = CM_CANDID;
    } else {
        mm.type = CM_ORPHAN;
    }
}

// TransExtension::calc_middle_ed (src/extend.cpp:878-920)
CM_HDN int cm_calc_middle_ed(ExtCtx &x, const ChainView &ch, int edth, const uint8_t *qseq, int qseq_len) {
    const cm_params &p = x.v->p;
    int mid_err = 0;
    if (ch.chain_len == 0) return 0;
    if (ch.aux && ch.aux->has_mid && edth == p.max_ed) return ch.aux->mid_ed;
    for (uint32_t i = 0; i + 1 < ch.chain_len; i++) {
        if (ch.frags[i + 1].qpos > ch.frags[i].qpos + p.kmer) {
            int diff = (int) (ch.frags[i + 1].rpos - ch.frags[i].rpos) - (ch.frags[i + 1].qpos - ch.frags[i].qpos);
            int qspos = ch.frags[i].qpos + p.kmer;
            int qlen = ch.frags[i + 1].qpos - qspos;
            uint32_t rspos = ch.frags[i].rpos + (uint32_t) p.kmer;
            int rlen = qlen + diff;
            if (rlen < 0) rlen = 0;
            if ((diff >= 0 && diff <= p.band_width) || (diff < 0 && diff >= -p.band_width)) {
                // the reference ignores pac2char's result here; out-of-range text is stale memory there
                const uint8_t *rseq = cm_pac2char(*x.v, rspos, rlen);
                if (!rseq) rseq = x.v->c.genome;
                if (diff >= 0) mid_err += cm_one_side_banded(qseq + qspos, qlen, rseq, rlen, diff, &x.st);
                else mid_err += cm_one_side_banded(rseq, rlen, qseq + qspos, qlen, -diff, &x.st);
            }
            if (mid_err > edth) return edth + 1;
        }
    }
    return mid_err;
}

// The two DPs of the transcript walk, behind the seed / capture lists.
CM_HD const SpecEntry *cm_spec_find(const ExtCtx &x, const uint8_t *s, int n, const uint8_t *t, int m, int kind) {
    if (!x.seed) return nullptr;
    const uint32_t so = (uint32_t) (s - x.v->c.genome);
    const uint16_t to = (uint16_t) (t - x.tbase);
    const uint32_t cnt = x.seed->n < CM_SPEC_CAP ? x.seed->n : CM_SPEC_CAP;
    for (uint32_t k = 0; k < cnt; k++) {
        const SpecEntry &e = x.seed->e[k];
        if (e.s_off == so && e.t_off == to && e.n == (uint16_t) n && e.m == (uint16_t) m && e.kind == (uint8_t) kind) return &e;
    }
    return nullptr;
}
CM_HD void cm_spec_put(ExtCtx &x, const uint8_t *s, int n, const uint8_t *t, int m, int kind, int ed, int sclen, int indel, int score) {
    if (!x.capture || x.capture->n >= CM_SPEC_CAP) return;
    SpecEntry &e = x.capture->e[x.capture->n++];
    e.s_off = (uint32_t) (s - x.v->c.genome);
    e.t_off = (uint16_t) (t - x.tbase);
    e.n = (uint16_t) n; e.m = (uint16_t) m; e.kind = (uint8_t) kind;
    e.ed = (int8_t) ed; e.sclen = (int16_t) sclen; e.indel = (int16_t) indel; e.score = score;
}
// DropAlignment::local_alignment_{right,left}_sc of reference text s against read text t
// EDIT: the TransExtension was made with EDIT_ALIGNMENT (stage 2, src/process_circ.cpp:25): alignment-> is virtual
// (src/align.h:170-171) and the soft-clipping aligner is EditDistAlignment's.  A template parameter of the whole extension tree
// (default false), so that stage 1's kernels carry none of it.
template <bool EDIT = false>
CM_HD int cm_x_drop(ExtCtx &x, const uint8_t *s, int n, const uint8_t *t, int m, bool rev, int &sclen, int &indel, int &score) {
    if (EDIT) return cm_edit_alignment_sc(s, n, t, m, rev, x.v->p, sclen, indel, score, &x.st);
    const DropHint &h = x.hint;
    const bool hinted = h.valid && h.s == s && h.t == t && h.n == n && h.m == m && h.rev == (int) rev;
    if (!hinted) {
        if (const SpecEntry *e = cm_spec_find(x, s, n, t, m, rev ? 1 : 0)) {
            sclen = e->sclen; indel = e->indel; score = e->score;
            return e->ed;
        }
    }
    int ed = cm_drop_alignment_sc(s, n, t, m, rev, x.v->p, sclen, indel, score, &x.st, &x.hint);
    if (!hinted) cm_spec_put(x, s, n, t, m, rev ? 1 : 0, ed, sclen, indel, score);
    return ed;
}
// Alignment::local_alignment_{right,left} of read text t (first argument of the reference call) against exon text s
CM_HD int cm_x_local(ExtCtx &x, const uint8_t *t, int m, const uint8_t *s, int n, bool rev, int &indel, int &score) {
    if (const SpecEntry *e = cm_spec_find(x, s, n, t, m, rev ? 3 : 2)) {
        indel = e->indel; score = e->score;
        return e->ed;
    }
    int ed = cm_local_alignment(t, m, s, n, rev, x.v->p, indel, score, &x.st);
    cm_spec_put(x, s, n, t, m, rev ? 3 : 2, ed, 0, indel, score);
    return ed;
}

// ---- right side --------------------------------------------------------------------------
// extend_right_middle (src/extend.cpp:435-461)
CM_HDN bool cm_extend_right_middle(ExtCtx &x, uint32_t pos, uint32_t exon_len, const uint8_t *qseq, uint32_t qseq_len,
                                   int ed_th, AlignResD &best, AlignResD &curr, AlignResD &exon_res) {
    const cm_params &p = x.v->p;
    const uint8_t *ref = cm_pac2char(*x.v, pos + 1, (int) exon_len);
    if (!ref) return false;
    int indel, score;
    uint32_t seq_remain = exon_len + (uint32_t) p.band_width < qseq_len ? exon_len + (uint32_t) p.band_width : qseq_len;
    int edit_dist = cm_x_local(x, qseq, (int) seq_remain, ref, (int) exon_len, false, indel, score);
    uint32_t new_rmpos = pos + exon_len;
    ar_set(exon_res, new_rmpos, edit_dist, 0, -indel, (int) exon_len - indel, score);
    if (curr.ed + edit_dist <= ed_th) {
        ar_update(curr, edit_dist, 0, new_rmpos, -indel, (int) exon_len - indel, score);
        ar_update_side(best, curr, true, p);
        return true;
    }
    return false;
}

// extend_right_end (src/extend.cpp:463-487)
template <bool EDIT = false>
CM_HDN void cm_extend_right_end(ExtCtx &x, uint32_t pos, uint32_t ref_len, const uint8_t *qseq, int qseq_len, int ed_th,
                                AlignResD &best, AlignResD &curr, AlignResD &exon_res) {
    const cm_params &p = x.v->p;
    const uint8_t *ref = cm_pac2char(*x.v, pos + 1, (int) ref_len);
    if (!ref) return;
    int sclen, indel, score;
    int edit_dist = cm_x_drop<EDIT>(x, ref, (int) ref_len, qseq, qseq_len, false, sclen, indel, score);
    uint32_t new_rmpos = pos + (uint32_t) qseq_len - (uint32_t) indel;
    ar_set(exon_res, new_rmpos, edit_dist, sclen, indel, qseq_len, score);
    int actual_mapped_bp = qseq_len - sclen;
    if ((curr.ed + edit_dist <= ed_th) && (sclen <= p.max_sc) && (actual_mapped_bp >= sclen)) {
        ar_update(curr, edit_dist, sclen, new_rmpos, indel, qseq_len, score);
        ar_update_by_score(best, curr, true);
    }
}

// shared by both sides: apply a memoised middle result; false = caller returns
CM_HD bool cm_apply_memo_middle(const AlignResD &m, int ed_th, AlignResD &best, AlignResD &curr, bool right, const cm_params &p) {
    if (curr.ed + m.ed > ed_th) return false;
    ar_update(curr, m.ed, m.sclen, m.pos, m.indel, m.qcovlen, m.score);
    ar_update_side(best, curr, right, p);
    return true;
}
CM_HD void cm_apply_memo_end(const AlignResD &m, int ed_th, AlignResD &best, AlignResD &curr, bool right, const cm_params &p) {
    int actual_mapped_bp = m.qcovlen - m.sclen;
    if ((curr.ed + m.ed > ed_th) || (m.sclen > p.max_sc) || (actual_mapped_bp < m.sclen)) return;
    ar_update(curr, m.ed, m.sclen, m.pos, m.indel, m.qcovlen, m.score);
    ar_update_by_score(best, curr, right);
}

// extend_right_trans (src/extend.cpp:491-650): [pos + 1, pos + len]
template <bool EDIT = false>
CM_HDN void cm_extend_right_trans(ExtCtx &x, uint32_t tid, uint32_t pos, int ref_len, const uint8_t *qseq, int qseq_len,
                                  int ed_th, uint32_t ub, AlignResD &best, bool &consecutive, AlignMemo &memo) {
    const AnnotDev &a = x.v->a;
    const cm_params &p = x.v->p;
    consecutive = false;
    AlignResD curr = ar_init(ub), exon_res = ar_init(ub);
    int it_ind;
    int it_seg = cm_location_overlap_ind(a, pos, it_ind);
    int covered = 0;
    if (it_seg < 0) return;   // probably wrong chaining to intron
    int it_ind_start = a.trans_start_ind[tid];
    int rel_ind = it_ind - it_ind_start;
    uint32_t rspos = pos;
    int exon_len = (int) (a.iv_epos[it_seg] - pos);
    int remain_ref_len = ref_len;
    int indel;
    const uint32_t tsz = a.t2s_off[tid + 1] - a.t2s_off[tid];
    const uint8_t *t2s = a.t2s + a.t2s_off[tid];
    for (unsigned int i = (unsigned int) (rel_ind + 1); i < tsz; i++) {
        if (exon_len >= qseq_len - covered) break;
        if (t2s[i] == 1) {
            indel = 0;
            if (exon_len > 0) {
                if (rspos + (uint32_t) exon_len > ub) return;
                int rq = exon_len + p.band_width < qseq_len - covered ? exon_len + p.band_width : qseq_len - covered;
                uint32_t remain_qseq_len = (uint32_t) rq;
                int mi = memo_find(memo, rspos, (uint32_t) exon_len, (uint32_t) covered, remain_qseq_len);
                if (mi >= 0) {
                    if (!cm_apply_memo_middle(memo.e[mi].res, ed_th, best, curr, true, p)) return;
                    indel = memo.e[mi].res.indel;
                } else {
                    bool success = cm_extend_right_middle(x, rspos, (uint32_t) exon_len, qseq + covered, remain_qseq_len, ed_th,
                                                          best, curr, exon_res);
                    memo_insert(memo, rspos, (uint32_t) exon_len, (uint32_t) covered, remain_qseq_len, exon_res);
                    if (!success) return;
                    indel = exon_res.indel;
                }
            }
            remain_ref_len -= exon_len;
            covered += exon_len + indel;
            exon_len = 0;
            it_seg = (int) (i + it_ind_start);
            rspos = a.iv_spos[it_seg] - 1;
        }
        if (t2s[i] != 0) {
            it_seg = (int) (i + it_ind_start);
            exon_len += (int) (a.iv_epos[it_seg] - a.iv_spos[it_seg] + 1);
        }
    }
    // reached end of transcript and read remains
    if ((exon_len > 0) && (exon_len < qseq_len - covered) && (rspos + (uint32_t) exon_len <= ub)) {
        int rq = exon_len + p.band_width < qseq_len - covered ? exon_len + p.band_width : qseq_len - covered;
        uint32_t remain_qseq_len = (uint32_t) rq;
        int mi = memo_find(memo, rspos, (uint32_t) exon_len, (uint32_t) covered, remain_qseq_len);
        if (mi >= 0) {
            cm_apply_memo_middle(memo.e[mi].res, ed_th, best, curr, true, p);
        } else {
            cm_extend_right_middle(x, rspos, (uint32_t) exon_len, qseq + covered, remain_qseq_len, ed_th, best, curr, exon_res);
            memo_insert(memo, rspos, (uint32_t) exon_len, (uint32_t) covered, remain_qseq_len, exon_res);
        }
        return;
    }
    if (covered >= qseq_len || (rspos + (uint32_t) (qseq_len - covered) > ub) || (exon_len < qseq_len - covered)) return;
    consecutive = (rspos == pos);
    remain_ref_len = remain_ref_len < exon_len ? remain_ref_len : exon_len;
    int mi = memo_find(memo, rspos, (uint32_t) remain_ref_len, (uint32_t) covered, (uint32_t) (qseq_len - covered));
    if (mi >= 0) {
        cm_apply_memo_end(memo.e[mi].res, ed_th, best, curr, true, p);
    } else {
        cm_extend_right_end<EDIT>(x, rspos, (uint32_t) remain_ref_len, qseq + covered, qseq_len - covered, ed_th, best, curr, exon_res);
        memo_insert(memo, rspos, (uint32_t) remain_ref_len, (uint32_t) covered, (uint32_t) (qseq_len - covered), exon_res);
    }
}

// does interval iv list transcript t?  (inner loop of intersect_trans, src/utils.cpp:322-333)
CM_HD bool cm_iv_has_tid(const AnnotDev &a, int iv, uint32_t t) {
    uint32_t b = a.seg_tid_off[a.iv_seg_off[iv]], e = a.seg_tid_off[a.iv_seg_off[iv + 1]];
    for (uint32_t k = b; k < e; k++)
        if (a.tid[k] == t) return true;
    return false;
}

// extend_right (src/extend.cpp:285-357): pos exclusive, [pos+1, pos+len]
template <bool EDIT = false>
CM_HDN bool cm_extend_right(ExtCtx &x, TidSet ts, const uint8_t *seq, uint32_t &pos, int len, int ed_th, uint32_t ub,
                            AlignResD &best) {
    const AnnotDev &a = x.v->a;
    const cm_params &p = x.v->p;
    int seq_len = len;
    int ref_len = len + p.band_width;
    uint32_t orig_pos = pos;
    int indel;
    bool consecutive = false;
    AlignResD curr = ar_init(ub);
    ar_set(best, pos, ed_th + 1, len + 1, p.band_width + 1, 0, 0);
    AlignMemo memo;
    memo.n = 0;
    if (ts.iv_a >= 0 && ts.iv_b >= 0) {
        uint32_t b = a.seg_tid_off[a.iv_seg_off[ts.iv_a]], e = a.seg_tid_off[a.iv_seg_off[ts.iv_a + 1]];
        for (uint32_t k = b; k < e; k++) {
            uint32_t t = a.tid[k];
            if (cm_iv_has_tid(a, ts.iv_b, t) && (!EDIT || x.ts_c < 0 || (cm_iv_has_tid(a, x.ts_c, t) && cm_iv_has_tid(a, x.ts_d, t))))
                cm_extend_right_trans<EDIT>(x, t, pos, ref_len, seq, seq_len, ed_th, ub, best, consecutive, memo);
        }
    } else if (EDIT && ts.iv_a <= -2) {     // one transcript: stage 2 only
        cm_extend_right_trans<EDIT>(x, (uint32_t) (-2 - ts.iv_a), pos, ref_len, seq, seq_len, ed_th, ub, best, consecutive, memo);
    }
    uint32_t best_rmpos = best.pos;
    int min_ed = best.ed;
    int sclen_best = best.sclen;
    if (min_ed <= ed_th) {
        pos = best_rmpos - (uint32_t) sclen_best;
        if (best.qcovlen >= seq_len && sclen_best <= p.max_sc) return true;
    }
    // intron retention
    const uint8_t *ref;
    if (!consecutive && (ref = cm_pac2char(*x.v, orig_pos + 1, ref_len)) != nullptr) {
        int score;
        min_ed = EDIT ? cm_edit_alignment_sc(ref, ref_len, seq, seq_len, false, p, sclen_best, indel, score, &x.st)
                             : cm_drop_alignment_sc(ref, ref_len, seq, seq_len, false, p, sclen_best, indel, score, &x.st, &x.hint);
        if (min_ed <= ed_th && sclen_best <= p.max_sc) {
            ar_set(curr, orig_pos + (uint32_t) seq_len - (uint32_t) indel, min_ed, sclen_best, indel, seq_len, score);
            if (ar_update_by_score(best, curr, true)) {
                pos = orig_pos + (uint32_t) seq_len - (uint32_t) indel - (uint32_t) sclen_best;
                return true;
            }
        }
    }
    // no extension was possible: roll back
    if (best.qcovlen <= 0) {
        pos = orig_pos;
        ar_set(best, pos, 0, 0, 0, 0, CM_NEGINF);
    }
    int qremain = seq_len - best.qcovlen;
    if (qremain + best.sclen <= p.max_sc) {
        ar_set(best, pos, best.ed, best.sclen + qremain, best.indel, seq_len, best.score);
        return true;
    }
    return (best.qcovlen >= seq_len && best.ed <= ed_th);
}

// ---- left side ---------------------------------------------------------------------------
// extend_left_middle (src/extend.cpp:653-679)
CM_HDN bool cm_extend_left_middle(ExtCtx &x, uint32_t pos, uint32_t exon_len, const uint8_t *qseq, uint32_t qseq_len,
                                  int ed_th, AlignResD &best, AlignResD &curr, AlignResD &exon_res) {
    const cm_params &p = x.v->p;
    const uint8_t *ref = cm_pac2char(*x.v, pos - exon_len, (int) exon_len);
    if (!ref) return false;
    int indel, score;
    uint32_t seq_remain = exon_len + (uint32_t) p.band_width < qseq_len ? exon_len + (uint32_t) p.band_width : qseq_len;
    int edit_dist = cm_x_local(x, qseq, (int) seq_remain, ref, (int) exon_len, true, indel, score);
    uint32_t new_lmpos = pos - exon_len;
    ar_set(exon_res, new_lmpos, edit_dist, 0, -indel, (int) exon_len - indel, score);
    if (curr.ed + edit_dist <= ed_th) {
        ar_update(curr, edit_dist, 0, new_lmpos, -indel, (int) exon_len - indel, score);
        ar_update_side(best, curr, false, p);
        return true;
    }
    return false;
}

// extend_left_end (src/extend.cpp:681-705)
template <bool EDIT = false>
CM_HDN void cm_extend_left_end(ExtCtx &x, uint32_t pos, uint32_t ref_len, const uint8_t *qseq, int qseq_len, int ed_th,
                               AlignResD &best, AlignResD &curr, AlignResD &exon_res) {
    const cm_params &p = x.v->p;
    const uint8_t *ref = cm_pac2char(*x.v, pos - ref_len, (int) ref_len);
    if (!ref) return;
    int sclen, indel, score;
    int edit_dist = cm_x_drop<EDIT>(x, ref, (int) ref_len, qseq, qseq_len, true, sclen, indel, score);
    uint32_t new_lmpos = pos - (uint32_t) qseq_len + (uint32_t) indel;
    ar_set(exon_res, new_lmpos, edit_dist, sclen, indel, qseq_len, score);
    int actual_mapped_bp = qseq_len - sclen;
    if ((curr.ed + edit_dist <= ed_th) && (sclen <= p.max_sc) && (actual_mapped_bp >= sclen)) {
        ar_update(curr, edit_dist, sclen, new_lmpos, indel, qseq_len, score);
        ar_update_by_score(best, curr, false);
    }
}

// extend_left_trans (src/extend.cpp:708-875): [pos - len, pos - 1]
template <bool EDIT = false>
CM_HDN void cm_extend_left_trans(ExtCtx &x, uint32_t tid, uint32_t pos, int ref_len, const uint8_t *qseq, int qseq_len,
                                 int ed_th, uint32_t lb, AlignResD &best, bool &consecutive, AlignMemo &memo) {
    const AnnotDev &a = x.v->a;
    const cm_params &p = x.v->p;
    consecutive = false;
    AlignResD curr = ar_init(lb), exon_res = ar_init(lb);
    int it_ind;
    int covered = 0;
    int it_seg = cm_location_overlap_ind(a, pos, it_ind);
    if (it_seg < 0) return;
    int it_ind_start = a.trans_start_ind[tid];
    int rel_ind = it_ind - it_ind_start;
    uint32_t lepos = pos;
    int exon_len = 0;
    int remain_ref_len = ref_len;
    int indel;
    bool first_seg = true;
    const int tsz = (int) (a.t2s_off[tid + 1] - a.t2s_off[tid]);
    const uint8_t *t2s = a.t2s + a.t2s_off[tid];
    for (int i = rel_ind; i >= 0; i--) {
        uint8_t state = i < tsz ? t2s[i] : 0;   // the reference would index past the vector here
        if (state != 0) {
            it_seg = i + it_ind_start;
            if (first_seg) {
                exon_len = (int) (pos - a.iv_spos[it_seg]);
                first_seg = false;
            } else {
                if (exon_len == 0) lepos = a.iv_epos[it_seg] + 1;
                exon_len += (int) (a.iv_epos[it_seg] - a.iv_spos[it_seg] + 1);
            }
        }
        if (exon_len >= qseq_len - covered) break;
        if (state == 1) {
            indel = 0;
            if (exon_len > 0) {
                if (lepos < lb + (uint32_t) exon_len) return;
                int rq = exon_len + p.band_width < qseq_len - covered ? exon_len + p.band_width : qseq_len - covered;
                uint32_t remain_qseq_len = (uint32_t) rq;
                int mi = memo_find(memo, lepos, (uint32_t) exon_len, (uint32_t) covered, remain_qseq_len);
                if (mi >= 0) {
                    if (!cm_apply_memo_middle(memo.e[mi].res, ed_th, best, curr, false, p)) return;
                    indel = memo.e[mi].res.indel;
                } else {
                    bool success = cm_extend_left_middle(x, lepos, (uint32_t) exon_len, qseq + qseq_len - covered - rq,
                                                         remain_qseq_len, ed_th, best, curr, exon_res);
                    memo_insert(memo, lepos, (uint32_t) exon_len, (uint32_t) covered, remain_qseq_len, exon_res);
                    if (!success) return;
                    indel = exon_res.indel;
                }
            }
            remain_ref_len -= exon_len;
            covered += exon_len + indel;
            exon_len = 0;
        }
    }
    // reached end of transcript and read remains
    if ((exon_len > 0) && (exon_len < qseq_len - covered) && (lepos >= lb + (uint32_t) exon_len)) {
        int rq = exon_len + p.band_width < qseq_len - covered ? exon_len + p.band_width : qseq_len - covered;
        uint32_t remain_qseq_len = (uint32_t) rq;
        int mi = memo_find(memo, lepos, (uint32_t) exon_len, (uint32_t) covered, remain_qseq_len);
        if (mi >= 0) {
            cm_apply_memo_middle(memo.e[mi].res, ed_th, best, curr, false, p);
        } else {
            cm_extend_left_middle(x, lepos, (uint32_t) exon_len, qseq + qseq_len - covered - rq, remain_qseq_len, ed_th, best,
                                  curr, exon_res);
            memo_insert(memo, lepos, (uint32_t) exon_len, (uint32_t) covered, remain_qseq_len, exon_res);
        }
        return;
    }
    if (covered >= qseq_len || (lepos < lb + (uint32_t) (qseq_len - covered)) || (exon_len < qseq_len - covered)) return;
    consecutive = (lepos == pos);
    remain_ref_len = remain_ref_len < exon_len ? remain_ref_len : exon_len;
    int mi = memo_find(memo, lepos, (uint32_t) remain_ref_len, (uint32_t) covered, (uint32_t) (qseq_len - covered));
    if (mi >= 0) {
        cm_apply_memo_end(memo.e[mi].res, ed_th, best, curr, false, p);
    } else {
        cm_extend_left_end<EDIT>(x, lepos, (uint32_t) remain_ref_len, qseq, qseq_len - covered, ed_th, best, curr, exon_res);
        memo_insert(memo, lepos, (uint32_t) remain_ref_len, (uint32_t) covered, (uint32_t) (qseq_len - covered), exon_res);
    }
}

// extend_left (src/extend.cpp:361-432): pos exclusive, [pos-len, pos-1]
template <bool EDIT = false>
CM_HDN bool cm_extend_left(ExtCtx &x, TidSet ts, const uint8_t *seq, uint32_t &pos, int len, int ed_th, uint32_t lb,
                           AlignResD &best) {
    const AnnotDev &a = x.v->a;
    const cm_params &p = x.v->p;
    int seq_len = len;
    int ref_len = len + p.band_width;
    uint32_t orig_pos = pos;
    int indel;
    bool consecutive = false;
    AlignResD curr = ar_init(lb);
    ar_set(best, pos, ed_th + 1, len + 1, p.band_width + 1, 0, 0);
    AlignMemo memo;
    memo.n = 0;
    if (ts.iv_a >= 0 && ts.iv_b >= 0) {
        uint32_t b = a.seg_tid_off[a.iv_seg_off[ts.iv_a]], e = a.seg_tid_off[a.iv_seg_off[ts.iv_a + 1]];
        for (uint32_t k = b; k < e; k++) {
            uint32_t t = a.tid[k];
            if (cm_iv_has_tid(a, ts.iv_b, t) && (!EDIT || x.ts_c < 0 || (cm_iv_has_tid(a, x.ts_c, t) && cm_iv_has_tid(a, x.ts_d, t))))
                cm_extend_left_trans<EDIT>(x, t, pos, ref_len, seq, seq_len, ed_th, lb, best, consecutive, memo);
        }
    } else if (EDIT && ts.iv_a <= -2) {
        cm_extend_left_trans<EDIT>(x, (uint32_t) (-2 - ts.iv_a), pos, ref_len, seq, seq_len, ed_th, lb, best, consecutive, memo);
    }
    uint32_t lmpos_best = best.pos;
    int min_ed = best.ed;
    int sclen_best = best.sclen;
    if (min_ed <= ed_th) {
        pos = lmpos_best + (uint32_t) sclen_best;
        if (best.qcovlen >= seq_len && sclen_best <= p.max_sc) return true;
    }
    // intron retention
    const uint8_t *ref;
    if (!consecutive && (ref = cm_pac2char(*x.v, orig_pos - (uint32_t) ref_len, ref_len)) != nullptr) {
        int score;
        min_ed = EDIT ? cm_edit_alignment_sc(ref, ref_len, seq, seq_len, true, p, sclen_best, indel, score, &x.st)
                             : cm_drop_alignment_sc(ref, ref_len, seq, seq_len, true, p, sclen_best, indel, score, &x.st, &x.hint);
        if (min_ed <= ed_th && sclen_best <= p.max_sc) {
            ar_set(curr, orig_pos - (uint32_t) seq_len + (uint32_t) indel, min_ed, sclen_best, indel, seq_len, score);
            if (ar_update_by_score(best, curr, false)) {
                pos = orig_pos - (uint32_t) seq_len + (uint32_t) indel + (uint32_t) sclen_best;
                return true;
            }
        }
    }
    if (best.qcovlen <= 0) {
        pos = orig_pos;
        ar_set(best, pos, 0, 0, 0, 0, CM_NEGINF);
    }
    int qremain = seq_len - best.qcovlen;
    if (qremain + best.sclen <= p.max_sc) {
        ar_set(best, pos, best.ed, best.sclen + qremain, best.indel, seq_len, best.score);
        return true;
    }
    return (best.qcovlen >= seq_len && best.ed <= ed_th);
}

// ---- chain level -------------------------------------------------------------------------
// extend_chain_right (src/extend.cpp:215-246)
template <bool EDIT = false>
CM_HDN bool cm_extend_chain_right(ExtCtx &x, TidSet ts, const ChainView &ch, const uint8_t *seq, int seq_len, uint32_t ub,
                                  MateD &mr, int &err) {
    const cm_params &p = x.v->p;
    const FragDev &fl = ch.frags[ch.chain_len - 1];
    uint32_t rm_pos = fl.rpos + (uint32_t) p.kmer - 1;
    int remain_end = seq_len - (fl.qpos + p.kmer);
    bool right_ok = (remain_end <= 0);
    AlignResD best = ar_init(ub);
    if (remain_end > 0) {
        cm_set_hint(x, ch, seq, seq_len, true);
        if (x.hint.valid && ch.spec) { x.seed = ch.spec + 1; x.tbase = seq; }
        right_ok = cm_extend_right<EDIT>(x, ts, seq + seq_len - remain_end, rm_pos, remain_end, p.max_ed - err, ub, best);
        x.hint.valid = 0;
        x.seed = nullptr;
    }
    int sclen_right = best.sclen;
    int err_right = best.ed;
    remain_end -= best.qcovlen;
    mr.epos = rm_pos;
    mr.matched_len -= (uint32_t) (right_ok ? sclen_right : remain_end);
    mr.qepos -= (uint32_t) (right_ok ? sclen_right : remain_end);
    mr.sclen_right = sclen_right;
    mr.right_ed = best.ed;
    err += err_right;
    return right_ok;
}

// extend_chain_left (src/extend.cpp:248-280)
template <bool EDIT = false>
CM_HDN bool cm_extend_chain_left(ExtCtx &x, TidSet ts, const ChainView &ch, const uint8_t *seq, int32_t qspos, uint32_t lb,
                                 MateD &mr, int &err) {
    const cm_params &p = x.v->p;
    uint32_t lm_pos = ch.frags[0].rpos;
    int remain_beg = ch.frags[0].qpos - qspos;
    bool left_ok = (remain_beg <= 0);
    AlignResD best = ar_init(lb);
    if (remain_beg > 0) {
        if (qspos == 0) cm_set_hint(x, ch, seq, 0, false);
        if (x.hint.valid && ch.spec) { x.seed = ch.spec; x.tbase = seq; }
        left_ok = cm_extend_left<EDIT>(x, ts, seq, lm_pos, remain_beg, p.max_ed - err, lb, best);
        x.hint.valid = 0;
        x.seed = nullptr;
    }
    int sclen_left = best.sclen;
    int err_left = best.ed;
    remain_beg -= best.qcovlen;
    mr.spos = lm_pos;
    mr.matched_len -= (uint32_t) (left_ok ? sclen_left : remain_beg);
    mr.qspos += (uint32_t) (left_ok ? sclen_left : remain_beg);
    mr.sclen_left = sclen_left;
    mr.left_ed = best.ed;
    err += err_left;
    return left_ok;
}

// extend_both_mates (src/extend.cpp:37-125)
template <bool EDIT = false>
CM_HDN bool cm_extend_both_mates(ExtCtx &x, const ChainView &lch, const ChainView &rch, TidSet ts, const uint8_t *lseq,
                                 const uint8_t *rseq, int lqspos, int rqspos, int lseq_len, int rseq_len, MateD &lmm,
                                 MateD &rmm) {
    const cm_params &p = x.v->p;
    lmm.middle_ed = cm_calc_middle_ed(x, lch, p.max_ed, lseq, lseq_len);
    rmm.middle_ed = cm_calc_middle_ed(x, rch, p.max_ed, rseq, rseq_len);
    if (lmm.middle_ed <= p.max_ed) cm_is_concord(lch, (uint32_t) lseq_len, lmm, p, true);
    if (rmm.middle_ed <= p.max_ed) cm_is_concord(rch, (uint32_t) rseq_len, rmm, p, true);
    if (lmm.middle_ed > p.max_ed || rmm.middle_ed > p.max_ed) return false;
    bool l_extend = true;
    lmm.is_concord = false;
    if (lch.chain_len <= 0) { lmm.type = CM_ORPHAN; lmm.matched_len = 0; l_extend = false; }
    bool r_extend = true;
    rmm.is_concord = false;
    if (rch.chain_len <= 0) { rmm.type = CM_ORPHAN; rmm.matched_len = 0; r_extend = false; }
    bool llok = false, lrok = false, rlok = false, rrok = false;
    int lerr = lmm.middle_ed;
    int rerr = rmm.middle_ed;
    if (l_extend) {
        lmm.matched_len = (uint32_t) (lseq_len - lqspos + 1);
        lmm.qspos = (uint32_t) lqspos;
        lmm.qepos = (uint32_t) lseq_len;
        llok = cm_extend_chain_left<EDIT>(x, ts, lch, lseq, lqspos - 1, CM_MINLB, lmm, lerr);
    }
    if (r_extend) {
        rmm.matched_len = (uint32_t) (rseq_len - rqspos + 1);
        rmm.qspos = (uint32_t) rqspos;
        rmm.qepos = (uint32_t) rseq_len;
        rlok = cm_extend_chain_left<EDIT>(x, ts, rch, rseq, rqspos - 1, l_extend ? lmm.spos : CM_MINLB, rmm, rerr);
    }
    if (r_extend) rrok = cm_extend_chain_right<EDIT>(x, ts, rch, rseq, rseq_len, CM_MAXUB, rmm, rerr);
    if (l_extend) lrok = cm_extend_chain_right<EDIT>(x, ts, lch, lseq, lseq_len, r_extend ? rmm.epos : CM_MAXUB, lmm, lerr);
    if (l_extend) cm_update_match_mate_info(llok, lrok, lerr, lmm, p);
    if (r_extend) cm_update_match_mate_info(rlok, rrok, rerr, rmm, p);
    return true;
}

// extend_chain_both_sides (src/extend.cpp:131-213)
CM_HDN int cm_extend_chain_both_sides(ExtCtx &x, const ChainView &ch, const uint8_t *seq, int seq_len, MateD &mr, int dir) {
    const cm_params &p = x.v->p;
    mr.is_concord = false;
    if (ch.chain_len <= 0) {
        mr.type = CM_ORPHAN;
        return mr.type;
    }
    mr.middle_ed = cm_estimate_middle_error(ch, p);
    if (cm_is_concord(ch, (uint32_t) seq_len, mr, p, false)) {
        mr.dir = dir;
        return mr.type;
    }
    TidSet empty;
    empty.iv_a = -1;
    empty.iv_b = -1;
    uint32_t lm_pos = ch.frags[0].rpos;
    int remain_beg = ch.frags[0].qpos;
    bool left_ok = (remain_beg <= 0);
    AlignResD best_left = ar_init(CM_MINLB);
    if (remain_beg > 0) {
        cm_set_hint(x, ch, seq, seq_len, false);
        left_ok = cm_extend_left(x, empty, seq, lm_pos, remain_beg, p.max_ed - mr.middle_ed, CM_MINLB, best_left);
        x.hint.valid = 0;
    }
    int err_left = best_left.ed;
    int sclen_left = best_left.sclen;
    remain_beg -= best_left.qcovlen;
    const FragDev &fl = ch.frags[ch.chain_len - 1];
    uint32_t rm_pos = fl.rpos + (uint32_t) p.kmer - 1;
    int remain_end = seq_len - (fl.qpos + p.kmer);
    bool right_ok = (remain_end <= 0);
    AlignResD best = ar_init(CM_MAXUB);
    if (remain_end > 0) {
        cm_set_hint(x, ch, seq, seq_len, true);
        right_ok = cm_extend_right(x, empty, seq + seq_len - remain_end, rm_pos, remain_end, p.max_ed - mr.middle_ed - err_left,
                                   CM_MAXUB, best);
        x.hint.valid = 0;
    }
    int err_right = best.ed;
    int sclen_right = best.sclen;
    remain_end -= best.qcovlen;
    mr.spos = lm_pos;
    mr.epos = rm_pos;
    mr.matched_len = (uint32_t) seq_len;
    mr.matched_len -= (uint32_t) (left_ok ? sclen_left : remain_beg);
    mr.matched_len -= (uint32_t) (right_ok ? sclen_right : remain_end);
    mr.qspos = (uint32_t) (1 + (left_ok ? sclen_left : remain_beg));
    mr.qepos = (uint32_t) (seq_len - (right_ok ? sclen_right : remain_end));
    mr.right_ed = best.ed;
    mr.left_ed = best_left.ed;
    mr.dir = dir;
    if (left_ok && right_ok && (err_left + err_right <= p.max_ed) && sclen_left <= p.max_sc && sclen_right <= p.max_sc) {
        mr.is_concord = true;
        mr.type = CM_CONCRD;
    } else if (left_ok || right_ok) {
        mr.type = CM_CANDID;
    } else {
        mr.type = CM_ORPHAN;
    }
    return mr.type;
}


// ---- dense pre-alignment -------------------------------------------------------------------
// kind 0: calc_middle_ed(ch, maxEd); kind 1 / 2: the genome-text X-drop of the left / right read end.
// Each kind writes its own fields of ChainAux, so the three may run in different threads.
// For an end with something to align, kinds 1 / 2 return true when the end's transcript walk may pose further DPs
// (cm_chain_walk records those); spec[0] / spec[1] is emptied either way.
CM_HDN bool cm_chain_spec(ExtCtx &x, const ChainView &ch, const uint8_t *seq, int seq_len, int kind, ChainAux &aux, EndSpec *spec) {
    const cm_params &p = x.v->p;
    if (kind == 0) {
        ChainView c = ch;
        c.aux = nullptr;
        c.spec = nullptr;
        aux.mid_ed = cm_calc_middle_ed(x, c, p.max_ed, seq, seq_len);
        aux.has_mid = 1;
        return false;
    }
    const bool right = kind == 2;
    const uint8_t *s, *t;
    int n, m;
    EndRes r;
    r.ed = r.sclen = r.indel = r.score = 0;
    uint8_t has = 0;
    bool walk = false;
    if (cm_end_problem(*x.v, ch, seq, seq_len, right, s, n, t, m)) {
        r.ed = cm_drop_alignment_sc(s, n, t, m, !right, p, r.sclen, r.indel, r.score, &x.st, nullptr);
        has = 1;
        if (spec) {
            spec[right ? 1 : 0].n = 0;
            uint32_t pos = right ? ch.frags[ch.chain_len - 1].rpos + (uint32_t) p.kmer - 1 : ch.frags[0].rpos;
            int ind;
            const int iv = cm_location_overlap_ind(x.v->a, pos, ind);
            // When the exon piece under the end holds m + w more bases, every transcript's walk stops at once and poses the
            // genome-text problem solved above (extend_*_trans: "exon_len >= qseq_len" break, remain_ref_len == ref_len).
            walk = iv >= 0 && (right ? (int) (x.v->a.iv_epos[iv] - pos) : (int) (pos - x.v->a.iv_spos[iv])) < m + p.band_width;
        }
    }
    if (right) { aux.r = r; aux.has_r = has; } else { aux.l = r; aux.has_l = has; }
    return walk;
}

// The transcript walk of one chain end: extend_left / extend_right over every transcript of the end's exon interval
// with the most permissive thresholds, recording the DPs met on the way in spec[0] (left) / spec[1] (right).
CM_HDN void cm_chain_walk(ExtCtx &x, const ChainView &ch, const uint8_t *seq, int seq_len, bool right, const ChainAux &aux, EndSpec *spec) {
    const cm_params &p = x.v->p;
    const uint8_t *s, *t;
    int n, m;
    if (!spec || !(right ? aux.has_r : aux.has_l) || !cm_end_problem(*x.v, ch, seq, seq_len, right, s, n, t, m)) return;
    EndSpec &out = spec[right ? 1 : 0];
    uint32_t pos = right ? ch.frags[ch.chain_len - 1].rpos + (uint32_t) p.kmer - 1 : ch.frags[0].rpos;
    int ind;
    const int iv = cm_location_overlap_ind(x.v->a, pos, ind);
    if (iv < 0) return;
    const EndRes &r = right ? aux.r : aux.l;
    x.hint.s = s; x.hint.t = t; x.hint.n = n; x.hint.m = m; x.hint.rev = right ? 0 : 1;
    x.hint.ed = r.ed; x.hint.sclen = r.sclen; x.hint.indel = r.indel; x.hint.score = r.score;
    x.hint.valid = 1;
    x.seed = nullptr;
    x.capture = &out;
    x.tbase = seq;
    TidSet ts;
    ts.iv_a = iv;
    ts.iv_b = iv;
    AlignResD best;
    if (right) cm_extend_right(x, ts, seq + seq_len - m, pos, m, p.max_ed, CM_MAXUB, best);
    else cm_extend_left(x, ts, seq, pos, m, p.max_ed, CM_MINLB, best);
    x.capture = nullptr;
    x.hint.valid = 0;
}
