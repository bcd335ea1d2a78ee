// Per-pair classification: FilterRead::process_read / process_mates / pair_chains
// (src/filter.cpp:86-395, 484-551), the explanation checks of src/utils.cpp (calc_tlen :53,
// concordant_explanation :157, check_chimeric :215, check_bsj :235, check_2bsj :270, same_gene
// :616-639, is_left_chain :827, overlap_to_* :667-683) and MatchedRead::update / go_for_update /
// update_type (src/common.cpp:286-411), all on cm_mread.
#pragma once
#include "cm_types.h"
#include "cm_annot.h"
#include "cm_extend.h"

// chains of one strand as the chaining stage leaves them
struct ChainListView {
    int count;                 // best_chain_count
    const float *score;        // [C]
    const uint32_t *len;       // [C]
    const FragDev *frags;      // [C][F]
    int F;
    const ChainAux *aux;       // [n_aux] or nullptr: pre-alignment results of the first n_aux chains
    const EndSpec *spec;       // [n_aux][2] or nullptr: transcript-walk DPs of the two ends
    int n_aux;
};
CM_HD ChainView cl_chain(const ChainListView &l, int i) {
    ChainView c;
    c.frags = l.frags + (size_t) i * l.F;
    c.chain_len = l.len[i];
    c.aux = (l.aux && i < l.n_aux) ? l.aux + i : nullptr;
    c.spec = (c.aux && l.spec) ? l.spec + 2 * (size_t) i : nullptr;
    return c;
}

// MatchedRead::go_for_update (src/common.cpp:362-411)
CM_HD bool mr_go_for_update(const cm_mread &m, const MateD &r1, const MateD &r2, int32_t tlen, bool gm, int type) {
    if (type < m.type) return true;
    if (type > m.type) return false;
    if (gm && !m.gm_compatible) return true;
    if (!gm && m.gm_compatible) return false;
    int edit_dist = r1.left_ed + r1.middle_ed + r1.right_ed + r2.left_ed + r2.middle_ed + r2.right_ed;
    uint32_t mlen_old = m.mlen_r1 + m.mlen_r2, mlen_new = r1.matched_len + r2.matched_len;
    if (type < CM_CHIBSJ) {
        if ((m.ed_r1 + m.ed_r2) > edit_dist) return true;
        if ((m.ed_r1 + m.ed_r2) < edit_dist) return false;
        if (m.tlen > tlen) return true;
        if (m.tlen < tlen) return false;
        if (mlen_old < mlen_new) return true;
        if (mlen_old > mlen_new) return false;
    } else {
        if (mlen_old < mlen_new) return true;
        if (mlen_old > mlen_new) return false;
        if ((m.ed_r1 + m.ed_r2) > edit_dist) return true;
        if ((m.ed_r1 + m.ed_r2) < edit_dist) return false;
    }
    return false;
}

// chromosome entry of con2chr[contig] (GTFParser::get_shift); chr index + shift
struct ConShiftD {
    int32_t chr;
    uint32_t shift;
};

// MatchedRead::update (src/common.cpp:286-351)
CM_HD bool mr_update(cm_mread &m, const MateD &r1, const MateD &r2, ConShiftD cs, int32_t tlen, uint16_t jun_between, bool gm,
                     int type, bool r1_first, int contig_id) {
    if (!mr_go_for_update(m, r1, r2, tlen, gm, type)) return false;
    m.type = type;
    m.chr_r1 = cs.chr;
    m.chr_r2 = cs.chr;
    const MateD &a = r1_first ? r1 : r2, &b = r1_first ? r2 : r1;
    m.spos_r1 = a.spos - cs.shift; m.epos_r1 = a.epos - cs.shift;
    m.qspos_r1 = a.qspos; m.qepos_r1 = a.qepos;
    m.mlen_r1 = a.matched_len;
    m.ed_r1 = a.left_ed + a.middle_ed + a.right_ed;
    m.spos_r2 = b.spos - cs.shift; m.epos_r2 = b.epos - cs.shift;
    m.qspos_r2 = b.qspos; m.qepos_r2 = b.qepos;
    m.mlen_r2 = b.matched_len;
    m.ed_r2 = b.left_ed + b.middle_ed + b.right_ed;
    m.r1_forward = a.dir > 0;
    m.r2_forward = b.dir > 0;
    m.tlen = tlen;
    m.junc_num = jun_between;
    m.gm_compatible = gm;
    m.contig_num = contig_id;
    return true;
}
CM_HD bool mr_update_type(cm_mread &m, int type) {
    if (type < m.type) { m.type = type; return true; }
    return false;
}

CM_HD void cm_overlap_to_epos(const AnnotDev &a, MateD &m) {
    if (m.looked_up_epos || m.exons_epos >= 0) return;
    m.exons_epos = cm_location_overlap_ind(a, m.epos, m.exon_ind_epos);
    m.looked_up_epos = true;
}
CM_HD void cm_overlap_to_spos(const AnnotDev &a, MateD &m) {
    if (m.looked_up_spos || m.exons_spos >= 0) return;
    m.exons_spos = cm_location_overlap_ind(a, m.spos, m.exon_ind_spos);
    m.looked_up_spos = true;
}

// calc_tlen (src/utils.cpp:53-113); uint32 index arithmetic as in the reference
CM_HDN int cm_calc_tlen(const AnnotDev &a, const MateD &sm, const MateD &lm, int &intron_num) {
    int min_tlen = CM_INF;
    int iv = sm.exons_epos;
    for (uint32_t s = a.iv_seg_off[iv]; s < a.iv_seg_off[iv + 1]; s++)
        for (uint32_t k = a.seg_tid_off[s]; k < a.seg_tid_off[s + 1]; k++) {
            uint32_t tid = a.tid[k];
            int start_ind = a.trans_start_ind[tid];
            uint32_t start_table_ind = (uint32_t) (sm.exon_ind_epos - start_ind);
            uint32_t end_table_ind = (uint32_t) (lm.exon_ind_spos - start_ind);
            uint32_t tsz = a.t2s_off[tid + 1] - a.t2s_off[tid];
            if (lm.exon_ind_spos < start_ind || end_table_ind >= tsz || a.t2s[a.t2s_off[tid] + end_table_ind] == 0) continue;
            int in, tlen;
            if (start_table_ind == end_table_ind) {
                in = 0;
                tlen = (int) (lm.spos - sm.epos + 1);
            } else {
                bool pre_zero = false;
                in = 0;
                tlen = (int) (a.iv_epos[sm.exons_epos] - sm.epos + 1);
                int this_it_ind = sm.exon_ind_epos;
                for (uint32_t q = start_table_ind + 1; q < end_table_ind; q++) {
                    this_it_ind++;
                    if (a.t2s[a.t2s_off[tid] + q] != 0) {
                        tlen += (int) (a.iv_epos[this_it_ind] - a.iv_spos[this_it_ind] + 1);
                        pre_zero = false;
                    } else {
                        if (!pre_zero) in++;
                        pre_zero = true;
                    }
                }
                tlen += (int) (lm.spos - a.iv_spos[lm.exons_spos] + 1);
            }
            if (tlen < min_tlen) {
                intron_num = in;
                min_tlen = tlen;
            }
        }
    return (min_tlen == CM_INF) ? -1 : (int) ((uint32_t) min_tlen + sm.matched_len - 1 + lm.matched_len - 1);
}

// same_gene(const MatchedMate &mm, const MatchedMate &other) (src/utils.cpp:628-639)
CM_HD bool cm_same_gene_mm(const AnnotDev &a, const MateD &mm, const MateD &other) {
    int iv = mm.exons_spos;
    for (uint32_t s = a.iv_seg_off[iv]; s < a.iv_seg_off[iv + 1]; s++) {
        uint32_t g = a.seg_gene_id[s];
        if (a.gene_start[g] <= other.spos && other.epos <= a.gene_end[g]) return true;
    }
    return false;
}
// same_gene(const IntervalInfo<UniqSeg> *mate, uint32_t s, uint32_t e) (src/utils.cpp:616-626)
CM_HD bool cm_same_gene_range(const AnnotDev &a, int iv, uint32_t s0, uint32_t e0) {
    for (uint32_t s = a.iv_seg_off[iv]; s < a.iv_seg_off[iv + 1]; s++) {
        uint32_t g = a.seg_gene_id[s];
        if (a.gene_start[g] <= s0 && e0 <= a.gene_end[g]) return true;
    }
    return false;
}
// any seg of iv_a sharing gene_id with any seg of iv_b
CM_HD bool cm_any_same_gene(const AnnotDev &a, int iv_a, int iv_b) {
    for (uint32_t i = a.iv_seg_off[iv_a]; i < a.iv_seg_off[iv_a + 1]; i++)
        for (uint32_t j = a.iv_seg_off[iv_b]; j < a.iv_seg_off[iv_b + 1]; j++)
            if (a.seg_gene_id[i] == a.seg_gene_id[j]) return true;
    return false;
}

struct FilterCtx {
    ExtCtx x;
    int contig_id;
};

CM_HD ConShiftD cm_con_shift(const AnnotDev &a, uint32_t loc) {
    int i = cm_get_shift(a, loc);
    ConShiftD c;
    c.chr = a.chr_id[i];
    c.shift = a.chr_shift[i];
    return c;
}

// UniqSeg::same_exon over two intervals' seg lists, visiting (i, j) in the reference's order
// concordant_explanation (src/utils.cpp:157-213); sm should start before lm
CM_HDN bool cm_concordant_explanation(FilterCtx &f, const MateD &sm, const MateD &lm, cm_mread &mr, ConShiftD cs, bool r1_sm,
                                      int pair_type) {
    const AnnotDev &a = f.x.v->a;
    const cm_params &p = f.x.v->p;
    if (sm.spos > lm.spos) return false;
    int32_t tlen;
    bool on_cdna = (sm.exons_spos >= 0) && (sm.exons_epos >= 0) && (lm.exons_spos >= 0) && (lm.exons_epos >= 0);
    if (sm.exons_spos < 0 || lm.exons_spos < 0) {
        tlen = (int32_t) (lm.spos - sm.epos - 1 + lm.matched_len + sm.matched_len);
        if (tlen <= p.max_tlen) mr_update(mr, sm, lm, cs, tlen, 0, false, CM_CONGNM, r1_sm, f.contig_id);
        else if (tlen <= CM_MAXDISCRDTLEN) mr_update(mr, sm, lm, cs, tlen, 0, false, CM_CONGNM, r1_sm, f.contig_id);
    } else {
        for (uint32_t i = a.iv_seg_off[sm.exons_spos]; i < a.iv_seg_off[sm.exons_spos + 1]; i++)
            for (uint32_t j = a.iv_seg_off[lm.exons_spos]; j < a.iv_seg_off[lm.exons_spos + 1]; j++)
                if (a.seg_start[i] == a.seg_start[j] && a.seg_end[i] == a.seg_end[j]) {
                    tlen = (int32_t) (lm.spos + lm.matched_len - sm.spos);
                    if (tlen <= p.max_tlen)
                        mr_update(mr, sm, lm, cs, tlen, 0, on_cdna, (pair_type == 0) ? CM_CONCRD : CM_CONGEN, r1_sm, f.contig_id);
                    else
                        mr_update(mr, sm, lm, cs, tlen, 0, on_cdna, CM_DISCRD, r1_sm, f.contig_id);
                }
    }
    if (sm.exons_epos < 0 || lm.exons_spos < 0) {
        tlen = (int32_t) (lm.spos - sm.epos - 1 + sm.matched_len + lm.matched_len);
        if (tlen <= p.max_tlen) mr_update(mr, sm, lm, cs, tlen, 0, false, CM_CONGNM, r1_sm, f.contig_id);
        else if (tlen <= CM_MAXDISCRDTLEN) mr_update(mr, sm, lm, cs, tlen, 0, false, CM_CONGNM, r1_sm, f.contig_id);
    } else {
        int intron_num = 0;
        tlen = cm_calc_tlen(a, sm, lm, intron_num);
        if (tlen >= 0 && tlen <= p.max_tlen) {
            mr_update(mr, sm, lm, cs, tlen, (uint16_t) intron_num, on_cdna, (pair_type == 0) ? CM_CONCRD : CM_CONGEN, r1_sm,
                      f.contig_id);
        } else {
            if (tlen < 0) {
                tlen = (int32_t) (lm.spos - sm.epos - 1 + sm.matched_len + lm.matched_len);
                intron_num = 0;
            }
            mr_update(mr, sm, lm, cs, tlen, (uint16_t) intron_num, on_cdna, CM_DISCRD, r1_sm, f.contig_id);
        }
    }
    return mr.type == CM_CONCRD;
}

// check_chimeric (src/utils.cpp:215-231)
CM_HD bool cm_check_chimeric(FilterCtx &f, const MateD &sm, const MateD &lm, cm_mread &mr, ConShiftD cs, bool r1_sm) {
    const AnnotDev &a = f.x.v->a;
    if (mr.type == CM_CONCRD) return false;
    if (sm.exons_spos < 0 || lm.exons_spos < 0) return false;
    if (cm_any_same_gene(a, sm.exons_spos, lm.exons_spos) && sm.spos < lm.spos) {
        mr_update(mr, sm, lm, cs, (int32_t) (lm.epos - sm.spos + 1), 0, false, CM_CHIORF, r1_sm, f.contig_id);
        return true;
    }
    return false;
}

// the "ciRNA" test shared by check_bsj / check_2bsj (src/utils.cpp:249-256, 299-306)
CM_HD bool cm_cirna(const AnnotDev &a, const MateD &sm, const MateD &lm) {
    return cm_bit(a.intronic, a.n_bits, sm.spos) && cm_bit(a.intronic, a.n_bits, lm.spos) && (sm.exon_ind_spos >= 0) &&
           (lm.exon_ind_epos >= 0) && (sm.exon_ind_spos == lm.exon_ind_epos) &&
           (sm.spos - a.iv_epos[sm.exon_ind_spos] <= (uint32_t) CM_LARIAT2BEGTH);
}

CM_HD bool cm_bsj_tail(FilterCtx &f, const MateD &sm, const MateD &lm, cm_mread &mr, ConShiftD cs, bool r1_sm, int type) {
    const AnnotDev &a = f.x.v->a;
    int32_t tlen = (int32_t) (lm.epos - sm.spos + 1);
    if (sm.exons_spos < 0 || lm.exons_spos < 0) {
        if ((sm.exons_spos >= 0 && cm_same_gene_mm(a, sm, lm)) || (lm.exons_spos >= 0 && cm_same_gene_mm(a, lm, sm))) {
            mr_update(mr, sm, lm, cs, tlen, 0, false, type, r1_sm, f.contig_id);
            return true;
        }
        if (cm_cirna(a, sm, lm)) {
            mr_update(mr, sm, lm, cs, tlen, 0, false, type, r1_sm, f.contig_id);
            return true;
        }
        return false;
    }
    if (cm_any_same_gene(a, sm.exons_spos, lm.exons_spos)) {
        mr_update(mr, sm, lm, cs, tlen, 0, false, type, r1_sm, f.contig_id);
        return true;
    }
    return false;
}

// check_bsj (src/utils.cpp:235-266)
CM_HD bool cm_check_bsj(FilterCtx &f, const MateD &sm, const MateD &lm, cm_mread &mr, ConShiftD cs, bool r1_sm) {
    if (mr.type == CM_CONCRD || mr.type == CM_DISCRD) return false;
    if (!sm.right_ok || !lm.left_ok) return false;
    return cm_bsj_tail(f, sm, lm, mr, cs, r1_sm, CM_CHIBSJ);
}

// check_2bsj (src/utils.cpp:270-318)
CM_HD bool cm_check_2bsj(FilterCtx &f, const MateD &sm, const MateD &lm, cm_mread &mr, ConShiftD cs, bool r1_sm) {
    if (mr.type < CM_CHI2BSJ) return false;
    if (sm.spos > lm.spos) return false;
    if (sm.right_ok && lm.right_ok && (sm.spos != lm.spos)) return false;
    if (sm.left_ok && lm.left_ok && (sm.epos != lm.epos)) return false;
    if (sm.left_ok && lm.right_ok) return false;
    return cm_bsj_tail(f, sm, lm, mr, cs, r1_sm, CM_CHI2BSJ);
}

// is_left_chain (src/utils.cpp:827-887)
CM_HDN bool cm_is_left_chain(const ChainView &a, const ChainView &b, int read_length, int kmer) {
    uint32_t a_beg = a.frags[0].rpos, b_beg = b.frags[0].rpos;
    uint32_t a_end = a.frags[a.chain_len - 1].rpos + kmer - 1;
    uint32_t b_end = b.frags[b.chain_len - 1].rpos + kmer - 1;
    bool non_overlaping = (b_beg > a_end) || (a_beg > b_end);
    if (non_overlaping) return a_beg < b_beg;
    uint32_t i = 0, j = 0;
    int best_distance = CM_INF;
    int best_i = -1, best_j = -1;
    while (i < a.chain_len && j < b.chain_len) {
        uint32_t bj_beg = b.frags[j].rpos;
        uint32_t ai_end = a.frags[i].rpos + kmer - 1;
        if (ai_end < bj_beg) {
            int distance = (int) (bj_beg - ai_end);
            if (distance < best_distance) { best_distance = distance; best_i = (int) i; best_j = (int) j; }
            ++i;
            continue;
        }
        uint32_t ai_beg = a.frags[i].rpos;
        uint32_t bj_end = b.frags[j].rpos + kmer - 1;
        if (bj_end < ai_beg) {
            int distance = (int) (ai_beg - bj_end);
            if (distance < best_distance) { best_distance = distance; best_i = (int) i; best_j = (int) j; }
            ++j;
            continue;
        }
        best_i = (int) i;
        best_j = (int) j;
        break;
    }
    if (best_i < 0 || best_j < 0) return a_beg < b_beg;   // unreachable: overlapping spans always enter the loop
    uint32_t common_bp = a.frags[best_i].rpos > b.frags[best_j].rpos ? a.frags[best_i].rpos : b.frags[best_j].rpos;
    int32_t a_ov_qpos = a.frags[best_i].qpos + (int32_t) (common_bp - a.frags[best_i].rpos);
    int32_t b_ov_qpos = b.frags[best_j].qpos + (int32_t) (common_bp - b.frags[best_j].rpos);
    if (a_ov_qpos < read_length && b_ov_qpos < read_length) return a_ov_qpos >= b_ov_qpos;
    return a_beg < b_beg;
}

// get_location_overlap (src/gene_annotation.cpp:538-551)
CM_HD int cm_location_overlap_ne(const AnnotDev &a, uint32_t pos) {
    int ind;
    return cm_location_overlap_ind(a, pos, ind);
}

// same_transcript(s, r, common_tid) != empty (src/utils.cpp:335-354)
CM_HD bool cm_same_transcript(const AnnotDev &a, int iv_s, int iv_r) {
    if (iv_s < 0 || iv_r < 0) return false;
    uint32_t b = a.seg_tid_off[a.iv_seg_off[iv_s]], e = a.seg_tid_off[a.iv_seg_off[iv_s + 1]];
    for (uint32_t k = b; k < e; k++)
        if (cm_iv_has_tid(a, iv_r, a.tid[k])) return true;
    return false;
}

// one strand of a pair as process_mates sees it
struct StrandView {
    ChainListView chains;
    const uint8_t *seq;   // forward_rec->seq or backward_rec->rcseq
    int seq_len;
};

// One MatePair of process_mates' main loop (src/filter.cpp:272-356): extend both chains, classify, update mr.
// Returns true when process_mates returns CONCRD at once (scan-lev 0).
struct MatesAcc {
    int min_ret1, min_ret2;
    bool r1_genic, r2_genic;
};
CM_HDN bool cm_mates_candidate(FilterCtx &f, const StrandView &fw, const StrandView &bw, int i, int j, int pair_type, int f_iv_i,
                               int b_iv_j, cm_mread &mr, bool r1_forward, MatesAcc &acc) {
    const AnnotDev &a = f.x.v->a;
    const cm_params &p = f.x.v->p;
    ChainView cf = cl_chain(fw.chains, i), cb = cl_chain(bw.chains, j);
    TidSet ts;
    ts.iv_a = pair_type == 0 ? f_iv_i : -1;
    ts.iv_b = pair_type == 0 ? b_iv_j : -1;
    MateD r1_mm, r2_mm;
    mate_init(r1_mm, p);
    mate_init(r2_mm, p);
    r1_mm.dir = 1;
    r2_mm.dir = -1;
    bool success;
    bool is_forward_left = cm_is_left_chain(cf, cb, fw.seq_len, p.kmer);
    if (is_forward_left) {
        success = cm_extend_both_mates(f.x, cf, cb, ts, fw.seq, bw.seq, 1, 1, fw.seq_len, bw.seq_len, r1_mm, r2_mm);
        if (success) {
            ConShiftD cs = cm_con_shift(a, r1_mm.spos);
            cm_overlap_to_epos(a, r1_mm);
            cm_overlap_to_spos(a, r1_mm);
            cm_overlap_to_epos(a, r2_mm);
            cm_overlap_to_spos(a, r2_mm);
            if (r1_mm.type == CM_CONCRD && r2_mm.type == CM_CONCRD) {
                if (cm_concordant_explanation(f, r1_mm, r2_mm, mr, cs, r1_forward, pair_type) && p.scan_level == 0) return true;
            } else if ((r1_mm.type == CM_CANDID && r2_mm.type == CM_CONCRD) || (r1_mm.type == CM_CONCRD && r2_mm.type == CM_CANDID)) {
                cm_check_bsj(f, r1_mm, r2_mm, mr, cs, r1_forward);
            } else if (r1_mm.type == CM_CANDID && r2_mm.type == CM_CANDID) {
                cm_check_2bsj(f, r1_mm, r2_mm, mr, cs, r1_forward);
            }
        }
    } else {
        success = cm_extend_both_mates(f.x, cb, cf, ts, bw.seq, fw.seq, 1, 1, bw.seq_len, fw.seq_len, r2_mm, r1_mm);
        if (success) {
            ConShiftD cs = cm_con_shift(a, r2_mm.spos);
            cm_overlap_to_epos(a, r1_mm);
            cm_overlap_to_spos(a, r1_mm);
            cm_overlap_to_epos(a, r2_mm);
            cm_overlap_to_spos(a, r2_mm);
            if (r1_mm.type == CM_CONCRD && r2_mm.type == CM_CONCRD) {
                cm_check_chimeric(f, r2_mm, r1_mm, mr, cs, !r1_forward);
            } else if ((r1_mm.type == CM_CANDID && r2_mm.type == CM_CONCRD) || (r1_mm.type == CM_CONCRD && r2_mm.type == CM_CANDID)) {
                cm_check_bsj(f, r2_mm, r1_mm, mr, cs, !r1_forward);
            } else if (r1_mm.type == CM_CANDID && r2_mm.type == CM_CANDID) {
                cm_check_2bsj(f, r2_mm, r1_mm, mr, cs, !r1_forward);
            }
        }
    }
    acc.min_ret1 = r1_mm.type < acc.min_ret1 ? r1_mm.type : acc.min_ret1;
    acc.min_ret2 = r2_mm.type < acc.min_ret2 ? r2_mm.type : acc.min_ret2;
    acc.r1_genic = (r1_mm.exons_spos >= 0) || (r1_mm.exons_epos >= 0);
    acc.r2_genic = (r2_mm.exons_spos >= 0) || (r2_mm.exons_epos >= 0);
    return false;
}

// pair_chains (src/filter.cpp:484-551) for one (i, j): 0 = not paired, else MatePair::type + 1
CM_HD int cm_pair_code(const AnnotDev &a, const cm_params &p, const ChainView &ci, const ChainView &cj, int f_iv_i, int b_iv_j,
                       int saved_type) {
    uint32_t fs = ci.frags[0].rpos, fe = ci.frags[ci.chain_len - 1].rpos + (uint32_t) p.kmer;
    uint32_t rs = cj.frags[0].rpos, re = cj.frags[cj.chain_len - 1].rpos + (uint32_t) p.kmer;
    int tlen = (fs < rs) ? (int) (re - fs) : (int) (fe - rs);
    bool same_tr = false, same_gen = false;
    if (f_iv_i >= 0 && b_iv_j >= 0) same_tr = cm_same_transcript(a, f_iv_i, b_iv_j);
    if (!same_tr && f_iv_i >= 0 && ((p.scan_level == 0 && saved_type > CM_CONGEN) || (p.scan_level > 0 && saved_type >= CM_CONGEN)))
        same_gen = cm_same_gene_range(a, f_iv_i, rs, re);
    if (!same_gen && b_iv_j >= 0 && (saved_type >= CM_CONGEN)) same_gen = cm_same_gene_range(a, b_iv_j, fs, fe);
    if (same_tr || same_gen || ((tlen <= CM_MAXDISCRDTLEN) && (saved_type >= CM_CONGNM))) return same_tr ? 1 : (same_gen ? 2 : 3);
    return 0;
}

// The rest of process_mates after the MatePair loop (src/filter.cpp:358-394): unpaired chains, fallback type.
CM_HDN int cm_mates_leftovers(FilterCtx &f, const StrandView &fw, const StrandView &bw, cm_mread &mr, uint32_t f_paired,
                              uint32_t b_paired, MatesAcc acc) {
    const AnnotDev &a = f.x.v->a;
    const cm_params &p = f.x.v->p;
    const int nf = fw.chains.count, nb = bw.chains.count;
    int min_ret1 = acc.min_ret1, min_ret2 = acc.min_ret2;
    bool r1_genic = acc.r1_genic, r2_genic = acc.r2_genic;
    if (mr.type == CM_CONCRD || mr.type == CM_DISCRD || mr.type == CM_CHIORF || mr.type == CM_CHIBSJ || mr.type == CM_CHI2BSJ)
        return mr.type;
    MateD mm1;
    mate_init(mm1, p);
    if (min_ret1 != CM_CONCRD)
        for (int i = 0; i < nf; i++)
            if (!((f_paired >> i) & 1)) {
                int ex_ret = cm_extend_chain_both_sides(f.x, cl_chain(fw.chains, i), fw.seq, fw.seq_len, mm1, 1);
                min_ret1 = ex_ret < min_ret1 ? ex_ret : min_ret1;
                cm_overlap_to_spos(a, mm1);
                cm_overlap_to_epos(a, mm1);
                r1_genic = (mm1.exons_spos >= 0) || (mm1.exons_epos >= 0);
            }
    MateD mm2;
    mate_init(mm2, p);
    if (min_ret2 != CM_CONCRD)
        for (int j = 0; j < nb; j++)
            if (!((b_paired >> j) & 1)) {
                int ex_ret = cm_extend_chain_both_sides(f.x, cl_chain(bw.chains, j), bw.seq, bw.seq_len, mm2, -1);
                min_ret2 = ex_ret < min_ret2 ? ex_ret : min_ret2;
                cm_overlap_to_spos(a, mm2);
                cm_overlap_to_epos(a, mm2);
                r2_genic = (mm2.exons_spos >= 0) || (mm2.exons_epos >= 0);
            }
    int new_type = (((min_ret1 == CM_ORPHAN) && (min_ret2 == CM_CONCRD)) || ((min_ret1 == CM_CONCRD) && (min_ret2 == CM_ORPHAN)))
                       ? CM_OEANCH
                       : ((min_ret1 == CM_ORPHAN) || (min_ret2 == CM_ORPHAN))
                             ? CM_ORPHAN
                             : ((min_ret1 == CM_CONCRD) && (min_ret2 == CM_CONCRD) && (r1_genic && r2_genic))
                                   ? CM_CHIFUS
                                   : ((min_ret1 == CM_CONCRD) && (min_ret2 == CM_CONCRD)) ? CM_OEA2 : CM_CANDID;
    mr_update_type(mr, new_type);
    return mr.type;
}


// FilterRead::process_mates (src/filter.cpp:244-395) with pair_chains (:484-551) folded in: the
// MatePair vector is replaced by a 2-bit code per (i, j) visited in the same order.
CM_HDN int cm_process_mates(FilterCtx &f, const StrandView &fw, const StrandView &bw, cm_mread &mr, bool r1_forward) {
    const AnnotDev &a = f.x.v->a;
    const cm_params &p = f.x.v->p;
    const int nf = fw.chains.count, nb = bw.chains.count;
    const int saved_type = mr.type;
    int f_iv[CM_MAX_CHAINS], b_iv[CM_MAX_CHAINS];
    for (int i = 0; i < nf; i++) f_iv[i] = cm_location_overlap_ne(a, cl_chain(fw.chains, i).frags[0].rpos);
    for (int j = 0; j < nb; j++) b_iv[j] = cm_location_overlap_ne(a, cl_chain(bw.chains, j).frags[0].rpos);
    uint64_t rowc[CM_MAX_CHAINS];   // row i of the code matrix, 2 bits per j
    uint32_t f_paired = 0, b_paired = 0;
    for (int i = 0; i < nf; i++) {
        ChainView ci = cl_chain(fw.chains, i);
        uint64_t r = 0;
        for (int j = 0; j < nb; j++) {
            int c = cm_pair_code(a, p, ci, cl_chain(bw.chains, j), f_iv[i], b_iv[j], saved_type);
            if (c) {
                r |= (uint64_t) c << (2 * j);
                b_paired |= 1u << j;
            }
        }
        rowc[i] = r;
        if (r) f_paired |= 1u << i;
    }
    MatesAcc acc;
    acc.min_ret1 = acc.min_ret2 = CM_ORPHAN;
    acc.r1_genic = acc.r2_genic = false;
    for (int i = 0; i < nf; i++) {
        uint64_t r = rowc[i];
        for (int j = 0; r; j++, r >>= 2) {
            const int c = (int) (r & 3u);
            if (!c) continue;
            if (cm_mates_candidate(f, fw, bw, i, j, c - 1, f_iv[i], b_iv[j], mr, r1_forward, acc)) return CM_CONCRD;
        }
    }
    return cm_mates_leftovers(f, fw, bw, mr, f_paired, b_paired, acc);
}

// one pair: the four strands' chain lists + high_hits (get_best_chains results), R1 fwd, R1 rc, R2 fwd, R2 rc
struct PairView {
    StrandView s[4];
    int high_hits[4];
};

// FilterRead::process_read, PE (src/filter.cpp:124-241)
CM_HDN int cm_process_read_pe(FilterCtx &f, const PairView &pv, cm_mread &mr) {
    const cm_params &p = f.x.v->p;
    const int nf1 = pv.s[0].chains.count, nb1 = pv.s[1].chains.count, nf2 = pv.s[2].chains.count, nb2 = pv.s[3].chains.count;
    if (nf1 + nb1 + nf2 + nb2 <= 0) {
        if ((pv.high_hits[0] + pv.high_hits[1] > 0) && (pv.high_hits[2] + pv.high_hits[3] > 0)) {
            mr_update_type(mr, CM_NOPROC_MANYHIT);
            return CM_NOPROC_MANYHIT;
        }
        mr_update_type(mr, CM_NOPROC_NOMATCH);
        return CM_NOPROC_NOMATCH;
    }
    if ((nf1 + nb1 <= 0) || (nf2 + nb2 <= 0)) {
        mr_update_type(mr, CM_OEANCH);
        return CM_OEANCH;
    }
    float fc1 = nf1 > 0 ? pv.s[0].chains.score[0] : 0.f;
    float bc1 = nb1 > 0 ? pv.s[1].chains.score[0] : 0.f;
    float fc2 = nf2 > 0 ? pv.s[2].chains.score[0] : 0.f;
    float bc2 = nb2 > 0 ? pv.s[3].chains.score[0] : 0.f;
#if defined(__CUDA_ARCH__)
    bool first = __fadd_rn(fc1, bc2) >= __fadd_rn(fc2, bc1);
#else
    volatile float lhs = fc1 + bc2, rhs = fc2 + bc1;
    bool first = lhs >= rhs;
#endif
    // attempt A: forward R1 / backward R2 ; attempt B: forward R2 / backward R1
    for (int k = 0; k < 2; k++) {
        bool attempt_a = (k == 0) == first;
        int r = attempt_a ? cm_process_mates(f, pv.s[0], pv.s[3], mr, true) : cm_process_mates(f, pv.s[2], pv.s[1], mr, false);
        if (p.scan_level == 0 && r == CM_CONCRD) return CM_CONCRD;
    }
    return mr.type;
}

// FilterRead::process_read, SE (src/filter.cpp:86-121): s[0] forward, s[1] reverse complement
CM_HDN int cm_process_read_se(FilterCtx &f, const StrandView &fw, const StrandView &bw) {
    const cm_params &p = f.x.v->p;
    MateD mr;
    mate_init(mr, p);
    int min_ret = CM_ORPHAN;
    for (int i = 0; i < fw.chains.count; i++) {
        int ex_ret = cm_extend_chain_both_sides(f.x, cl_chain(fw.chains, i), fw.seq, fw.seq_len, mr, 1);
        if (ex_ret == CM_CONCRD) return CM_CONCRD;
        if (ex_ret < min_ret) min_ret = ex_ret;
    }
    for (int i = 0; i < bw.chains.count; i++) {
        int ex_ret = cm_extend_chain_both_sides(f.x, cl_chain(bw.chains, i), bw.seq, bw.seq_len, mr, -1);
        if (ex_ret == CM_CONCRD) return CM_CONCRD;
        if (ex_ret < min_ret) min_ret = ex_ret;
    }
    return min_ret;
}
