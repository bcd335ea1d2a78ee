// Round semantics of the reference's mapping loop (src/circminer.cpp:229-308), as pure functions shared by the
// device kernels (k_round_commit / k_round_compact) and the g++ build in tests/hostsim.
#pragma once
#include "cm_types.h"

CM_HD bool cm_type_mapped(int t) {   // the types whose coordinates the header carries (src/filter.cpp:422-423)
    return t == CM_CONCRD || t == CM_DISCRD || t == CM_CHIORF || t == CM_CHIBSJ || t == CM_CHI2BSJ || t == CM_CONGNM || t == CM_CONGEN;
}
// what FASTQParser::fill_map_info rebuilds from the header write_read_category wrote
CM_HD cm_mread cm_header_round_trip(const cm_mread &m, const cm_params &p) {
    if (cm_type_mapped(m.type)) return m;
    cm_mread r;
    r.spos_r1 = r.spos_r2 = r.epos_r1 = r.epos_r2 = 0;
    r.qspos_r1 = r.qspos_r2 = r.qepos_r1 = r.qepos_r2 = 0;
    r.mlen_r1 = r.mlen_r2 = 0;
    r.ed_r1 = r.ed_r2 = p.max_ed + 1;
    r.type = m.type;
    r.tlen = CM_INF;
    r.contig_num = 0;
    r.chr_r1 = r.chr_r2 = -1;
    r.junc_num = 0;
    r.r1_forward = r.r2_forward = 1;
    r.gm_compatible = 0;
    r.pad_[0] = r.pad_[1] = r.pad_[2] = 0;
    return r;
}

// the skip rule of map_reads (src/circminer.cpp:386-390): true = the pair is printed in this round and never sees the next contig
CM_HD bool cm_skip_rule(const cm_params &p, int state, const cm_mread &m, int len_sum) {
    if (p.scan_level == 0) return state == CM_CONCRD;
    if (p.scan_level == 1)
        return state == CM_CONCRD && m.gm_compatible && (m.ed_r1 + m.ed_r2 == 0) && ((int) (m.mlen_r1 + m.mlen_r2) == len_sum);
    return false;
}
