// Per-strand and per-pair drivers shared by the CUDA kernels (cm_gpu.cu) and the g++ build in
// tests/hostsim (which exists only to debug this logic against the reference without a GPU).
#pragma once
#include "cm_types.h"
#include "cm_seed.h"
#include "cm_annot.h"
#include "cm_chain.h"
#include "cm_align.h"
#include "cm_extend.h"
#include "cm_filter.h"
#include "cm_rounds.h"

// Device-resident batch layout (all arrays indexed by strand; PE pair i owns strands 4i..4i+3 =
// R1 fwd, R1 rc, R2 fwd, R2 rc; SE read i owns 2i, 2i+1).
struct BatchDev {
    const uint8_t *strands;      // [n_strands][stride] ASCII
    const uint16_t *strand_len;  // [n_strands]
    uint32_t n_strands;
    uint32_t stride;
    int F;                       // fragment slots per chain in this batch
    int C;                       // chain slots per strand (max_chain_len)
    uint32_t *seed_off;          // [n_strands][n_slots]
    uint32_t *seed_cnt;          // [n_strands][n_slots]  raw equal-range sizes
    int32_t *chain_cnt;          // [n_strands]
    int32_t *chain_hh;           // [n_strands] high_hits
    float *chain_score;          // [n_strands][C]
    uint32_t *chain_len;         // [n_strands][C]
    FragDev *chain_frags;        // [n_strands][C][F]
    const uint32_t *chain_base;  // [n_strands + 1] exclusive prefix of min(chain_cnt, spec_k) (nullptr: no pre-alignment)
    ChainAux *aux;               // [pre-aligned chains]
    EndSpec *spec;               // [pre-aligned chains][2] transcript-walk DPs (nullptr: none)
    int spec_k;                  // leading chains of each strand that k_chain_spec pre-aligns
};

// GenomeSeeder::split_match_hash for one strand (src/match_read.cpp:270-286): fills n_slots (off, cnt)
CM_HD void cm_strand_seed(const DevView &v, const uint8_t *seq, int len, uint32_t *seed_off, uint32_t *seed_cnt) {
    for (int t = 0; t < v.n_slots; t++) {
        SeedHit h;
        h.off = 0;
        h.cnt = 0;
        int q = t * v.p.kmer;
        if (q + v.p.kmer <= len) {
            int hv, cv;
            cm_hash_kmer(seq + q, v.p.kmer, hv, cv);
            h = cm_seed_lookup(v.c, hv, cv, nullptr);
        }
        seed_off[t] = h.off;
        seed_cnt[t] = h.cnt;
    }
}

CM_HD ChainListView cm_chain_list(const BatchDev &b, uint32_t strand) {
    ChainListView l;
    l.count = b.chain_cnt[strand];
    l.score = b.chain_score + (size_t) strand * b.C;
    l.len = b.chain_len + (size_t) strand * b.C;
    l.frags = b.chain_frags + (size_t) strand * b.C * b.F;
    l.F = b.F;
    l.aux = (b.aux && b.chain_base) ? b.aux + b.chain_base[strand] : nullptr;
    l.spec = (l.aux && b.spec) ? b.spec + 2 * (size_t) b.chain_base[strand] : nullptr;
    l.n_aux = l.count < b.spec_k ? l.count : b.spec_k;
    return l;
}

CM_HD StrandView cm_strand_view(const BatchDev &b, uint32_t strand) {
    StrandView s;
    s.chains = cm_chain_list(b, strand);
    s.seq = b.strands + (size_t) strand * b.stride;
    s.seq_len = b.strand_len[strand];
    return s;
}

// FilterRead::process_read (PE) for pair i of the batch
CM_HD int cm_pair_process(const DevView &v, const BatchDev &b, uint32_t pair, cm_mread &mr, AlignStats &st) {
    FilterCtx f;
    ext_ctx_init(f.x, &v);
    f.contig_id = v.contig_id;
    PairView pv;
    for (int k = 0; k < 4; k++) {
        pv.s[k] = cm_strand_view(b, 4 * pair + k);
        pv.high_hits[k] = b.chain_hh[4 * pair + k];
    }
    int r = cm_process_read_pe(f, pv, mr);
    st = f.x.st;
    return r;
}

// SE overload for read i
CM_HD int cm_read_process_se(const DevView &v, const BatchDev &b, uint32_t read, AlignStats &st) {
    FilterCtx f;
    ext_ctx_init(f.x, &v);
    f.contig_id = v.contig_id;
    StrandView fw = cm_strand_view(b, 2 * read), bw = cm_strand_view(b, 2 * read + 1);
    int r = cm_process_read_se(f, fw, bw);
    st = f.x.st;
    return r;
}
