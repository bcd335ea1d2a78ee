"""TEST INFRASTRUCTURE ONLY.  ctypes wrapper over oracle/_ref/libcm_ref.so (the unmodified
reference + oracle/ref_harness.cpp).  The reference keeps everything in process-global
singletons, so one process can open exactly one (reference, GTF) pair; use `python -m
oracle.refdump` (a subprocess) to produce stage dumps for a dataset.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libcm_ref.so")
REF_BIN = os.path.join(HERE, "_ref", "circminer")


class CmParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "kmer", "max_read_len", "max_ed", "max_sc", "band_width", "seed_lim", "max_tlen", "max_intron_len",
        "max_chain_len", "scan_level", "mat_sc", "mis_sc", "ind_sc", "xdrop")]


def default_params(**kw) -> CmParams:
    p = CmParams(20, 300, 4, 7, 3, 500, 500, 2000000, 30, 0, 1, -3, -3, 8)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


MREAD_DTYPE = np.dtype([
    ("spos_r1", "<u4"), ("spos_r2", "<u4"), ("epos_r1", "<u4"), ("epos_r2", "<u4"),
    ("qspos_r1", "<u4"), ("qspos_r2", "<u4"), ("qepos_r1", "<u4"), ("qepos_r2", "<u4"),
    ("mlen_r1", "<u4"), ("mlen_r2", "<u4"), ("ed_r1", "<i4"), ("ed_r2", "<i4"), ("type", "<i4"),
    ("tlen", "<i4"), ("contig_num", "<i4"), ("chr_r1", "<i4"), ("chr_r2", "<i4"), ("junc_num", "<u2"),
    ("r1_forward", "u1"), ("r2_forward", "u1"), ("gm_compatible", "u1"), ("pad_", "u1", (3,))])
assert MREAD_DTYPE.itemsize == 76


MATE_DTYPE = np.dtype([("spos", "<u4"), ("epos", "<u4"), ("qspos", "<u4"), ("qepos", "<u4"), ("matched_len", "<u4"),
                       ("left_ed", "<i4"), ("middle_ed", "<i4"), ("right_ed", "<i4"), ("sclen_left", "<i4"), ("sclen_right", "<i4"),
                       ("dir", "<i4"), ("type", "<i4"), ("is_concord", "u1"), ("left_ok", "u1"), ("right_ok", "u1"), ("pad_", "u1")])
assert MATE_DTYPE.itemsize == 52


def fresh_mreads(n: int, max_ed: int = 4) -> np.ndarray:
    """MatchedRead::MatchedRead(void), /root/reference/src/common.cpp:268-271 (unset fields 0)."""
    m = np.zeros(n, MREAD_DTYPE)
    m["r1_forward"] = 1
    m["r2_forward"] = 1
    m["ed_r1"] = max_ed + 1
    m["ed_r2"] = max_ed + 1
    m["type"] = 13
    m["tlen"] = 1000000000
    m["chr_r1"] = -1
    m["chr_r2"] = -1
    return m


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class RefHarness:
    def __init__(self, params: CmParams, threads: int = 1):
        if not os.path.exists(REF_SO):
            raise RuntimeError(f"{REF_SO} missing: run `make -C oracle ref` where /root/reference exists")
        self.lib = C.CDLL(REF_SO)
        L = self.lib
        L.cmref_hash_table.restype = C.c_void_p
        L.cmref_hash_mem.restype = C.c_void_p
        L.cmref_hash_mem_size.restype = C.c_uint64
        L.cmref_cref_gen.restype = C.c_void_p
        L.cmref_ref_gen_len.restype = C.c_uint32
        L.cmref_chr_name.restype = C.c_char_p
        self.params = params
        L.cmref_set_params(C.byref(params), threads)

    def build_index(self, ref_fa: str, contig_size: int = 0, compact: bool = False):
        r = self.lib.cmref_build_index(ref_fa.encode(), C.c_uint(contig_size), int(compact))
        if r != 0:
            raise RuntimeError("reference index build failed")

    def open(self, ref_fa: str, gtf: str, tmp_prefix: str):
        r = self.lib.cmref_open(ref_fa.encode(), gtf.encode(), tmp_prefix.encode())
        if r != 0:
            raise RuntimeError(f"cmref_open failed: {r}")
        self.kmer = self.lib.cmref_kmer()
        self.S = self.lib.cmref_max_seg_cnt()
        self.C_ = self.params.max_chain_len
        self.F = (self.S + 1) // 2
        self.chr_names = [self.lib.cmref_chr_name(i).decode() for i in range(self.lib.cmref_n_chr())]

    def chr_info(self):
        out = []
        for i in range(len(self.chr_names)):
            a, b, c = C.c_uint32(), C.c_uint32(), C.c_uint32()
            self.lib.cmref_chr_info(i, C.byref(a), C.byref(b), C.byref(c))
            out.append((self.chr_names[i], a.value, b.value, c.value))
        return out

    def next_contig(self) -> int:
        r = self.lib.cmref_next_contig()
        if r < 0:
            raise RuntimeError("cmref_next_contig failed")
        self.contig = self.lib.cmref_contig_num()
        return r

    # raw in-memory index, as a maintainer's shim would pass it
    def index_pointers(self):
        L = self.lib
        return (L.cmref_hash_table(), L.cmref_hash_mem(), L.cmref_hash_mem_size(), L.cmref_cref_gen(),
                L.cmref_ref_gen_len())

    def genome_chars(self, start: int, length: int) -> bytes:
        buf = C.create_string_buffer(length + 1)
        r = self.lib.cmref_genome_chars(C.c_uint32(start), length, buf)
        if r != 0:
            raise RuntimeError("pac2char out of range")
        return buf.raw[:length]

    def annotation(self) -> dict:
        L = self.lib
        n = [C.c_uint32() for _ in range(7)]
        L.cmref_annot_sizes(*[C.byref(x) for x in n])
        n_iv, n_seg, n_tid, n_trans, n_t2s, n_genes, n_chr = [x.value for x in n]
        u4 = lambda k: np.zeros(max(k, 1), np.uint32)
        a = dict(iv_spos=u4(n_iv), iv_epos=u4(n_iv), iv_max_end=u4(n_iv), iv_min_end=u4(n_iv),
                 iv_max_next_exon=u4(n_iv), iv_seg_off=u4(n_iv + 1), seg_start=u4(n_seg), seg_end=u4(n_seg),
                 seg_next_exon_beg=u4(n_seg), seg_gene_id=u4(n_seg), seg_tid_off=u4(n_seg + 1), tid=u4(n_tid),
                 trans_start_ind=np.zeros(max(n_trans, 1), np.int32), t2s_off=u4(n_trans + 1),
                 t2s=np.zeros(max(n_t2s, 1), np.uint8), gene_start=u4(n_genes), gene_end=u4(n_genes),
                 chr_shift=u4(n_chr), chr_id=np.zeros(max(n_chr, 1), np.int32))
        order = ["iv_spos", "iv_epos", "iv_max_end", "iv_min_end", "iv_max_next_exon", "iv_seg_off", "seg_start",
                 "seg_end", "seg_next_exon_beg", "seg_gene_id", "seg_tid_off", "tid", "trans_start_ind", "t2s_off",
                 "t2s", "gene_start", "gene_end", "chr_shift", "chr_id"]
        L.cmref_annot_fill(*[a[k].ctypes.data_as(C.c_void_p) for k in order])
        n_bits = int(L.cmref_ref_gen_len()) + 64
        nw = (n_bits + 63) // 64
        nb = np.zeros(nw, np.uint64)
        ib = np.zeros(nw, np.uint64)
        L.cmref_annot_masks(C.c_uint64(n_bits), nb.ctypes.data_as(C.c_void_p), ib.ctypes.data_as(C.c_void_p))
        a.update(near_border=nb, intronic=ib, n_bits=n_bits, n_iv=n_iv, n_seg=n_seg, n_tid=n_tid, n_trans=n_trans,
                 n_t2s=n_t2s, n_genes=n_genes, n_chr=n_chr)
        # trim the max(k,1) padding
        for k, cnt in (("iv_spos", n_iv), ("iv_epos", n_iv), ("iv_max_end", n_iv), ("iv_min_end", n_iv),
                       ("iv_max_next_exon", n_iv), ("seg_start", n_seg), ("seg_end", n_seg),
                       ("seg_next_exon_beg", n_seg), ("seg_gene_id", n_seg), ("tid", n_tid),
                       ("trans_start_ind", n_trans), ("t2s", n_t2s), ("gene_start", n_genes), ("gene_end", n_genes),
                       ("chr_shift", n_chr), ("chr_id", n_chr)):
            a[k] = a[k][:cnt]
        return a

    def annot_lookup(self, pos, mlen, rlen) -> dict:
        pos = np.ascontiguousarray(pos, np.uint32)
        n = len(pos)
        out = dict(ub=np.zeros(n, np.uint32), max_end=np.zeros(n, np.uint32), ol_iv=np.zeros(n, np.int32), loc_iv=np.zeros(n, np.int32),
                   loc_ind=np.zeros(n, np.int32), shift_idx=np.zeros(n, np.int32), intronic=np.zeros(n, np.uint8))
        self.lib.cmref_annot_lookup(C.c_uint32(n), pos.ctypes.data_as(C.c_void_p), C.c_uint32(mlen), C.c_uint32(rlen),
                                    *[out[k].ctypes.data_as(C.c_void_p) for k in
                                      ("ub", "max_end", "ol_iv", "loc_iv", "loc_ind", "shift_idx", "intronic")])
        return out

    def extend(self, mode, prob: dict) -> dict:
        """TransExtension::extend_both_mates (mode 0) / extend_chain_both_sides (mode 1) on the problems of `prob`
        (arrays named as the arguments of cm_gpu_extend_batch)."""
        n = len(prob["llen"])
        F = prob["l_rpos"].shape[1]
        res = dict(lmm=np.zeros(n, MATE_DTYPE), rmm=np.zeros(n, MATE_DTYPE), ret=np.zeros(n, np.int32))
        vp = lambda a: None if a is None else np.ascontiguousarray(a).ctypes.data_as(C.c_void_p)
        keep = {k: (None if v is None else np.ascontiguousarray(v)) for k, v in prob.items()}
        g = lambda k: None if keep.get(k) is None else keep[k].ctypes.data_as(C.c_void_p)
        self.lib.cmref_extend(mode, C.c_uint32(n), g("lseq"), g("rseq"), C.c_uint32(keep["lseq"].shape[1]), g("llen"), g("rlen"),
                              C.c_uint32(F), g("l_chain_len"), g("l_rpos"), g("l_qpos"), g("r_chain_len"), g("r_rpos"), g("r_qpos"),
                              g("iv_a"), g("iv_b"), g("dir"), vp(res["lmm"]), vp(res["rmm"]), vp(res["ret"]))
        return res

    def circ_chain(self, prob: dict, ws: int = 8, step: int = 3) -> dict:
        """ProcessCirc::chaining (src/process_circ.cpp:678) per row of `prob` (seqs, qspos, qepos, gene_start, gene_end) on the
        RegionalHashTable of that gene."""
        a = np.ascontiguousarray(prob["seqs"], np.uint8)
        n, Cn = a.shape[0], self.params.max_chain_len
        F = (self.params.max_read_len - ws) // step + 1
        u = {k: np.ascontiguousarray(prob[k], np.uint32) for k in ("qspos", "qepos", "gene_start", "gene_end")}
        out = dict(chain_count=np.zeros(n, np.int32), score=np.zeros((n, Cn), np.float32), chain_len=np.zeros((n, Cn), np.uint32),
                   frag_rpos=np.zeros((n, Cn, F), np.uint32), frag_qpos=np.zeros((n, Cn, F), np.int32))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.cmref_circ_chain(C.c_uint32(n), vp(a), C.c_uint32(a.shape[1]), vp(u["qspos"]), vp(u["qepos"]), vp(u["gene_start"]),
                                      vp(u["gene_end"]), ws, step, F, *[vp(out[k]) for k in ("chain_count", "score", "chain_len", "frag_rpos", "frag_qpos")])
        if r != 0:
            raise RuntimeError(f"cmref_circ_chain failed: {r}")
        return out


    def circ_exact(self, prob: dict, window_size=8) -> dict:
        """ProcessCirc::find_exact_coord per row (src/process_circ.cpp:739); `prob` holds arrays named as the C arguments
        (seqs, r1_spos, r1_epos, r2_spos, r2_epos, dir, qspos, rlen, whole_len, chain_len, frag_rpos, frag_qpos)."""
        pass
        a = np.ascontiguousarray(prob["seqs"], np.uint8)
        n = a.shape[0]
        u = {k: np.ascontiguousarray(prob[k], np.uint32) for k in ("r1_spos", "r1_epos", "r2_spos", "r2_epos", "qspos", "chain_len", "frag_rpos")}
        s = {k: np.ascontiguousarray(prob[k], np.int32) for k in ("dir", "rlen", "whole_len", "frag_qpos")}
        res = dict(pmm=np.zeros(n, MATE_DTYPE), ret=np.zeros(n, np.int32))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.cmref_circ_exact(C.c_uint32(n), vp(a), C.c_uint32(a.shape[1]), vp(u["r1_spos"]), vp(u["r1_epos"]), vp(u["r2_spos"]), vp(u["r2_epos"]),
               vp(s["dir"]), vp(u["qspos"]), vp(s["rlen"]), vp(s["whole_len"]), C.c_uint32(u["frag_rpos"].shape[1]), vp(u["chain_len"]),
               vp(u["frag_rpos"]), vp(s["frag_qpos"]), window_size, vp(res["pmm"]), vp(res["ret"]))
        assert r == 0, r
        return res

    def circ_single(self, s1, s2, l1, l2, mreads, ws: int = 8, cap: int = 20000) -> dict:
        """ProcessCirc::call_circ_single_split's path for every CHIBSJ pair of the current contig with every check_split_map call
        recorded (cmref_circ_single)."""
        from circminer_b200.api import CIRC_MATE_DTYPE, CIRC_RES_DTYPE
        a1, a2 = np.ascontiguousarray(s1, np.uint8), np.ascontiguousarray(s2, np.uint8)
        n, stride = a1.shape
        l1, l2 = np.ascontiguousarray(l1, np.uint16), np.ascontiguousarray(l2, np.uint16)
        mr = np.ascontiguousarray(mreads)
        rec = dict(call_pair=np.zeros(cap, np.uint32), r1_partial=np.zeros(cap, np.uint8), remain_rows=np.zeros((cap, stride), np.uint8),
                   fullmap_rows=np.zeros((cap, stride), np.uint8), remain_len=np.zeros(cap, np.int32), fullmap_len=np.zeros(cap, np.int32),
                   in_r1=np.zeros(cap, CIRC_MATE_DTYPE), in_r2=np.zeros(cap, CIRC_MATE_DTYPE), in_pm=np.zeros(cap, CIRC_MATE_DTYPE),
                   ret=np.zeros(cap, np.int32), cr=np.zeros(cap, CIRC_RES_DTYPE), out_r1=np.zeros(cap, CIRC_MATE_DTYPE),
                   out_r2=np.zeros(cap, CIRC_MATE_DTYPE), out_pm=np.zeros(cap, CIRC_MATE_DTYPE))
        best = np.zeros(n, np.int32)
        gene_cap = 16
        mr_con = np.zeros(n, mr.dtype)
        gene_cnt = np.zeros(n, np.uint32)
        gene_se = np.zeros((n, gene_cap, 2), np.uint32)
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        order = ("call_pair", "r1_partial", "remain_rows", "fullmap_rows", "remain_len", "fullmap_len", "in_r1", "in_r2", "in_pm", "ret", "cr",
                 "out_r1", "out_r2", "out_pm")
        k = self.lib.cmref_circ_single(C.c_uint32(n), vp(a1), vp(a2), C.c_uint32(stride), vp(l1), vp(l2), vp(mr), ws, C.c_uint32(cap),
                                       *[vp(rec[x]) for x in order], vp(best), vp(mr_con), C.c_uint32(gene_cap), vp(gene_cnt), vp(gene_se))
        if k < 0:
            raise RuntimeError(f"cmref_circ_single failed: {k}")
        out = {x: rec[x][:k] for x in order}
        out["best_type"] = best
        out["mr_con"], out["gene_cnt"], out["gene_se"] = mr_con, gene_cnt, gene_se
        return out

    def circ_single_real(self, s1, s2, l1, l2, mreads, ws: int = 8, type_cap: int = 64) -> dict:
        """The unmodified ProcessCirc::call_circ on every CHIBSJ pair of the current contig: the types it writes to candidates.pam
        and what it pushes to circ_res (cmref_circ_single_real)."""
        a1, a2 = np.ascontiguousarray(s1, np.uint8), np.ascontiguousarray(s2, np.uint8)
        n, stride = a1.shape
        l1, l2 = np.ascontiguousarray(l1, np.uint16), np.ascontiguousarray(l2, np.uint16)
        mr = np.ascontiguousarray(mreads)
        out = dict(n_types=np.zeros(n, np.uint32), types=np.zeros((n, type_cap), np.int32), pushed=np.zeros(n, np.uint8),
                   push_type=np.zeros(n, np.int32), push_spos=np.zeros(n, np.uint32), push_epos=np.zeros(n, np.uint32))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.cmref_circ_single_real(C.c_uint32(n), vp(a1), vp(a2), C.c_uint32(stride), vp(l1), vp(l2), vp(mr), ws, C.c_uint32(type_cap),
                                            *[vp(out[k]) for k in ("n_types", "types", "pushed", "push_type", "push_spos", "push_epos")])
        assert r == 0, r
        return out

    def circ_call_real(self, s1, s2, l1, l2, mreads, want_type: int, ws: int = 8, push_cap: int = 4) -> dict:
        """The unmodified ProcessCirc::call_circ on every pair of type want_type of the current contig: its candidates.pam text per pair
        and everything it pushes to circ_res (cmref_circ_call_real)."""
        a1, a2 = np.ascontiguousarray(s1, np.uint8), np.ascontiguousarray(s2, np.uint8)
        n, stride = a1.shape
        l1, l2 = np.ascontiguousarray(l1, np.uint16), np.ascontiguousarray(l2, np.uint16)
        mr = np.ascontiguousarray(mreads)
        cap = 64 << 20
        text = np.zeros(cap, np.uint8)
        out = dict(text_off=np.zeros(n + 1, np.uint64), n_push=np.zeros(n, np.uint32), push_type=np.zeros((n, push_cap), np.int32),
                   push_spos=np.zeros((n, push_cap), np.uint32), push_epos=np.zeros((n, push_cap), np.uint32),
                   push_sig=np.zeros((n, push_cap, 8), np.uint8))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.cmref_circ_call_real(C.c_uint32(n), vp(a1), vp(a2), C.c_uint32(stride), vp(l1), vp(l2), vp(mr), ws, want_type, C.c_uint64(cap),
                                          vp(out["text_off"]), vp(text), C.c_uint32(push_cap),
                                          *[vp(out[k]) for k in ("n_push", "push_type", "push_spos", "push_epos", "push_sig")])
        assert r == 0, r
        out["text"] = text[:int(out["text_off"][-1])].copy()
        return out

    def circ_check(self, rec: dict, ws: int = 8) -> dict:
        """ProcessCirc::check_split_map on given inputs (same record layout as Mapper.circ_check)."""
        from circminer_b200.api import CIRC_MATE_DTYPE, CIRC_RES_DTYPE
        rr, fr = np.ascontiguousarray(rec["remain_rows"], np.uint8), np.ascontiguousarray(rec["fullmap_rows"], np.uint8)
        n = rr.shape[0]
        rl, fl = np.ascontiguousarray(rec["remain_len"], np.int32), np.ascontiguousarray(rec["fullmap_len"], np.int32)
        rp = np.ascontiguousarray(rec["r1_partial"], np.uint8)
        out = dict(r1=np.array(rec["in_r1"], CIRC_MATE_DTYPE), r2=np.array(rec["in_r2"], CIRC_MATE_DTYPE), pm=np.array(rec["in_pm"], CIRC_MATE_DTYPE),
                   ret=np.zeros(n, np.int32), cr=np.zeros(n, CIRC_RES_DTYPE))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.cmref_circ_check(C.c_uint32(n), vp(rr), vp(fr), C.c_uint32(rr.shape[1]), vp(rl), vp(fl), vp(rp), vp(out["r1"]), vp(out["r2"]),
                                      vp(out["pm"]), ws, vp(out["ret"]), vp(out["cr"]))
        assert r == 0, r
        return out

    def circ_check4(self, rec: dict, ws: int = 8) -> dict:
        """ProcessCirc::check_split_map on four mates (process_circ.cpp:922) on given inputs: r1_rows, r2_rows, l1, l2, fullmap_sel,
        r1_1, r2_1, r1_2, r2_2 -> ret, cr."""
        from circminer_b200.api import CIRC_MATE_DTYPE, CIRC_RES_DTYPE
        a, b = np.ascontiguousarray(rec["r1_rows"], np.uint8), np.ascontiguousarray(rec["r2_rows"], np.uint8)
        n = a.shape[0]
        l1, l2 = np.ascontiguousarray(rec["l1"], np.int32), np.ascontiguousarray(rec["l2"], np.int32)
        fs = np.ascontiguousarray(rec["fullmap_sel"], np.uint8)
        m = [np.ascontiguousarray(np.array(rec[k], CIRC_MATE_DTYPE)) for k in ("r1_1", "r2_1", "r1_2", "r2_2")]
        out = dict(ret=np.zeros(n, np.int32), cr=np.zeros(n, CIRC_RES_DTYPE))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.cmref_circ_check4(C.c_uint32(n), vp(a), vp(b), C.c_uint32(a.shape[1]), vp(l1), vp(l2), vp(fs), *[vp(x) for x in m], ws,
                                       vp(out["ret"]), vp(out["cr"]))
        assert r == 0, r
        return out

    def seed(self, seqs: np.ndarray, lens: np.ndarray) -> dict:
        n, S = len(lens), self.S
        out = dict(qpos=np.zeros((n, S), np.int32), frag_count=np.zeros((n, S), np.uint32),
                   has_frags=np.zeros((n, S), np.uint8), first_info=np.zeros((n, S), np.uint32),
                   last_info=np.zeros((n, S), np.uint32), sum_info=np.zeros((n, S), np.uint64))
        seqs = np.ascontiguousarray(seqs)
        for i in range(n):
            self.lib.cmref_seed(seqs[i].ctypes.data_as(C.c_char_p), int(lens[i]),
                                *[out[k][i].ctypes.data_as(C.c_void_p) for k in
                                  ("qpos", "frag_count", "has_frags", "first_info", "last_info", "sum_info")])
        return out

    def chain(self, seqs: np.ndarray, lens: np.ndarray) -> dict:
        n, Cn, F = len(lens), self.C_, self.F
        out = dict(chain_count=np.zeros(n, np.int32), high_hits=np.zeros(n, np.int32),
                   score=np.zeros((n, Cn), np.float32), chain_len=np.zeros((n, Cn), np.uint32),
                   frag_rpos=np.zeros((n, Cn, F), np.uint32), frag_qpos=np.zeros((n, Cn, F), np.int32))
        seqs = np.ascontiguousarray(seqs)
        for i in range(n):
            self.lib.cmref_chain(seqs[i].ctypes.data_as(C.c_char_p), int(lens[i]),
                                 out["chain_count"][i:].ctypes.data_as(C.c_void_p),
                                 out["high_hits"][i:].ctypes.data_as(C.c_void_p),
                                 *[out[k][i].ctypes.data_as(C.c_void_p) for k in
                                   ("score", "chain_len", "frag_rpos", "frag_qpos")])
        return out

    def process_pe(self, s1, s2, l1, l2, mreads):
        n = len(l1)
        s1 = np.ascontiguousarray(s1)
        s2 = np.ascontiguousarray(s2)
        assert s1.shape[1] == s2.shape[1]
        out = np.zeros(n, MREAD_DTYPE)
        state = np.zeros(n, np.int32)
        l1 = np.ascontiguousarray(l1, np.uint16)
        l2 = np.ascontiguousarray(l2, np.uint16)
        mreads = np.ascontiguousarray(mreads)
        self.lib.cmref_process_pe(C.c_uint32(n), s1.ctypes.data_as(C.c_char_p), s2.ctypes.data_as(C.c_char_p),
                                  C.c_uint32(s1.shape[1]), l1.ctypes.data_as(C.c_void_p),
                                  l2.ctypes.data_as(C.c_void_p), mreads.ctypes.data_as(C.c_void_p),
                                  out.ctypes.data_as(C.c_void_p), state.ctypes.data_as(C.c_void_p))
        return out, state

    def process_se(self, s, l):
        n = len(l)
        s = np.ascontiguousarray(s)
        l = np.ascontiguousarray(l, np.uint16)
        state = np.zeros(n, np.int32)
        self.lib.cmref_process_se(C.c_uint32(n), s.ctypes.data_as(C.c_char_p), C.c_uint32(s.shape[1]),
                                  l.ctypes.data_as(C.c_void_p), state.ctypes.data_as(C.c_void_p))
        return state

    def align(self, kind, s, ns, t, nt, aux=None):
        n = len(ns)
        s = np.ascontiguousarray(s)
        t = np.ascontiguousarray(t)
        assert s.shape[1] == t.shape[1]
        ns = np.ascontiguousarray(ns, np.uint16)
        nt = np.ascontiguousarray(nt, np.uint16)
        aux = np.zeros(n, np.int32) if aux is None else np.ascontiguousarray(aux, np.int32)
        o = [np.zeros(n, np.int32) for _ in range(4)]
        self.lib.cmref_align(kind, C.c_uint32(n), s.ctypes.data_as(C.c_char_p), ns.ctypes.data_as(C.c_void_p),
                             t.ctypes.data_as(C.c_char_p), nt.ctypes.data_as(C.c_void_p), C.c_uint32(s.shape[1]),
                             aux.ctypes.data_as(C.c_void_p), *[x.ctypes.data_as(C.c_void_p) for x in o])
        return dict(ed=o[0], sclen=o[1], indel=o[2], score=o[3])
