"""Host-side mirror of the reference interface over the C-ABI (include/cm_gpu.h).

The method names follow the reference's own seam (src/filter.h:46-62, src/match_read.h:45,
src/chain.h:19): `split_match_hash`, `get_best_chains`, `process_read`.  This module only
marshals numpy buffers into `libcm_gpu.so`; there is no CPU implementation behind it and
it raises if the CUDA library is missing or no device is present.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libcm_gpu.so")

CATEGORY = ["CONCRD", "DISCRD", "CHIORF", "CHIBSJ", "CHI2BSJ", "CONGEN", "CHIFUS", "CONGNM", "OEA2", "CANDID",
            "OEANCH", "ORPHAN", "NOPROC_MANYHIT", "NOPROC_NOMATCH"]
MAPPED_TYPES = (0, 1, 2, 3, 4, 5, 7)   # carry coordinates (src/filter.cpp:422-423)


class CmParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "kmer", "max_read_len", "max_ed", "max_sc", "band_width", "seed_lim", "max_tlen", "max_intron_len",
        "max_chain_len", "scan_level", "mat_sc", "mis_sc", "ind_sc", "xdrop")]


class CmCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "seeds", "seed_sectors", "seed_hits", "chain_links", "xdrop_cells", "edit_cells", "xdrop_calls",
        "edit_calls", "pairs", "kernel_launches")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


_U4P = C.POINTER(C.c_uint32)


class CmAnnotFlat(C.Structure):
    _fields_ = [
        ("n_iv", C.c_uint32),
        ("iv_spos", C.c_void_p), ("iv_epos", C.c_void_p), ("iv_max_end", C.c_void_p), ("iv_min_end", C.c_void_p),
        ("iv_max_next_exon", C.c_void_p), ("iv_seg_off", C.c_void_p),
        ("n_seg", C.c_uint32),
        ("seg_start", C.c_void_p), ("seg_end", C.c_void_p), ("seg_next_exon_beg", C.c_void_p),
        ("seg_gene_id", C.c_void_p), ("seg_tid_off", C.c_void_p),
        ("n_tid", C.c_uint32), ("tid", C.c_void_p),
        ("n_trans", C.c_uint32), ("trans_start_ind", C.c_void_p), ("t2s_off", C.c_void_p), ("t2s", C.c_void_p),
        ("n_genes", C.c_uint32), ("gene_start", C.c_void_p), ("gene_end", C.c_void_p),
        ("n_chr", C.c_uint32), ("chr_shift", C.c_void_p), ("chr_id", C.c_void_p),
        ("n_bits", C.c_uint64), ("near_border", C.c_void_p), ("intronic", C.c_void_p)]


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
# cm_circ_mate / cm_circ_res (include/cm_gpu.h): stage 2's MatchedMate with its cached exon look-ups, and CircRes::set_bp
CIRC_MATE_DTYPE = np.dtype([("m", MATE_DTYPE), ("exons_spos", "<i4"), ("exons_epos", "<i4"), ("exon_ind_spos", "<i4"), ("exon_ind_epos", "<i4"),
                            ("looked_up_spos", "u1"), ("looked_up_epos", "u1"), ("pad_", "u1", (2,))])
assert CIRC_MATE_DTYPE.itemsize == 72
CIRC_RES_DTYPE = np.dtype([("spos", "<u4"), ("epos", "<u4"), ("start_signal", "u1", (2,)), ("end_signal", "u1", (2,)), ("start_bp_ref", "u1", (2,)),
                           ("end_bp_ref", "u1", (2,)), ("start_signal_n", "u1"), ("end_signal_n", "u1"), ("ref_n", "u1"), ("was_set", "u1")])
assert CIRC_RES_DTYPE.itemsize == 20

MREAD_COMPARE_FIELDS = [f for f in MREAD_DTYPE.names if f != "pad_"]


def default_params(**kw) -> CmParams:
    p = CmParams(20, 300, 4, 7, 3, 500, 500, 2000000, 30, 0, 1, -3, -3, 8)
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def fresh_mreads(n: int, max_ed: int = 4) -> np.ndarray:
    """MatchedRead::MatchedRead(void) (src/common.cpp:268-271)."""
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


class CmGpuError(RuntimeError):
    pass


_ANNOT_U4 = ("iv_spos", "iv_epos", "iv_max_end", "iv_min_end", "iv_max_next_exon", "iv_seg_off", "seg_start", "seg_end",
             "seg_next_exon_beg", "seg_gene_id", "seg_tid_off", "tid", "t2s_off", "gene_start", "gene_end", "chr_shift")
_ANNOT_OTHER = (("trans_start_ind", np.int32), ("chr_id", np.int32), ("t2s", np.uint8), ("near_border", np.uint64),
                ("intronic", np.uint64))


def annot_struct(a: dict):
    """cm_annot_flat over the arrays of `a` (+ the list that keeps them alive)."""
    f = CmAnnotFlat()
    keep = []
    for k in ("n_iv", "n_seg", "n_tid", "n_trans", "n_genes", "n_chr", "n_bits"):
        setattr(f, k, int(a[k]))
    for nm, dt in [(n, np.uint32) for n in _ANNOT_U4] + list(_ANNOT_OTHER):
        arr = np.ascontiguousarray(a[nm], dt)
        if arr.size == 0:
            arr = np.zeros(1, dt)
        keep.append(arr)
        setattr(f, nm, arr.ctypes.data)
    return f, keep


def write_annotation_file(path, a: dict, max_read_len: int = 300, tag: int = 0):
    """One packed contig's flat annotation as a build-once file (cm_annot_file_write; host only)."""
    lib = load_library()
    f, keep = annot_struct(a)
    r = lib.cm_annot_file_write(os.fsencode(path), C.byref(f), C.c_uint32(max_read_len), C.c_uint64(tag))
    if r != 0:
        raise CmGpuError(f"cm_annot_file_write({path}): status {r}")


def read_annotation_file(path) -> dict:
    """The arrays of a file written by write_annotation_file, as numpy copies (cm_annot_file_read; host only)."""
    lib = load_library()
    f = CmAnnotFlat()
    mrl, tag, st = C.c_uint32(), C.c_uint64(), C.c_void_p()
    r = lib.cm_annot_file_read(os.fsencode(path), C.byref(f), C.byref(mrl), C.byref(tag), C.byref(st))
    if r != 0:
        raise CmGpuError(f"cm_annot_file_read({path}): status {r}")
    try:
        a = {k: int(getattr(f, k)) for k in ("n_iv", "n_seg", "n_tid", "n_trans", "n_genes", "n_chr", "n_bits")}
        nw = (a["n_bits"] + 63) // 64
        counts = dict(iv_spos=a["n_iv"], iv_epos=a["n_iv"], iv_max_end=a["n_iv"], iv_min_end=a["n_iv"], iv_max_next_exon=a["n_iv"],
                      iv_seg_off=a["n_iv"] + 1, seg_start=a["n_seg"], seg_end=a["n_seg"], seg_next_exon_beg=a["n_seg"],
                      seg_gene_id=a["n_seg"], seg_tid_off=a["n_seg"] + 1, tid=a["n_tid"], t2s_off=a["n_trans"] + 1,
                      gene_start=a["n_genes"], gene_end=a["n_genes"], chr_shift=a["n_chr"], trans_start_ind=a["n_trans"],
                      chr_id=a["n_chr"], near_border=nw, intronic=nw)

        def arr(nm, dt, cnt):
            buf = (C.c_char * (cnt * np.dtype(dt).itemsize)).from_address(getattr(f, nm)) if cnt else b""
            return np.frombuffer(buf, dt, cnt).copy()
        for nm in _ANNOT_U4:
            a[nm] = arr(nm, np.uint32, counts[nm])
        for nm, dt in _ANNOT_OTHER:
            if nm != "t2s":
                a[nm] = arr(nm, dt, counts[nm])
        a["t2s"] = arr("t2s", np.uint8, int(a["t2s_off"][-1]) if a["n_trans"] else 0)
        a["max_read_len"], a["tag"] = mrl.value, tag.value
        return a
    finally:
        lib.cm_annot_file_free(st)


_lib = None


def load_library():
    """dlopen the CUDA library; fail loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CmGpuError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'`. "
                         "There is no CPU fallback.")
    lib = C.CDLL(os.environ.get("CM_GPU_LIB", LIB_PATH))   # CM_GPU_LIB: an alternative build of the same library (A/B runs)
    lib.cm_gpu_last_error.restype = C.c_char_p
    lib.cm_gpu_last_error.argtypes = [C.c_void_p]
    _lib = lib
    return lib


def _vp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _rows(seqs, lens=None):
    """Accept a 2-D uint8/S1 array (fixed rows) and return (contiguous uint8 2-D, uint16 lens)."""
    a = np.ascontiguousarray(seqs)
    if a.dtype != np.uint8:
        a = a.view(np.uint8)
    if a.ndim != 2:
        raise ValueError("reads must be a 2-D byte array [n, stride]")
    if lens is None:
        lens = np.full(a.shape[0], a.shape[1], np.uint16)
    lens = np.ascontiguousarray(lens, np.uint16)
    if len(lens) != a.shape[0]:
        raise ValueError("len(lens) != number of rows")
    return a, lens


class Mapper:
    """One context per GPU (threading contract of SURVEY.md 8(b))."""

    def __init__(self, params: CmParams | None = None, device: int = 0, parent: "Mapper | None" = None):
        """parent: share that Mapper's resident contigs and annotation (cm_gpu_create_shared); own stream and buffers."""
        self.lib = load_library()
        self.params = parent.params if parent is not None else (params or default_params())
        self.ctx = C.c_void_p()
        self.parent = parent
        if parent is not None:
            r = self.lib.cm_gpu_create_shared(parent.ctx, C.byref(self.ctx))
        else:
            r = self.lib.cm_gpu_create(device, C.byref(self.params), C.byref(self.ctx))
        if r != 0:
            raise CmGpuError(f"cm_gpu_create: {self.lib.cm_gpu_last_error(None).decode()} (status {r})")
        k = self.params.kmer
        self.S = 2 * ((self.params.max_read_len + k - 1) // k) - 1
        self.F = (self.S + 1) // 2
        self.Cn = self.params.max_chain_len
        self._keep = []

    def close(self):
        if self.ctx:
            self.lib.cm_gpu_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, r, what):
        if r != 0:
            raise CmGpuError(f"{what}: {self.lib.cm_gpu_last_error(self.ctx).decode()} (status {r})")

    # ---- index / annotation ------------------------------------------------------------
    def load_contig_from_reference(self, contig_id, bucket_lists, table_mem, table_mem_entries, cref_gen,
                                   ref_gen_len):
        """Raw pointers of the reference's in-memory index (see cm_gpu_load_contig)."""
        self._chk(self.lib.cm_gpu_load_contig(self.ctx, contig_id, C.c_void_p(bucket_lists), C.c_void_p(table_mem),
                                              C.c_uint64(table_mem_entries), C.c_void_p(cref_gen),
                                              C.c_uint32(ref_gen_len)), "cm_gpu_load_contig")

    def build_contig(self, contig_id, seq: np.ndarray):
        seq = np.ascontiguousarray(seq, np.uint8)
        self._chk(self.lib.cm_gpu_build_contig(self.ctx, contig_id, _vp(seq), C.c_uint32(len(seq))),
                  "cm_gpu_build_contig")

    def build_contig_packed(self, contig_id, cref_gen, ref_gen_len):
        """Device build from the reference's 3-bit packed contig (raw pointer to _ih_crefGen)."""
        self._chk(self.lib.cm_gpu_build_contig_packed(self.ctx, contig_id, C.c_void_p(cref_gen), C.c_uint32(ref_gen_len)),
                  "cm_gpu_build_contig_packed")

    def write_index(self, path, n_contigs, compact=False, contig_max_size=1_300_000_000):
        """The reference's <ref>.packed.fa.index from the resident contigs (cm_gpu_write_index)."""
        self._chk(self.lib.cm_gpu_write_index(self.ctx, os.fsencode(path), n_contigs, int(compact), C.c_uint32(contig_max_size)),
                  "cm_gpu_write_index")

    def load_index_file(self, path, first_id=0):
        n = C.c_int()
        self._chk(self.lib.cm_gpu_load_index_file(self.ctx, os.fsencode(path), first_id, C.byref(n)), "cm_gpu_load_index_file")
        return n.value

    def contig_stats(self, contig_id):
        n = C.c_uint64()
        ln = C.c_uint32()
        self._chk(self.lib.cm_gpu_contig_stats(self.ctx, contig_id, C.byref(n), C.byref(ln)), "cm_gpu_contig_stats")
        return n.value, ln.value

    def read_bucket(self, contig_id, hv, cap=4096):
        cks = np.zeros(cap, np.uint16)
        pos = np.zeros(cap, np.uint32)
        n = C.c_uint32()
        self._chk(self.lib.cm_gpu_read_bucket(self.ctx, contig_id, C.c_uint32(hv), C.c_uint32(cap), _vp(cks), _vp(pos),
                                              C.byref(n)), "cm_gpu_read_bucket")
        m = min(n.value, cap)
        return cks[:m], pos[:m], n.value

    def load_annotation(self, contig_id, a: dict):
        """`a` holds the flat arrays named as in cm_annot_flat (annotation.flatten / RefHarness.annotation)."""
        f, keep = annot_struct(a)
        self._chk(self.lib.cm_gpu_load_annotation(self.ctx, contig_id, C.byref(f)), "cm_gpu_load_annotation")

    def load_annotation_file(self, contig_id, path, tag=0):
        """Upload a file written by write_annotation_file (cm_gpu_load_annotation_file)."""
        self._chk(self.lib.cm_gpu_load_annotation_file(self.ctx, contig_id, os.fsencode(path), C.c_uint64(tag)),
                  "cm_gpu_load_annotation_file")

    def extend(self, contig_id, mode, prob: dict) -> dict:
        """TransExtension::extend_both_mates (mode 0) / extend_chain_both_sides (mode 1) per row (cm_gpu_extend_batch);
        `prob` holds arrays named as the C arguments (lseq, rseq, llen, rlen, l_chain_len, l_rpos, l_qpos, r_*, iv_a, iv_b, dir)."""
        keep = {k: (None if v is None else np.ascontiguousarray(v)) for k, v in prob.items()}
        n = len(keep["llen"])
        F = keep["l_rpos"].shape[1]
        res = dict(lmm=np.zeros(n, MATE_DTYPE), rmm=np.zeros(n, MATE_DTYPE), ret=np.zeros(n, np.int32))
        g = lambda k: _vp(keep.get(k))
        self._chk(self.lib.cm_gpu_extend_batch(self.ctx, contig_id, mode, C.c_uint32(n), g("lseq"), g("rseq"),
                                               C.c_uint32(keep["lseq"].shape[1]), g("llen"), g("rlen"), C.c_uint32(F), g("l_chain_len"),
                                               g("l_rpos"), g("l_qpos"), g("r_chain_len"), g("r_rpos"), g("r_qpos"), g("iv_a"), g("iv_b"),
                                               g("dir"), _vp(res["lmm"]), _vp(res["rmm"]), _vp(res["ret"])), "cm_gpu_extend_batch")
        return res

    def circ_chaining(self, contig_id, seqs, qspos, qepos, gene_start, gene_end, window_size=8, step=3, F=None) -> dict:
        """ProcessCirc::chaining per row (src/process_circ.cpp:678; cm_gpu_circ_chain_batch): rows = remain_seq as ASCII codes."""
        a = np.ascontiguousarray(seqs, np.uint8)
        n, Cn = a.shape[0], self.Cn
        u = lambda x: np.ascontiguousarray(x, np.uint32)
        qspos, qepos, gene_start, gene_end = u(qspos), u(qepos), u(gene_start), u(gene_end)
        if F is None:   # fragment slots: what the longest range of the batch can need (the reference sizes by maxReadLength)
            span = int((qepos.astype(np.int64) - qspos.astype(np.int64) + 1).max()) if n else 0
            F = max(1, (max(span, window_size) - window_size) // step + 1)
        out = dict(chain_count=np.zeros(n, np.int32), kmer_cnt=np.zeros(n, np.int32), score=np.zeros((n, Cn), np.float32),
                   chain_len=np.zeros((n, Cn), np.uint32), frag_rpos=np.zeros((n, Cn, F), np.uint32), frag_qpos=np.zeros((n, Cn, F), np.int32))
        self._chk(self.lib.cm_gpu_circ_chain_batch(self.ctx, contig_id, C.c_uint32(n), _vp(a), C.c_uint32(a.shape[1]), _vp(qspos), _vp(qepos),
                                                   _vp(gene_start), _vp(gene_end), window_size, step, F,
                                                   *[_vp(out[k]) for k in ("chain_count", "kmer_cnt", "score", "chain_len", "frag_rpos", "frag_qpos")]),
                  "cm_gpu_circ_chain_batch")
        return out


    def circ_exact(self, contig_id, prob: dict, window_size=8) -> dict:
        """ProcessCirc::find_exact_coord per row (src/process_circ.cpp:739); `prob` holds arrays named as the C arguments
        (seqs, r1_spos, r1_epos, r2_spos, r2_epos, dir, qspos, rlen, whole_len, chain_len, frag_rpos, frag_qpos)."""
        a = np.ascontiguousarray(prob["seqs"], np.uint8)
        n = a.shape[0]
        u = {k: np.ascontiguousarray(prob[k], np.uint32) for k in ("r1_spos", "r1_epos", "r2_spos", "r2_epos", "qspos", "chain_len", "frag_rpos")}
        s = {k: np.ascontiguousarray(prob[k], np.int32) for k in ("dir", "rlen", "whole_len", "frag_qpos")}
        res = dict(pmm=np.zeros(n, MATE_DTYPE), ret=np.zeros(n, np.int32))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.cm_gpu_circ_exact_batch(self.ctx, contig_id, C.c_uint32(n), vp(a), C.c_uint32(a.shape[1]), vp(u["r1_spos"]), vp(u["r1_epos"]), vp(u["r2_spos"]), vp(u["r2_epos"]),
               vp(s["dir"]), vp(u["qspos"]), vp(s["rlen"]), vp(s["whole_len"]), C.c_uint32(u["frag_rpos"].shape[1]), vp(u["chain_len"]),
               vp(u["frag_rpos"]), vp(s["frag_qpos"]), window_size, vp(res["pmm"]), vp(res["ret"]))
        self._chk(r, "cm_gpu_circ_exact_batch")
        return res


    def circ_check(self, contig_id, rec: dict, window_size=8) -> dict:
        """ProcessCirc::check_split_map (three mates) per row (src/process_circ.cpp:892; cm_gpu_circ_check_batch).  rec: remain_rows,
        fullmap_rows, remain_len, fullmap_len, r1_partial, in_r1, in_r2, in_pm -> ret, cr and the mates as the call leaves them."""
        rr = np.ascontiguousarray(rec["remain_rows"], np.uint8)
        fr = np.ascontiguousarray(rec["fullmap_rows"], np.uint8)
        n = rr.shape[0]
        rl, fl = np.ascontiguousarray(rec["remain_len"], np.int32), np.ascontiguousarray(rec["fullmap_len"], np.int32)
        rp = np.ascontiguousarray(rec["r1_partial"], np.uint8)
        out = dict(r1=np.array(rec["in_r1"], CIRC_MATE_DTYPE), r2=np.array(rec["in_r2"], CIRC_MATE_DTYPE), pm=np.array(rec["in_pm"], CIRC_MATE_DTYPE),
                   ret=np.zeros(n, np.int32), cr=np.zeros(n, CIRC_RES_DTYPE))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.cm_gpu_circ_check_batch(self.ctx, contig_id, C.c_uint32(n), vp(rr), vp(fr), C.c_uint32(rr.shape[1] if n else 0), vp(rl), vp(fl), vp(rp), vp(out["r1"]), vp(out["r2"]),
               vp(out["pm"]), window_size, vp(out["ret"]), vp(out["cr"]))
        self._chk(r, "cm_gpu_circ_check_batch")
        return out

    def circ_check4(self, contig_id, rec: dict, window_size=8) -> dict:
        """ProcessCirc::check_split_map (four mates) per row (src/process_circ.cpp:922; cm_gpu_circ_check4_batch).  rec: r1_rows, r2_rows,
        l1, l2, fullmap_sel, r1_1, r2_1, r1_2, r2_2 -> ret, cr."""
        a, b = np.ascontiguousarray(rec["r1_rows"], np.uint8), np.ascontiguousarray(rec["r2_rows"], np.uint8)
        n = a.shape[0]
        l1, l2 = np.ascontiguousarray(rec["l1"], np.int32), np.ascontiguousarray(rec["l2"], np.int32)
        fs = np.ascontiguousarray(rec["fullmap_sel"], np.uint8)
        m = [np.ascontiguousarray(np.array(rec[k], CIRC_MATE_DTYPE)) for k in ("r1_1", "r2_1", "r1_2", "r2_2")]
        out = dict(ret=np.zeros(n, np.int32), cr=np.zeros(n, CIRC_RES_DTYPE))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.cm_gpu_circ_check4_batch(self.ctx, contig_id, C.c_uint32(n), vp(a), vp(b), C.c_uint32(a.shape[1] if n else 0), vp(l1), vp(l2), vp(fs),
                                              *[vp(x) for x in m], window_size, vp(out["ret"]), vp(out["cr"]))
        self._chk(r, "cm_gpu_circ_check4_batch")
        return out

    def circ_single(self, contig_id, s1, s2, l1, l2, mr_con, gene_cnt, gene_se, window_size=8, step=3, cap=20000) -> dict:
        """ProcessCirc::call_circ_single_split for a batch (src/process_circ.cpp:360; cm_gpu_circ_single_batch).  mr_con: MatchedReads
        with contig coordinates; gene_cnt[n], gene_se[n][G][2]: the genes get_gene_overlap(mm_r1.spos) lists for each pair."""
        a1, a2 = np.ascontiguousarray(s1, np.uint8), np.ascontiguousarray(s2, np.uint8)
        n, stride = a1.shape
        l1, l2 = np.ascontiguousarray(l1, np.uint16), np.ascontiguousarray(l2, np.uint16)
        mr = np.ascontiguousarray(mr_con)
        gene_cnt = np.asarray(gene_cnt, np.int64)
        goff = np.zeros(n + 1, np.uint32)
        goff[1:] = np.cumsum(gene_cnt)
        sel = np.arange(gene_se.shape[1])[None, :] < gene_cnt[:, None]
        gs = np.ascontiguousarray(gene_se[:, :, 0][sel], np.uint32)
        ge = np.ascontiguousarray(gene_se[:, :, 1][sel], np.uint32)
        if len(gs) == 0:
            gs, ge = np.zeros(1, np.uint32), np.zeros(1, np.uint32)
        rec = dict(call_pair=np.zeros(cap, np.uint32), r1_partial=np.zeros(cap, np.uint8), in_r1=np.zeros(cap, CIRC_MATE_DTYPE),
                   in_r2=np.zeros(cap, CIRC_MATE_DTYPE), in_pm=np.zeros(cap, CIRC_MATE_DTYPE), ret=np.zeros(cap, np.int32),
                   cr=np.zeros(cap, CIRC_RES_DTYPE), out_r1=np.zeros(cap, CIRC_MATE_DTYPE), out_r2=np.zeros(cap, CIRC_MATE_DTYPE),
                   out_pm=np.zeros(cap, CIRC_MATE_DTYPE))
        best, best_cr = np.zeros(n, np.int32), np.zeros(n, CIRC_RES_DTYPE)
        k = C.c_uint32(0)
        self._chk(self.lib.cm_gpu_circ_single_batch(self.ctx, contig_id, C.c_uint32(n), _vp(a1), _vp(a2), C.c_uint32(stride), _vp(l1), _vp(l2), _vp(mr),
                                                    _vp(goff), _vp(gs), _vp(ge), window_size, step, C.c_uint32(cap), C.byref(k),
                                                    *[_vp(rec[x]) for x in ("call_pair", "r1_partial", "in_r1", "in_r2", "in_pm", "ret", "cr",
                                                                            "out_r1", "out_r2", "out_pm")], _vp(best), _vp(best_cr)),
                  "cm_gpu_circ_single_batch")
        out = {x: v[:min(k.value, cap)] for x, v in rec.items()}
        out["n_attempts"], out["best_type"], out["best_cr"] = k.value, best, best_cr
        return out

    def circ_double(self, contig_id, s1, s2, l1, l2, mr_con, gene_cnt, gene_se, window_size=8, step=3, cap=50000) -> dict:
        """ProcessCirc::call_circ_double_split for a batch of CHI2BSJ pairs (src/process_circ.cpp:485; cm_gpu_circ_double_batch)."""
        a1, a2 = np.ascontiguousarray(s1, np.uint8), np.ascontiguousarray(s2, np.uint8)
        n, stride = a1.shape
        l1, l2 = np.ascontiguousarray(l1, np.uint16), np.ascontiguousarray(l2, np.uint16)
        mr = np.ascontiguousarray(mr_con)
        gene_cnt = np.asarray(gene_cnt, np.int64)
        goff = np.zeros(n + 1, np.uint32)
        goff[1:] = np.cumsum(gene_cnt)
        sel = np.arange(gene_se.shape[1])[None, :] < gene_cnt[:, None]
        gs = np.ascontiguousarray(gene_se[:, :, 0][sel], np.uint32)
        ge = np.ascontiguousarray(gene_se[:, :, 1][sel], np.uint32)
        if len(gs) == 0:
            gs, ge = np.zeros(1, np.uint32), np.zeros(1, np.uint32)
        ev = dict(pair=np.zeros(cap, np.uint32), kind=np.zeros(cap, np.uint8), r1=np.zeros(cap, CIRC_MATE_DTYPE), r2=np.zeros(cap, CIRC_MATE_DTYPE),
                  p1=np.zeros(cap, CIRC_MATE_DTYPE), p2=np.zeros(cap, CIRC_MATE_DTYPE), ret=np.zeros(cap, np.int32), cr=np.zeros(cap, CIRC_RES_DTYPE))
        push_type, push_cr = np.zeros(n, np.int32), np.zeros(n, CIRC_RES_DTYPE)
        k = C.c_uint32(0)
        self._chk(self.lib.cm_gpu_circ_double_batch(self.ctx, contig_id, C.c_uint32(n), _vp(a1), _vp(a2), C.c_uint32(stride), _vp(l1), _vp(l2), _vp(mr),
                                                    _vp(goff), _vp(gs), _vp(ge), window_size, step, C.c_uint32(cap), C.byref(k),
                                                    *[_vp(ev[x]) for x in ("pair", "kind", "r1", "r2", "p1", "p2", "ret", "cr")], _vp(push_type),
                                                    _vp(push_cr)), "cm_gpu_circ_double_batch")
        assert k.value <= cap
        out = {x: v[:k.value] for x, v in ev.items()}
        out["push_type"], out["push_cr"] = push_type, push_cr
        return out

    def annot_lookup(self, contig_id, pos, mlen, rlen) -> dict:
        """GTFParser look-ups per position (cm_gpu_annot_lookup_batch)."""
        pos = np.ascontiguousarray(pos, np.uint32)
        n = len(pos)
        out = dict(ub=np.zeros(n, np.uint32), max_end=np.zeros(n, np.uint32), ol_iv=np.zeros(n, np.int32), loc_iv=np.zeros(n, np.int32),
                   loc_ind=np.zeros(n, np.int32), shift_idx=np.zeros(n, np.int32), intronic=np.zeros(n, np.uint8))
        self._chk(self.lib.cm_gpu_annot_lookup_batch(self.ctx, contig_id, C.c_uint32(n), _vp(pos), C.c_uint32(mlen), C.c_uint32(rlen),
                                                     *[_vp(out[k]) for k in ("ub", "max_end", "ol_iv", "loc_iv", "loc_ind", "shift_idx", "intronic")]),
                  "cm_gpu_annot_lookup_batch")
        return out

    # ---- stages ---------------------------------------------------------------------------
    def split_match_hash(self, contig_id, seqs, lens=None) -> dict:
        """GenomeSeeder::split_match_hash for every row (src/match_read.cpp:270)."""
        a, lens = _rows(seqs, lens)
        n, S = a.shape[0], self.S
        out = dict(qpos=np.zeros((n, S), np.int32), frag_count=np.zeros((n, S), np.uint32),
                   has_frags=np.zeros((n, S), np.uint8), first_info=np.zeros((n, S), np.uint32),
                   last_info=np.zeros((n, S), np.uint32), sum_info=np.zeros((n, S), np.uint64))
        self._chk(self.lib.cm_gpu_seed_batch(self.ctx, contig_id, C.c_uint32(n), _vp(a), C.c_uint32(a.shape[1]),
                                             _vp(lens), _vp(out["qpos"]), _vp(out["frag_count"]),
                                             _vp(out["has_frags"]), _vp(out["first_info"]), _vp(out["last_info"]),
                                             _vp(out["sum_info"])), "cm_gpu_seed_batch")
        return out

    def get_best_chains(self, contig_id, seqs, lens=None) -> dict:
        """FilterRead::get_best_chains for every row (src/filter.cpp:469)."""
        a, lens = _rows(seqs, lens)
        n, Cn, F = a.shape[0], self.Cn, self.F
        out = dict(chain_count=np.zeros(n, np.int32), high_hits=np.zeros(n, np.int32),
                   score=np.zeros((n, Cn), np.float32), chain_len=np.zeros((n, Cn), np.uint32),
                   frag_rpos=np.zeros((n, Cn, F), np.uint32), frag_qpos=np.zeros((n, Cn, F), np.int32))
        self._chk(self.lib.cm_gpu_chain_batch(self.ctx, contig_id, C.c_uint32(n), _vp(a), C.c_uint32(a.shape[1]),
                                              _vp(lens), _vp(out["chain_count"]), _vp(out["high_hits"]),
                                              _vp(out["score"]), _vp(out["chain_len"]), _vp(out["frag_rpos"]),
                                              _vp(out["frag_qpos"])), "cm_gpu_chain_batch")
        return out

    def align(self, kind, s, ns, t, nt, aux=None) -> dict:
        s, ns = _rows(s, ns)
        t, nt = _rows(t, nt)
        if s.shape != t.shape:
            raise ValueError("s and t must share shape")
        n = s.shape[0]
        aux = np.zeros(n, np.int32) if aux is None else np.ascontiguousarray(aux, np.int32)
        o = [np.zeros(n, np.int32) for _ in range(4)]
        self._chk(self.lib.cm_gpu_align_batch(self.ctx, kind, C.c_uint32(n), _vp(s), _vp(ns), _vp(t), _vp(nt),
                                              C.c_uint32(s.shape[1]), _vp(aux), *[_vp(x) for x in o]),
                  "cm_gpu_align_batch")
        return dict(ed=o[0], sclen=o[1], indel=o[2], score=o[3])

    def align_noalloc(self, kind, s, ns, t, nt, aux=None) -> dict:
        """cm_gpu_align_batch_noalloc: same results as align(), staged in the library's fixed 4 MiB device array."""
        s, ns = _rows(s, ns)
        t, nt = _rows(t, nt)
        if s.shape != t.shape:
            raise ValueError("s and t must share shape")
        n = s.shape[0]
        aux = np.zeros(n, np.int32) if aux is None else np.ascontiguousarray(aux, np.int32)
        o = [np.zeros(n, np.int32) for _ in range(4)]
        self._chk(self.lib.cm_gpu_align_batch_noalloc(self.ctx, kind, C.c_uint32(n), _vp(s), _vp(ns), _vp(t), _vp(nt),
                                                      C.c_uint32(s.shape[1]), _vp(aux), *[_vp(x) for x in o]),
                  "cm_gpu_align_batch_noalloc")
        return dict(ed=o[0], sclen=o[1], indel=o[2], score=o[3])

    def process_read(self, contig_id, seqs1, seqs2, mreads, lens1=None, lens2=None, out=None, state=None):
        """FilterRead::process_read, PE overload, for a batch (src/filter.cpp:124).  out / state: caller-owned result
        buffers (pinned memory makes the device->host copy several times faster)."""
        a1, l1 = _rows(seqs1, lens1)
        a2, l2 = _rows(seqs2, lens2)
        if a1.shape != a2.shape:
            raise ValueError("mate arrays must share shape")
        n = a1.shape[0]
        mreads = np.ascontiguousarray(mreads)
        assert mreads.dtype == MREAD_DTYPE and len(mreads) == n
        if out is None:
            out = np.zeros(n, MREAD_DTYPE)
        if state is None:
            state = np.zeros(n, np.int32)
        assert out.dtype == MREAD_DTYPE and len(out) == n and state.dtype == np.int32 and len(state) == n
        self._chk(self.lib.cm_gpu_process_batch_pe(self.ctx, contig_id, C.c_uint32(n), _vp(a1), _vp(a2),
                                                   C.c_uint32(a1.shape[1]), _vp(l1), _vp(l2), _vp(mreads),
                                                   _vp(out), _vp(state)), "cm_gpu_process_batch_pe")
        return out, state

    def map_rounds(self, seqs1, seqs2, mreads, first=0, n_contigs=1, lens1=None, lens2=None, snapshots=False,
                   out=None, state=None):
        """The reference's round loop for a batch over resident contigs first .. first + n_contigs - 1
        (src/circminer.cpp:229-308; cm_gpu_map_rounds_pe).  Returns a dict: out / state / final_round per pair and,
        with snapshots, round_out / round_state [n_contigs, n] (state -1 = the pair was already final)."""
        a1, l1 = _rows(seqs1, lens1)
        a2, l2 = _rows(seqs2, lens2)
        if a1.shape != a2.shape:
            raise ValueError("mate arrays must share shape")
        n = a1.shape[0]
        mreads = np.ascontiguousarray(mreads)
        assert mreads.dtype == MREAD_DTYPE and len(mreads) == n
        res = dict(out=np.zeros(n, MREAD_DTYPE) if out is None else out,
                   state=np.zeros(n, np.int32) if state is None else state, final_round=np.zeros(n, np.int32))
        if snapshots:
            res["round_out"] = np.zeros((n_contigs, n), MREAD_DTYPE)
            res["round_state"] = np.zeros((n_contigs, n), np.int32)
        self._chk(self.lib.cm_gpu_map_rounds_pe(self.ctx, first, n_contigs, C.c_uint32(n), _vp(a1), _vp(a2),
                                                C.c_uint32(a1.shape[1]), _vp(l1), _vp(l2), _vp(mreads), _vp(res["out"]),
                                                _vp(res["state"]), _vp(res["final_round"]), _vp(res.get("round_out")),
                                                _vp(res.get("round_state"))), "cm_gpu_map_rounds_pe")
        return res

    def run_staged_rounds(self, first=0, n_contigs=1):
        self._chk(self.lib.cm_gpu_run_staged_rounds_pe(self.ctx, first, n_contigs), "cm_gpu_run_staged_rounds_pe")

    def fetch_rounds(self):
        out = np.zeros(self._staged_n, MREAD_DTYPE)
        state = np.zeros(self._staged_n, np.int32)
        fr = np.zeros(self._staged_n, np.int32)
        self._chk(self.lib.cm_gpu_fetch_rounds_pe(self.ctx, _vp(out), _vp(state), _vp(fr)), "cm_gpu_fetch_rounds_pe")
        return out, state, fr

    def rounds_ms(self, cap=8):
        """(ms[8], pairs_per_round): see cm_gpu_last_rounds_ms."""
        a = (C.c_float * 8)()
        pr = (C.c_uint32 * cap)()
        k = self.lib.cm_gpu_last_rounds_ms(self.ctx, a, pr, cap)
        return list(a), list(pr)[:max(0, k)]

    def process_read_se(self, contig_id, seqs, lens=None):
        a, l = _rows(seqs, lens)
        state = np.zeros(a.shape[0], np.int32)
        self._chk(self.lib.cm_gpu_process_batch_se(self.ctx, contig_id, C.c_uint32(a.shape[0]), _vp(a),
                                                   C.c_uint32(a.shape[1]), _vp(l), _vp(state)),
                  "cm_gpu_process_batch_se")
        return state

    # ---- device-resident batch (kernel-only timing) -------------------------------------
    def stage(self, seqs1, seqs2, mreads, lens1=None, lens2=None):
        a1, l1 = _rows(seqs1, lens1)
        a2, l2 = _rows(seqs2, lens2)
        mreads = np.ascontiguousarray(mreads)
        self._staged_n = a1.shape[0]
        self._chk(self.lib.cm_gpu_stage_batch_pe(self.ctx, C.c_uint32(a1.shape[0]), _vp(a1), _vp(a2),
                                                 C.c_uint32(a1.shape[1]), _vp(l1), _vp(l2), _vp(mreads)),
                  "cm_gpu_stage_batch_pe")

    def run_staged(self, contig_id, upto_stage=3):
        self._chk(self.lib.cm_gpu_run_staged_pe(self.ctx, contig_id, upto_stage), "cm_gpu_run_staged_pe")

    def fetch_staged(self):
        out = np.zeros(self._staged_n, MREAD_DTYPE)
        state = np.zeros(self._staged_n, np.int32)
        self._chk(self.lib.cm_gpu_fetch_staged_pe(self.ctx, _vp(out), _vp(state)), "cm_gpu_fetch_staged_pe")
        return out, state

    def stage_ms(self):
        a = (C.c_float * 5)()
        self.lib.cm_gpu_last_stage_ms(self.ctx, a)
        return list(a)

    def align_ms(self):
        """[k_chain_spec, k_chain_walk] device ms of the last chain stage."""
        a = (C.c_float * 2)()
        self._chk(self.lib.cm_gpu_last_align_ms(self.ctx, a), "cm_gpu_last_align_ms")
        return list(a)

    def int32_peak(self) -> float:
        v = C.c_double()
        self._chk(self.lib.cm_gpu_int32_peak(self.ctx, C.byref(v)), "cm_gpu_int32_peak")
        return v.value

    def counters(self, reset=False) -> dict:
        c = CmCounters()
        self._chk(self.lib.cm_gpu_get_counters(self.ctx, C.byref(c), int(reset)), "cm_gpu_get_counters")
        return c.as_dict()
