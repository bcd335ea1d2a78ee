"""TEST INFRASTRUCTURE ONLY: ctypes wrapper over tests/hostsim/libhostsim.so (g++ build of the device
headers).  Mirrors circminer_b200.api.Mapper so the same comparison code serves CPU debugging and
the GPU parity tests."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from circminer_b200.api import MREAD_DTYPE, CmAnnotFlat, CmCounters, CmParams, _rows, _vp, default_params

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "hostsim", "libhostsim.so")


def build(target="libhostsim.so"):
    subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "hostsim"), target])


class HostSim:
    def __init__(self, params: CmParams | None = None, variant: str = ""):
        """variant "smallcap": the build with tiny fixed-size lists in the stage-2 classification (tests/hostsim/Makefile)."""
        so = f"libhostsim_{variant}.so" if variant else "libhostsim.so"
        build(so)
        self.lib = C.CDLL(os.path.join(HERE, "hostsim", so))
        self.lib.hs_create.restype = C.c_void_p
        self.params = params or default_params()
        self.ctx = C.c_void_p(self.lib.hs_create(C.byref(self.params)))
        k = self.params.kmer
        self.S = 2 * ((self.params.max_read_len + k - 1) // k) - 1
        self.F = (self.S + 1) // 2
        self.Cn = self.params.max_chain_len
        self._keep = []

    def load_contig_from_reference(self, contig_id, bucket_lists, table_mem, genome_ascii: np.ndarray):
        g = np.ascontiguousarray(genome_ascii, np.uint8)
        self._keep.append(g)
        self.lib.hs_load_contig(self.ctx, contig_id, C.c_void_p(bucket_lists), C.c_void_p(table_mem), _vp(g),
                                C.c_uint32(len(g)))

    def build_contig(self, contig_id, seq: np.ndarray):
        """Index of one packed contig from its text, (bucket, checksum, position) order
        (src/mrsfast/HashTable.c:769-839, Sort.c:110-176) -- numpy restatement for the CPU tests."""
        g = np.ascontiguousarray(seq, np.uint8)
        k = self.params.kmer
        code = np.full(256, 4, np.int64)
        for i, ch in enumerate(b"ACGT"):
            code[ch] = i
        v = code[g]
        n = len(g) - k + 1
        if n <= 0:
            key = np.zeros(0, np.int64); pos = np.zeros(0, np.uint32)
        else:
            bad = np.concatenate([[0], np.cumsum(v == 4)])
            ok = (bad[k:k + n] - bad[:n]) == 0
            key = np.zeros(n, np.int64)
            for j in range(k):
                key = key * 4 + (v[j:j + n] & 3)
            pos = np.nonzero(ok)[0]
            key = key[pos]
            order = np.argsort(key, kind="stable")
            key = key[order]
            pos = (pos[order] + 1).astype(np.uint32)
        cs = 2 * (k - 14)
        hv = key >> cs
        cks = (key & ((1 << cs) - 1)).astype(np.uint16)
        dirv = np.zeros(4 ** 14 + 1, np.uint32)
        dirv[1:] = np.cumsum(np.bincount(hv, minlength=4 ** 14)).astype(np.uint32)
        self._keep += [g]
        self.lib.hs_set_index(self.ctx, contig_id, _vp(dirv), _vp(cks), _vp(np.ascontiguousarray(pos)), C.c_uint64(len(pos)),
                              _vp(g), C.c_uint32(len(g)))

    def seed_x_check(self, contig_id, big_th=96, n_probe=200000, seed=1):
        self.lib.hs_seed_x_check.restype = C.c_long
        return self.lib.hs_seed_x_check(self.ctx, contig_id, C.c_uint32(big_th), C.c_uint32(n_probe), C.c_uint64(seed))

    def load_annotation(self, contig_id, a: dict):
        f = CmAnnotFlat()
        for k in ("n_iv", "n_seg", "n_tid", "n_trans", "n_genes", "n_chr", "n_bits"):
            setattr(f, k, int(a[k]))

        def put(name, dtype):
            arr = np.ascontiguousarray(a[name], dtype)
            if arr.size == 0:
                arr = np.zeros(1, dtype)
            self._keep.append(arr)
            setattr(f, name, arr.ctypes.data)
        for nm in ("iv_spos", "iv_epos", "iv_max_end", "iv_min_end", "iv_max_next_exon", "iv_seg_off", "seg_start",
                   "seg_end", "seg_next_exon_beg", "seg_gene_id", "seg_tid_off", "tid", "t2s_off", "gene_start",
                   "gene_end", "chr_shift"):
            put(nm, np.uint32)
        put("trans_start_ind", np.int32)
        put("chr_id", np.int32)
        put("t2s", np.uint8)
        put("near_border", np.uint64)
        put("intronic", np.uint64)
        self.lib.hs_load_annotation(self.ctx, contig_id, C.byref(f))

    def get_best_chains(self, contig_id, seqs, lens=None):
        a, lens = _rows(seqs, lens)
        n, Cn, F = a.shape[0], self.Cn, self.F
        out = dict(chain_count=np.zeros(n, np.int32), high_hits=np.zeros(n, np.int32),
                   score=np.zeros((n, Cn), np.float32), chain_len=np.zeros((n, Cn), np.uint32),
                   frag_rpos=np.zeros((n, Cn, F), np.uint32), frag_qpos=np.zeros((n, Cn, F), np.int32))
        self.lib.hs_chain_batch(self.ctx, contig_id, C.c_uint32(n), _vp(a), C.c_uint32(a.shape[1]), _vp(lens),
                                _vp(out["chain_count"]), _vp(out["high_hits"]), _vp(out["score"]),
                                _vp(out["chain_len"]), _vp(out["frag_rpos"]), _vp(out["frag_qpos"]))
        return out

    def circ_chaining(self, contig_id, genome_ascii, seqs, qspos, qepos, gene_start, gene_end, window_size=8, step=3):
        """ProcessCirc::chaining per row through cm_circ.h; the contig's ws-mer index is made here with numpy
        (genome_ascii[p - 1] = base at 1-based p)."""
        pos, off = self._circ_index(genome_ascii, window_size)
        ws = window_size
        a = np.ascontiguousarray(seqs, np.uint8)
        n, Cn = a.shape[0], self.Cn
        F = (self.params.max_read_len - ws) // step + 1
        u = lambda x: np.ascontiguousarray(x, np.uint32)
        qspos, qepos, gene_start, gene_end = u(qspos), u(qepos), u(gene_start), u(gene_end)
        out = dict(chain_count=np.zeros(n, np.int32), kmer_cnt=np.zeros(n, np.int32), score=np.zeros((n, Cn), np.float32),
                   chain_len=np.zeros((n, Cn), np.uint32), frag_rpos=np.zeros((n, Cn, F), np.uint32), frag_qpos=np.zeros((n, Cn, F), np.int32))
        self.lib.hs_circ_chain(self.ctx, contig_id, _vp(pos), _vp(off), C.c_uint32(n), _vp(a), C.c_uint32(a.shape[1]), _vp(qspos), _vp(qepos),
                               _vp(gene_start), _vp(gene_end), ws, step, F,
                               *[_vp(out[k]) for k in ("chain_count", "kmer_cnt", "score", "chain_len", "frag_rpos", "frag_qpos")])
        return out

    def _circ_index(self, genome_ascii, ws):
        """The contig's ws-mer index (positions sorted by (hash, position), run offsets), cached per genome array."""
        key = (id(genome_ascii), ws)
        if getattr(self, "_circ_ix", None) and self._circ_ix[0] == key:
            return self._circ_ix[1], self._circ_ix[2]
        g = np.ascontiguousarray(genome_ascii, np.uint8)
        code = np.full(256, 4, np.int64)
        for i, ch in enumerate(b"ACGT"):
            code[ch] = i
            code[ch + 32] = i
        v = code[g]
        nk = len(g) - ws + 1
        bad = np.concatenate([[0], np.cumsum(v == 4)])
        ok = (bad[ws:ws + nk] - bad[:nk]) == 0
        key = np.zeros(nk, np.int64)
        for j in range(ws):
            key = key * 4 + (v[j:j + nk] & 3)
        p0 = np.nonzero(ok)[0]
        order = np.argsort(key[p0], kind="stable")
        pos = np.ascontiguousarray(p0[order] + 1, np.uint32)
        off = np.zeros(4 ** ws + 1, np.uint32)
        off[1:] = np.cumsum(np.bincount(key[p0], minlength=4 ** ws))
        self._circ_ix = ((id(genome_ascii), ws), pos, off)
        return pos, off

    def circ_single(self, contig_id, genome_ascii, s1, s2, l1, l2, mr_con, gene_cnt, gene_se, window_size=8, step=3, cap=20000):
        """call_circ_single_split's loop (csrc/cm_circ_orch.h) over the host build of chaining / find_exact_coord / check_split_map."""
        from circminer_b200.api import CIRC_MATE_DTYPE, CIRC_RES_DTYPE
        pos, off = self._circ_index(genome_ascii, window_size)
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
        k = self.lib.hs_circ_single(self.ctx, contig_id, _vp(pos), _vp(off), C.c_uint32(n), _vp(a1), _vp(a2), C.c_uint32(stride), _vp(l1), _vp(l2),
                                    _vp(mr), _vp(goff), _vp(gs), _vp(ge), window_size, step, C.c_uint32(cap),
                                    *[_vp(rec[x]) for x in ("call_pair", "r1_partial", "in_r1", "in_r2", "in_pm", "ret", "cr", "out_r1", "out_r2", "out_pm")],
                                    _vp(best), _vp(best_cr))
        assert k >= 0, k
        out = {x: v[:k] for x, v in rec.items()}
        out["best_type"], out["best_cr"] = best, best_cr
        return out


    def circ_exact(self, contig_id, prob: dict, window_size=8) -> dict:
        """ProcessCirc::find_exact_coord per row (src/process_circ.cpp:739); `prob` holds arrays named as the C arguments
        (seqs, r1_spos, r1_epos, r2_spos, r2_epos, dir, qspos, rlen, whole_len, chain_len, frag_rpos, frag_qpos)."""
        from circminer_b200.api import MATE_DTYPE
        a = np.ascontiguousarray(prob["seqs"], np.uint8)
        n = a.shape[0]
        u = {k: np.ascontiguousarray(prob[k], np.uint32) for k in ("r1_spos", "r1_epos", "r2_spos", "r2_epos", "qspos", "chain_len", "frag_rpos")}
        s = {k: np.ascontiguousarray(prob[k], np.int32) for k in ("dir", "rlen", "whole_len", "frag_qpos")}
        res = dict(pmm=np.zeros(n, MATE_DTYPE), ret=np.zeros(n, np.int32))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.hs_circ_exact(self.ctx, contig_id, C.c_uint32(n), vp(a), C.c_uint32(a.shape[1]), vp(u["r1_spos"]), vp(u["r1_epos"]), vp(u["r2_spos"]), vp(u["r2_epos"]),
               vp(s["dir"]), vp(u["qspos"]), vp(s["rlen"]), vp(s["whole_len"]), C.c_uint32(u["frag_rpos"].shape[1]), vp(u["chain_len"]),
               vp(u["frag_rpos"]), vp(s["frag_qpos"]), window_size, vp(res["pmm"]), vp(res["ret"]))
        assert r == 0
        return res


    def circ_check(self, contig_id, rec: dict, window_size=8) -> dict:
        """ProcessCirc::check_split_map (three mates) per row (src/process_circ.cpp:892; cm_gpu_circ_check_batch).  rec: remain_rows,
        fullmap_rows, remain_len, fullmap_len, r1_partial, in_r1, in_r2, in_pm -> ret, cr and the mates as the call leaves them."""
        from circminer_b200.api import CIRC_MATE_DTYPE, CIRC_RES_DTYPE
        rr = np.ascontiguousarray(rec["remain_rows"], np.uint8)
        fr = np.ascontiguousarray(rec["fullmap_rows"], np.uint8)
        n = rr.shape[0]
        rl, fl = np.ascontiguousarray(rec["remain_len"], np.int32), np.ascontiguousarray(rec["fullmap_len"], np.int32)
        rp = np.ascontiguousarray(rec["r1_partial"], np.uint8)
        out = dict(r1=np.array(rec["in_r1"], CIRC_MATE_DTYPE), r2=np.array(rec["in_r2"], CIRC_MATE_DTYPE), pm=np.array(rec["in_pm"], CIRC_MATE_DTYPE),
                   ret=np.zeros(n, np.int32), cr=np.zeros(n, CIRC_RES_DTYPE))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.hs_circ_check(self.ctx, contig_id, C.c_uint32(n), vp(rr), vp(fr), C.c_uint32(rr.shape[1] if n else 0), vp(rl), vp(fl), vp(rp), vp(out["r1"]), vp(out["r2"]),
               vp(out["pm"]), window_size, vp(out["ret"]), vp(out["cr"]))
        assert r == 0
        return out

    def circ_check4(self, contig_id, rec: dict, window_size=8) -> dict:
        """ProcessCirc::check_split_map (four mates) per row (src/process_circ.cpp:922; cm_circ_check4_row, the function k_circ_check4 runs).
        rec as RefHarness.circ_check4."""
        from circminer_b200.api import CIRC_MATE_DTYPE, CIRC_RES_DTYPE
        a, b = np.ascontiguousarray(rec["r1_rows"], np.uint8), np.ascontiguousarray(rec["r2_rows"], np.uint8)
        n = a.shape[0]
        l1, l2 = np.ascontiguousarray(rec["l1"], np.int32), np.ascontiguousarray(rec["l2"], np.int32)
        fs = np.ascontiguousarray(rec["fullmap_sel"], np.uint8)
        m = [np.ascontiguousarray(np.array(rec[k], CIRC_MATE_DTYPE)) for k in ("r1_1", "r2_1", "r1_2", "r2_2")]
        out = dict(ret=np.zeros(n, np.int32), cr=np.zeros(n, CIRC_RES_DTYPE))
        vp = lambda x: x.ctypes.data_as(C.c_void_p)
        r = self.lib.hs_circ_check4(self.ctx, contig_id, C.c_uint32(n), vp(a), vp(b), C.c_uint32(a.shape[1] if n else 0), vp(l1), vp(l2), vp(fs),
                                    *[vp(x) for x in m], window_size, vp(out["ret"]), vp(out["cr"]))
        assert r == 0
        return out

    def circ_double(self, contig_id, genome_ascii, s1, s2, l1, l2, mr_con, gene_cnt, gene_se, window_size=8, step=3, cap=50000):
        """call_circ_double_split for a batch (csrc/cm_circ_orch.h: circ_orch_double) over the host build of the device functions."""
        from circminer_b200.api import CIRC_MATE_DTYPE, CIRC_RES_DTYPE
        pos, off = self._circ_index(genome_ascii, window_size)
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
        k = self.lib.hs_circ_double(self.ctx, contig_id, _vp(pos), _vp(off), C.c_uint32(n), _vp(a1), _vp(a2), C.c_uint32(stride), _vp(l1), _vp(l2),
                                    _vp(mr), _vp(goff), _vp(gs), _vp(ge), window_size, step, C.c_uint32(cap),
                                    *[_vp(ev[x]) for x in ("pair", "kind", "r1", "r2", "p1", "p2", "ret", "cr")], _vp(push_type), _vp(push_cr))
        assert 0 <= k <= cap, k
        out = {x: v[:k] for x, v in ev.items()}
        out["push_type"], out["push_cr"] = push_type, push_cr
        return out

    def process_read(self, contig_id, seqs1, seqs2, mreads, lens1=None, lens2=None):
        a1, l1 = _rows(seqs1, lens1)
        a2, l2 = _rows(seqs2, lens2)
        n = a1.shape[0]
        mreads = np.ascontiguousarray(mreads)
        out = np.zeros(n, MREAD_DTYPE)
        state = np.zeros(n, np.int32)
        self.ctr = CmCounters()
        self.lib.hs_process_batch_pe(self.ctx, contig_id, C.c_uint32(n), _vp(a1), _vp(a2), C.c_uint32(a1.shape[1]),
                                     _vp(l1), _vp(l2), _vp(mreads), _vp(out), _vp(state), C.byref(self.ctr))
        return out, state

    def map_rounds(self, seqs1, seqs2, mreads, first=0, n_contigs=1, lens1=None, lens2=None, snapshots=False):
        a1, l1 = _rows(seqs1, lens1)
        a2, l2 = _rows(seqs2, lens2)
        n = a1.shape[0]
        res = dict(out=np.zeros(n, MREAD_DTYPE), state=np.zeros(n, np.int32), final_round=np.zeros(n, np.int32))
        if snapshots:
            res["round_out"] = np.zeros((n_contigs, n), MREAD_DTYPE)
            res["round_state"] = np.zeros((n_contigs, n), np.int32)
        self.lib.hs_map_rounds_pe(self.ctx, first, n_contigs, C.c_uint32(n), _vp(a1), _vp(a2), C.c_uint32(a1.shape[1]), _vp(l1),
                                  _vp(l2), _vp(np.ascontiguousarray(mreads)), _vp(res["out"]), _vp(res["state"]),
                                  _vp(res["final_round"]), _vp(res.get("round_out")), _vp(res.get("round_state")))
        return res

    def process_read_se(self, contig_id, seqs, lens=None):
        a, l = _rows(seqs, lens)
        state = np.zeros(a.shape[0], np.int32)
        self.lib.hs_process_batch_se(self.ctx, contig_id, C.c_uint32(a.shape[0]), _vp(a), C.c_uint32(a.shape[1]),
                                     _vp(l), _vp(state))
        return state

    def align(self, kind, s, ns, t, nt, aux=None):
        s, ns = _rows(s, ns)
        t, nt = _rows(t, nt)
        n = s.shape[0]
        aux = np.zeros(n, np.int32) if aux is None else np.ascontiguousarray(aux, np.int32)
        o = [np.zeros(n, np.int32) for _ in range(4)]
        self.lib.hs_align_batch(self.ctx, kind, C.c_uint32(n), _vp(s), _vp(ns), _vp(t), _vp(nt),
                                C.c_uint32(s.shape[1]), _vp(aux), *[_vp(x) for x in o])
        return dict(ed=o[0], sclen=o[1], indel=o[2], score=o[3])
