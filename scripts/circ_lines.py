"""Timed lines for the stage-2 pieces on the device at GRCh38 scale: the per-contig 8-mer index build,
cm_gpu_circ_chain_batch (ProcessCirc::chaining) and cm_gpu_circ_exact_batch (ProcessCirc::find_exact_coord) on split-read
problems over the genes of the first packed contig of C3 (1.06 Gbp), with the reference's own functions timed beside them on a
bounded sample of the same problems (oracle/_ref/libcm_ref.so: ProcessCirc::chaining on its RegionalHashTable, one table build
per gene as get_hash_table_smart does; ProcessCirc::find_exact_coord), whose results must equal the device's.

usage (GPU box): python scripts/circ_lines.py [--problems N] [--ref-sample M] [--data DIR] [--out FILE.json]
TEST / BENCH INFRASTRUCTURE: executes oracle/_ref as the checker and the timed baseline."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def exact_problems(prob, res, flat, L, seed, top=10):
    """One find_exact_coord problem per chain ProcessCirc would try; the pair's mates sit in exon intervals near the chain."""
    rng = np.random.default_rng(seed)
    ivs, ive = flat["iv_spos"].astype(np.int64), flat["iv_epos"].astype(np.int64)
    cnt = np.minimum(res["chain_count"], top)
    rows = np.repeat(np.arange(len(cnt)), np.maximum(cnt, 0))
    js = np.concatenate([np.arange(c) for c in cnt if c > 0]) if len(rows) else np.zeros(0, np.int64)
    n = len(rows)
    anchor = res["frag_rpos"][rows, js, 0].astype(np.int64)
    k = np.clip(np.searchsorted(ivs, anchor, side="right") - 1 + rng.integers(-2, 3, size=n), 0, len(ivs) - 1)
    k2 = np.clip(np.searchsorted(ivs, anchor, side="right") - 1 + rng.integers(-2, 3, size=n), 0, len(ivs) - 1)
    s1 = ivs[k] + (rng.random(n) * (ive[k] - ivs[k] + 1)).astype(np.int64)
    s2 = ivs[k2] + (rng.random(n) * (ive[k2] - ivs[k2] + 1)).astype(np.int64)
    return dict(seqs=prob["seqs"][rows], r1_spos=s1.astype(np.uint32), r1_epos=(s1 + rng.integers(20, L, size=n)).astype(np.uint32),
                r2_spos=s2.astype(np.uint32), r2_epos=(s2 + rng.integers(20, L, size=n)).astype(np.uint32),
                dir=np.where(rng.random(n) < 0.5, 1, -1).astype(np.int32), qspos=prob["qspos"][rows],
                rlen=(prob["qepos"][rows].astype(np.int64) - prob["qspos"][rows].astype(np.int64) + 1).astype(np.int32),
                whole_len=np.full(n, L, np.int32), chain_len=res["chain_len"][rows, js], frag_rpos=res["frag_rpos"][rows, js],
                frag_qpos=res["frag_qpos"][rows, js])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--problems", type=int, default=100_000)
    ap.add_argument("--ref-sample", type=int, default=3000)
    ap.add_argument("--data", default="/tmp/cm_bench")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    from circminer_b200 import synth, workload
    from circminer_b200.api import Mapper
    w = workload.prepare(a.workload, os.path.join(a.data, f"{a.workload}_r0"), n_pairs=1000)
    p = workload.params_for(a.workload)
    L = w["s1"].shape[1]
    m = Mapper(p)
    m.build_contig(0, np.ascontiguousarray(w["contigs"][0]))
    m.load_annotation(0, w["flat"][0])
    flat = w["flat"][0]
    t = time.time()
    prob = synth.split_read_problems(w["contigs"][0], flat["gene_start"], flat["gene_end"], L, a.problems, 77)
    t_prob = time.time() - t
    t = time.time()
    m._chk(m.lib.cm_gpu_build_circ_index(m.ctx, 0, 8), "cm_gpu_build_circ_index")
    t_index = time.time() - t
    m.circ_chaining(0, **{k: v[:2000] for k, v in prob.items()})          # warm-up
    t = time.time()
    res = m.circ_chaining(0, **prob)
    t_chain = time.time() - t
    ex = exact_problems(prob, res, flat, L, 78)
    m.circ_exact(0, {k: v[:2000] for k, v in ex.items()})
    t = time.time()
    xr = m.circ_exact(0, ex)
    t_exact = time.time() - t
    out = {"workload": a.workload, "contig_bp": int(len(w["contigs"][0])), "genes": int(len(flat["gene_start"])), "read_len": int(L),
           "index_build_s": t_index, "index_bytes": 4 * int(len(w["contigs"][0])),
           "chaining": {"problems": int(a.problems), "wall_s_with_copies": t_chain, "problems_per_s": a.problems / t_chain,
                        "with_chains": int((res["chain_count"] > 0).sum()), "kmer_rows": float(res["kmer_cnt"].mean())},
           "find_exact_coord": {"problems": int(len(ex["dir"])), "wall_s_with_copies": t_exact, "problems_per_s": len(ex["dir"]) / t_exact,
                                "concordant": int((xr["ret"] == 1).sum())},
           "problem_generation_s": t_prob}
    m.close()
    # the reference's own functions on a sample of the same problems (its genome and GTF through the harness)
    if a.ref_sample:
        from oracle.cpu_baseline import ensure_ref_index
        from oracle.refharness import RefHarness
        ensure_ref_index(w, a.workload, p.kmer)
        h = RefHarness(p)
        h.open(w["fa"], w["gtf"], os.path.join(a.data, f"{a.workload}_r0", "circ_ref"))
        h.next_contig()
        # sorted by gene, as do_process meets the candidates (one table build per gene)
        order = np.argsort(prob["gene_start"][:a.ref_sample], kind="stable")
        sp = {k: v[:a.ref_sample][order] for k, v in prob.items()}
        t = time.time()
        rr = h.circ_chain(sp)
        t_ref = time.time() - t
        Fd = res["frag_rpos"].shape[2]   # the device call sizes the fragment slots by the batch, the reference by maxReadLength
        same = all(np.array_equal(rr[k], res[k][:a.ref_sample][order]) for k in ("chain_count", "score", "chain_len"))
        same = same and all(np.array_equal(rr[k][:, :, :Fd], res[k][:a.ref_sample][order]) and not rr[k][:, :, Fd:].any() for k in ("frag_rpos", "frag_qpos"))
        nx = min(len(ex["dir"]), 10 * a.ref_sample)
        sx = {k: v[:nx] for k, v in ex.items()}
        t = time.time()
        rx = h.circ_exact(sx)
        t_refx = time.time() - t
        same_x = bool((rx["ret"] == xr["ret"][:nx]).all() and (rx["pmm"] == xr["pmm"][:nx]).all())
        out["reference"] = {"cores": 1, "chaining": {"problems": int(a.ref_sample), "genes": int(len(np.unique(sp["gene_start"]))), "seconds": t_ref,
                                                      "problems_per_s": a.ref_sample / t_ref, "equal_to_device": bool(same)},
                            "find_exact_coord": {"problems": int(nx), "seconds": t_refx, "problems_per_s": nx / t_refx, "equal_to_device": same_x}}
    line = json.dumps(out)
    print(line)
    if a.out:
        with open(a.out, "w") as f:
            f.write(line + "\n")
    if a.ref_sample and not (out["reference"]["chaining"]["equal_to_device"] and out["reference"]["find_exact_coord"]["equal_to_device"]):
        raise SystemExit("device and reference differ")


if __name__ == "__main__":
    main()
