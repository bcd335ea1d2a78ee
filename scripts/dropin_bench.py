"""Whole-binary comparison at bench scale: the unmodified reference (`oracle/_ref/circminer -t <all cores>`)
against the drop-in (`oracle/_ref/circminer_gpu`, the same objects with the mapping threads routed to
libcm_gpu.so) on the first --pairs pairs of a bench workload (default C2, 1 M pairs 2x150; C3 = GRCh38 scale, three
rounds), the drop-in once per --devices setting (CM_GPU_DEVICES: which GPUs the batches are dealt to).

Reports, per binary, the mapping time (sum of the reference's own "Completed round" real-time log lines,
src/circminer.cpp:302: FASTQ parse, mapping and PAM / remain writes; index and GTF load excluded), the whole
process wall time, and whether the outputs agree: candidates.pam and circ_report byte for byte, mapping.pam
as a sorted multiset of lines (the reference's line order depends on thread timing).

usage (GPU box): python scripts/dropin_bench.py [--workload C2] [--pairs N] [--data DIR] [--out FILE.json]
TEST / BENCH INFRASTRUCTURE: executes oracle/_ref as the checker and the timed baseline."""
import argparse
import hashlib
import json
import os
import re
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(binary, d, fq1, fq2, out, threads, env=None, extra=()):
    cmd = [binary, "-r", d["fa"], "-g", d["gtf"], "-1", fq1, "-2", fq2, "-k", "20", "-t", str(threads), "-P", "-o", out, *extra]
    e = dict(os.environ, LC_ALL="C")
    e.update(env or {})
    t0 = time.time()
    p = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=e)
    timeline, lines = [], []
    for ln in p.stdout:                       # every log line with the time it appeared at: where the wall time goes
        lines.append(ln)
        if len(timeline) < 400 and ln.strip():
            timeline.append([round(time.time() - t0, 2), ln.rstrip()[:150]])
    p.wait()
    wall = time.time() - t0
    txt = "".join(lines)
    if p.returncode != 0:
        raise SystemExit(f"{binary} failed:\n{txt[-2000:]}")
    rounds = [float(x) for x in re.findall(r"Completed round \S+ \(CPU time: [\d.]+s; Real time: ([\d.]+)s\)", txt)]
    phases = [ln for ln in txt.splitlines() if ln.startswith(("circminer_gpu:", "cm_gpu:"))]
    circ = re.search(r"CircRNA detection Completed! \(CPU time: [\d.]+s; Real time: ([\d.]+)s\)", txt)   # do_process (src/process_circ.cpp:325)
    sort = re.findall(r"Sorting remaining read mappings.*?\n.*?Completed! \(CPU time: [\d.]+s; Real time: ([\d.]+)s\)", txt, re.S)
    return {"wall_s": wall, "round_s": rounds, "mapping_s": sum(rounds), "phases": phases, "timeline": timeline,
            "stage2_do_process_s": float(circ.group(1)) if circ else None, "stage2_sort_s": float(sort[0]) if sort else None}


def digest(path, sort_lines=False):
    if not os.path.exists(path):
        return None
    h = hashlib.sha256()
    if sort_lines:
        with open(path, "rb") as f:
            for ln in sorted(f.readlines()):
                h.update(ln)
    else:
        with open(path, "rb") as f:
            for blk in iter(lambda: f.read(1 << 24), b""):
                h.update(blk)
    return h.hexdigest()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--pairs", type=int, default=0)
    ap.add_argument("--data", default="/tmp/cm_bench")
    ap.add_argument("--out", default="")
    ap.add_argument("--batch", type=int, default=524288)
    ap.add_argument("--devices", nargs="+", default=["0"], help="CM_GPU_DEVICES settings to time the drop-in under, e.g. 0 0,1 all")
    ap.add_argument("--stage2", action="store_true", help="also run the reference with -q 0 (mapping only): whole wall minus that = stage 2 (process_circ)")
    a = ap.parse_args()
    from circminer_b200 import synth, workload
    from oracle.cpu_baseline import ensure_ref_index
    from oracle.refharness import REF_BIN
    gpu_bin = os.path.join(ROOT, "oracle", "_ref", "circminer_gpu")
    cfg = workload.config(a.workload)
    datadir = os.path.join(a.data, f"{a.workload}_r0")
    d = workload.prepare(a.workload, datadir, n_pairs=a.pairs or cfg.n_pairs)
    ensure_ref_index(d, a.workload, workload.params_for(a.workload).kmer)
    n = len(d["r1"])
    fq1 = os.path.join(datadir, f"dropin_{n}_1.fastq")
    fq2 = os.path.join(datadir, f"dropin_{n}_2.fastq")
    if not os.path.exists(fq2):
        synth.write_fastq(fq1, d["r1"], 1)
        synth.write_fastq(fq2, d["r2"], 2)
    ncores = os.cpu_count() or 1
    ref_out = os.path.join(datadir, "dropin_ref")
    gpu_out = os.path.join(datadir, "dropin_gpu")
    res = {"workload": a.workload, "pairs": n, "read_len": int(d["r1"].shape[1]), "host_cores": ncores}
    res["reference"] = run(REF_BIN, d, fq1, fq2, ref_out, ncores)
    res["reference"]["threads"] = ncores
    res["reference"]["reads_per_s_mapping"] = 2.0 * n / res["reference"]["mapping_s"]
    if a.stage2:
        q0 = run(REF_BIN, d, fq1, fq2, ref_out + "_q0", ncores, extra=("-q", "0"))
        res["reference"]["wall_s_mapping_only"] = q0["wall_s"]
        res["reference"]["stage2_s"] = res["reference"]["wall_s"] - q0["wall_s"]
        res["reference"]["load_s"] = q0["wall_s"] - q0["mapping_s"]
    same = {}
    res["dropin"] = []
    for dev in a.devices:
        r = run(gpu_bin, d, fq1, fq2, gpu_out, ncores, env={"CM_GPU_BATCH": str(a.batch), "CM_GPU_VERBOSE": "1", "CM_GPU_DEVICES": dev})
        r.update(batch=a.batch, devices=dev, reads_per_s_mapping=2.0 * n / r["mapping_s"] if r["mapping_s"] else None,
                 speedup_mapping=res["reference"]["mapping_s"] / r["mapping_s"] if r["mapping_s"] else None)
        for suffix, sort_lines in ((".mapping.pam", True), (".candidates.pam", False), (".circ_report", False)):
            ra, rb = digest(ref_out + suffix, sort_lines), digest(gpu_out + suffix, sort_lines)
            same[f"{dev}:{suffix}"] = {"identical": ra is not None and ra == rb, "bytes": os.path.getsize(ref_out + suffix) if ra else None,
                                       "compared": "sorted lines" if sort_lines else "bytes"}
        res["dropin"].append(r)
    res["outputs"] = same
    line = json.dumps(res)
    print(line)
    if a.out:
        with open(a.out, "w") as f:
            f.write(line + "\n")
    if not all(v["identical"] for v in same.values()):
        raise SystemExit("outputs differ")


if __name__ == "__main__":
    main()
