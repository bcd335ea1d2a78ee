"""TEST / BENCH INFRASTRUCTURE ONLY (bench.py's cpu_baseline and --impl reference legs).

Times the UNMODIFIED reference (oracle/_ref, built from /root/reference by oracle/Makefile) on the
host cores of this box, on a bounded sample of the bench workload:
  --threads 1 : FilterRead::process_read called per pair through oracle/_ref/libcm_ref.so, every packed contig in
                order with the reference's skip rule and header round trip between them (oracle/rounds.py): the hot
                path alone, reads already in memory, index loads outside the clock;
  --threads N : oracle/_ref/circminer -t N, mapping stage only (-q 0), time = sum of the
                "Completed round" real-time log lines (src/circminer.cpp:302), index/GTF load excluded.
Prints one JSON object on the last line.

The GRCh38-scale configs need <ref>.packed.fa.index (25 GB).  If it is not there yet it is made here: from the device
index with cm_gpu_write_index when a GPU and libcm_gpu.so are present (13 s; the file equals the reference builder's,
tests/test_gpu_index_io.py), else by the reference's own `--index` (about 8 minutes).  CM_REF_INDEX=reference forces
the latter."""
from __future__ import annotations

import argparse
import fcntl
import json
import os
import re
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def ensure_ref_index(w, workload_name, kmer, compact=False):
    """Reference index next to the FASTA; built once per data dir."""
    from circminer_b200 import workload
    from oracle.refharness import REF_BIN
    fa = w["fa"]
    path = fa + ".packed.fa.index"
    with open(path + ".lock", "w") as lk:
        fcntl.flock(lk, fcntl.LOCK_EX)
        try:
            if os.path.exists(path + ".done") or (not workload.is_scale(workload_name) and os.path.exists(path)):
                return "cached"
            how = None
            if workload.is_scale(workload_name) and os.environ.get("CM_REF_INDEX") != "reference":
                try:
                    from circminer_b200.api import Mapper
                    m = Mapper(workload.params_for(workload_name))
                    for c, seq in enumerate(w["contigs"]):
                        m.build_contig(c, np.ascontiguousarray(seq))
                    m.write_index(path, len(w["contigs"]), compact=compact)
                    m.close()
                    how = "cm_gpu_write_index"
                except Exception as e:  # noqa: BLE001
                    sys.stderr.write(f"cpu_baseline: device index writer unavailable ({e!r}); using the reference's builder\n")
            if how is None:
                if not os.path.exists(fa + ".done") and workload.is_scale(workload_name):
                    from circminer_b200 import synth_scale
                    synth_scale.write_fasta(fa, w["g"], w["off"], w["names"])
                    open(fa + ".done", "w").close()
                csz = getattr(workload.config(workload_name), "contig_size", 1_100_000_000)
                if csz != 1_100_000_000:   # the CLI always packs at 1.1 Gbp; the scaled-down configs pack smaller contigs
                    code = ("import sys; sys.path.insert(0, %r)\n"
                            "from oracle.refharness import RefHarness, default_params\n"
                            "RefHarness(default_params(kmer=%d), threads=%d).build_index(%r, contig_size=%d, compact=%r)"
                            % (ROOT, kmer, os.cpu_count() or 1, fa, csz, bool(compact)))
                    subprocess.run([sys.executable, "-c", code], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                else:
                    cmd = [REF_BIN, "--index", "-r", fa, "-k", str(kmer), "-t", str(os.cpu_count() or 1)] + (["-m"] if compact else [])
                    subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                how = "circminer --index"
            open(path + ".done", "w").write(how + "\n")
            return how
        finally:
            fcntl.flock(lk, fcntl.LOCK_UN)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--data", required=True)
    ap.add_argument("--pairs", type=int, default=0)
    ap.add_argument("--threads", type=int, default=1)
    ap.add_argument("--target-s", type=float, default=12.0)
    a = ap.parse_args()
    from circminer_b200 import synth, workload
    from oracle import rounds
    from oracle.refharness import REF_BIN, RefHarness, default_params, fresh_mreads
    cfg = workload.config(a.workload)
    scale = workload.is_scale(a.workload)
    k_bin = 200_000 if scale else 400_000
    n_want = a.pairs or cfg.n_pairs
    if a.threads > 1:
        n_want = min(n_want, k_bin)
    elif scale:
        n_want = min(n_want, 250_000)
    w = workload.prepare(a.workload, a.data, n_pairs=n_want)
    p = workload.params_for(a.workload)
    index_how = ensure_ref_index(w, a.workload, p.kmer, compact=bool(getattr(cfg, "compact_index", False)))
    s1, s2 = w["s1"], w["s2"]
    n, L = s1.shape
    n_contigs = len(w["contigs"])
    if a.threads <= 1:
        h = RefHarness(default_params(**{f: getattr(p, f) for f, _ in p._fields_}))
        h.open(w["fa"], w["gtf"], os.path.join(a.data, "cpu_base"))
        lens = np.full(n, L, np.uint16)
        # size the sample on the first contig: a small slice, then ~target seconds for all rounds together
        more = h.next_contig()
        k0 = min(n, 4000 if scale else 20000)
        t0 = time.time()
        h.process_pe(s1[:k0], s2[:k0], lens[:k0], lens[:k0], fresh_mreads(k0, p.max_ed))
        dt0 = time.time() - t0
        k = int(min(n, max(k0, a.target_s / (1.8 if n_contigs > 1 else 1.0) / max(dt0, 1e-6) * k0)))
        ids = np.arange(k)
        mr = fresh_mreads(k, p.max_ed)
        dt = 0.0
        visits = []
        c = 0
        while True:
            t0 = time.time()
            out, st = h.process_pe(s1[ids], s2[ids], lens[ids], lens[ids], mr)
            dt += time.time() - t0
            visits.append(len(ids))
            if not more:
                break
            keep = ~rounds.skip_rule(p.scan_level, st, out, 2 * L)
            ids, mr = ids[keep], rounds.header_round_trip(out[keep], p.max_ed)
            more = h.next_contig()
            c += 1
            if not len(ids):
                break
        print(json.dumps({"value": 2.0 * k / dt, "unit": "reads/s", "cores": 1, "kind": "reference",
                          "seconds": dt, "index": index_how,
                          "sample": f"first {k} of {n_want} pairs of {a.workload}, {n_contigs} packed contig(s) in order (pairs per round {visits}), "
                                    "FilterRead::process_read per pair via oracle/_ref/libcm_ref.so, index loads outside the clock"}))
        sys.stdout.flush()
        os._exit(0)   # the reference's global destructors are not worth running
    # multi-threaded: the reference binary on FASTQ files
    k = min(n, k_bin)
    fq1 = os.path.join(a.data, f"base_{k}_1.fastq")
    fq2 = os.path.join(a.data, f"base_{k}_2.fastq")
    if not os.path.exists(fq2):
        synth.write_fastq(fq1, w["r1"][:k], 1)
        synth.write_fastq(fq2, w["r2"][:k], 2)
    out = os.path.join(a.data, "base_out")
    extra = []
    for flag, f, dflt in (("-a", "scan_level", 0), ("-e", "max_ed", 4), ("-w", "band_width", 3), ("-l", "max_read_len", 300), ("-S", "seed_lim", 500)):
        if getattr(p, f) != dflt:
            extra += [flag, str(getattr(p, f))]
    pr = subprocess.run([REF_BIN, "-r", w["fa"], "-g", w["gtf"], "-1", fq1, "-2", fq2, "-k", str(p.kmer), "-t", str(a.threads),
                         "-o", out, "-q", "0", "-P", *extra], capture_output=True, text=True)
    txt = pr.stdout + pr.stderr
    secs = [float(x) for x in re.findall(r"Completed round \S+ \(CPU time: [\d.]+s; Real time: ([\d.]+)s\)", txt)]
    if pr.returncode != 0 or not secs:
        print(json.dumps({"value": None, "error": txt[-300:]}))
        return
    dt = sum(secs)
    print(json.dumps({"value": 2.0 * k / dt, "unit": "reads/s", "cores": a.threads, "kind": "reference",
                      "seconds": dt, "round_seconds": secs, "index": index_how,
                      "sample": f"first {k} pairs of {a.workload}, all {len(secs)} rounds, oracle/_ref/circminer -t {a.threads} "
                                "-q 0 (mapping stage; FASTQ parse and PAM / remain writing inside the timed rounds as in the reference)"}))


if __name__ == "__main__":
    main()
