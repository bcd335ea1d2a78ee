"""TEST / BENCH INFRASTRUCTURE ONLY (bench.py's cpu_baseline and --impl reference legs).

Times the UNMODIFIED reference (oracle/_ref, built from /root/reference by oracle/Makefile) on the
host cores of this box, on a bounded sample of the bench workload:
  --threads 1 : FilterRead::process_read called per pair through oracle/_ref/libcm_ref.so
                (the hot path alone, reads already in memory);
  --threads N : oracle/_ref/circminer -t N, mapping stage only (-q 0), time = sum of the
                "Completed round" real-time log lines (src/circminer.cpp:302), index/GTF load excluded.
Prints one JSON object on the last line."""
from __future__ import annotations

import argparse
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


def ensure_ref_index(fa):
    """Reference index next to the FASTA (GenomePacker::build_index); built once per data dir."""
    from oracle.refharness import REF_BIN
    if os.path.exists(fa + ".packed.fa.index.info") and os.path.exists(fa + ".packed.fa.index"):
        return
    subprocess.run([REF_BIN, "--index", "-r", fa, "-k", "20", "-t", str(os.cpu_count() or 1)], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--data", required=True)
    ap.add_argument("--pairs", type=int, default=0)
    ap.add_argument("--threads", type=int, default=1)
    ap.add_argument("--target-s", type=float, default=12.0)
    a = ap.parse_args()
    from circminer_b200 import synth
    from oracle.refharness import REF_BIN, RefHarness, default_params, fresh_mreads
    from dataclasses import replace
    cfg = synth.CONFIGS[a.workload]
    if a.pairs:
        cfg = replace(cfg, n_pairs=a.pairs)
    d = synth.generate(cfg, a.data, write_reads=False)
    ensure_ref_index(d["fa"])
    s1 = synth.codes_to_ascii(d["r1"])
    s2 = synth.codes_to_ascii(d["r2"])
    n = len(s1)
    L = s1.shape[1]
    if a.threads <= 1:
        h = RefHarness(default_params())
        h.open(d["fa"], d["gtf"], os.path.join(a.data, "cpu_base"))
        more = h.next_contig()
        assert more == 0 or True
        # calibrate on a small slice, then size the sample for ~target seconds
        k0 = min(n, 20000)
        t0 = time.time()
        h.process_pe(s1[:k0], s2[:k0], np.full(k0, L, np.uint16), np.full(k0, L, np.uint16), fresh_mreads(k0))
        dt = time.time() - t0
        k = int(min(n, max(k0, a.target_s / max(dt, 1e-6) * k0)))
        t0 = time.time()
        h.process_pe(s1[:k], s2[:k], np.full(k, L, np.uint16), np.full(k, L, np.uint16), fresh_mreads(k))
        dt = time.time() - t0
        print(json.dumps({"value": 2.0 * k / dt, "unit": "reads/s", "cores": 1, "kind": "reference",
                          "seconds": dt,
                          "sample": f"first {k} of {n} pairs of {a.workload}, first packed contig, "
                                    "FilterRead::process_read per pair via oracle/_ref/libcm_ref.so"}))
        sys.stdout.flush()
        os._exit(0)   # the reference's global destructors are not worth running
    # multi-threaded: the reference binary on FASTQ files
    k = min(n, 400000)
    fq1 = os.path.join(a.data, f"base_{k}_1.fastq")
    fq2 = os.path.join(a.data, f"base_{k}_2.fastq")
    if not os.path.exists(fq2):
        synth.write_fastq(fq1, d["r1"][:k], 1)
        synth.write_fastq(fq2, d["r2"][:k], 2)
    out = os.path.join(a.data, "base_out")
    p = subprocess.run([REF_BIN, "-r", d["fa"], "-g", d["gtf"], "-1", fq1, "-2", fq2, "-k", "20", "-t", str(a.threads),
                        "-o", out, "-q", "0", "-P"], capture_output=True, text=True)
    txt = p.stdout + p.stderr
    secs = [float(x) for x in re.findall(r"Completed round \S+ \(CPU time: [\d.]+s; Real time: ([\d.]+)s\)", txt)]
    if p.returncode != 0 or not secs:
        print(json.dumps({"value": None, "error": txt[-300:]}))
        return
    dt = sum(secs)
    print(json.dumps({"value": 2.0 * k / dt, "unit": "reads/s", "cores": a.threads, "kind": "reference",
                      "seconds": dt,
                      "sample": f"first {k} of {n} pairs of {a.workload}, all rounds, oracle/_ref/circminer -t {a.threads} "
                                "-q 0 (mapping stage; FASTQ parse inside the timed rounds as in the reference)"}))


if __name__ == "__main__":
    main()
