"""Debug helper: run the T1 workload through the C-ABI (no checker), for compute-sanitizer / ncu."""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from circminer_b200 import workload
from circminer_b200.api import Mapper, default_params, fresh_mreads
name = sys.argv[1] if len(sys.argv) > 1 else "T1"
npairs = int(sys.argv[2]) if len(sys.argv) > 2 else None
w = workload.prepare(name, tempfile.mkdtemp(prefix="cm_run_"), contig_size=3_500_000 if name == "T1" else 1_100_000_000, n_pairs=npairs)
m = Mapper(default_params())
workload.load_into(m, w)
mr = fresh_mreads(len(w["s1"]))
for c in range(len(w["contigs"])):
    mr, st = m.process_read(c, w["s1"], w["s2"], mr)
    print("contig", c, "ok", m.counters())
m.close()
