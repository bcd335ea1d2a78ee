"""Debug helper: raw work counters after each stage of the staged pipeline (T1, contig 0)."""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from circminer_b200 import workload
from circminer_b200.api import Mapper, default_params, fresh_mreads
w = workload.prepare("T1", tempfile.mkdtemp(prefix="cm_run_"), contig_size=3_500_000)
m = Mapper(default_params())
workload.load_into(m, w)
n = len(w["s1"])
mr = fresh_mreads(n)
m.stage(w["s1"], w["s2"], mr)
for rep in range(3):
    for upto in (1, 2, 3):
        m.counters(reset=True)
        m.run_staged(0, upto)
        c = m.counters()
        print(os.environ.get("CM_GPU_NO_SPEC"), os.environ.get("CM_GPU_DBG"), "rep", rep, "upto", upto, "pairs", hex(int(c["pairs"])), int(c["pairs"]),
              {k: int(v) for k, v in c.items() if k != "pairs"}, flush=True)
m.close()
