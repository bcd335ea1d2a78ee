import sys, os, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from parity import live_dump
from circminer_b200 import workload
from circminer_b200.api import Mapper, default_params
d = "/tmp/dbgcs"; os.makedirs(d, exist_ok=True)
w = workload.prepare("T1", d + "/t1", contig_size=3_500_000, n_pairs=20000)
dump = live_dump(d + "/cs.npz", workload="T1", contig_size=3_500_000, pairs=20000, circ_single=True, no_stage_dumps=True)
m = Mapper(default_params()); workload.load_into(m, w)
c = 0
rec = {k[len(f"c{c}_circs_"):]: dump[k] for k in dump.files if k.startswith(f"c{c}_circs_")}
got = m.circ_check(c, rec)
bad = np.nonzero(got["cr"] != rec["cr"])[0]
print("bad", bad)
for i in bad[:3]:
    print("rec", i, "ret", got["ret"][i], rec["ret"][i], "lens", rec["remain_len"][i], rec["fullmap_len"][i], "r1p", rec["r1_partial"][i])
    print(" got", got["cr"][i]); print(" want", rec["cr"][i])
    print(" in_r1", rec["in_r1"][i]); print(" in_r2", rec["in_r2"][i]); print(" in_pm", rec["in_pm"][i])
    print(" remain", bytes(rec["remain_rows"][i][:rec["remain_len"][i] + 4])); print(" full  ", bytes(rec["fullmap_rows"][i][:rec["fullmap_len"][i] + 4]))
    one = {k: v[i:i + 1] for k, v in rec.items() if k != "best_type"}
    g1 = m.circ_check(c, one)
    print(" alone", g1["cr"][0], g1["ret"][0])
    for rep in range(3):
        g = m.circ_check(c, rec)
        print(" again", g["cr"][i])
