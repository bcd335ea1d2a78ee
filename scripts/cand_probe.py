"""Distribution of candidate chain pairs per read pair on the bench workload (who dominates k_process_pe?)."""
import sys, numpy as np
sys.path.insert(0, ".")
from circminer_b200 import workload
from circminer_b200.api import Mapper, default_params, fresh_mreads

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
w = workload.prepare("C2", "/tmp/cm_cand", n_pairs=n)
m = Mapper(default_params())
workload.load_into(m, w)
rcs = lambda a: None
cnt = {}
from circminer_b200 import synth
def rc(a):
    comp = np.zeros(256, np.uint8); comp[list(b"ACGTN")] = list(b"TGCAN")
    return comp[a[:, ::-1]]
for name, arr in (("f1", w["s1"]), ("b1", rc(w["s1"])), ("f2", w["s2"]), ("b2", rc(w["s2"]))):
    cnt[name] = m.get_best_chains(0, arr)["chain_count"].astype(np.int64)
cand = cnt["f1"] * cnt["b2"] + cnt["f2"] * cnt["b1"]
print("pairs", n, "sum candidates", cand.sum(), "mean", cand.mean())
for th in (1, 2, 4, 8, 16, 64, 256, 512, 1024):
    sel = cand > th
    print(f"cand > {th:5d}: {sel.mean()*100:6.2f}% of pairs, {cand[sel].sum()/cand.sum()*100:6.2f}% of candidates")
mr = fresh_mreads(n)
out, st = m.process_read(0, w["s1"], w["s2"], mr)
types = np.bincount(out["type"], minlength=14)
print("types", types)
for th in (16, 256):
    sel = cand > th
    print("types of pairs with cand >", th, np.bincount(out["type"][sel], minlength=14))
