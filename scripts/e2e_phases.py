"""Where the end-to-end time of one batch call goes: stage (H2D + strand prep), kernels, fetch (D2H)."""
import os, sys, time, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from circminer_b200 import workload
from circminer_b200.api import Mapper, default_params, fresh_mreads
w = workload.prepare("C2", "/tmp/cm_bench/C2_r0")
m = Mapper(default_params()); workload.load_into(m, w)
n = len(w["s1"])
def pinned(arr):
    t = torch.empty(arr.nbytes, dtype=torch.uint8, pin_memory=True)
    v = t.numpy().view(arr.dtype).reshape(arr.shape); v[...] = arr
    return t, v
keep = [pinned(w["s1"]), pinned(w["s2"]), pinned(fresh_mreads(n))]
s1, s2, mr = keep[0][1], keep[1][1], keep[2][1]
for rep in range(4):
    t0 = time.time(); m.stage(s1, s2, mr); t1 = time.time(); m.run_staged(0, 3); t2 = time.time(); out = m.fetch_staged(); t3 = time.time()
    o2, st = m.process_read(0, s1, s2, mr); t4 = time.time()
    print(f"stage {1e3*(t1-t0):.1f} ms  run {1e3*(t2-t1):.1f} ms  fetch {1e3*(t3-t2):.1f} ms | process_read {1e3*(t4-t3):.1f} ms  device {m.stage_ms()}", flush=True)
m.close()
