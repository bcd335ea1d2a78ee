"""A/B of the seed look-up kernel on contig 0 of a GRCh38-scale workload: one Mapper per environment variant (the directory
layout switches are read when the contig is built).  CM_GPU_LIB picks an alternative build of the library."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from circminer_b200 import workload
    from circminer_b200.api import Mapper, fresh_mreads
    name, n = "C3", 1_000_000
    w = workload.prepare(name, f"/tmp/cm_bench/{name}_r0", n_pairs=n)
    seq = np.ascontiguousarray(w["contigs"][0])
    for var in sys.argv[1:] or [""]:
        kv = dict(x.split("=", 1) for x in var.split(",") if x)
        os.environ.update(kv)
        m = Mapper(workload.params_for(name))
        m.build_contig(0, seq)
        for k in kv:
            os.environ.pop(k, None)
        m.stage(w["s1"], w["s2"], fresh_mreads(n))
        ms = []
        for _ in range(6):
            m.run_staged(0, 1)
            ms.append(m.stage_ms()[1])
        print(json.dumps({"lib": os.environ.get("CM_GPU_LIB", "default"), "variant": var, "seed_ms": ms[1:], "best": min(ms[1:])}), flush=True)
        m.close()


if __name__ == "__main__":
    main()
