#!/bin/bash
# candidate-count threshold above which a pair goes to the warp-per-pair kernel, at the bench batch size
run() { python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['value'])"; }
echo "== default (off at this batch size)"; run
for t in ${TH_LIST:-32 64 128 256 512}; do echo "== CM_GPU_HEAVY_PAIR_TH=$t"; CM_GPU_HEAVY_PAIR_TH=$t run; done
