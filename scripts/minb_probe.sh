#!/bin/bash
# compare builds of the per-pair kernel with different launch bounds (libs under _variants/)
run() { python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['e2e']['value'])"; }
echo "== default"; run
for f in _variants/*.so; do echo "== $f"; CM_GPU_LIB=$PWD/$f run; done
