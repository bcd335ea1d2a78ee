#!/bin/bash
# compare the pair work-signature variants (CM_GPU_PAIR_SIG = 0 / 1 / 2) on the bench workload
for m in 0 1 2; do
  echo "== CM_GPU_PAIR_SIG=$m"
  CM_GPU_PAIR_SIG=$m python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d.get('stages_ms'), d['e2e']['value'])"
done
