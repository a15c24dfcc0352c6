#!/bin/bash
# first-radius sweep of the fast path: expected candidates a*k + b (PM_FAST_EK_A / _B)
mkdir -p gpurun_out
for ab in "1.25 10" "1.2 8" "1.15 8" "1.15 4" "1.1 6"; do
  set -- $ab
  PM_FAST_EK_A=$1 PM_FAST_EK_B=$2 timeout 600 python bench.py --steps 20 --warmup 3 --no-pipelines --no-cpu-baseline --no-stage-extras 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=j['roofline']
print('ek', '$1', '$2', 'step_ms', round(j['ms_per_step'],4), 'kernel_ms', round(r['kernel_ms'],4), 'handovers', r['general_grid_kernel_queries'], 'e2e_gather', round(j['e2e_gather']['value']/1e6,1))"
done
