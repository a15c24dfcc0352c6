#!/bin/bash
# on the GPU box: kNN / gather tests + a short C2 bench line; prints the figures. usage: bash tools/q.sh [tag] [pytest -k expr]
tag=${1:-q}; mkdir -p gpurun_out
python -m pytest tests/test_gpu_kdknn.py tests/test_zz_no_alloc.py -m gpu -x -q ${2:+-k "$2"} > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -n 2 gpurun_out/${tag}_pytest.log
python bench.py --steps 20 --warmup 3 --no-pipelines --no-cpu-baseline --no-stage-extras > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
python - <<PY
import json
j=json.loads(open('gpurun_out/${tag}_bench.json').read().strip().splitlines()[-1]); r=j['roofline']
print('ms_per_step',round(j['ms_per_step'],4),'Mq/s',round(j['value']/1e6,1),'frac',round(r['frac'],4),'kernel_ms',round(r['kernel_ms'],4),'ho',r['general_grid_kernel_queries'],'fb',r['kd_tree_fallback_queries'],'build',round(j['stage2_build']['ms'],3),'e2e',round(j['e2e']['value']/1e6,1), 'e2e_gather', round(j.get('e2e_gather',{}).get('value',0)/1e6,1))
PY
