#!/bin/bash
# build, quick GPU batch (tests + bench + ncu of k_knn_fast), print the figures. usage: tools/quick.sh <tag>
cd /root/repo
tag=$1
(cd photon-mapping_b200/csrc && make -j8 2>&1 | grep -v "^nvcc\|^mkdir\|warning" | head -5)
tools/gpurun_retry.sh 2000 gpurun_out/${tag}_call.log 1 "bash tools/gpu_batch.sh ${tag} quick"
tail -2 gpurun_out/${tag}_call.log
python - <<PY
import json
t=open('/root/repo/gpurun_out/${tag}_bench.json').read().strip()
if t:
    j=json.loads(t.splitlines()[-1]); r=j['roofline']
    print('ms_per_step',j['ms_per_step'],'value',j['value']/1e6,'frac',r['frac'],'kernel_ms',r['kernel_ms'],'ho',r['general_grid_kernel_queries'],'fb',r['kd_tree_fallback_queries'], 'build', j['stage2_build']['ms'])
PY
tail -2 gpurun_out/${tag}_pytest.log
python tools/summarize_profile.py ${tag} nonexistent.csv gpurun_out/${tag}_knn_step.ncu-rep 2>&1 | grep -v "^smsp__average" | head -22
