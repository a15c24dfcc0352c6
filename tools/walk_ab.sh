#!/bin/bash
# stage-1 parity tests + A/B of the two walk kernels (PM_WALK_OLD=1: one path per lane)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_scene.py tests/test_gpu_multirank.py tests/test_ppm.py tests/test_gpu_host.py tests/test_view.py -m gpu -q -x 2>&1 | tail -5
for v in new old; do
  if [ $v = old ]; then export PM_WALK_OLD=1; else unset PM_WALK_OLD; fi
  echo "== walk $v: cornell 4M"; timeout 300 python tools/prof_emit.py cornell 4000000 3 2>&1 | tail -3
  echo "== walk $v: 1M-triangle room 16M"; timeout 300 python tools/prof_emit.py big 16000000 3 2>&1 | tail -3
done
