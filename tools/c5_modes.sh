#!/bin/bash
# experiment: radius-gather variants on the C5 hit points. usage: tools/c5_modes.sh <tag>
tag=$1
o=gpurun_out/${tag}_c5modes.log
python -m pytest tests/test_gpu_kdknn.py tests/test_ppm.py -q -x -k "radius or ppm" > gpurun_out/${tag}_pytest.log 2>&1
tail -3 gpurun_out/${tag}_pytest.log > $o
for ranks in 1 2 4 8; do
  echo "== warp" >> $o
  PM_RADIUS_MODE=0 C5_FRACS=0.25 python tools/c5_cellfrac.py $ranks 2>&1 | grep ranks >> $o
  echo "== thread" >> $o
  PM_RADIUS_DENSE=1000000 C5_FRACS=0.25,0.5 python tools/c5_cellfrac.py $ranks 2>&1 | grep ranks >> $o
  echo "== auto" >> $o
  C5_FRACS=0.25 python tools/c5_cellfrac.py $ranks 2>&1 | grep ranks >> $o
done
