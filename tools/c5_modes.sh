#!/bin/bash
# pm_gather_radius on the C5 hit points for the photon share of 1, 2, 4, 8 ranks. usage: tools/c5_modes.sh <tag>
tag=$1
o=gpurun_out/${tag}_c5modes.log
python -m pytest tests/test_gpu_kdknn.py tests/test_ppm.py tests/test_gpu_multirank.py -q -x -k "radius or ppm" > gpurun_out/${tag}_pytest.log 2>&1
tail -3 gpurun_out/${tag}_pytest.log > $o
for ranks in 1 2 4 8; do
  C5_FRACS=0.25 python tools/c5_cellfrac.py $ranks 2>&1 | grep ranks >> $o
done
