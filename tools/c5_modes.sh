#!/bin/bash
tag=$1
o=gpurun_out/${tag}_c5modes.log
: > $o
for ranks in 1 8; do
  C5_FRACS=0.25 timeout 300 python tools/c5_cellfrac.py $ranks 2>&1 | grep ranks >> $o
done
