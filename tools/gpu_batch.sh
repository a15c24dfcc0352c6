#!/bin/bash
# One gpurun call of the round-2 loop: GPU tests, bench line, eK sweep, launch list + full ncu capture of the top kernel.
# usage: tools/gpu_batch.sh [tag]   (outputs under gpurun_out/<tag>_*)
tag=${1:-b}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${tag}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -rP > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 3 --no-pipelines > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
tail -c 1500 gpurun_out/${tag}_bench.json
for ab in "1.15 8" "1.1 4" "1.35 12"; do
  set -- $ab
  PM_FAST_EK_A=$1 PM_FAST_EK_B=$2 timeout 600 python bench.py --steps 10 --warmup 3 --no-pipelines --no-cpu-baseline --no-stage-extras \
     > gpurun_out/${tag}_bench_ek_$1_$2.json 2>> gpurun_out/${tag}_bench.err
done
timeout 600 python bench.py --steps 3 --warmup 3 --no-pipelines --no-cpu-baseline --no-stage-extras > gpurun_out/${tag}_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py --steps 3 --warmup 3 --no-pipelines --no-cpu-baseline --no-stage-extras > gpurun_out/${tag}_ncu_l.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_knn_fast -s 3 -c 1 -o gpurun_out/${tag}_knn_fast \
  python bench.py --steps 3 --warmup 3 --no-pipelines --no-cpu-baseline --no-stage-extras > gpurun_out/${tag}_ncu_f.log 2>&1
ls -la gpurun_out | tail -20
