#!/bin/bash
# One gpurun call of the round-2 loop: GPU tests, bench line, launch list + full ncu captures.
# usage: tools/gpu_batch.sh [tag] [quick]   (outputs under gpurun_out/<tag>_*)
tag=${1:-b}
quick=${2:-}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${tag}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -rP > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
B="python bench.py --steps 3 --warmup 3 --no-pipelines --no-cpu-baseline --no-stage-extras"
if [ -n "$quick" ]; then
  timeout 600 python bench.py --steps 20 --warmup 3 --no-pipelines --no-cpu-baseline --no-stage-extras > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
  tail -c 1200 gpurun_out/${tag}_bench.json; tail -5 gpurun_out/${tag}_bench.err
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_knn_fast' -s 4 -c 1 \
    -o gpurun_out/${tag}_knn_step $B > gpurun_out/${tag}_ncu_f.log 2>&1
  du -sh gpurun_out; exit 0
fi
timeout 1200 python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"
tail -c 600 gpurun_out/${tag}_bench.json; tail -5 gpurun_out/${tag}_bench.err
timeout 600 $B > gpurun_out/${tag}_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_launches.csv $B > gpurun_out/${tag}_ncu_l.log 2>&1
# one warm pm_knn step: its six kernels (3 binning + fast + general + kd), skipping the 3 builds' tuning samples and 3 warm-up steps
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_qbin|k_knn_fast|k_grid_query|k_kdw' -s 28 -c 6 \
  -o gpurun_out/${tag}_knn_step $B > gpurun_out/${tag}_ncu_f.log 2>&1
timeout 300 python tools/prof_emit.py cornell 4000000 3 > gpurun_out/${tag}_plain_emit.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_walk -s 1 -c 1 -o gpurun_out/${tag}_walk \
  python tools/prof_emit.py cornell 4000000 3 > gpurun_out/${tag}_ncu_w.log 2>&1
timeout 300 python tools/prof_emit.py big 16000000 2 > gpurun_out/${tag}_plain_emit_big.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_walk -s 1 -c 1 -o gpurun_out/${tag}_walk_big \
  python tools/prof_emit.py big 16000000 2 > gpurun_out/${tag}_ncu_wb.log 2>&1
timeout 300 python tools/prof_build.py 10000000 2 > gpurun_out/${tag}_plain_build.log 2>&1 && \
timeout 1500 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed.avg.per_cycle_active,l1tex__throughput.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread \
  --clock-control none -o gpurun_out/${tag}_build python tools/prof_build.py 10000000 1 > gpurun_out/${tag}_ncu_b.log 2>&1
du -sh gpurun_out
ls -la gpurun_out | tail -30
