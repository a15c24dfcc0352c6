#!/bin/bash
# bench.py over N ranks of one box (the driver's launch line). usage: tools/gpu_multi.sh N tag [extra bench args]
n=$1; tag=$2; shift 2
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/${tag}_topo.txt 2>&1
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29531 \
  bench.py --gpus $n --steps 10 --warmup 3 "$@" > gpurun_out/${tag}_bench_${n}gpu.json 2> gpurun_out/${tag}_bench_${n}gpu.err
echo "exit $?"; tail -c 3000 gpurun_out/${tag}_bench_${n}gpu.json; tail -20 gpurun_out/${tag}_bench_${n}gpu.err
