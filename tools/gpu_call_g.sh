#!/bin/bash
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
N=${1:-8}
nvidia-smi -L > gpurun_out/g_smi.txt 2>&1
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus $N --steps 20 --warmup 3 \
  > gpurun_out/g_bench_n$N.json 2> gpurun_out/g_bench_n$N.err; echo "bench n$N rc=$?"
tail -c 500 gpurun_out/g_bench_n$N.err
head -c 600 gpurun_out/g_bench_n$N.json
