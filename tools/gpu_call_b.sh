#!/bin/bash
# Round-2 GPU call B (2+ GPUs of one box): multi-GPU tests, then the scaling bench line.
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
N=${1:-2}
nvidia-smi -L > gpurun_out/b_smi.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_dist.py "tests/test_gpu_parity.py::test_two_contexts_in_one_process_run_tensor_core_sweeps" -m gpu -q -s > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -8 gpurun_out/b_pytest.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $N --steps 20 --warmup 3 \
  > gpurun_out/b_bench_n$N.json 2> gpurun_out/b_bench_n$N.err; echo "bench n$N rc=$?"
tail -c 600 gpurun_out/b_bench_n$N.err
ls -la gpurun_out | tail -8
