#!/bin/bash
# final check on a 2-GPU box: the whole GPU suite (the multi-GPU tests run here), smoke(), the N = 2 bench line
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
nvidia-smi -L > gpurun_out/p_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/p_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/p_pytest.log
tail -4 gpurun_out/p_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/p_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/p_smoke.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 --steps 20 --warmup 3 \
  > gpurun_out/p_bench_n2.json 2> gpurun_out/p_bench_n2.err; echo "bench n2 rc=$?"
tail -c 300 gpurun_out/p_bench_n2.err
head -c 400 gpurun_out/p_bench_n2.json
