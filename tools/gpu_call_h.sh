#!/bin/bash
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/h_pytest.log
tail -3 gpurun_out/h_pytest.log
for c in 2 3 3c; do
  timeout 900 python bench.py --config $c --steps $([ $c = 2 ] && echo 100 || echo 5) --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/h_bench_cfg$c.json 2> gpurun_out/h_bench_cfg$c.err; echo "bench cfg$c rc=$?"
  python -c "import json; d=json.load(open('gpurun_out/h_bench_cfg$c.json')); print('cfg$c ms %.3f share %.3f'%(d['ms_per_step'], d['roofline']['kernel_share_of_step']))"
done
