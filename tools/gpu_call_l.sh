#!/bin/bash
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -q -s -k "cfg4 or multiclass or ragged or config_shapes or class_weight or zz_no_alloc or no_alloc or static" > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/l_pytest.log
tail -4 gpurun_out/l_pytest.log
timeout 900 python bench.py --config 4 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/l_bench_cfg4.json 2> gpurun_out/l_bench_cfg4.err; echo "bench cfg4 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/l_bench_cfg4.json')); print('cfg4 ms %.1f frac %.3f'%(d['ms_per_step'], d['roofline']['frac']))"
