#!/bin/bash
# config 4 pinned at full size: the new GPU test, then the bench line with its parity record
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -q -s -k "config4_full" > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/q_pytest.log
tail -5 gpurun_out/q_pytest.log | cut -c1-1500
timeout 900 python bench.py --config 4 --steps 3 --warmup 3 > gpurun_out/q_bench_cfg4.json 2> gpurun_out/q_bench_cfg4.err; echo "bench cfg4 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/q_bench_cfg4.json')); print('cfg4 ms %.1f frac %.3f'%(d['ms_per_step'], d['roofline']['frac'])); print(d.get('parity'))"
