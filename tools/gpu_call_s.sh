#!/bin/bash
# Gram kernel, diagonal blocks: Ridge tests, then configs 3 / 3c
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -q -s -k "ridge or config3 or linear or collinear or rank_deficient or scoring" > gpurun_out/s_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s_pytest.log
tail -3 gpurun_out/s_pytest.log
for c in 3 3c; do
  timeout 600 python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/s_bench_cfg$c.json 2> gpurun_out/s_bench_cfg$c.err; echo "bench cfg$c rc=$?"
  python -c "import json; d=json.load(open('gpurun_out/s_bench_cfg$c.json')); print('cfg$c ms %.3f frac %.3f'%(d['ms_per_step'], d['roofline']['frac']), d.get('parity'))"
done
