#!/bin/bash
# config 5 against the full-size cache: the bench line with its parity record
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1200 python bench.py --config 5 --steps 3 --warmup 3 > gpurun_out/r_bench_cfg5.json 2> gpurun_out/r_bench_cfg5.err; echo "bench cfg5 rc=$?"
tail -c 400 gpurun_out/r_bench_cfg5.err
python -c "import json; d=json.load(open('gpurun_out/r_bench_cfg5.json')); print('cfg5 ms %.1f frac %.3f'%(d['ms_per_step'], d['roofline']['frac'])); print(d.get('parity'))"
