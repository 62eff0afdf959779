#!/bin/bash
# default bench line of the final tree
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
SECONDS=0
timeout 900 python bench.py > gpurun_out/x_bench.json 2> gpurun_out/x_bench.err; echo "bench rc=$? in ${SECONDS}s"
python -c "import json; d=json.load(open('gpurun_out/x_bench.json')); print('cfg2 ms %.3f value %.0f e2e %.0f frac %.3f'%(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac']), d['parity']['best_index_equal'], d['clocks']['reasons'], d['cpu_baseline']['value'])"
