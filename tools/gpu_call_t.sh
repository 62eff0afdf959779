#!/bin/bash
# the driver's round-end sequence on one GPU: GPU tests, smoke(), default bench line, reference arm
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/t_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_pytest.log
tail -3 gpurun_out/t_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/t_smoke.log 2>&1; echo "smoke rc=$?"
/usr/bin/time -v timeout 900 python bench.py > gpurun_out/t_bench.json 2> gpurun_out/t_bench.err; echo "bench rc=$?"
grep -E "Elapsed|Maximum resident" gpurun_out/t_bench.err
python -c "import json; d=json.load(open('gpurun_out/t_bench.json')); print('cfg2 ms %.3f value %.0f e2e %.0f frac %.3f'%(d['ms_per_step'], d['value'], d['e2e']['value'], d['roofline']['frac']), d['parity']['best_index_equal'], d['clocks'])"
