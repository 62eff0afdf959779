#!/bin/bash
# multinomial epilogue + 88-register forward kernel: all GPU tests, then config 4 and config 2 lines
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/m_pytest.log
tail -4 gpurun_out/m_pytest.log
timeout 900 python bench.py --config 4 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/m_bench_cfg4.json 2> gpurun_out/m_bench_cfg4.err; echo "bench cfg4 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/m_bench_cfg4.json')); print('cfg4 ms %.1f frac %.3f'%(d['ms_per_step'], d['roofline']['frac']))"
for r in 1 2; do
timeout 600 python bench.py --config 2 --steps 100 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/m_bench_cfg2_$r.json 2> gpurun_out/m_bench_cfg2_$r.err; echo "bench cfg2 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/m_bench_cfg2_$r.json')); print('cfg2 ms %.3f frac %.3f parity %s'%(d['ms_per_step'], d['roofline']['frac'], d.get('parity')))"
done
