#!/bin/bash
# Round-2 GPU call C (one B200): tests + bench lines after the scoring-by-group / Cholesky changes.
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c_pytest.log
tail -6 gpurun_out/c_pytest.log
for c in 2 3 3c 5; do
  timeout 900 python bench.py --config $c --steps $([ $c = 2 ] && echo 20 || echo 3) --warmup 3 --no-cpu-baseline > gpurun_out/c_bench_cfg$c.json 2> gpurun_out/c_bench_cfg$c.err; echo "bench cfg$c rc=$?"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c_launches_cfg3.csv \
  python bench.py --config 3 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/c_ncu3.log 2>&1; echo "ncu launches cfg3 rc=$?"
ls -la gpurun_out | tail -12
