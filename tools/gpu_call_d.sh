#!/bin/bash
# Round-2 GPU call D (one B200): Ridge after the Cholesky rewrite, scoring-by-group A/B on config 2.
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests -m gpu -q -s -k "ridge or Ridge or linear_regression or refit or api_end_to_end or holdout or config_shapes" > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest.log
tail -4 gpurun_out/d_pytest.log
for c in 3 3c; do
  timeout 900 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/d_bench_cfg$c.json 2> gpurun_out/d_bench_cfg$c.err; echo "bench cfg$c rc=$?"
done
for v in ; do
  DML_SCORE_BY_GROUP=$v timeout 600 python bench.py --steps 100 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import sys,json; d=json.load(sys.stdin); print('by_group=$v ms_per_step %.3f value %.1f'%(d['ms_per_step'], d['value']))" >> gpurun_out/d_ab_score.log
done
cat gpurun_out/d_ab_score.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/d_launches_cfg3.csv \
  python bench.py --config 3 --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/d_ncu3.log 2>&1; echo "ncu launches cfg3 rc=$?"
