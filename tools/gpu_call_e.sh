#!/bin/bash
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 600 python __graft_entry__.py smoke > gpurun_out/e_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/e_smoke.log
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:ridge_cholesky_blocked" -c 1 \
  -o gpurun_out/e_chol -f python bench.py --config 3 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/e_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/e_chol.ncu-rep --page raw --csv > gpurun_out/e_chol_raw.csv 2>/dev/null
ncu -i gpurun_out/e_chol.ncu-rep --page source --csv > gpurun_out/e_chol_src.csv 2>/dev/null
rm -f gpurun_out/e_chol.ncu-rep
