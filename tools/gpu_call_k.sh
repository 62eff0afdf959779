#!/bin/bash
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 2400 ncu --set full --clock-control none --import-source on -k "regex:tc_fwd_kernel" -s 15 -c 1 \
  -o gpurun_out/k_src_cfg4 -f python bench.py --config 4 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/k_ncu4.log 2>&1; echo "ncu cfg4 rc=$?"
ncu -i gpurun_out/k_src_cfg4.ncu-rep --page source --csv > gpurun_out/k_src_cfg4.csv 2>/dev/null
rm -f gpurun_out/k_src_cfg4.ncu-rep
ls -la gpurun_out/k_*
