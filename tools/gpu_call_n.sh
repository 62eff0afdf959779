#!/bin/bash
# config 4: bench line, then one full-width forward + gradient launch under ncu (raw + source pages)
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python bench.py --config 4 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/n_bench_cfg4.json 2> gpurun_out/n_bench_cfg4.err; echo "bench cfg4 rc=$?"
python -c "import json; d=json.load(open('gpurun_out/n_bench_cfg4.json')); print('cfg4 ms %.1f frac %.3f'%(d['ms_per_step'], d['roofline']['frac']))"
timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:tc_fwd_kernel|tc_grad2_kernel" -s 30 -c 2 \
  -o gpurun_out/n_cfg4 -f python bench.py --config 4 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/n_ncu4.log 2>&1; echo "ncu cfg4 rc=$?"
ncu -i gpurun_out/n_cfg4.ncu-rep --page raw --csv > gpurun_out/n_full_cfg4.csv 2>/dev/null
ncu -i gpurun_out/n_cfg4.ncu-rep --page source --csv --kernel-name regex:tc_fwd_kernel > gpurun_out/n_src_fwd_cfg4.csv 2>/dev/null
rm -f gpurun_out/n_cfg4.ncu-rep
ls -la gpurun_out/n_*
