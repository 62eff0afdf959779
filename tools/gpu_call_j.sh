#!/bin/bash
# Round-2 GPU call J (one B200): config-3 full-size parity, ncu of the multinomial kernels at config 4.
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "config3_full_size" > gpurun_out/j_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j_pytest.log
tail -4 gpurun_out/j_pytest.log
timeout 900 python bench.py --config 3 --steps 5 --warmup 3 > gpurun_out/j_bench_cfg3.json 2> gpurun_out/j_bench_cfg3.err; echo "bench cfg3 rc=$?"
timeout 1500 ncu --set full --clock-control none -k "regex:tc_fwd_kernel|tc_grad2_kernel" -s 2 -c 4 \
  -o gpurun_out/j_full_cfg4 -f python bench.py --config 4 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/j_ncu4.log 2>&1; echo "ncu cfg4 rc=$?"
ncu -i gpurun_out/j_full_cfg4.ncu-rep --page raw --csv > gpurun_out/j_full_cfg4.csv 2>/dev/null
rm -f gpurun_out/j_full_cfg4.ncu-rep
ls -la gpurun_out/j_*
