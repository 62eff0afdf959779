#!/bin/bash
# Round-2 GPU call A (one B200): tests, bench lines of every config, reference arm, ncu launch list + full captures.
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
nvidia-smi --query-gpu=name,power.limit,clocks.max.sm --format=csv > gpurun_out/a_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
timeout 600 python bench.py > gpurun_out/a_bench_cfg2.json 2> gpurun_out/a_bench_cfg2.err; echo "bench cfg2 rc=$?"
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/a_bench_ref.json 2> gpurun_out/a_bench_ref.err; echo "bench ref rc=$?"
for c in 3 3c 4 5; do
  timeout 900 python bench.py --config $c --steps 3 --warmup 3 > gpurun_out/a_bench_cfg$c.json 2> gpurun_out/a_bench_cfg$c.err; echo "bench cfg$c rc=$?"
done
# ncu: launch list of the default command, then full captures of one whole sweep (config 2) and of the Ridge kernels
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/a_launches_cfg2.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/a_ncu1.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:zero_round_kernel|tc_fwd_kernel|tc_grad2_kernel" -c 22 \
  -o gpurun_out/a_full_cfg2 -f python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/a_ncu2.log 2>&1; echo "ncu full cfg2 rc=$?"
ncu -i gpurun_out/a_full_cfg2.ncu-rep --page raw --csv > gpurun_out/a_full_cfg2.csv 2>/dev/null
ncu -i gpurun_out/a_full_cfg2.ncu-rep --page source --csv --kernel-name regex:tc_fwd_kernel --launch-skip 2 --launch-count 1 > gpurun_out/a_src_fwd.csv 2>/dev/null
ncu -i gpurun_out/a_full_cfg2.ncu-rep --page source --csv --kernel-name regex:tc_grad2_kernel --launch-skip 2 --launch-count 1 > gpurun_out/a_src_grad.csv 2>/dev/null
rm -f gpurun_out/a_full_cfg2.ncu-rep
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:tc_gram_kernel|xty_kernel|tc_fwd_kernel|ridge_cholesky" -c 5 \
  -o gpurun_out/a_full_cfg3 -f python bench.py --config 3 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/a_ncu3.log 2>&1; echo "ncu full cfg3 rc=$?"
ncu -i gpurun_out/a_full_cfg3.ncu-rep --page raw --csv > gpurun_out/a_full_cfg3.csv 2>/dev/null
rm -f gpurun_out/a_full_cfg3.ncu-rep
ls -la gpurun_out | tail -30
# config 4: one full-width multinomial forward + gradient launch
timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:tc_fwd_kernel|tc_grad2_kernel" -s 30 -c 2 \
  -o gpurun_out/a_full_cfg4 -f python bench.py --config 4 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/a_ncu4.log 2>&1; echo "ncu full cfg4 rc=$?"
ncu -i gpurun_out/a_full_cfg4.ncu-rep --page raw --csv > gpurun_out/a_full_cfg4.csv 2>/dev/null
ncu -i gpurun_out/a_full_cfg4.ncu-rep --page source --csv --kernel-name regex:tc_fwd_kernel > gpurun_out/a_src_fwd_cfg4.csv 2>/dev/null
rm -f gpurun_out/a_full_cfg4.ncu-rep
ls -la gpurun_out | grep " a_"
