#!/bin/bash
# the driver's GPU test command (host page-lock now ends with the array)
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/ -x -q -m gpu > gpurun_out/w_pytest_1.log 2>&1; rc=$?
echo "run 1 rc=$rc: $(tail -1 gpurun_out/w_pytest_1.log)"
if [ $rc -ne 0 ]; then grep -n "Error\|FAILED" gpurun_out/w_pytest_1.log | head -10; fi
