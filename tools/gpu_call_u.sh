#!/bin/bash
# the driver's GPU test command, three times in a row, with the worker's log shown on failure
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
for r in 1 2 3; do
  timeout 900 python -m pytest tests/ -x -q -m gpu -o log_cli=false > gpurun_out/u_pytest_$r.log 2>&1; rc=$?
  echo "run $r rc=$rc: $(tail -1 gpurun_out/u_pytest_$r.log)"
  if [ $rc -ne 0 ]; then grep -n "error\|Error\|failed" gpurun_out/u_pytest_$r.log | head -20; fi
done
