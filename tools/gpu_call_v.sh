#!/bin/bash
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 900 python -m pytest tests/ -x -q -m gpu -s > gpurun_out/v_pytest.log 2>&1; echo "rc=$? $(tail -1 gpurun_out/v_pytest.log)"
grep -n "dml_b200\]" gpurun_out/v_pytest.log | head -20
