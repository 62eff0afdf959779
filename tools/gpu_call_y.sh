#!/bin/bash
# the three tests that need a second GPU, on the final tree
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 150 python -m pytest tests/test_gpu_dist.py tests/test_gpu_parity.py -m gpu -q -k "row_shard or two_contexts or page_lock" > gpurun_out/y_pytest.log 2>&1; echo "rc=$? $(tail -1 gpurun_out/y_pytest.log)"
grep -n "Error\|FAILED" gpurun_out/y_pytest.log | head
