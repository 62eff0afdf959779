#!/bin/bash
# last sanity of the rebuilt library: smoke() and the fixture sweeps
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 150 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "fixture or class_weight_vs" > gpurun_out/z_pytest.log 2>&1; echo "rc=$? $(tail -1 gpurun_out/z_pytest.log)"
