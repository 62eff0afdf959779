"""Multi-GPU parity on one box (needs >= 2 GPUs): row shards + NCCL all-reduce == single GPU."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on the box")
def test_row_sharded_sweeps_match_single_gpu():
    n = min(4, torch.cuda.device_count())
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29641", os.path.join(ROOT, "tests", "dist_gpu_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    line = next(l for l in res.stdout.splitlines() if l.startswith("DIST_CHECK "))
    out = json.loads(line[len("DIST_CHECK "):])
    for tc in (0, 1):
        r = out[f"logreg_tc{tc}"]
        assert r["n_test_equal"] and r["n_iter_equal"], out
        assert r["coef_relerr"] <= 1e-4, out      # different summation order across shards
        assert r["counts_equal"], out
    assert out["ridge"]["coef_relerr"] <= 1e-5 and out["ridge"]["sse_relerr"] <= 1e-5 and out["ridge"]["n_test_equal"]
    assert out["ranks_identical"]
