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


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on the box")
def test_row_shard_blocks_through_master_and_workers_match_one_gpu():
    """partition="rows" behind the subtask contract: the master emits one block per GPU naming its row shard,
    each worker uploads only its rows, the workers' contexts join one NCCL communicator (ids exchanged through the
    scheduler) and every sweep all-reduces per round.  Result = the one-GPU job's."""
    import numpy as np
    from sklearn.linear_model import LogisticRegression, Ridge
    from sklearn.model_selection import GridSearchCV
    import cs230_distributed_machine_learning_b200 as pkg
    from cs230_distributed_machine_learning_b200.client import LocalBackend, MLTaskManager
    from tests.synth import make_dataset
    n_gpus = min(4, torch.cuda.device_count())
    X, y = make_dataset("binary", 50_000, 64, seed=21)
    pkg.register_dataset("rows-bin", X, y)
    Xr, yr = make_dataset("regression", 50_000, 64, seed=22)
    pkg.register_dataset("rows-reg", Xr, yr)
    est = GridSearchCV(LogisticRegression(), {"C": [float(c) for c in np.logspace(-4, 1, 6)]}, cv=5)
    estw = GridSearchCV(LogisticRegression(class_weight="balanced"), {"C": [0.01, 1.0]}, cv=5, return_train_score=True)
    estr = GridSearchCV(Ridge(), {"alpha": [0.1, 10.0, 1000.0]}, cv=5)
    got = {}
    for label, gpus, part in (("one", [0], "candidates"), ("rows", list(range(n_gpus)), "rows"), ("cand", list(range(n_gpus)), "candidates")):
        m = MLTaskManager(LocalBackend(gpus=gpus, partition=part, monitor_interval=None))
        try:
            r = m.train(est, "rows-bin", train_params={"test_size": 0.2, "random_state": 3}, wait_for_completion=True)
            assert r["job_status"] == "completed", r
            rr = m.train(estr, "rows-reg", wait_for_completion=True)
            assert rr["job_status"] == "completed", rr
            rw = m.train(estw, "rows-bin", wait_for_completion=True)      # class weights + training scores, sharded
            assert rw["job_status"] == "completed", rw
            got[label] = (r["job_result"], rr["job_result"], rw["job_result"])
        finally:
            m.close()
    one, rows, cand = got["one"], got["rows"], got["cand"]
    for label, (a, b, c) in got.items():       # a failed candidate carries its error: show it instead of a NaN score
        assert not [x["error"] for x in a["results"] + b["results"] + c["results"] if x.get("error")], label
    assert all(x["executor"] == "b200" and x["row_shards"] == n_gpus for x in rows[0]["results"])
    for other in (rows, cand):
        assert other[0]["cv_results_"]["mean_test_score"] == one[0]["cv_results_"]["mean_test_score"]     # counts are exact
        assert other[0]["best_params_"] == one[0]["best_params_"]
        assert [x["n_iter"] for x in other[0]["results"]] == [x["n_iter"] for x in one[0]["results"]]
        assert [x["accuracy"] for x in other[0]["results"]] == [x["accuracy"] for x in one[0]["results"]]
        np.testing.assert_allclose(other[1]["cv_results_"]["mean_test_score"], one[1]["cv_results_"]["mean_test_score"], rtol=1e-6)
        assert other[1]["best_params_"] == one[1]["best_params_"]
        assert other[2]["cv_results_"]["mean_test_score"] == one[2]["cv_results_"]["mean_test_score"]
        np.testing.assert_allclose(other[2]["cv_results_"]["mean_train_score"], one[2]["cv_results_"]["mean_train_score"], atol=1e-6)
    gsw = GridSearchCV(LogisticRegression(class_weight="balanced"), {"C": [0.01, 1.0]}, cv=5, return_train_score=True).fit(X.astype(np.float64), y)
    np.testing.assert_allclose(rows[2]["cv_results_"]["mean_test_score"], gsw.cv_results_["mean_test_score"], atol=1e-4)
    np.testing.assert_allclose(rows[2]["cv_results_"]["mean_train_score"], gsw.cv_results_["mean_train_score"], atol=1e-4)
