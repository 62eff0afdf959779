"""bench.py's host-side pieces (no GPU): candidate lists of the BASELINE.json configs, the parity record."""
import json
import os

import numpy as np

import bench
from cs230_distributed_machine_learning_b200 import master
from cs230_distributed_machine_learning_b200.client import MLTaskManager, wire


def test_config4_candidates_are_the_masters_randomized_search_draw():
    """Config 4 = RandomizedSearchCV over list-valued distributions: the candidates bench.py sweeps must be the ones
    the master draws for that job (ParameterSampler, seed 42, task_handler.py:213), in draw order."""
    from sklearn.linear_model import LogisticRegression
    from sklearn.model_selection import RandomizedSearchCV
    cands = bench.candidates_of(bench.CONFIGS["4"], 512)
    assert len(cands) == 512 and len({json.dumps(c, sort_keys=True) for c in cands}) == 512     # without replacement
    dists = {"C": [float(v) for v in np.logspace(-5, 2, 64)], "tol": [float(v) for v in np.logspace(-5, -3, 8)],
             "fit_intercept": [True, False]}
    est = RandomizedSearchCV(LogisticRegression(), dists, n_iter=512, cv=5)
    details = MLTaskManager._extract_model_details(None, est)
    payload = wire({"job_id": "j", "session_id": "s", "dataset_id": "d", "model_details": details, "train_params": {}})
    subtasks = master.create_subtasks(payload)
    assert [{k: st["parameters"][k] for k in ("C", "tol", "fit_intercept")} for st in subtasks] == \
           [{k: c[k] for k in ("C", "tol", "fit_intercept")} for c in cands]


def test_config_candidate_grids():
    c2 = bench.candidates_of(bench.CONFIGS["2"], 64)
    assert [c["C"] for c in c2] == [float(v) for v in np.logspace(-6, 2, 64)]
    c3 = bench.candidates_of(bench.CONFIGS["3"], 128)
    assert [c["alpha"] for c in c3] == [float(v) for v in np.logspace(-3, 5, 128)]
    assert bench.CONFIGS["5"]["cv"] == 10 and bench.CONFIGS["5"]["n"] == 16_000_000


def test_parity_report_counts_what_differs(tmp_path):
    rng = np.random.default_rng(0)
    n, d, n_cand, cv = 400, 6, 3, 2
    X = rng.standard_normal((n, d)).astype(np.float32)
    y = (X[:, 0] > 0).astype(np.int32)
    fold = (np.arange(n) % cv).astype(np.int8)
    coef = rng.standard_normal((n_cand, cv, d)); icpt = rng.standard_normal((n_cand, cv)) * 0.1
    correct = np.zeros((n_cand, cv), np.int64); n_test = np.zeros((n_cand, cv), np.int64)
    for i in range(n_cand):
        for k in range(cv):
            rows = fold == k
            z = X[rows].astype(np.float64) @ coef[i, k] + icpt[i, k]
            correct[i, k] = ((z > 0) == y[rows]).sum(); n_test[i, k] = rows.sum()
    scores = correct / n_test
    golden = tmp_path / "g.npz"
    np.savez(golden, coef=coef, intercept=icpt, n_iter=np.full((n_cand, cv), 7), nfev=np.full((n_cand, cv), 8),
             correct=correct, n_test=n_test, mean_test_score=scores.mean(axis=1), best_index=np.int64(bench.search_best_index(scores)))
    out = dict(coef=coef[:, :, None, :].astype(np.float32), intercept=icpt[:, :, None].astype(np.float32),
               n_iter=np.full((n_cand, cv), 7), nfev=np.full((n_cand, cv), 8), correct=correct.copy(), n_test=n_test)
    rep = bench.parity_report(out, X, y, fold, cv, str(golden))
    assert rep["n_iter_equal"] == 6 and rep["count_diffs"] == 0 and rep["best_index_equal"] and rep["max_coef_relerr"] < 1e-6
    out["correct"][1, 0] += 5          # five rows off, none of them a tie: must be reported, not excused
    out["n_iter"][2, 1] = 9
    rep = bench.parity_report(out, X, y, fold, cv, str(golden))
    assert rep["count_diffs"] == 1 and rep["max_count_diff"] == 5 and rep["count_diffs_beyond_ties"] == 1
    assert rep["n_iter_equal"] == 5
    assert "unavailable" in bench.parity_report(out, X, y, fold, cv, str(tmp_path / "missing.npz"))


def test_committed_golden_file_is_the_headline_config():
    z = np.load(bench.GOLDEN_CFG2)
    meta = json.loads(str(z["meta"]))
    cfg = bench.CONFIGS["2"]
    assert (meta["n"], meta["d"], meta["cv"], meta["seed"]) == (cfg["n"], cfg["d"], cfg["cv"], cfg["seed"])
    assert z["coef"].shape == (64, 5, 256) and z["n_iter"].shape == (64, 5) and int(z["best_index"]) == 21
    assert meta["C"] == [c["C"] for c in bench.candidates_of(cfg, 64)]
    assert bench.golden_nfev_sum() == int(z["nfev"].sum())
