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


def test_parity_report_multi_counts_what_differs(tmp_path):
    """The config-4 record: a count may differ only on rows whose two largest decision values tie; anything else is
    reported as beyond ties; the picked candidates are looked up by their index in the full candidate list."""
    rng = np.random.default_rng(1)
    n, d, T, cv, pick = 600, 5, 4, 2, [1, 3]
    X = rng.standard_normal((n, d)).astype(np.float32)
    y = rng.integers(0, T, n).astype(np.int32)
    fold = (np.arange(n) % cv).astype(np.int8)
    coef = rng.standard_normal((len(pick), cv, T, d)); icpt = rng.standard_normal((len(pick), cv, T)) * 0.1
    correct = np.zeros((len(pick), cv), np.int64); n_test = np.zeros((len(pick), cv), np.int64)
    for i in range(len(pick)):
        for k in range(cv):
            rows = fold == k
            Z = X[rows].astype(np.float64) @ coef[i, k].T + icpt[i, k]
            correct[i, k] = (Z.argmax(axis=1) == y[rows]).sum(); n_test[i, k] = rows.sum()
    golden = tmp_path / "g4.npz"
    np.savez(golden, index=np.array(pick), coef=coef, intercept=icpt, n_iter=np.full((2, cv), 6), nfev=np.full((2, cv), 7),
             correct=correct, n_test=n_test, meta=np.array(json.dumps({"index": pick})))
    n_all = 5                                                   # the sweep holds 5 candidates; the golden ones are 1 and 3
    out = dict(coef=np.zeros((n_all, cv, T, d), np.float32), intercept=np.zeros((n_all, cv, T), np.float32),
               n_iter=np.full((n_all, cv), 6), nfev=np.full((n_all, cv), 7), correct=np.zeros((n_all, cv), np.int64),
               n_test=np.zeros((n_all, cv), np.int64))
    out["coef"][pick] = coef; out["intercept"][pick] = icpt; out["correct"][pick] = correct; out["n_test"][pick] = n_test
    rep = bench.parity_report_multi(out, X, fold, cv, str(golden))
    assert rep["fits_checked"] == 4 and rep["n_iter_equal"] == 4 and rep["nfev_equal"] == 4
    assert rep["count_diffs"] == 0 and rep["n_test_equal"] and rep["max_coef_relerr"] < 1e-6 and rep["best_of_checked_equal"]
    out["correct"][3, 1] -= 3; out["n_iter"][1, 0] = 9
    rep = bench.parity_report_multi(out, X, fold, cv, str(golden))
    assert rep["count_diffs"] == 1 and rep["max_count_diff"] == 3 and rep["count_diffs_beyond_ties"] == 1 and rep["n_iter_equal"] == 3
    sub = {k: v[pick] for k, v in out.items()}                  # the GPU test sweeps only the picked candidates
    assert bench.parity_report_multi(sub, X, fold, cv, str(golden), pos=range(2))["count_diffs"] == 1
    assert "unavailable" in bench.parity_report_multi(out, X, fold, cv, str(tmp_path / "missing.npz"))


def test_committed_config4_golden_file_is_the_bench_config():
    z = np.load(bench.GOLDEN_CFG4)
    meta = json.loads(str(z["meta"]))
    cfg = bench.CONFIGS["4"]
    assert (meta["kind"], meta["n"], meta["d"], meta["cv"], meta["seed"]) == (cfg["kind"], cfg["n"], cfg["d"], cfg["cv"], cfg["seed"])
    cands = bench.candidates_of(cfg, cfg["n_cand"])
    assert [cands[i] for i in z["index"]] == meta["params"] and meta["n_candidates"] == cfg["n_cand"]
    assert z["coef"].shape == (8, 5, 10, 512) and z["n_iter"].shape == (8, 5) and (z["n_test"].sum(axis=1) == cfg["n"]).all()
    assert (z["nfev"] >= z["n_iter"]).all() and (z["intercept"][[not p["fit_intercept"] for p in meta["params"]]] == 0).all()


def test_parity_report_cached_ties(tmp_path):
    """The config-5 record: the cache carries the near-tie row counts, so no X is needed on the host."""
    rng = np.random.default_rng(2)
    m, cv, d, pick = 2, 3, 7, [0, 4]
    coef = rng.standard_normal((m, cv, d)); icpt = rng.standard_normal((m, cv))
    correct = rng.integers(100, 200, (m, cv)); n_test = np.full((m, cv), 250); ties = np.array([[0, 2, 0], [1, 0, 0]])
    golden = tmp_path / "g5.npz"
    np.savez(golden, index=np.array(pick), coef=coef, intercept=icpt, n_iter=np.full((m, cv), 5), nfev=np.full((m, cv), 6),
             correct=correct, n_test=n_test, ties=ties)
    out = dict(coef=np.zeros((6, cv, 1, d), np.float32), intercept=np.zeros((6, cv, 1), np.float32), n_iter=np.full((6, cv), 5),
               nfev=np.full((6, cv), 6), correct=np.zeros((6, cv), np.int64), n_test=np.full((6, cv), 250))
    out["coef"][pick] = coef[:, :, None, :]; out["intercept"][pick] = icpt[:, :, None]; out["correct"][pick] = correct
    rep = bench.parity_report_cached_ties(out, str(golden))
    assert rep["fits_checked"] == 6 and rep["n_iter_equal"] == 6 and rep["count_diffs"] == 0 and rep["max_coef_relerr"] < 1e-6
    out["correct"][0, 1] += 2          # two rows, two near ties there: excused, but counted
    out["correct"][4, 2] -= 1          # one row where the oracle saw no tie: beyond ties
    rep = bench.parity_report_cached_ties(out, str(golden))
    assert rep["count_diffs"] == 2 and rep["rows_excused_as_ties"] == 2 and rep["count_diffs_beyond_ties"] == 1
    assert "unavailable" in bench.parity_report_cached_ties({k: v[:3] for k, v in out.items()}, str(golden))


def test_committed_config5_golden_file_is_the_bench_config():
    z = np.load(bench.GOLDEN_CFG5)
    meta = json.loads(str(z["meta"]))
    cfg = bench.CONFIGS["5"]
    assert (meta["kind"], meta["n"], meta["d"], meta["cv"], meta["seed"]) == (cfg["kind"], cfg["n"], cfg["d"], cfg["cv"], cfg["seed"])
    cands = bench.candidates_of(cfg, cfg["n_cand"])
    assert [cands[i]["C"] for i in z["index"]] == meta["C"] and meta["n_candidates"] == cfg["n_cand"]
    assert z["coef"].shape == (8, 10, 512) and z["ties"].shape == (8, 10) and (z["n_test"].sum(axis=1) == cfg["n"]).all()
    assert meta["pin"]["n_iter_nfev_equal"] and meta["pin"]["max_coef_reldiff"] < 1e-9 and (z["status"] == 0).all()


def test_committed_config3_golden_file_is_the_bench_config():
    z = np.load(bench.GOLDEN_CFG3)
    meta = json.loads(str(z["meta"]))
    cfg = bench.CONFIGS["3"]
    assert (meta["kind"], meta["n"], meta["d"], meta["cv"], meta["seed"]) == (cfg["kind"], cfg["n"], cfg["d"], cfg["cv"], cfg["seed"])
    cands = bench.candidates_of(cfg, cfg["n_cand"])
    assert [cands[i]["alpha"] for i in meta["alpha_index"]] == meta["alpha"]
    assert z["coef"].shape == (8, 5, 512) and int(z["n_test"].sum()) == cfg["n"] and meta["sklearn_pin_coef_reldiff"] < 1e-9
