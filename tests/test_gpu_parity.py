"""CUDA path vs the oracle, through the C ABI (solver.SweepSolver -> ctypes -> libdml_b200.so).

Bars (BASELINE.json north_star): identical fold labels, per-fold correct counts
bit-exact except on decision ties, coefficients / scores within 1e-4 relative,
best candidate identical under sklearn's rule.

The comparison target is scikit-learn fed the same values as float64 -- the
reference worker's actual path (pd.read_csv -> float64, worker.py:425-429).
scikit-learn's float32 path is NOT a usable target: it sums the per-sample loss in
float32, and that noise moves More'-Thuente interpolation steps enough to land up
to 1e-3 away from scikit-learn's own float64 result on identical values
("precision").
"""
import json
import os
import warnings

import numpy as np
import pytest
import torch

from cs230_distributed_machine_learning_b200 import folds
from oracle import folds as ofolds
from oracle import lbfgs as olbfgs
from oracle import reference_path as rp
from oracle import ridge as oridge
from tests.synth import make_dataset

pytestmark = pytest.mark.gpu
RTOL = 1e-4   # north_star tolerance for coefficients and scores


def rel_err(got, want):
    return float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-30))


def near_ties(X, coef, icpt, rows, eps=2e-5):
    """rows whose sklearn decision is within eps of flipping (binary: |z|, multi: top-2 gap)."""
    z = X[rows].astype(np.float64) @ np.atleast_2d(coef).T + icpt
    if z.shape[1] == 1:
        return int((np.abs(z[:, 0]) < eps * max(1.0, np.abs(z).max())).sum())
    zs = np.sort(z, axis=1)
    return int(((zs[:, -1] - zs[:, -2]) < eps * max(1.0, np.abs(z).max())).sum())


def load_case(golden_dir, name):
    z = np.load(os.path.join(golden_dir, "sweep_small.npz"))
    meta = json.loads(str(z[f"{name}_meta"]))
    X, y = make_dataset(meta["kind"], meta["n"], meta["d"], seed=meta["seed"])
    return z, meta, X, y


@pytest.mark.parametrize("tc", [1, 0], ids=["tcgen05", "ffma"])
@pytest.mark.parametrize("name", ["cfg2", "cfg5", "cfg4"])
def test_logreg_sweep_vs_sklearn_fixture(solver, golden_dir, name, tc):
    z, meta, X, y = load_case(golden_dir, name)
    cv, cands = meta["cv"], meta["candidates"]
    fold = folds.stratified_labels(y, cv)
    assert (fold == ofolds.stratified_fold_ids(y, cv)).all()
    ds = solver.upload_dataset(X, y, classifier=True)
    out = solver.logreg_sweep(ds, fold, cv, [c["C"] for c in cands], [c.get("tol", 1e-4) for c in cands],
                              [c.get("fit_intercept", True) for c in cands], [100] * len(cands), use_tensor_cores=tc)
    want_coef, want_icpt = z[f"{name}_coef"], z[f"{name}_intercept"]
    iter_diff = excused = 0
    for i in range(len(cands)):
        for k in range(cv):
            if out["n_iter"][i, k] != z[f"{name}_n_iter"][i, k]:
                iter_diff += 1      # a column stopping one iterate apart is reported, not hidden
                continue
            assert rel_err(out["coef"][i, k], want_coef[i, k]) <= RTOL, (i, k)
            assert np.abs(out["intercept"][i, k] - want_icpt[i, k]).max() <= RTOL * max(1.0, np.abs(want_coef[i, k]).max())
            ties = near_ties(X, want_coef[i, k], want_icpt[i, k], fold == k)
            diff = abs(int(out["correct"][i, k]) - int(z[f"{name}_correct"][i, k]))
            assert diff <= ties, (i, k)
            excused += diff
    print(f"{name} tc={tc}: {iter_diff} fits at another iterate, {excused} rows excused as decision ties")
    assert iter_diff <= max(1, len(cands) * cv // 20), f"{iter_diff} fits stopped at a different iterate"
    assert (out["n_test"] == z[f"{name}_ntest"]).all()
    assert (out["status"] == 0).all()
    scores = out["correct"] / out["n_test"]
    np.testing.assert_allclose(scores, z[f"{name}_scores"], atol=RTOL)
    mean = np.average(scores, axis=1)
    gold_mean = np.average(z[f"{name}_scores"], axis=1)
    if np.sort(gold_mean)[-1] - np.sort(gold_mean)[-2] > 2.0 / meta["n"]:   # winner separated by > 1 row
        assert rp.search_rules(scores)[3] == int(z[f"{name}_best_index"])


@pytest.mark.parametrize("name,classifier", [("cfg3r", False), ("cfg3c", True)])
def test_ridge_sweep_vs_sklearn_fixture(solver, golden_dir, name, classifier):
    z, meta, X, y = load_case(golden_dir, name)
    cv, cands = meta["cv"], meta["candidates"]
    fold = folds.stratified_labels(y, cv) if classifier else folds.kfold_labels(len(y), cv)
    ds = solver.upload_dataset(X, y, classifier=classifier)
    out = solver.ridge_sweep(ds, fold, cv, [c["alpha"] for c in cands], [True] * len(cands), classifier=classifier)
    assert (out["info"] == 0).all()
    for i in range(len(cands)):
        for k in range(cv):
            assert rel_err(out["coef"][i, k], z[f"{name}_coef"][i, k]) <= RTOL, (i, k)
            assert np.abs(out["intercept"][i, k] - z[f"{name}_intercept"][i, k]).max() <= RTOL
    if classifier:
        for i in range(len(cands)):
            for k in range(cv):
                ties = near_ties(X, z[f"{name}_coef"][i, k], z[f"{name}_intercept"][i, k], fold == k)
                assert abs(int(out["correct"][i, k]) - int(z[f"{name}_correct"][i, k])) <= ties
        assert (out["n_test"] == z[f"{name}_ntest"]).all()
    else:
        sst = out["syy"] - out["sy"] ** 2 / out["n_test"]
        np.testing.assert_allclose(out["sse"], z[f"{name}_sse"], rtol=RTOL)
        np.testing.assert_allclose(sst, z[f"{name}_sst"], rtol=RTOL)
        np.testing.assert_allclose(1.0 - out["sse"] / sst, z[f"{name}_scores"], rtol=RTOL, atol=RTOL)


def test_score_kernel_counts_bit_exact(solver):
    """Oracle-supplied coefficients: counts equal the numpy predictions exactly."""
    X, y = make_dataset("binary", 5000, 24, seed=31)
    fold = folds.stratified_labels(y, 5)
    rng = np.random.default_rng(0)
    W = (rng.standard_normal((10, 24)) / 5).astype(np.float32)
    b = rng.standard_normal(10).astype(np.float32) * 0.1
    sel, val = np.zeros(10, dtype=np.int32), (np.arange(10) % 5).astype(np.int32)
    ds = solver.upload_dataset(X, y, classifier=True)
    got = solver.score(ds, fold, W, b, sel, val, T=1)
    z = X.astype(np.float64) @ W.T.astype(np.float64) + b
    for s in range(10):
        rows = fold == val[s]
        exact = int(((z[rows, s] > 0).astype(int) == y[rows]).sum())
        slack = int((np.abs(z[rows, s]) < 1e-5).sum())
        assert abs(int(got[s]) - exact) <= slack
    # multiclass arg-max, ragged tile (n % 128 != 0, d % 4 != 0)
    X, y = make_dataset("multiclass5", 3333, 13, seed=32)
    fold = folds.stratified_labels(y, 3)
    W = (rng.standard_normal((15, 13)) / 3).astype(np.float32)
    b = np.zeros(15, dtype=np.float32)
    sel, val = np.zeros(15, dtype=np.int32), np.repeat(np.arange(3), 5).astype(np.int32)
    ds = solver.upload_dataset(X, y, classifier=True)
    got = solver.score(ds, fold, W, b, sel, val, T=5)
    for g in range(3):
        rows = fold == g
        zz = X[rows].astype(np.float64) @ W[g * 5:(g + 1) * 5].T.astype(np.float64)
        zs = np.sort(zz, axis=1)
        slack = int(((zs[:, -1] - zs[:, -2]) < 1e-5).sum())
        assert abs(int(got[g]) - int((zz.argmax(1) == y[rows]).sum())) <= slack


@pytest.mark.parametrize("n,d,k", [(1000, 13, 2), (777, 5, 3), (4096, 32, 2), (130, 3, 2)])
def test_logreg_ragged_shapes_and_options(solver, n, d, k):
    """Edge shapes (n % 128, d % 4), fit_intercept=False, tight tol, iteration cap."""
    from sklearn.linear_model import LogisticRegression
    X, y = make_dataset("binary" if k == 2 else f"multiclass{k}", n, d, seed=40 + d)
    cv = 3
    fold = folds.stratified_labels(y, cv)
    cands = [dict(C=1.0, tol=1e-4, fit_intercept=True, max_iter=100),
             dict(C=0.01, tol=1e-6, fit_intercept=False, max_iter=100),
             dict(C=100.0, tol=1e-4, fit_intercept=True, max_iter=3)]
    ds = solver.upload_dataset(X, y, classifier=True)
    out = solver.logreg_sweep(ds, fold, cv, [c["C"] for c in cands], [c["tol"] for c in cands],
                              [c["fit_intercept"] for c in cands], [c["max_iter"] for c in cands])
    compared = skipped = 0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for i, c in enumerate(cands):
            for f in range(cv):
                tr, te = fold != f, fold == f
                m = LogisticRegression(**c).fit(X[tr].astype(np.float64), y[tr])
                if out["n_iter"][i, f] != m.n_iter_[0]:
                    # stopped at another iterate (candidate 1 asks for tol = 1e-6, at the noise level of fp32 gradients
                    # on a few hundred rows): counted and capped below, and still held to a bound -- both runs are
                    # then within the stopping tolerance of the same optimum
                    skipped += 1
                    assert abs(int(out["n_iter"][i, f]) - int(m.n_iter_[0])) <= 3, (i, f, out["n_iter"][i, f], m.n_iter_[0])
                    assert rel_err(out["coef"][i, f], m.coef_) <= 2e-2, (i, f)
                    continue
                compared += 1
                assert rel_err(out["coef"][i, f], m.coef_) <= 5 * RTOL, (i, f)
                ties = near_ties(X, m.coef_, m.intercept_, te, eps=1e-4)
                assert abs(int(out["correct"][i, f]) - int((m.predict(X[te]) == y[te]).sum())) <= ties
                assert out["n_test"][i, f] == te.sum()
    print(f"ragged {n}x{d} k={k}: {compared} fits compared at equal n_iter, {skipped} at another iterate")
    assert compared >= 6 and skipped <= 3, f"{skipped} of {compared + skipped} fits stopped at a different iterate"
    assert (out["n_iter"][2] <= 3).all() and (out["status"][2] == 1).all()   # cap reached -> status 1
    assert (out["intercept"][1] == 0).all()


def test_holdout_fit_equals_train_test_split(solver):
    from sklearn.linear_model import LogisticRegression, Ridge
    from sklearn.model_selection import train_test_split
    X, y = make_dataset("binary", 3000, 16, seed=50)
    fold = folds.stratified_labels(y, 5)
    hold = folds.holdout_labels(len(y), 0.25, 42)
    ds = solver.upload_dataset(X, y, classifier=True)
    out = solver.logreg_sweep(ds, fold, 5, [0.5], [1e-4], [True], [100], hold_id=hold)
    Xtr, Xte, ytr, yte = train_test_split(X.astype(np.float64), y, test_size=0.25, random_state=42)
    m = LogisticRegression(C=0.5).fit(Xtr, ytr)
    assert out["coef"].shape == (1, 6, 1, 16)
    assert rel_err(out["coef"][0, 5], m.coef_) <= RTOL
    assert out["n_test"][0, 5] == len(yte)
    assert abs(int(out["correct"][0, 5]) - int((m.predict(Xte) == yte).sum())) <= near_ties(Xte, m.coef_, m.intercept_, slice(None), 1e-4)
    Xr, yr = make_dataset("regression", 3000, 16, seed=51)
    dr = solver.upload_dataset(Xr, yr, classifier=False)
    outr = solver.ridge_sweep(dr, folds.kfold_labels(3000, 5), 5, [2.0], [True], classifier=False, hold_id=hold)
    Xtr, Xte, ytr, yte = train_test_split(Xr.astype(np.float64), yr.astype(np.float64), test_size=0.25, random_state=42)
    mr = Ridge(alpha=2.0).fit(Xtr, ytr)
    assert rel_err(outr["coef"][0, 5], mr.coef_) <= RTOL
    assert outr["sse"][0, 5] == pytest.approx(float(((yte - mr.predict(Xte)) ** 2).sum()), rel=RTOL)


def test_restatement_tracks_gpu_trajectory(solver):
    """Same rule set => same number of iterations and function evaluations as the fp64 restatement."""
    X, y = make_dataset("binary", 20000, 32, seed=60)
    fold = folds.stratified_labels(y, 5)
    Cs = [1e-5, 1e-3, 1e-1, 10.0]
    ds = solver.upload_dataset(X, y, classifier=True)
    out = solver.logreg_sweep(ds, fold, 5, Cs, [1e-4] * 4, [True] * 4, [100] * 4)
    same = 0
    for i, C in enumerate(Cs):
        for k in range(5):
            r = olbfgs.fit_logreg(X[fold != k].astype(np.float64), y[fold != k], 2, C=C)
            same += int(r["n_iter"] == out["n_iter"][i, k] and r["nfev"] == out["nfev"][i, k])
            if r["n_iter"] == out["n_iter"][i, k]:
                assert rel_err(out["coef"][i, k], r["coef"]) <= RTOL
    assert same >= 18, f"only {same}/20 fits followed the oracle's trajectory"


def test_api_end_to_end_best_params_equal_gridsearchcv():
    """MLTaskManager.train on float32 data -> b200 executor; compare with sklearn GridSearchCV."""
    from sklearn.linear_model import LogisticRegression, Ridge
    from sklearn.model_selection import GridSearchCV
    import cs230_distributed_machine_learning_b200 as pkg
    X, y = make_dataset("binary", 8000, 32, seed=70)
    pkg.register_dataset("syn-bin", X, y)
    m = pkg.MLTaskManager("local://1")
    try:
        grid = {"C": [float(c) for c in np.logspace(-5, 1, 7)]}
        r = m.train(GridSearchCV(LogisticRegression(), grid, cv=5), "syn-bin", wait_for_completion=True)
        assert r["job_status"] == "completed", r
        gs = GridSearchCV(LogisticRegression(), grid, cv=5).fit(X.astype(np.float64), y)
        jr = r["job_result"]
        assert all(res["executor"] == "b200" for res in jr["results"])
        np.testing.assert_allclose(jr["cv_results_"]["mean_test_score"], gs.cv_results_["mean_test_score"], atol=RTOL)
        top2 = np.sort(gs.cv_results_["mean_test_score"])[-2:]
        if top2[1] - top2[0] > 2.0 / len(y):
            assert jr["best_params_"] == gs.best_params_
        Xr, yr = make_dataset("regression", 8000, 32, seed=71)
        pkg.register_dataset("syn-reg", Xr, yr)
        gridr = {"alpha": [float(a) for a in np.logspace(-2, 4, 7)]}
        r = m.train(GridSearchCV(Ridge(), gridr, cv=5), "syn-reg", train_params={"test_size": 0.2, "random_state": 1},
                    wait_for_completion=True)
        gsr = GridSearchCV(Ridge(), gridr, cv=5).fit(Xr.astype(np.float64), yr.astype(np.float64))
        jr = r["job_result"]
        np.testing.assert_allclose(jr["cv_results_"]["mean_test_score"], gsr.cv_results_["mean_test_score"], rtol=RTOL)
        assert jr["best_params_"] == gsr.best_params_
        assert "r2_score" in jr["results"][0] and os.path.exists(jr["results"][0]["model_path"])
    finally:
        m.close()


def test_refit_on_gpu_equals_gridsearchcv_best_estimator(tmp_path, monkeypatch):
    """train_params {"refit": True}: the best candidate is fitted on ALL rows by the sweep library
    (one column, empty hold-out set) and pickled as a scikit-learn estimator."""
    import pickle
    from sklearn.linear_model import LogisticRegression, Ridge
    from sklearn.model_selection import GridSearchCV
    import cs230_distributed_machine_learning_b200 as pkg
    monkeypatch.chdir(tmp_path)
    X, y = make_dataset("binary", 12_000, 40, seed=90)
    pkg.register_dataset("rf-bin", X, y)
    Xr, yr = make_dataset("regression", 12_000, 40, seed=91)
    pkg.register_dataset("rf-reg", Xr, yr)
    m = pkg.MLTaskManager("local://1")
    try:
        grid = {"C": [1e-4, 1e-2, 1.0]}
        r = m.train(GridSearchCV(LogisticRegression(), grid, cv=5), "rf-bin", train_params={"refit": True},
                    wait_for_completion=True)
        assert r["job_status"] == "completed", r
        jr = r["job_result"]
        gs = GridSearchCV(LogisticRegression(), grid, cv=5).fit(X.astype(np.float64), y)
        be = jr["best_estimator_"]
        assert be["executor"] == "b200" and be["parameters"]["C"] == gs.best_params_["C"]
        assert rel_err(np.asarray(be["coef"]), gs.best_estimator_.coef_) < RTOL
        assert rel_err(np.asarray(be["intercept"]), gs.best_estimator_.intercept_) < 10 * RTOL
        with open(be["model_path"], "rb") as fh:
            est = pickle.load(fh)
        agree = (est.predict(X.astype(np.float64)) == gs.best_estimator_.predict(X.astype(np.float64))).mean()
        assert agree > 0.9995                                # decision ties only
        gridr = {"alpha": [0.1, 10.0, 1000.0]}
        r = m.train(GridSearchCV(Ridge(), gridr, cv=5), "rf-reg", train_params={"refit": True}, wait_for_completion=True)
        gsr = GridSearchCV(Ridge(), gridr, cv=5).fit(Xr.astype(np.float64), yr.astype(np.float64))
        be = r["job_result"]["best_estimator_"]
        assert be["parameters"]["alpha"] == gsr.best_params_["alpha"]
        assert rel_err(np.asarray(be["coef"]), gsr.best_estimator_.coef_) < RTOL
        with open(be["model_path"], "rb") as fh:
            est = pickle.load(fh)
        np.testing.assert_allclose(est.predict(Xr[:100].astype(np.float64)),
                                   gsr.best_estimator_.predict(Xr[:100].astype(np.float64)), rtol=1e-3, atol=1e-3)
    finally:
        m.close()


def test_linear_regression_is_ridge_with_alpha_zero():
    """LinearRegression on the b200 executor: the Ridge normal equations with alpha = 0."""
    from sklearn.linear_model import LinearRegression
    from sklearn.model_selection import GridSearchCV
    import cs230_distributed_machine_learning_b200 as pkg
    X, y = make_dataset("regression", 20_000, 48, seed=95)
    pkg.register_dataset("lin-reg", X, y)
    m = pkg.MLTaskManager("local://1")
    try:
        grid = {"fit_intercept": [True, False]}
        r = m.train(GridSearchCV(LinearRegression(), grid, cv=5), "lin-reg", train_params={"refit": True},
                    wait_for_completion=True)
        assert r["job_status"] == "completed", r
        jr = r["job_result"]
        assert all(res["executor"] == "b200" for res in jr["results"])
        gs = GridSearchCV(LinearRegression(), grid, cv=5).fit(X.astype(np.float64), y.astype(np.float64))
        np.testing.assert_allclose(jr["cv_results_"]["mean_test_score"], gs.cv_results_["mean_test_score"], rtol=RTOL)
        assert jr["best_params_"] == gs.best_params_
        assert rel_err(np.asarray(jr["best_estimator_"]["coef"]), gs.best_estimator_.coef_) < RTOL
    finally:
        m.close()


def test_queued_jobs_prefetch_their_dataset_and_match_blocking_runs():
    """Jobs submitted without waiting: the worker starts the next job's host->device copy on its
    copy stream while the current sweep runs.  Results equal the one-job-at-a-time results exactly."""
    from sklearn.linear_model import LogisticRegression
    from sklearn.model_selection import GridSearchCV
    import cs230_distributed_machine_learning_b200 as pkg
    from cs230_distributed_machine_learning_b200 import worker as wk
    data = {}
    for nm, seed in (("pf-a", 80), ("pf-b", 81)):
        data[nm] = make_dataset("binary", 20_000, 64, seed=seed)
        pkg.register_dataset(nm, *data[nm])
    m = pkg.MLTaskManager("local://1")
    try:
        est = GridSearchCV(LogisticRegression(), {"C": [0.001, 0.1, 10.0]}, cv=3)
        want = {}
        for nm in data:
            r = m.train(est, nm, wait_for_completion=True)
            assert r["job_status"] == "completed", r
            want[nm] = r["job_result"]["cv_results_"]["mean_test_score"]
        ex = wk.gpu_executor(0)
        old_budget, ex.max_resident_bytes = ex.max_resident_bytes, int(1.5 * data["pf-a"][0].nbytes)
        before = ex.prefetched
        order = ["pf-a", "pf-b", "pf-a", "pf-b", "pf-a", "pf-b"]
        jobs = []
        with ex.lock:                    # hold the first sweep back until every job is queued
            for nm in order:
                assert m.train(est, nm, wait_for_completion=False)["status"] == "Training started"
                jobs.append(m.job_id)
        for nm, j in zip(order, jobs):
            r = m._wait_for_completion(j)
            assert r["job_status"] == "completed", r
            assert r["job_result"]["cv_results_"]["mean_test_score"] == want[nm]
        assert ex.prefetched - before >= 4      # every job after the first found its copy already started
        assert len(ex.resident) <= 2
        ex.max_resident_bytes = old_budget
    finally:
        m.close()


@pytest.mark.parametrize("env", [{"DML_FOLD_SKIP": "0"}, {"DML_GRAD_SKIP": "1"}, {"DML_FIRST_ROUND_ONCE": "0"},
                                 {"DML_ROUND_AHEAD": "0"}, {"DML_GB_IN_GRAD": "0"}, {"DML_GRAD_ZERO_FILL": "0"}],
                         ids=["no-fold-skip", "grad-skip", "full-round-0", "no-round-ahead", "gb-in-forward", "no-zero-fill"])
def test_skipping_knobs_do_not_change_results(solver, monkeypatch, env):
    """Fold-aware column skipping and the round-0 shortcut only remove work whose result is zero or
    known: every switch setting gives the same iterates (compared with the default context)."""
    from cs230_distributed_machine_learning_b200.solver import SweepSolver
    X, y = make_dataset("binary", 60_000, 256, seed=12)
    fold = folds.stratified_labels(y, 5)
    # 64 candidates: every fold's group of fits is 64 columns, so the default context runs the fold-major
    # padded layout with column-range skipping (and falls back to the dense list as fits finish), while
    # DML_FOLD_SKIP=0 runs the dense fit-major list throughout
    Cs = np.logspace(-5, 1, 64)
    args = ([1e-4] * 64, [True] * 64, [100] * 64)
    ds = solver.upload_dataset(X, y, classifier=True)
    want = solver.logreg_sweep(ds, fold, 5, Cs, *args)
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    other = SweepSolver(0)
    try:
        ds2 = other.upload_dataset(X, y, classifier=True)
        got = other.logreg_sweep(ds2, fold, 5, Cs, *args)
    finally:
        other.close()
    # same work in a different order: the iterates agree far inside the parity tolerance (knobs that move the
    # sum of residuals between kernels change its rounding: coefficients to 1e-5 of their scale, counts up to ties;
    # the others reproduce coefficients to 1e-6 and counts exactly)
    loose = "DML_GB_IN_GRAD" in env or "DML_GRAD_SKIP" in env
    assert (got["n_iter"] == want["n_iter"]).all()
    assert np.abs(got["correct"] - want["correct"]).max() <= (2 if loose else 0)
    if loose:
        for i in range(got["coef"].shape[0]):
            assert rel_err(got["coef"][i], want["coef"][i]) < 1e-5, i
    else:
        np.testing.assert_allclose(got["coef"], want["coef"], rtol=1e-6, atol=1e-9)
    if "DML_FOLD_SKIP" in env:       # and the grouped layout itself against scikit-learn (float64), three fits
        from sklearn.linear_model import LogisticRegression
        X64 = X.astype(np.float64)
        compared = 0
        for i, f in ((5, 0), (33, 2), (60, 4)):
            m = LogisticRegression(C=float(Cs[i])).fit(X64[fold != f], y[fold != f])
            if m.n_iter_[0] == want["n_iter"][i, f]:
                compared += 1
                assert rel_err(want["coef"][i, f], m.coef_) < RTOL, (i, f)
        assert compared >= 2


def test_grouped_layout_with_holdout_column_group(solver, monkeypatch):
    """Fold-major layout when every candidate also has a hold-out fit: the hold-out group (rows chosen by
    a permutation, never a whole tile) must not be skipped anywhere.  Compared with the dense layout."""
    from cs230_distributed_machine_learning_b200.solver import SweepSolver
    X, y = make_dataset("binary", 30_000, 128, seed=14)
    fold = folds.stratified_labels(y, 3)
    hold = folds.holdout_labels(len(y), 0.25, 7)
    Cs = np.logspace(-4, 1, 32)
    args = ([1e-4] * 32, [True] * 32, [100] * 32)
    ds = solver.upload_dataset(X, y, classifier=True)
    want = solver.logreg_sweep(ds, fold, 3, Cs, *args, hold_id=hold)
    monkeypatch.setenv("DML_FOLD_SKIP", "0")
    other = SweepSolver(0)
    try:
        got = other.logreg_sweep(other.upload_dataset(X, y, classifier=True), fold, 3, Cs, *args, hold_id=hold)
    finally:
        other.close()
    assert want["coef"].shape == (32, 4, 1, 128)
    assert (got["n_iter"] == want["n_iter"]).all() and (got["correct"] == want["correct"]).all()
    assert (got["n_test"] == want["n_test"]).all() and want["n_test"][0, 3] == hold.sum()
    np.testing.assert_allclose(got["coef"], want["coef"], rtol=1e-6, atol=1e-9)


def test_medium_size_properties(solver):
    """Size-independent properties at a size the oracle cannot sweep in seconds (200k x 256, 16 C x 5 folds)."""
    X, y = make_dataset("binary", 200_000, 256, seed=2)
    fold = folds.stratified_labels(y, 5)
    Cs = np.logspace(-6, 2, 16)
    ds = solver.upload_dataset(X, y, classifier=True)
    a = solver.logreg_sweep(ds, fold, 5, Cs, [1e-4] * 16, [True] * 16, [100] * 16)
    b = solver.logreg_sweep(ds, fold, 5, Cs[::-1].copy(), [1e-4] * 16, [True] * 16, [100] * 16)
    assert (a["n_test"].sum(axis=1) == len(y)).all()                       # folds partition the rows
    assert ((a["correct"] >= 0) & (a["correct"] <= a["n_test"])).all()
    assert (a["status"] == 0).all() and (a["n_iter"] < 100).all()
    np.testing.assert_array_equal(a["correct"], b["correct"][::-1])       # column order does not matter
    np.testing.assert_array_equal(a["n_iter"], b["n_iter"][::-1])
    np.testing.assert_array_equal(a["coef"], b["coef"][::-1])             # deterministic reductions
    nrm = np.linalg.norm(a["coef"][:, 0, 0], axis=1)
    assert (np.diff(nrm) > 0).all()                                        # ||w|| grows with C
    # recount the winner's predictions on the host from the returned coefficients
    i = int(np.argmax(a["correct"].sum(axis=1)))
    for k in range(5):
        rows = fold == k
        zz = X[rows] @ a["coef"][i, k, 0] + a["intercept"][i, k, 0]
        assert abs(int(((zz > 0) == y[rows]).sum()) - int(a["correct"][i, k])) <= int((np.abs(zz) < 1e-5).sum())


def test_config_shapes_at_moderate_size(solver):
    """The shapes of BASELINE.json configs 3-5 (d = 512, cv = 10, 10 classes, per-candidate tol /
    fit_intercept) at sizes the oracle can still spot-check."""
    from sklearn.linear_model import LogisticRegression, Ridge
    # config 5 shape: binary, d = 512, cv = 10 (tensor-core path, 4 feature blocks)
    X, y = make_dataset("binary", 60_000, 512, seed=5)
    fold = folds.stratified_labels(y, 10)
    Cs = np.logspace(-6, 2, 16)
    ds = solver.upload_dataset(X, y, classifier=True)
    out = solver.logreg_sweep(ds, fold, 10, Cs, [1e-4] * 16, [True] * 16, [100] * 16)
    assert (out["status"] == 0).all() and (out["n_test"].sum(axis=1) == len(y)).all()
    for i, k in [(0, 0), (7, 3), (15, 9)]:
        r = olbfgs.fit_logreg(X[fold != k].astype(np.float64), y[fold != k], 2, C=Cs[i])
        assert r["n_iter"] == out["n_iter"][i, k]
        assert rel_err(out["coef"][i, k], r["coef"]) <= RTOL
        z = X[fold == k].astype(np.float64) @ r["coef"][0] + r["intercept"][0]
        assert abs(int(((z > 0) == y[fold == k]).sum()) - int(out["correct"][i, k])) <= int((np.abs(z) < 2e-5).sum())
    # config 4 shape: 10 classes, d = 512, candidates differing in C, tol and fit_intercept
    X, y = make_dataset("multiclass10", 30_000, 512, seed=4)
    fold = folds.stratified_labels(y, 5)
    cands = [(1e-3, 1e-4, True), (1.0, 1e-5, False), (10.0, 1e-3, True), (0.05, 1e-4, False)]
    ds = solver.upload_dataset(X, y, classifier=True)
    out = solver.logreg_sweep(ds, fold, 5, [c[0] for c in cands], [c[1] for c in cands], [c[2] for c in cands], [100] * 4)
    assert out["coef"].shape == (4, 5, 10, 512)
    compared = 0
    for i, k in [(0, 1), (1, 4), (3, 0)]:
        C_, tol_, fi_ = cands[i]
        m = LogisticRegression(C=C_, tol=tol_, fit_intercept=fi_).fit(X[fold != k].astype(np.float64), y[fold != k])
        if m.n_iter_[0] == out["n_iter"][i, k]:
            compared += 1
            assert rel_err(out["coef"][i, k], m.coef_) <= RTOL
            ties = near_ties(X, m.coef_, m.intercept_, fold == k)
            assert abs(int((m.predict(X[fold == k].astype(np.float64)) == y[fold == k]).sum()) - int(out["correct"][i, k])) <= ties
    assert compared >= 2, f"only {compared} of 3 multinomial fits stopped at scikit-learn's iterate"
    # config 3 shape: Ridge, d = 512, many alphas, one pass over X for all folds
    X, y = make_dataset("regression", 100_000, 512, seed=3)
    kf = folds.kfold_labels(len(y), 5)
    alphas = np.logspace(-3, 5, 16)
    ds = solver.upload_dataset(X, y, classifier=False)
    out = solver.ridge_sweep(ds, kf, 5, alphas, [True] * 16, classifier=False)
    assert (out["info"] == 0).all()
    for i, k in [(0, 0), (9, 2), (15, 4)]:
        m = Ridge(alpha=alphas[i]).fit(X[kf != k].astype(np.float64), y[kf != k].astype(np.float64))
        assert rel_err(out["coef"][i, k], m.coef_) <= RTOL
        sse = float(((y[kf == k] - m.predict(X[kf == k].astype(np.float64))) ** 2).sum())
        assert out["sse"][i, k] == pytest.approx(sse, rel=RTOL)


def full_size_parity(out, z, X, y, fold, cv, rtol=RTOL, eps=2e-5):
    """Compare a sweep result with a cached full-oracle record (tests/golden/make_cfg_full.py):
    returns the numbers north_star asks to be reported, never hiding a fit that differs."""
    n_cand = out["n_iter"].shape[0]
    same_iter = out["n_iter"] == z["n_iter"]
    coef_err = np.zeros((n_cand, cv)); icpt_err = np.zeros((n_cand, cv))
    diffs = np.zeros((n_cand, cv), np.int64); ties = np.zeros((n_cand, cv), np.int64)
    for k in range(cv):
        rows = fold == k
        Xk = X[rows].astype(np.float64)
        Z = Xk @ z["coef"][:, k, :].T + z["intercept"][:, k][None, :]          # sklearn's decision values
        zmax = np.maximum(1.0, np.abs(Z).max(axis=0))
        ties[:, k] = (np.abs(Z) < eps * zmax[None, :]).sum(axis=0)
        for i in range(n_cand):
            w = z["coef"][i, k]
            coef_err[i, k] = np.abs(out["coef"][i, k, 0] - w).max() / np.abs(w).max()
            icpt_err[i, k] = abs(out["intercept"][i, k, 0] - z["intercept"][i, k]) / max(1.0, np.abs(w).max())
        diffs[:, k] = np.abs(out["correct"][:, k] - z["correct"][:, k])
    scores = out["correct"] / out["n_test"]
    best = rp.search_rules(scores)[3]
    return dict(fits_checked=int(same_iter.size), n_iter_equal=int(same_iter.sum()),
                nfev_equal=int((out["nfev"] == z["nfev"]).sum()),
                max_coef_relerr=float(coef_err[same_iter].max()), max_intercept_err=float(icpt_err[same_iter].max()),
                count_diffs=int((diffs > 0).sum()), max_count_diff=int(diffs.max()),
                count_diffs_beyond_ties=int((diffs > ties).sum()), rows_excused_as_ties=int(np.minimum(diffs, ties).sum()),
                rows_near_ties=int(ties.sum()), n_test_equal=bool((out["n_test"] == z["n_test"]).all()),
                best_index=int(best), best_index_oracle=int(z["best_index"]),
                best_index_equal=bool(best == int(z["best_index"])),
                max_mean_score_diff=float(np.abs(np.average(scores, axis=1) - z["mean_test_score"]).max()))


def test_config2_full_size_vs_cached_oracle(solver, golden_dir):
    """BASELINE.json config 2 at FULL size (1M x 256, 64 C x cv 5) against the cached scikit-learn (float64)
    answers of tests/golden/make_cfg_full.py: n_iter, coefficients within 1e-4, per-fold correct counts
    exact except rows within 2e-5 of a decision tie, best_params_ identical."""
    z = np.load(os.path.join(golden_dir, "cfg2_full.npz"))
    meta = json.loads(str(z["meta"]))
    X, y = make_dataset(meta["kind"], meta["n"], meta["d"], seed=meta["seed"])
    cv, Cs = meta["cv"], np.asarray(meta["C"])
    fold = folds.stratified_labels(y, cv)
    ds = solver.upload_dataset(X, y, classifier=True)
    n = len(Cs)
    out = solver.logreg_sweep(ds, fold, cv, Cs, [1e-4] * n, [True] * n, [100] * n)
    rep = full_size_parity(out, z, X, y, fold, cv)
    print("cfg2 full-size parity:", json.dumps(rep))
    if os.path.isdir(os.path.join(os.path.dirname(golden_dir), "..", "gpurun_out")):
        json.dump(rep, open(os.path.join(os.path.dirname(golden_dir), "..", "gpurun_out", "parity_cfg2_full.json"), "w"))
    assert rep["n_test_equal"] and (out["status"] == 0).all()
    assert rep["n_iter_equal"] >= 0.95 * rep["fits_checked"], rep     # a fit one iterate apart is counted, not hidden
    assert rep["max_coef_relerr"] <= RTOL and rep["max_intercept_err"] <= RTOL, rep
    assert rep["count_diffs_beyond_ties"] == 0, rep
    assert rep["max_mean_score_diff"] <= RTOL, rep
    assert rep["best_index_equal"], rep


@pytest.mark.parametrize("kind,n,d", [("multiclass3", 9000, 24), ("multiclass10", 20000, 64)])
def test_ridge_classifier_multiclass_vs_sklearn(solver, kind, n, d):
    """RidgeClassifier with more than two classes: one-vs-rest +-1 target columns from ONE Gram pass, arg-max
    prediction (sklearn/linear_model/_ridge.py:1291-1348).  Against scikit-learn on float64."""
    from sklearn.linear_model import RidgeClassifier
    X, y = make_dataset(kind, n, d, seed=33)
    cv = 5
    fold = folds.stratified_labels(y, cv)
    alphas = [1e-2, 1.0, 100.0, 1e4]
    ds = solver.upload_dataset(X, y, classifier=True)
    out = solver.ridge_sweep(ds, fold, cv, alphas, [True] * len(alphas), classifier=True)
    T = int(kind[len("multiclass"):])
    assert out["coef"].shape == (len(alphas), cv, T, d) and (out["info"] == 0).all()
    X64 = X.astype(np.float64)
    excused = 0
    for i, a in enumerate(alphas):
        for k in range(cv):
            tr, te = fold != k, fold == k
            m = RidgeClassifier(alpha=a).fit(X64[tr], y[tr])
            assert rel_err(out["coef"][i, k], m.coef_) <= RTOL, (i, k)
            assert np.abs(out["intercept"][i, k] - m.intercept_).max() <= RTOL
            want = int((m.predict(X64[te]) == y[te]).sum())
            ties = near_ties(X, m.coef_, m.intercept_, te)
            diff = abs(int(out["correct"][i, k]) - want)
            assert diff <= ties, (i, k, diff, ties)
            excused += diff
            assert out["n_test"][i, k] == te.sum()
    print(f"ridge classifier {kind}: {excused} rows excused as arg-max ties")


def test_ridge_rank_deficient_system_is_flagged_and_rerouted():
    """An exactly collinear column: scikit-learn returns the min-norm lstsq solution for LinearRegression and
    a well-defined Ridge solution for alpha > 0.  The fp32 Gram matrix cannot resolve the zero pivot, so the
    library reports info != 0 for alpha = 0 and the worker re-runs that candidate on scikit-learn."""
    from sklearn.linear_model import LinearRegression, Ridge
    from sklearn.model_selection import GridSearchCV
    import cs230_distributed_machine_learning_b200 as pkg
    from cs230_distributed_machine_learning_b200.solver import SweepSolver
    X, y = make_dataset("regression", 8000, 16, seed=97)
    X = np.ascontiguousarray(X)
    X[:, 7] = X[:, 3]                      # duplicate feature
    s = SweepSolver(0)
    try:
        ds = s.upload_dataset(X, y, classifier=False)
        out = s.ridge_sweep(ds, folds.kfold_labels(len(y), 5), 5, [0.0, 10.0], [True, True], classifier=False)
        assert (out["info"][0] != 0).all(), out["info"]          # alpha = 0: singular, flagged
        assert (out["info"][1] == 0).all(), out["info"]          # alpha = 10: every pivot is at least about alpha
    finally:
        s.close()
    pkg.register_dataset("collinear-reg", X, y)
    m = pkg.MLTaskManager("local://1")
    try:
        r = m.train(GridSearchCV(LinearRegression(), {"fit_intercept": [True, False]}, cv=5), "collinear-reg",
                    wait_for_completion=True)
        assert r["job_status"] == "completed", r
        jr = r["job_result"]
        assert all(res["executor"] == "sklearn" and "rerouted" in res for res in jr["results"]), jr["results"]
        gs = GridSearchCV(LinearRegression(), {"fit_intercept": [True, False]}, cv=5).fit(X.astype(np.float64), y.astype(np.float64))
        np.testing.assert_allclose(jr["cv_results_"]["mean_test_score"], gs.cv_results_["mean_test_score"], rtol=RTOL)
        r = m.train(GridSearchCV(Ridge(), {"alpha": [5.0, 50.0]}, cv=5), "collinear-reg", wait_for_completion=True)
        jr = r["job_result"]
        assert all(res["executor"] == "b200" for res in jr["results"])
        gs = GridSearchCV(Ridge(), {"alpha": [5.0, 50.0]}, cv=5).fit(X.astype(np.float64), y.astype(np.float64))
        np.testing.assert_allclose(jr["cv_results_"]["mean_test_score"], gs.cv_results_["mean_test_score"], rtol=RTOL)
    finally:
        m.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on the box")
def test_two_contexts_in_one_process_run_tensor_core_sweeps():
    """One process, one ctx per GPU (worker.py runs a Worker thread per GPU): the shared-memory opt-ins of the
    tcgen05 kernels are per device, so the second GPU must get them too."""
    from cs230_distributed_machine_learning_b200.solver import SweepSolver
    X, y = make_dataset("binary", 20_000, 64, seed=98)
    Xr, yr = make_dataset("regression", 20_000, 64, seed=99)
    fold = folds.stratified_labels(y, 5)
    res = []
    for g in (0, 1):
        s = SweepSolver(g)
        try:
            ds = s.upload_dataset(X, y, classifier=True)
            a = s.logreg_sweep(ds, fold, 5, [1e-3, 1.0], [1e-4] * 2, [True] * 2, [100] * 2, use_tensor_cores=1)
            dr = s.upload_dataset(Xr, yr, classifier=False)
            b = s.ridge_sweep(dr, folds.kfold_labels(len(yr), 5), 5, [0.1, 10.0], [True, True], classifier=False, use_tensor_cores=1)
            res.append((a, b))
        finally:
            s.close()
    torch.cuda.set_device(0)
    np.testing.assert_array_equal(res[0][0]["coef"], res[1][0]["coef"])
    np.testing.assert_array_equal(res[0][0]["correct"], res[1][0]["correct"])
    np.testing.assert_array_equal(res[0][1]["coef"], res[1][1]["coef"])


@pytest.mark.parametrize("tc", [1, 0], ids=["tcgen05", "ffma"])
@pytest.mark.parametrize("kind,n,d", [("binary", 12_000, 32), ("multiclass4", 12_000, 32)])
def test_class_weight_vs_sklearn(solver, kind, n, d, tc):
    """LogisticRegression(class_weight="balanced" | dict): per-row weights by class inside the forward epilogue,
    1 / sum(weights) in the optimiser.  Against scikit-learn on float64 with the same class_weight."""
    from sklearn.linear_model import LogisticRegression
    from cs230_distributed_machine_learning_b200 import worker as wk
    X, y = make_dataset(kind, n, d, seed=61)
    if kind == "binary":                    # make the classes unbalanced so that "balanced" is not a no-op
        keep = (y == 0) | (np.arange(n) % 3 == 0)
        X, y = np.ascontiguousarray(X[keep]), y[keep]
    classes = np.unique(y)
    cv = 3
    fold = folds.stratified_labels(y, cv)
    dict_cw = {int(c): 1.0 + 0.7 * i for i, c in enumerate(classes)}
    cands = [dict(C=0.5, class_weight="balanced"), dict(C=0.01, class_weight=dict_cw), dict(C=1.0)]
    table = wk.class_weight_table(cands, classes, np.searchsorted(classes, y), fold, None, cv)
    ds = solver.upload_dataset(X, y, classifier=True)
    out = solver.logreg_sweep(ds, fold, cv, [c["C"] for c in cands], [1e-4] * 3, [True] * 3, [100] * 3,
                              class_weight=table, use_tensor_cores=tc)
    X64 = X.astype(np.float64)
    compared = 0
    for i, c in enumerate(cands):
        for k in range(cv):
            tr, te = fold != k, fold == k
            m = LogisticRegression(**c).fit(X64[tr], y[tr])
            if m.n_iter_[0] != out["n_iter"][i, k]:
                continue
            compared += 1
            assert rel_err(out["coef"][i, k], m.coef_) <= RTOL, (i, k)
            assert np.abs(out["intercept"][i, k] - m.intercept_).max() <= RTOL * max(1.0, np.abs(m.coef_).max())
            ties = near_ties(X, m.coef_, m.intercept_, te)
            assert abs(int(out["correct"][i, k]) - int((m.predict(X64[te]) == y[te]).sum())) <= ties, (i, k)
    assert compared >= 8, f"only {compared} of 9 fits stopped at scikit-learn's iterate"


def test_class_weight_through_the_api():
    from sklearn.linear_model import LogisticRegression
    from sklearn.model_selection import GridSearchCV
    import cs230_distributed_machine_learning_b200 as pkg
    X, y = make_dataset("binary", 9000, 24, seed=62)
    keep = (y == 1) | (np.arange(len(y)) % 4 == 0)
    X, y = np.ascontiguousarray(X[keep]), y[keep]
    pkg.register_dataset("cw-bin", X, y)
    m = pkg.MLTaskManager("local://1")
    try:
        grid = {"C": [0.01, 1.0], "class_weight": [None, "balanced", {0: 3.0, 1: 1.0}]}
        r = m.train(GridSearchCV(LogisticRegression(), grid, cv=5), "cw-bin", wait_for_completion=True)
        assert r["job_status"] == "completed", r
        jr = r["job_result"]
        assert all(res["executor"] == "b200" for res in jr["results"]), [res.get("executor_reason") for res in jr["results"]]
        gs = GridSearchCV(LogisticRegression(), grid, cv=5).fit(X.astype(np.float64), y)
        np.testing.assert_allclose(jr["cv_results_"]["mean_test_score"], gs.cv_results_["mean_test_score"], atol=RTOL)
    finally:
        m.close()


def test_reregistered_dataset_name_is_uploaded_again():
    """The executor keeps device copies and row labels per dataset NAME across jobs; registering the name again
    (new rows, new labels, another row count) must invalidate them -- the reference re-reads the CSV for every
    subtask (worker.py:425) and so always trains on current data."""
    from sklearn.linear_model import LogisticRegression
    from sklearn.model_selection import GridSearchCV
    import cs230_distributed_machine_learning_b200 as pkg
    m = pkg.MLTaskManager("local://1")
    try:
        grid = {"C": [0.01, 1.0]}
        for n, seed in ((6000, 101), (9000, 102)):
            X, y = make_dataset("binary", n, 24, seed=seed)
            pkg.register_dataset("same-name", X, y)
            r = m.train(GridSearchCV(LogisticRegression(), grid, cv=4), "same-name", train_params={"test_size": 0.25, "random_state": 0},
                        wait_for_completion=True)
            assert r["job_status"] == "completed", r
            gs = GridSearchCV(LogisticRegression(), grid, cv=4).fit(X.astype(np.float64), y)
            np.testing.assert_allclose(r["job_result"]["cv_results_"]["mean_test_score"], gs.cv_results_["mean_test_score"], atol=RTOL)
            assert all(res["executor"] == "b200" for res in r["job_result"]["results"])
    finally:
        m.close()


def test_page_lock_of_a_registered_array_ends_with_the_array():
    """The executor page-locks a registered float32 dataset in place (cudaHostRegister).  When the array is freed
    (its name re-registered, the process moving on) the registration has to go with it: the allocator hands the
    address range out again, and a host buffer overlapping a stale registration fails the next cudaMemcpy with
    "invalid argument"."""
    import gc
    from cs230_distributed_machine_learning_b200 import worker as W
    X = np.random.default_rng(5).standard_normal((300_000, 64), dtype=np.float32)        # 77 MB: its own mmap region
    key, ptr = id(X), X.ctypes.data
    W.GpuExecutor._pin(None, X)                              # (uses no executor state)
    assert W._PINNED.get(key) == ptr
    del X
    gc.collect()
    assert key not in W._PINNED
    Y = np.empty((300_000, 64), dtype=np.float32)            # very likely the same address range
    W.GpuExecutor._pin(None, Y)
    assert id(Y) in W._PINNED                                # would fail on a range that still overlaps a registration
    t = torch.ones(300_000, 64, device="cuda")
    Y[:] = t.cpu().numpy()                                    # D2H into fresh host memory
    torch.from_numpy(Y).cuda()                                # and H2D from it
    torch.cuda.synchronize()
    assert float(Y.sum()) == 300_000 * 64


def test_scoring_and_train_scores_on_the_b200_executor():
    """cv_params.scoring / return_train_score (core.py:123-130) on float32 data: the scores come from the sums of
    the GPU scoring pass (test rows) and of one more scoring call over all rows (training rows = all - test)."""
    from sklearn.linear_model import LogisticRegression, Ridge, RidgeClassifier
    from sklearn.model_selection import GridSearchCV
    import cs230_distributed_machine_learning_b200 as pkg
    Xr, yr = make_dataset("regression", 20_000, 40, seed=111)
    pkg.register_dataset("sc-reg", Xr, yr)
    X, y = make_dataset("binary", 20_000, 40, seed=112)
    pkg.register_dataset("sc-bin", X, y)
    Xm, ym = make_dataset("multiclass4", 12_000, 24, seed=113)
    pkg.register_dataset("sc-multi", Xm, ym)
    m = pkg.MLTaskManager("local://1")
    try:
        cases = [("sc-reg", Ridge(), {"alpha": [0.1, 100.0, 1e4]}, "neg_mean_squared_error", Xr, yr.astype(np.float64)),
                 ("sc-reg", Ridge(), {"alpha": [0.1, 100.0, 1e4]}, "r2", Xr, yr.astype(np.float64)),
                 ("sc-bin", LogisticRegression(), {"C": [1e-3, 1.0]}, "accuracy", X, y),
                 ("sc-bin", RidgeClassifier(), {"alpha": [1.0, 1e3]}, None, X, y),
                 ("sc-multi", LogisticRegression(), {"C": [1e-2, 1.0]}, None, Xm, ym)]
        for name, est, grid, scoring, Xh, yh in cases:
            gs = GridSearchCV(est, grid, cv=4, scoring=scoring, return_train_score=True)
            r = m.train(gs, name, wait_for_completion=True)
            assert r["job_status"] == "completed", r
            jr = r["job_result"]
            assert all(res.get("executor") == "b200" for res in jr["results"]), \
                [(res.get("executor"), res.get("executor_reason"), res.get("error")) for res in jr["results"]]
            want = gs.fit(Xh.astype(np.float64), yh).cv_results_
            for key in ("mean_test_score", "mean_train_score", "split1_train_score", "std_train_score"):
                np.testing.assert_allclose(jr["cv_results_"][key], want[key], rtol=2 * RTOL, atol=2 * RTOL, err_msg=f"{name} {scoring} {key}")
            assert jr["best_params_"] == gs.best_params_
    finally:
        m.close()


def test_config3_full_size_vs_cached_oracle(solver, golden_dir):
    """BASELINE.json config 3 at FULL size (4M x 512, KFold 5), 8 of the 128 alphas, against the float64 answers of
    tests/golden/make_cfg3_full.py (the oracle's Cholesky restatement from float64 moments, pinned at full size by one
    real scikit-learn Ridge fit): coefficients, intercepts, squared errors and R^2 within 1e-4."""
    z = np.load(os.path.join(golden_dir, "cfg3_full.npz"))
    meta = json.loads(str(z["meta"]))
    X, y = make_dataset(meta["kind"], meta["n"], meta["d"], seed=meta["seed"])
    kf = folds.kfold_labels(meta["n"], meta["cv"])
    ds = solver.upload_dataset(X, y, classifier=False)
    del X
    out = solver.ridge_sweep(ds, kf, meta["cv"], meta["alpha"], [True] * len(meta["alpha"]), classifier=False)
    assert (out["info"] == 0).all() and (out["n_test"] == z["n_test"][None, :]).all()
    worst = 0.0
    for i in range(len(meta["alpha"])):
        for k in range(meta["cv"]):
            worst = max(worst, rel_err(out["coef"][i, k, 0], z["coef"][i, k]))
            assert abs(out["intercept"][i, k, 0] - z["intercept"][i, k]) <= RTOL
    assert worst <= RTOL, worst
    np.testing.assert_allclose(out["sse"], z["sse"], rtol=RTOL)
    r2 = 1.0 - out["sse"] / (out["syy"] - out["sy"] ** 2 / out["n_test"])
    np.testing.assert_allclose(r2, z["r2"], atol=RTOL)
    assert int(np.argmax(r2.mean(axis=1))) == int(np.argmax(z["r2"].mean(axis=1)))
    print(f"cfg3 full-size parity: 40 fits, max coef rel. err {worst:.2e}, max R2 diff {np.abs(r2 - z['r2']).max():.2e}")


def test_config4_full_size_vs_cached_oracle(solver, golden_dir):
    """BASELINE.json config 4 at FULL size (10-class multinomial, 2M x 512, StratifiedKFold 5), 8 of the 512 sampled
    candidates (C, tol, fit_intercept as the sampler drew them), against scikit-learn on float64
    (tests/golden/make_cfg4_full.py): n_iter / nfev, coefficients and intercepts within 1e-4, per-fold correct counts
    exact except rows whose two largest decision values are within 2e-5 of each other."""
    import bench
    path = os.path.join(golden_dir, "cfg4_full.npz")
    z = np.load(path)
    meta = json.loads(str(z["meta"]))
    X, y = make_dataset(meta["kind"], meta["n"], meta["d"], seed=meta["seed"])
    cv, params = meta["cv"], meta["params"]
    fold = folds.stratified_labels(y, cv)
    ds = solver.upload_dataset(X, y, classifier=True)
    m = len(params)
    out = solver.logreg_sweep(ds, fold, cv, [p["C"] for p in params], [p["tol"] for p in params],
                              [p["fit_intercept"] for p in params], [100] * m)
    rep = bench.parity_report_multi(out, X, fold, cv, path, pos=range(m))
    print("cfg4 full-size parity:", json.dumps(rep))
    gout = os.path.join(os.path.dirname(golden_dir), "..", "gpurun_out")
    if os.path.isdir(gout):
        json.dump(rep, open(os.path.join(gout, "parity_cfg4_full.json"), "w"))
    assert rep["n_test_equal"] and (out["status"] == 0).all()
    assert rep["fits_checked"] == m * cv
    assert rep["n_iter_equal"] >= 0.95 * rep["fits_checked"], rep       # a fit one iterate apart is counted, not hidden
    assert rep["max_coef_relerr"] <= RTOL and rep["max_intercept_err"] <= RTOL, rep
    assert rep["count_diffs_beyond_ties"] == 0, rep
    assert rep["max_mean_score_diff"] <= RTOL, rep
