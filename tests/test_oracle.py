"""The oracle against scikit-learn / scipy and against the committed golden fixtures
(produced by importing the reference, tests/golden/make_golden.py)."""
import json
import os
import warnings

import numpy as np
import pytest
from sklearn.linear_model import LogisticRegression, Ridge, RidgeClassifier
from sklearn.model_selection import KFold, StratifiedKFold

from oracle import folds as ofolds
from oracle import lbfgs as olbfgs
from oracle import reference_path as rp
from oracle import ridge as oridge
from tests.synth import make_dataset


def test_dcsrch_matches_scipy_twin():
    from scipy.optimize._dcsrch import DCSRCH
    rng = np.random.default_rng(3)
    for trial in range(20):
        a, b, c = rng.uniform(0.5, 3), rng.uniform(-4, -0.5), rng.uniform(0.1, 2)
        phi = lambda t: a * t * t + b * t + c * np.cos(3 * t)
        dphi = lambda t: 2 * a * t + b - 3 * c * np.sin(3 * t)
        stp0 = rng.uniform(0.05, 5)
        want, *_ = DCSRCH(phi, dphi, 1e-3, 0.9, 0.1, 0.0, 1e10)(stp0, phi0=phi(0.0), derphi0=dphi(0.0), maxiter=50)
        ls = olbfgs.Dcsrch()
        stp, task = ls.start(stp0, phi(0.0), dphi(0.0))
        n = 0
        while task == "FG" and n < 50:
            stp, task = ls.step(stp, phi(stp), dphi(stp))
            n += 1
        if want is None:
            assert task == "WARN"
        else:
            assert stp == pytest.approx(want, rel=1e-14, abs=0)


@pytest.mark.parametrize("C", [1e-4, 1e-2, 1.0, 100.0])
def test_lbfgs_restatement_binary(C):
    X, y = make_dataset("binary", 3000, 16, seed=21)
    X = X.astype(np.float64)
    m = LogisticRegression(C=C).fit(X, y)
    r = olbfgs.fit_logreg(X, y, 2, C=C)
    assert r["n_iter"] == m.n_iter_[0]
    np.testing.assert_allclose(r["coef"], m.coef_, rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(r["intercept"], m.intercept_, rtol=1e-9, atol=1e-12)


@pytest.mark.parametrize("C,tol,fi", [(1.0, 1e-4, True), (0.01, 1e-5, False), (10.0, 1e-3, True)])
def test_lbfgs_restatement_multinomial(C, tol, fi):
    X, y = make_dataset("multiclass4", 3000, 12, seed=22)
    X = X.astype(np.float64)
    m = LogisticRegression(C=C, tol=tol, fit_intercept=fi).fit(X, y)
    r = olbfgs.fit_logreg(X, y, 4, C=C, tol=tol, fit_intercept=fi)
    assert r["n_iter"] == m.n_iter_[0]
    np.testing.assert_allclose(r["coef"], m.coef_, rtol=1e-8, atol=1e-11)
    np.testing.assert_allclose(r["intercept"], m.intercept_, rtol=1e-8, atol=1e-11)


def test_lbfgs_restatement_counts_nfev():
    from scipy import optimize
    X, y = make_dataset("binary", 2000, 8, seed=23)
    X = X.astype(np.float64)
    l2 = 1.0 / (1.0 * len(y))
    fg = lambda w: olbfgs.logistic_loss_grad(w, X, y.astype(float), 2, True, l2)
    ref = optimize.minimize(fg, np.zeros(9), method="L-BFGS-B", jac=True,
                            options={"maxiter": 100, "maxls": 50, "gtol": 1e-4, "ftol": 64 * np.finfo(float).eps})
    mine = olbfgs.lbfgs_minimize(fg, np.zeros(9))
    assert (mine["nit"], mine["nfev"]) == (ref.nit, ref.nfev)
    np.testing.assert_allclose(mine["x"], ref.x, rtol=1e-10)


def test_lbfgs_iteration_cap_is_trajectory_defined():
    from sklearn.datasets import load_iris
    X, y = load_iris(return_X_y=True)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = LogisticRegression(C=100.0).fit(X, y)
    r = olbfgs.fit_logreg(X, y, 3, C=100.0)
    assert r["n_iter"] == m.n_iter_[0] == 40 and r["nfev"] == 51   # SURVEY.md 8c hard case
    np.testing.assert_allclose(r["coef"], m.coef_, atol=1e-6)


@pytest.mark.parametrize("n,k,nf", [(1000, 2, 5), (1003, 3, 5), (997, 10, 10), (50, 4, 3)])
def test_fold_ids(n, k, nf):
    y = np.random.default_rng(n).integers(0, k, n)
    want = np.empty(n, dtype=int)
    for i, (_, te) in enumerate(StratifiedKFold(nf).split(y, y)):
        want[te] = i
    assert (ofolds.stratified_fold_ids(y, nf) == want).all()
    for i, (_, te) in enumerate(KFold(nf).split(y)):
        want[te] = i
    assert (ofolds.kfold_ids(n, nf) == want).all()


@pytest.mark.parametrize("alpha", [1e-3, 1.0, 1e3])
def test_ridge_restatement(alpha):
    X, y = make_dataset("regression", 2000, 10, seed=24)
    X = X.astype(np.float64); y = y.astype(np.float64)
    m = Ridge(alpha=alpha).fit(X, y)
    W, b = oridge.fit_ridge(X, y, alpha)
    np.testing.assert_allclose(W[0], m.coef_, rtol=1e-10)
    np.testing.assert_allclose(b[0], m.intercept_, rtol=1e-9, atol=1e-12)
    Xc, yc = make_dataset("multiclass3", 1500, 8, seed=25)
    mc = RidgeClassifier(alpha=alpha).fit(Xc.astype(np.float64), yc)
    W, b = oridge.fit_ridge(Xc, oridge.classifier_targets(yc, 3), alpha)
    np.testing.assert_allclose(W, mc.coef_, rtol=1e-9, atol=1e-12)


def test_reference_train_model_fixture(golden_dir):
    """oracle.reference_path.train_model == the imported reference worker.train_model."""
    from sklearn.datasets import load_iris
    g = json.load(open(os.path.join(golden_dir, "train_model_iris.json")))
    sets = {"iris": load_iris(return_X_y=True),
            "binary_2000x16_seed11": make_dataset("binary", 2000, 16, seed=11),
            "regression_1500x12_seed12": make_dataset("regression", 1500, 12, seed=12)}
    for name, rec in g.items():
        X, y = sets[name]
        for task, want in zip(rec["tasks"], rec["results"]):
            got = rp.train_model(task, X, y)
            # accuracies are ratios of integers -> exact; the fp32 regression set goes through a
            # pandas DataFrame in the reference (worker.py:425-429), which changes BLAS summation order
            rtol = 1e-12 if "accuracy" in want else 1e-6
            np.testing.assert_allclose(got["cv_scores"], want["cv_scores"], rtol=rtol)
            assert got["mean_cv_score"] == pytest.approx(want["mean_cv_score"], rel=rtol)
            for key in ("accuracy", "r2_score", "mean_squared_error"):
                if key in want:
                    assert got[key] == pytest.approx(want[key], rel=rtol)


def test_iris_known_answers():
    """SURVEY.md 8c table (sklearn 1.9.0, config 1 of BASELINE.json)."""
    from sklearn.datasets import load_iris
    X, y = load_iris(return_X_y=True)
    cands = [{"C": c, "solver": "lbfgs"} for c in (0.1, 1.0, 10.0, 100)]
    res = rp.sweep("LogisticRegression", {}, cands, X, y, cv=5)
    counts = [[f["correct"] for f in row] for row in res["fits"]]
    assert counts == [[26, 29, 28, 29, 30], [29, 30, 28, 29, 30], [30, 30, 29, 28, 30], [30, 30, 29, 28, 30]]
    assert res["best_index"] == 2 and res["best_score"] == pytest.approx(0.9800000000000001, abs=0)
    assert list(res["rank"]) == [4, 3, 1, 1]


def test_small_config_fixtures_vs_restatement(golden_dir):
    """fp64 restatement vs the committed sklearn (float64) answers: agreement to ~1e-9."""
    z = np.load(os.path.join(golden_dir, "sweep_small.npz"))
    meta = json.loads(str(z["cfg2_meta"]))
    X, y = make_dataset(meta["kind"], meta["n"], meta["d"], seed=meta["seed"])
    fold = ofolds.stratified_fold_ids(y, meta["cv"])
    for ci in range(8):
        for k in range(5):
            tr = fold != k
            r = olbfgs.fit_logreg(X[tr].astype(np.float64), y[tr], 2, C=meta["candidates"][ci]["C"])
            want = z["cfg2_coef"][ci, k]
            assert np.abs(r["coef"] - want).max() <= 1e-8 * np.abs(want).max() + 1e-12
            assert r["n_iter"] == z["cfg2_n_iter"][ci, k]
