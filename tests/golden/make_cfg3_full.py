"""Full-size oracle cache for BASELINE.json config 3 (Ridge, 4 000 000 x 512, cv=5), a subset of the alphas.

    python tests/golden/make_cfg3_full.py            # about 5 CPU-minutes, 30 GB of RAM

scikit-learn's Ridge(solver="cholesky") on 3.2M x 512 float64 rows needs a copy of X per fit and 640 fits; what is
cached here instead is (a) the oracle's restatement of that solver (oracle/ridge.py: centre, X^T X + alpha I,
Cholesky solve; checked against scikit-learn to 1e-16 in tests/test_oracle.py) evaluated at FULL size from float64
moments accumulated over row chunks -- for every fold and 8 of the 128 alphas -- and (b) one real
``sklearn.linear_model.Ridge`` fit at full size (fold 0, one alpha, X handed over as float64) that pins (a) there.
Test-fold sums (sse, sum y, sum y^2) are accumulated the same way, so the R^2 of every cached fit is known.

Inputs are the bytes the GPU gets: ``make_dataset('regression', 4_000_000, 512, seed=3)`` (float32), seen as float64.
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def main():
    from cs230_distributed_machine_learning_b200 import folds, make_dataset
    n, d, cv, seed = 4_000_000, 512, 5, 3
    if len(sys.argv) > 1:
        n = int(sys.argv[1])
    t0 = time.time()
    X, y = make_dataset("regression", n, d, seed=seed)
    y = y.astype(np.float32)
    fold = folds.kfold_labels(n, cv)
    alphas_all = np.logspace(-3, 5, 128)
    pick = [0, 18, 36, 54, 72, 90, 108, 127]
    # per-fold float64 moments over row chunks: Z = [X | y | 1]
    D = d + 2
    M = np.zeros((cv, D, D))
    chunk = 1 << 17
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        Z = np.empty((b - a, D))
        Z[:, :d] = X[a:b]; Z[:, d] = y[a:b]; Z[:, d + 1] = 1.0
        for k in np.unique(fold[a:b]):
            Zk = Z[fold[a:b] == k]
            M[k] += Zk.T @ Zk
    Mtot = M.sum(axis=0)
    print(f"moments {time.time() - t0:.0f} s", flush=True)
    coef = np.zeros((len(pick), cv, d)); icpt = np.zeros((len(pick), cv))
    for k in range(cv):
        Mtr = Mtot - M[k]
        n_tr = Mtr[D - 1, D - 1]
        mu = Mtr[:d, D - 1] / n_tr; ym = Mtr[d, D - 1] / n_tr
        A0 = Mtr[:d, :d] - n_tr * np.outer(mu, mu)            # centred X^T X   (oracle/ridge.py fit_ridge)
        rhs = Mtr[:d, d] - n_tr * mu * ym                      # centred X^T y
        for i, ai in enumerate(pick):
            A = A0.copy(); A.flat[:: d + 1] += alphas_all[ai]
            L = np.linalg.cholesky(A)
            w = np.linalg.solve(L.T, np.linalg.solve(L, rhs))
            coef[i, k] = w; icpt[i, k] = ym - w @ mu
    # test-fold sums in float64, chunked
    sse = np.zeros((len(pick), cv)); sy = np.zeros(cv); syy = np.zeros(cv); n_test = np.zeros(cv, np.int64)
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        Xc = X[a:b].astype(np.float64); yc = y[a:b].astype(np.float64)
        for k in np.unique(fold[a:b]):
            m = fold[a:b] == k
            pred = Xc[m] @ coef[:, k, :].T + icpt[:, k][None, :]
            sse[:, k] += ((yc[m][:, None] - pred) ** 2).sum(axis=0)
            sy[k] += yc[m].sum(); syy[k] += (yc[m] ** 2).sum(); n_test[k] += int(m.sum())
    r2 = 1.0 - sse / (syy - sy ** 2 / n_test)[None, :]
    print(f"solves + scores {time.time() - t0:.0f} s", flush=True)
    # (b) one scikit-learn fit at full size
    from sklearn.linear_model import Ridge
    tr = fold != 0
    m = Ridge(alpha=float(alphas_all[pick[3]])).fit(X[tr].astype(np.float64), y[tr].astype(np.float64))
    pin = float(np.abs(m.coef_ - coef[3, 0]).max() / np.abs(m.coef_).max())
    print(f"sklearn Ridge at full size vs restatement: coef rel. diff {pin:.2e}, intercept diff {abs(m.intercept_ - icpt[3, 0]):.2e}")
    assert pin < 1e-9
    meta = {"kind": "regression", "n": n, "d": d, "seed": seed, "cv": cv, "alpha_index": pick,
            "alpha": [float(alphas_all[i]) for i in pick], "sklearn_pin_coef_reldiff": pin,
            "sklearn": __import__("sklearn").__version__, "seconds": time.time() - t0}
    out = os.path.join(HERE, "cfg3_full.npz" if n == 4_000_000 else f"cfg3_n{n}.npz")
    np.savez_compressed(out, coef=coef, intercept=icpt, sse=sse, sy=sy, syy=syy, n_test=n_test, r2=r2,
                        meta=np.array(json.dumps(meta)))
    print("wrote", out, f"{time.time() - t0:.0f} s")


if __name__ == "__main__":
    main()
