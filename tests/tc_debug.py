"""Manual diagnostic (not a pytest file): tcgen05 path vs FFMA path vs oracle on small problems."""
import sys
import numpy as np
sys.path.insert(0, ".")
from cs230_distributed_machine_learning_b200 import folds, make_dataset
from cs230_distributed_machine_learning_b200.solver import SweepSolver
from oracle import lbfgs as olbfgs

s = SweepSolver(0)
for (n, d, nc, mi) in [(1024, 64, 4, 1), (1024, 64, 4, 3), (5000, 96, 8, 100), (20000, 256, 64, 100), (3333, 40, 3, 100)]:
    X, y = make_dataset("binary", n, d, seed=5)
    fold = folds.stratified_labels(y, 5)
    ds = s.upload_dataset(X, y, classifier=True)
    Cs = np.logspace(-4, 1, nc)
    res = {}
    for tc in (0, 1):
        res[tc] = s.logreg_sweep(ds, fold, 5, Cs, [1e-4] * nc, [True] * nc, [mi] * nc, use_tensor_cores=tc)
    a, b = res[0], res[1]
    den = np.abs(a["coef"]).max(axis=(2, 3)) + 1e-30
    err = np.abs(a["coef"] - b["coef"]).max(axis=(2, 3)) / den
    print(f"n={n} d={d} cand={nc} max_iter={mi}: coef relerr tc-vs-ffma max {err.max():.3e} median {np.median(err):.3e}; "
          f"n_iter equal {(a['n_iter'] == b['n_iter']).mean():.2f}; nfev equal {(a['nfev'] == b['nfev']).mean():.2f}; "
          f"correct equal {(a['correct'] == b['correct']).mean():.2f}; icpt err {np.abs(a['intercept'] - b['intercept']).max():.2e}; "
          f"ms ffma {a['timing']['total_ms']:.2f} tc {b['timing']['total_ms']:.2f}", flush=True)
    if mi == 100:
        k = 0
        for i in (0, nc - 1):
            r = olbfgs.fit_logreg(X[fold != k].astype(np.float64), y[fold != k], 2, C=Cs[i])
            for tc in (0, 1):
                e = np.abs(res[tc]["coef"][i, k] - r["coef"]).max() / np.abs(r["coef"]).max()
                print(f"   oracle C={Cs[i]:.1e}: tc={tc} relerr {e:.2e} n_iter {res[tc]['n_iter'][i, k]} vs {r['n_iter']}", flush=True)
s.close()
