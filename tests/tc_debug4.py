"""Manual diagnostic: multinomial sweep, tcgen05 vs FFMA vs sklearn."""
import sys, warnings
import numpy as np
sys.path.insert(0, ".")
from cs230_distributed_machine_learning_b200 import folds, make_dataset
from cs230_distributed_machine_learning_b200.solver import SweepSolver
from sklearn.linear_model import LogisticRegression
s = SweepSolver(0)
for (kind, n, d, nc) in [("multiclass10", 20000, 64, 6), ("multiclass5", 9000, 40, 4), ("multiclass3", 5000, 32, 3), ("multiclass10", 30000, 512, 4)]:
    X, y = make_dataset(kind, n, d, seed=8)
    fold = folds.stratified_labels(y, 5)
    ds = s.upload_dataset(X, y, classifier=True)
    Cs = np.logspace(-3, 1, nc); tols = [1e-4, 1e-5, 1e-3, 1e-4, 1e-4, 1e-4][:nc]; fis = [True, False, True, True, False, True][:nc]
    res = {}
    for tc in (0, 1):
        try:
            res[tc] = s.logreg_sweep(ds, fold, 5, Cs, tols, fis, [100] * nc, use_tensor_cores=tc)
        except Exception as e:
            print(kind, "tc", tc, "error", e); res[tc] = None
    a, b = res[0], res[1]
    if b is None: continue
    den = np.abs(a["coef"]).max(axis=(2, 3))
    err = np.abs(a["coef"] - b["coef"]).max(axis=(2, 3)) / den
    print(f"{kind} n={n} d={d}: coef relerr tc-vs-ffma max {err.max():.2e}; n_iter equal {(a['n_iter']==b['n_iter']).mean():.2f}; nfev equal {(a['nfev']==b['nfev']).mean():.2f}; "
          f"correct equal {(a['correct']==b['correct']).mean():.2f} maxdiff {np.abs(a['correct']-b['correct']).max()}; icpt {np.abs(a['intercept']-b['intercept']).max():.1e}; ms ffma {a['timing']['total_ms']:.1f} tc {b['timing']['total_ms']:.1f}", flush=True)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        i, k = 0, 1
        m = LogisticRegression(C=Cs[i], tol=tols[i], fit_intercept=fis[i]).fit(X[fold != k].astype(np.float64), y[fold != k])
        for tc in (0, 1):
            e = np.abs(res[tc]["coef"][i, k] - m.coef_).max() / np.abs(m.coef_).max()
            print(f"   sklearn: tc={tc} relerr {e:.2e} n_iter {res[tc]['n_iter'][i,k]} vs {m.n_iter_[0]} correct {res[tc]['correct'][i,k]} vs {(m.predict(X[fold==k].astype(np.float64))==y[fold==k]).sum()}")
s.close()
