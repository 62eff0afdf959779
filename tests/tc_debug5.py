"""Manual diagnostic: Ridge sweep, tcgen05 Gram vs FFMA vs sklearn."""
import sys
import numpy as np
sys.path.insert(0, ".")
from cs230_distributed_machine_learning_b200 import folds, make_dataset
from cs230_distributed_machine_learning_b200.solver import SweepSolver
from sklearn.linear_model import Ridge, RidgeClassifier
s = SweepSolver(0)
for (kind, n, d, clf) in [("regression", 5000, 64, False), ("regression", 40000, 200, False), ("binary", 30000, 128, True), ("multiclass4", 20000, 96, True), ("regression", 200000, 512, False)]:
    X, y = make_dataset(kind, n, d, seed=3)
    fold = folds.stratified_labels(y, 5) if clf else folds.kfold_labels(n, 5)
    ds = s.upload_dataset(X, y, classifier=clf)
    alphas = np.logspace(-2, 3, 6)
    res = {tc: s.ridge_sweep(ds, fold, 5, alphas, [True] * 6, classifier=clf, use_tensor_cores=tc) for tc in (0, 1)}
    a, b = res[0], res[1]
    err = np.abs(a["coef"] - b["coef"]).max() / np.abs(a["coef"]).max()
    extra = f"correct equal {(a['correct']==b['correct']).mean():.2f}" if clf else f"sse rel {np.abs(a['sse']/b['sse']-1).max():.1e}"
    print(f"{kind} n={n} d={d}: coef relerr tc-vs-ffma {err:.2e}; icpt {np.abs(a['intercept']-b['intercept']).max():.1e}; n_test eq {(a['n_test']==b['n_test']).all()}; info {b['info'].max()}; {extra}; ms ffma {a['timing']['total_ms']:.1f} tc {b['timing']['total_ms']:.1f}", flush=True)
    k = 2
    M = (RidgeClassifier if clf else Ridge)(alpha=alphas[3]).fit(X[fold != k].astype(np.float64), y[fold != k].astype(np.float64) if not clf else y[fold != k])
    for tc in (0, 1):
        e = np.abs(res[tc]["coef"][3, k] - np.atleast_2d(M.coef_)).max() / np.abs(M.coef_).max()
        print(f"   sklearn: tc={tc} relerr {e:.2e}")
s.close()
