"""Small end-to-end run for compute-sanitizer (manual): every kernel family once, ragged shapes."""
import sys
import numpy as np
sys.path.insert(0, ".")
from cs230_distributed_machine_learning_b200 import folds, make_dataset
from cs230_distributed_machine_learning_b200.solver import SweepSolver
s = SweepSolver(0)
for tc in (1, 0):
    X, y = make_dataset("binary", 1000, 40, seed=1)
    ds = s.upload_dataset(X, y, classifier=True)
    f = folds.stratified_labels(y, 3)
    h = folds.holdout_labels(1000, 0.25, 1)
    o = s.logreg_sweep(ds, f, 3, [0.01, 1.0], [1e-4] * 2, [True, False], [20, 20], hold_id=h, use_tensor_cores=tc)
    X, y = make_dataset("multiclass5", 900, 36, seed=2)
    ds = s.upload_dataset(X, y, classifier=True)
    o = s.logreg_sweep(ds, folds.stratified_labels(y, 3), 3, [0.1, 1.0], [1e-4] * 2, [True, True], [10, 10], use_tensor_cores=tc)
    o = s.ridge_sweep(ds, folds.stratified_labels(y, 3), 3, [0.1, 10.0], [True, True], classifier=True, use_tensor_cores=tc)
    X, y = make_dataset("regression", 1100, 44, seed=3)
    ds = s.upload_dataset(X, y, classifier=False)
    o = s.ridge_sweep(ds, folds.kfold_labels(1100, 4), 4, [0.1, 10.0], [True, False], classifier=False,
                      hold_id=folds.holdout_labels(1100, 0.2, 0), use_tensor_cores=tc)
    print("tc", tc, "ok", flush=True)
s.close()
