"""Full-size oracle cache for BASELINE.json config 2 (and, with --config, reduced-n slices of 3/4/5).

    python tests/golden/make_cfg_full.py            # config 2: 1 000 000 x 256, 64 C x cv 5 (about 10-15 CPU-minutes)

Runs the reference worker's call sequence (oracle/reference_path.py, i.e.
aws-prod/worker/worker.py:315-326 per candidate, driven by
aws-prod/master/task_handler.py:186-207) on scikit-learn with the SAME bytes the
GPU gets (``make_dataset('binary', n, d, seed=2)``), handed over as float64 like the
reference's ``pd.read_csv`` does, and commits what the parity tests and bench.py
compare against:

  coef [64,5,256] f64, intercept [64,5] f64, n_iter [64,5], nfev [64,5] (function
  evaluations = calls of LinearModelLoss.loss_gradient during the fit), correct /
  n_test [64,5] int64, mean_test_score [64], best_index (rankdata(-mean,'min').argmin()).

The GPU box never runs this: it has no /root/reference and the file is committed.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def count_nfev():
    """Wrap LinearModelLoss.loss_gradient so that every fit reports its function evaluations."""
    from sklearn.linear_model import _linear_loss
    counter = {"n": 0}
    orig = _linear_loss.LinearModelLoss.loss_gradient

    def wrapped(self, *a, **k):
        counter["n"] += 1
        return orig(self, *a, **k)
    _linear_loss.LinearModelLoss.loss_gradient = wrapped
    return counter


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--features", type=int, default=256)
    ap.add_argument("--candidates", type=int, default=64)
    ap.add_argument("--cv", type=int, default=5)
    ap.add_argument("--seed", type=int, default=2)
    ap.add_argument("--out", default=os.path.join(HERE, "cfg2_full.npz"))
    a = ap.parse_args()
    from cs230_distributed_machine_learning_b200 import make_dataset
    from oracle import reference_path as rp
    counter = count_nfev()
    X32, y = make_dataset("binary", a.rows, a.features, seed=a.seed)
    X = X32.astype(np.float64)
    del X32
    Cs = np.logspace(-6, 2, a.candidates)
    F = (a.candidates, a.cv)
    coef = np.zeros(F + (a.features,)); icpt = np.zeros(F)
    n_iter = np.zeros(F, np.int64); nfev = np.zeros(F, np.int64)
    correct = np.zeros(F, np.int64); n_test = np.zeros(F, np.int64)
    t0 = time.time()
    from sklearn.base import clone
    from sklearn.model_selection import StratifiedKFold
    splits = list(StratifiedKFold(a.cv).split(X, y))
    # fold-major loop so that one training-fold copy of X serves all candidates (same fits as cv_fits per candidate)
    for k, (tr, te) in enumerate(splits):
        Xtr, ytr, Xte, yte = X[tr], y[tr], X[te], y[te]
        for i, C in enumerate(Cs):
            model = rp.make_model("LogisticRegression", {"C": float(C)})
            counter["n"] = 0
            m = clone(model).fit(Xtr, ytr)
            nfev[i, k] = counter["n"]
            coef[i, k] = m.coef_[0]; icpt[i, k] = m.intercept_[0]; n_iter[i, k] = int(m.n_iter_[0])
            correct[i, k] = int((m.predict(Xte) == yte).sum()); n_test[i, k] = len(te)
        print(f"fold {k} done, {time.time() - t0:.0f} s", flush=True)
        del Xtr, Xte
    scores = correct / n_test
    mean, std, rank, best = rp.search_rules(scores)
    meta = {"kind": "binary", "n": a.rows, "d": a.features, "seed": a.seed, "cv": a.cv,
            "C": [float(c) for c in Cs], "sklearn": __import__("sklearn").__version__,
            "scipy": __import__("scipy").__version__, "numpy": np.__version__, "seconds": time.time() - t0}
    np.savez_compressed(a.out, coef=coef, intercept=icpt, n_iter=n_iter, nfev=nfev, correct=correct, n_test=n_test,
                        mean_test_score=mean, rank=rank, best_index=np.int64(best), meta=np.array(json.dumps(meta)))
    print("wrote", a.out, "best_index", best, "C", Cs[best], "in", meta["seconds"], "s")


if __name__ == "__main__":
    main()
