"""Full-size oracle cache for BASELINE.json config 4 (10-class multinomial, 2 000 000 x 512, cv=5), 8 of the 512 candidates.

    python tests/golden/make_cfg4_full.py            # about 20-30 CPU-minutes on 8 cores, 20 GB of RAM

Same recipe as make_cfg_full.py (config 2): the reference worker's call sequence
(oracle/reference_path.py, i.e. aws-prod/worker/worker.py:315-326 per candidate, driven by
aws-prod/master/task_handler.py:209-252 for a RandomizedSearchCV) on scikit-learn with the SAME bytes the
GPU gets (``make_dataset('multiclass10', 2_000_000, 512, seed=4)``), handed over as float64 like the
reference's ``pd.read_csv`` does.  The candidates are the ones bench.py sweeps (``bench.candidates_of``:
ParameterSampler over C x tol x fit_intercept, sampler seed 42); every 64th of the 512 in draw order is fitted
here, on all 5 folds:

  index [8], coef [8,5,10,512] f64, intercept [8,5,10] f64, n_iter [8,5], nfev [8,5] (calls of
  LinearModelLoss.loss_gradient during the fit), correct / n_test [8,5] int64.

The GPU box never runs this: it has no scikit-learn reference data and the file is committed.
"""
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
    d, cv, seed, n_cand = 512, 5, 4, 512
    from make_cfg_full import count_nfev
    import bench
    from cs230_distributed_machine_learning_b200 import make_dataset
    from oracle import reference_path as rp
    from sklearn.base import clone
    from sklearn.model_selection import StratifiedKFold
    counter = count_nfev()
    cands = bench.candidates_of(dict(bench.CONFIGS["4"]), n_cand)
    pick = list(range(0, n_cand, 64))
    X32, y = make_dataset("multiclass10", n, d, seed=seed)
    X = X32.astype(np.float64)
    del X32
    K = int(y.max()) + 1
    F = (len(pick), cv)
    coef = np.zeros(F + (K, d)); icpt = np.zeros(F + (K,))
    n_iter = np.zeros(F, np.int64); nfev = np.zeros(F, np.int64)
    correct = np.zeros(F, np.int64); n_test = np.zeros(F, np.int64)
    t0 = time.time()
    splits = list(StratifiedKFold(cv).split(X, y))
    for k, (tr, te) in enumerate(splits):            # fold-major: one training-fold copy of X serves all candidates
        Xtr, ytr, Xte, yte = X[tr], y[tr], X[te], y[te]
        for i, ci in enumerate(pick):
            model = rp.make_model("LogisticRegression", dict(cands[ci]))
            counter["n"] = 0
            m = clone(model).fit(Xtr, ytr)
            nfev[i, k] = counter["n"]
            coef[i, k] = m.coef_; icpt[i, k] = m.intercept_; n_iter[i, k] = int(np.max(m.n_iter_))
            correct[i, k] = int((m.predict(Xte) == yte).sum()); n_test[i, k] = len(te)
            print(f"fold {k} cand {ci} {cands[ci]} n_iter {n_iter[i, k]} nfev {nfev[i, k]} acc {correct[i, k] / n_test[i, k]:.5f} "
                  f"{time.time() - t0:.0f} s", flush=True)
        del Xtr, Xte
    meta = {"kind": "multiclass10", "n": n, "d": d, "seed": seed, "cv": cv, "n_candidates": n_cand, "index": pick,
            "params": [dict(cands[i]) for i in pick], "sklearn": __import__("sklearn").__version__,
            "scipy": __import__("scipy").__version__, "numpy": np.__version__, "seconds": time.time() - t0}
    out = os.path.join(HERE, "cfg4_full.npz" if n == 2_000_000 else f"cfg4_n{n}.npz")
    np.savez_compressed(out, index=np.array(pick), coef=coef, intercept=icpt, n_iter=n_iter, nfev=nfev, correct=correct,
                        n_test=n_test, meta=np.array(json.dumps(meta)))
    print("wrote", out, "in", meta["seconds"], "s")


if __name__ == "__main__":
    main()
