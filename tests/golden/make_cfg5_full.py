"""Full-size oracle cache for BASELINE.json config 5 (binary, 16 000 000 x 512, 256 C x cv 10), 8 of the 256 C values.

    python tests/golden/make_cfg5_full.py            # about 25 CPU-minutes on 8 cores, 45 GB of RAM

16M x 512 is 65 GB as float64, more than this container holds, so scikit-learn itself cannot be run on it.  What is
cached instead is the oracle's restatement of ``LogisticRegression(solver="lbfgs").fit`` (oracle/lbfgs.py: L-BFGS-B +
MINPACK-2 dcsrch; equal to scikit-learn to 1e-15 with identical n_iter / nfev in tests/test_oracle.py) with its
loss-and-gradient callback evaluated at FULL size in float64 over row chunks of the float32 data (each chunk widened
exactly, like the reference's ``pd.read_csv`` hands scikit-learn float64).  The 80 fits (8 C values x 10 folds) run
their own optimiser each, in lockstep: every pass over X serves all fits that are waiting for an evaluation.
(a) Before the full run the same machinery is pinned against REAL scikit-learn fits on the first 400 000 rows
(2 C x 2 folds, float64): n_iter and nfev equal, coefficients to 1e-9 -- recorded in ``meta``.
(b) Test-fold counts and the number of rows within 2e-5 of a decision tie are accumulated the same way, so the GPU's
counts can be judged without the 33 GB of X on the host.

Inputs are the bytes the GPU gets: ``datasets.make_dataset_rows('binary', 16_000_000, 512, seed=5, ...)`` (the
chunk-seeded generator bench.py uses for this config), folds = StratifiedKFold(10) labels (``folds.stratified_labels``).
"""
import json
import os
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

CHUNK = 1 << 16
EPS_TIE = 2e-5


class Lockstep:
    """Runs one oracle L-BFGS per fit in its own thread; their loss/gradient requests are answered together by one
    chunked float64 pass over X per round."""

    def __init__(self, X, y, fold, fit_fold, l2, n_tr, workers):
        self.X, self.y, self.fold = X, y.astype(np.float64), fold
        self.fit_fold, self.l2, self.n_tr = np.asarray(fit_fold), np.asarray(l2), np.asarray(n_tr, dtype=np.float64)
        self.F, self.d = len(fit_fold), X.shape[1]
        self.cv, self.want, self.answer = threading.Condition(), {}, {}
        self.running, self.passes, self.workers = self.F, 0, workers

    # ---- called from the fit threads
    def fg(self, i, w):
        with self.cv:
            self.want[i] = np.array(w, dtype=np.float64)
            self.cv.notify_all()
            while i not in self.answer:
                self.cv.wait()
            return self.answer.pop(i)

    def done(self, i):
        with self.cv:
            self.running -= 1
            self.cv.notify_all()

    # ---- one pass for the fits in `ids`
    def _chunk(self, a, b, ids, W, bias):
        Xc = self.X[a:b].astype(np.float64)
        yc = self.y[a:b]
        Z = Xc @ W.T + bias[None, :]
        tr = self.fold[a:b][:, None] != self.fit_fold[ids][None, :]
        e = np.exp(-np.abs(Z))
        loss = np.where(Z > 0, Z, 0.0) + np.log1p(e) - yc[:, None] * Z          # log(1 + e^z) - y z
        p = np.where(Z >= 0, 1.0 / (1.0 + e), e / (1.0 + e))
        R = np.where(tr, p - yc[:, None], 0.0)
        return np.where(tr, loss, 0.0).sum(axis=0), R.T @ Xc, R.sum(axis=0)

    def evaluate(self, ids, Wfull):
        d = self.d
        W, bias = np.ascontiguousarray(Wfull[:, :d]), Wfull[:, d].copy()
        n = self.X.shape[0]
        spans = [(a, min(n, a + CHUNK)) for a in range(0, n, CHUNK)]
        from threadpoolctl import threadpool_limits
        with threadpool_limits(limits=1), ThreadPoolExecutor(self.workers) as ex:
            parts = list(ex.map(lambda s: self._chunk(s[0], s[1], ids, W, bias), spans))
        fs = np.zeros(len(ids)); G = np.zeros((len(ids), d)); gb = np.zeros(len(ids))
        for f_, G_, gb_ in parts:                                                 # fixed order: reproducible sums
            fs += f_; G += G_; gb += gb_
        n_tr, l2 = self.n_tr[ids], self.l2[ids]
        f = fs / n_tr + 0.5 * l2 * (W * W).sum(axis=1)
        out = np.empty((len(ids), d + 1))
        out[:, :d] = G / n_tr[:, None] + l2[:, None] * W
        out[:, d] = gb / n_tr
        return f, out

    def serve(self, log=None):
        t0 = time.time()
        while True:
            with self.cv:
                while len(self.want) < self.running:
                    self.cv.wait()
                if self.running == 0:
                    return
                ids = np.array(sorted(self.want)); Wfull = np.stack([self.want[i] for i in ids])
                self.want.clear()
            f, g = self.evaluate(ids, Wfull)
            self.passes += 1
            if log:
                log(f"pass {self.passes}: {len(ids)} fits, {time.time() - t0:.0f} s")
            with self.cv:
                for j, i in enumerate(ids):
                    self.answer[int(i)] = (float(f[j]), g[j].copy())
                self.cv.notify_all()


def run_fits(X, y, fold, cv, Cs, tol=1e-4, max_iter=100, workers=8, log=None):
    """All len(Cs) x cv fits (fit_intercept=True); returns coef [m,cv,d], intercept, n_iter, nfev, status, passes."""
    from oracle import lbfgs as olbfgs
    m, d = len(Cs), X.shape[1]
    fits = [(i, k) for i in range(m) for k in range(cv)]
    n_tr = np.array([(fold != k).sum() for _, k in fits])
    l2 = np.array([1.0 / (Cs[i] * n_tr[j]) for j, (i, _) in enumerate(fits)])
    ls = Lockstep(X, y, fold, [k for _, k in fits], l2, n_tr, workers)
    res = [None] * len(fits)

    def one(j):
        try:
            res[j] = olbfgs.lbfgs_minimize(lambda w: ls.fg(j, w), np.zeros(d + 1), tol=tol, max_iter=max_iter)
        finally:
            ls.done(j)
    threads = [threading.Thread(target=one, args=(j,), daemon=True) for j in range(len(fits))]
    for t in threads:
        t.start()
    ls.serve(log)
    for t in threads:
        t.join()
    coef = np.zeros((m, cv, d)); icpt = np.zeros((m, cv))
    n_iter = np.zeros((m, cv), np.int64); nfev = np.zeros((m, cv), np.int64); status = np.zeros((m, cv), np.int64)
    for j, (i, k) in enumerate(fits):
        r = res[j]
        coef[i, k] = r["x"][:d]; icpt[i, k] = r["x"][d]
        n_iter[i, k] = min(r["nit"], max_iter); nfev[i, k] = r["nfev"]; status[i, k] = r["status"]
    return coef, icpt, n_iter, nfev, status, ls.passes


def score_fits(X, y, fold, cv, coef, icpt, workers=8):
    """Test-fold correct counts, n_test and rows within EPS_TIE of a decision tie, float64 over row chunks."""
    m, _, d = coef.shape
    W = coef.reshape(m * cv, d); b = icpt.reshape(m * cv)
    fit_fold = np.tile(np.arange(cv), m)
    n = X.shape[0]
    spans = [(a, min(n, a + CHUNK)) for a in range(0, n, CHUNK)]

    def zmax(s):
        Z = X[s[0]:s[1]].astype(np.float64) @ W.T + b[None, :]
        te = fold[s[0]:s[1]][:, None] == fit_fold[None, :]
        return np.where(te, np.abs(Z), 0.0).max(axis=0)

    def counts(s, lim):
        Z = X[s[0]:s[1]].astype(np.float64) @ W.T + b[None, :]
        te = fold[s[0]:s[1]][:, None] == fit_fold[None, :]
        ok = ((Z > 0) == (y[s[0]:s[1]][:, None] == 1)) & te
        return ok.sum(axis=0), te.sum(axis=0), ((np.abs(Z) < lim[None, :]) & te).sum(axis=0)
    from threadpoolctl import threadpool_limits
    with threadpool_limits(limits=1), ThreadPoolExecutor(workers) as ex:
        zm = np.max(list(ex.map(zmax, spans)), axis=0)
        lim = EPS_TIE * np.maximum(1.0, zm)
        parts = list(ex.map(lambda s: counts(s, lim), spans))
    correct = sum(p[0] for p in parts); n_test = sum(p[1] for p in parts); ties = sum(p[2] for p in parts)
    return (correct.reshape(m, cv).astype(np.int64), n_test.reshape(m, cv).astype(np.int64),
            ties.reshape(m, cv).astype(np.int64))


def pin_against_sklearn(X, y, log):
    """(a): the lockstep machinery against real scikit-learn fits (float64) on a slice."""
    from sklearn.linear_model import LogisticRegression
    from cs230_distributed_machine_learning_b200 import folds
    from make_cfg_full import count_nfev
    counter = count_nfev()
    n = min(400_000, X.shape[0])
    Xs, ys = X[:n], y[:n]
    cv, Cs = 2, [1e-4, 1.0]
    fold = folds.stratified_labels(ys, cv)
    coef, icpt, n_iter, nfev, _, _ = run_fits(Xs, ys, fold, cv, Cs)
    X64 = Xs.astype(np.float64)
    worst, same = 0.0, True
    for i, C in enumerate(Cs):
        for k in range(cv):
            counter["n"] = 0
            m = LogisticRegression(C=C).fit(X64[fold != k], ys[fold != k])
            same &= int(m.n_iter_[0]) == n_iter[i, k] and counter["n"] == nfev[i, k]
            worst = max(worst, float(np.abs(m.coef_[0] - coef[i, k]).max() / np.abs(m.coef_[0]).max()),
                        abs(float(m.intercept_[0]) - icpt[i, k]))
    log(f"pin against scikit-learn on {n} rows: n_iter/nfev equal {same}, worst coef / intercept diff {worst:.2e}")
    assert same and worst < 1e-9
    return {"rows": n, "C": Cs, "cv": cv, "n_iter_nfev_equal": bool(same), "max_coef_reldiff": worst}


def main():
    sys.path.insert(0, HERE)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000
    d, cv, seed, n_cand = 512, 10, 5, 256
    from cs230_distributed_machine_learning_b200 import folds
    from cs230_distributed_machine_learning_b200.datasets import CHUNK_ROWS, make_dataset_rows
    t0 = time.time()
    log = lambda s: print(f"[{time.time() - t0:6.0f} s] {s}", flush=True)
    X = np.empty((n, d), np.float32); y = np.empty(n, np.int32)
    step = 8 * CHUNK_ROWS
    for a0 in range(0, n, step):
        b0 = min(n, a0 + step)
        Xc, yc = make_dataset_rows("binary", n, d, seed, a0, b0, threads=8)
        X[a0:b0] = Xc; y[a0:b0] = yc
    log("data generated")
    pin = pin_against_sklearn(X, y, log)
    fold = folds.stratified_labels(y, cv)
    Cs_all = np.logspace(-6, 2, n_cand)
    pick = [0, 36, 73, 109, 146, 182, 219, 255]
    Cs = [float(Cs_all[i]) for i in pick]
    coef, icpt, n_iter, nfev, status, passes = run_fits(X, y, fold, cv, Cs, log=log)
    log(f"fits done in {passes} passes over X; n_iter {n_iter.min()}..{n_iter.max()}")
    correct, n_test, ties = score_fits(X, y, fold, cv, coef, icpt)
    log("scored")
    meta = {"kind": "binary", "n": n, "d": d, "seed": seed, "cv": cv, "n_candidates": n_cand, "index": pick, "C": Cs,
            "eps_tie": EPS_TIE, "passes": passes, "pin": pin, "sklearn": __import__("sklearn").__version__,
            "numpy": np.__version__, "seconds": time.time() - t0}
    out = os.path.join(HERE, "cfg5_full.npz" if n == 16_000_000 else f"cfg5_n{n}.npz")
    np.savez_compressed(out, index=np.array(pick), coef=coef, intercept=icpt, n_iter=n_iter, nfev=nfev, status=status,
                        correct=correct, n_test=n_test, ties=ties, meta=np.array(json.dumps(meta)))
    log(f"wrote {out}")


if __name__ == "__main__":
    main()
