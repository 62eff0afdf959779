"""Manual diagnostic: one batched loss/gradient evaluation, tcgen05 vs FFMA vs numpy fp64."""
import sys
import numpy as np
sys.path.insert(0, ".")
from cs230_distributed_machine_learning_b200 import folds, make_dataset
from cs230_distributed_machine_learning_b200.solver import SweepSolver

s = SweepSolver(0)
rng = np.random.default_rng(0)
for (n, d, K, wscale) in [(256, 32, 32, 0.0), (256, 32, 32, 0.3), (1000, 64, 20, 0.3), (4096, 256, 160, 0.1), (5000, 100, 37, 0.2), (8192, 256, 256, 0.1)]:
    X, y = make_dataset("binary", n, d, seed=9)
    fold = folds.stratified_labels(y, 5)
    ds = s.upload_dataset(X, y, classifier=True)
    W = (rng.standard_normal((K, d)) * wscale / np.sqrt(d)).astype(np.float32)
    b = (rng.standard_normal(K) * wscale).astype(np.float32)
    sel = np.zeros(K, np.int32); val = (np.arange(K) % 5).astype(np.int32)
    X64 = X.astype(np.float64)
    Z = X64 @ W.T.astype(np.float64) + b
    loss = np.logaddexp(0, Z) - y[:, None] * Z
    Rm = 1 / (1 + np.exp(-Z)) - y[:, None]
    tr = fold[:, None] != val[None, :]
    want = np.concatenate([(Rm * tr).T @ X64, (Rm * tr).sum(0)[:, None], (loss * tr).sum(0)[:, None]], axis=1)
    for tc in (0, 1):
        got = s.logreg_eval(ds, fold, W, b, sel, val, use_tensor_cores=tc)
        eG = np.abs(got[:, :d] - want[:, :d]).max() / np.abs(want[:, :d]).max()
        eb = np.abs(got[:, d] - want[:, d]).max() / np.abs(want[:, d]).max()
        el = np.abs(got[:, d + 1] - want[:, d + 1]).max() / np.abs(want[:, d + 1]).max()
        print(f"n={n} d={d} K={K} wscale={wscale} tc={tc}: relerr G {eG:.2e} gb {eb:.2e} loss {el:.2e}", flush=True)
        if tc == 1 and eG > 1e-3:
            bad = np.abs(got[:, :d] - want[:, :d]) > 1e-3 * np.abs(want[:, :d]).max()
            print("   bad G entries:", bad.sum(), "of", bad.size, "; slots with bad:", np.nonzero(bad.any(1))[0][:20],
                  "features with bad:", np.nonzero(bad.any(0))[0][:40])
            print("   got[0,:8]", got[0, :8], "\n   want[0,:8]", want[0, :8])
            ratio = got[:, :d] / want[:, :d]
            print("   ratio quantiles", np.nanquantile(ratio, [0.01, 0.25, 0.5, 0.75, 0.99]))
s.close()
