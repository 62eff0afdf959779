"""Manual timing (not a pytest file): BASELINE.json configs 4 and 5 at full size on one GPU (device-generated data)."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from cs230_distributed_machine_learning_b200 import folds
from cs230_distributed_machine_learning_b200.solver import SweepSolver, DeviceDataset
from sklearn.model_selection import ParameterSampler
which = sys.argv[1]
s = SweepSolver(0)
g = torch.Generator(device="cuda").manual_seed(4)
if which == "cfg5":
    n, d, nc, cv = int(sys.argv[2]) if len(sys.argv) > 2 else 16_000_000, 512, 256, 10
    X = torch.randn(n, d, device="cuda", generator=g)
    w = torch.randn(d, device="cuda", generator=g) / d ** 0.5
    y = ((X @ w + 0.5 * torch.randn(n, device="cuda", generator=g)) > 0).to(torch.int32)
    ds = DeviceDataset(X, y.contiguous(), None, np.array([0, 1]), n, d)
    fold = folds.stratified_labels(y.cpu().numpy(), cv)
    Cs = np.logspace(-6, 2, nc); tol = [1e-4] * nc; fi = [True] * nc
else:
    n, d, nc, cv = int(sys.argv[2]) if len(sys.argv) > 2 else 2_000_000, 512, 512, 5
    X = torch.randn(n, d, device="cuda", generator=g)
    W = torch.randn(d, 10, device="cuda", generator=g) / d ** 0.5
    u = torch.rand(n, 10, device="cuda", generator=g).clamp_(1e-9, 1 - 1e-9)
    y = torch.argmax(X @ W - torch.log(-torch.log(u)), dim=1).to(torch.int32)
    ds = DeviceDataset(X, y.contiguous(), None, np.arange(10), n, d)
    fold = folds.stratified_labels(y.cpu().numpy(), cv)
    grid = {"C": list(np.logspace(-5, 2, 64)), "tol": list(np.logspace(-5, -3, 8)), "fit_intercept": [True, False]}
    cands = list(ParameterSampler(grid, nc, random_state=42))
    Cs = [c["C"] for c in cands]; tol = [c["tol"] for c in cands]; fi = [c["fit_intercept"] for c in cands]
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = s.logreg_sweep(ds, fold, cv, Cs, tol, fi, [100] * nc)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    acc = (out["correct"] / out["n_test"]).mean(1)
    print(f"{which} n={n} d={d} cand={nc} cv={cv}: {dt:.3f} s wall ({nc*cv/dt:.0f} fits/s), rounds {out['rounds']}, column-rounds {out['timing']['column_rounds']}, "
          f"contraction {out['timing']['contraction_ms']/1e3:.3f} s, n_iter mean {out['n_iter'].mean():.2f} max {out['n_iter'].max()}, status!=0: {(out['status']!=0).sum()}, best acc {acc.max():.4f}", flush=True)
