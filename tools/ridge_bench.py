"""Manual timing (not a pytest file): BASELINE.json config 3 shape on one GPU, data generated on device."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from cs230_distributed_machine_learning_b200 import folds
from cs230_distributed_machine_learning_b200.solver import SweepSolver, DeviceDataset
n, d, na = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000, 512, 128
s = SweepSolver(0)
g = torch.Generator(device="cuda").manual_seed(3)
X = torch.randn(n, d, device="cuda", generator=g)
w = torch.randn(d, device="cuda", generator=g) / d ** 0.5
y = X @ w + 0.5 * torch.randn(n, device="cuda", generator=g)
ds = DeviceDataset(X, None, y.contiguous(), None, n, d)
kf = folds.kfold_labels(n, 5)
alphas = np.logspace(-3, 5, na)
for tc in (1, 1, 0):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = s.ridge_sweep(ds, kf, 5, alphas, [True] * na, classifier=False, use_tensor_cores=tc)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    r2 = 1 - out["sse"] / (out["syy"] - out["sy"] ** 2 / out["n_test"])
    print(f"ridge cfg3 n={n} d={d} alphas={na} tc={tc}: {dt*1e3:.1f} ms wall, lib {out['timing']['total_ms']:.1f} ms (Gram+score {out['timing']['contraction_ms']:.1f} ms) -> {na*5/dt:.0f} fits/s; best alpha {alphas[int(np.argmax(r2.mean(1)))]:.3g} info {out['info'].max()}", flush=True)
yc = (y > 0).to(torch.int32)
dsc = DeviceDataset(X, yc.contiguous(), None, np.array([0, 1]), n, d)
sf = folds.stratified_labels(yc.cpu().numpy(), 5)
for tc in (1,):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = s.ridge_sweep(dsc, sf, 5, alphas, [True] * na, classifier=True, use_tensor_cores=tc)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(f"ridge-classifier cfg3 tc={tc}: {dt*1e3:.1f} ms wall, lib {out['timing']['total_ms']:.1f} ms -> {na*5/dt:.0f} fits/s; best acc {(out['correct']/out['n_test']).mean(1).max():.4f}", flush=True)
