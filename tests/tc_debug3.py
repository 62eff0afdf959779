"""Manual diagnostic: look inside the scratch of one tcgen05 evaluation."""
import ctypes as C
import sys
import numpy as np
sys.path.insert(0, ".")
from cs230_distributed_machine_learning_b200 import folds, make_dataset
from cs230_distributed_machine_learning_b200.solver import SweepSolver

s = SweepSolver(0)
lib = s.lib
lib.dml_debug_fetch.restype = C.c_int
lib.dml_debug_fetch.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]

def fetch(which):
    av = C.c_size_t(0)
    lib.dml_debug_fetch(s.ctx.handle, which, None, 0, C.byref(av))
    buf = np.empty(av.value // 4, dtype=np.float32)
    lib.dml_debug_fetch(s.ctx.handle, which, buf.ctypes.data_as(C.c_void_p), av.value, None)
    return buf

rng = np.random.default_rng(0)
n, d, K = 256, 32, 32
X, y = make_dataset("binary", n, d, seed=9)
fold = folds.stratified_labels(y, 5)
ds = s.upload_dataset(X, y, classifier=True)
W = (rng.standard_normal((K, d)) * 0.3 / np.sqrt(d)).astype(np.float32)
b = (rng.standard_normal(K) * 0.3).astype(np.float32)
sel = np.zeros(K, np.int32); val = (np.arange(K) % 5).astype(np.int32)
got = s.logreg_eval(ds, fold, W, b, sel, val, use_tensor_cores=1)
Rhi, Rlo, Gp, Whi = fetch(0), fetch(1), fetch(2), fetch(3)
print("sizes", Rhi.size, Rlo.size, Gp.size, Whi.size)
Z = X.astype(np.float64) @ W.T.astype(np.float64) + b
Rm = (1 / (1 + np.exp(-Z)) - y[:, None]) * (fold[:, None] != val[None, :])
R = (Rhi.astype(np.float64) + Rlo)[: n * K].reshape(n, K)
print("R err", np.abs(R - Rm).max(), "Rhi nonzero", np.count_nonzero(Rhi), "Rlo nonzero", np.count_nonzero(Rlo))
print("Gpart nonzero", np.count_nonzero(Gp), "abs max", np.abs(Gp).max(), "first", Gp[:8])
want = Rm.T @ X.astype(np.float64)
print("want[0,:8]", want[0, :8])
G = Gp[: K * 36].reshape(K, 36)
print("Gpart[0,:8]", G[0, :8])
s.close()
