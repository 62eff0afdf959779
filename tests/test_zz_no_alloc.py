"""dml_score_static: scoring for callers that cannot allocate device memory.

Some users of the library run it where a cudaMalloc at call time is not an option
-- a worker whose device-memory reservation is already spent, or a scoring call on
a side stream while a sweep owns the GPU (cudaMalloc/cudaFree would serialise the
device).  dml_score_static keeps its scratch in a fixed array inside the library,
so it only launches kernels.  This file checks, through the C ABI, that it returns
what dml_score returns and what a float64 numpy recount of the same predictions
gives (integer counts: they may differ only by rows whose decision value is within
1e-5 of a tie, and the number of such rows is the tolerance -- dml_score runs on the
tensor cores in 3xTF32, this entry point on CUDA cores in fp32), twice in a row on
different problems -- the second call finds the first call's leftovers in the shared
array and must return the very same integers -- and that a request larger than the
fixed array is refused with a message instead of being run.  It is collected last (file name) so a failure
here cannot hide the result of any other test, and it needs nothing from the
tests before it: it passes alone in a fresh process as well.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from cs230_distributed_machine_learning_b200 import _cabi, folds
from tests.synth import make_dataset

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _score_static(ds, fold_d, W, b, sel, val, T, stream=None):
    lib = _cabi.load()
    dev = ds.X.device
    W_d = torch.as_tensor(np.ascontiguousarray(W, dtype=np.float32), device=dev)
    b_d = torch.as_tensor(np.ascontiguousarray(b, dtype=np.float32), device=dev)
    sel_d = torch.as_tensor(np.asarray(sel, dtype=np.int32), device=dev)
    val_d = torch.as_tensor(np.asarray(val, dtype=np.int32), device=dev)
    out = torch.full((W_d.shape[0] // T,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    rc = lib.dml_score_static(dev.index or 0, C.c_void_p(stream or 0), _ptr(ds.X), ds.n, ds.d, ds.X.stride(0),
                              _ptr(ds.y_cls), _ptr(fold_d), None, _ptr(W_d), _ptr(b_d), _ptr(sel_d), _ptr(val_d),
                              W_d.shape[0], T, _ptr(out))
    return rc, out.cpu().numpy()


def test_static_scratch_scoring_equals_dml_score_twice_in_a_row(solver):
    rng = np.random.default_rng(123)
    cases = [("binary", 20000, 24, 40, 1, 5),        # 157 row tiles x 40 columns
             ("multiclass4", 7777, 13, 24, 4, 3)]    # ragged n and d, arg-max over groups of 4 columns
    for kind, n, d, n_slots, T, cv in cases:
        X, y = make_dataset(kind, n, d, seed=n)
        fold = folds.stratified_labels(y, cv)
        ds = solver.upload_dataset(X, y, classifier=True)
        fold_d = solver.to_device(fold, torch.int8)
        W = (rng.standard_normal((n_slots, d)) / np.sqrt(d)).astype(np.float32)
        b = (0.1 * rng.standard_normal(n_slots)).astype(np.float32)
        sel = np.zeros(n_slots, np.int32)
        val = np.repeat(np.arange(n_slots // T) % cv, T).astype(np.int32)
        want = solver.score(ds, fold, W, b, sel, val, T=T)           # existing entry point (ctx arena)
        # float64 recount on the host: the anchor for both entry points
        z = X.astype(np.float64) @ W.T.astype(np.float64) + b.astype(np.float64)
        exact = np.zeros(n_slots // T, np.int64)
        slack = np.zeros(n_slots // T, np.int64)
        for g in range(n_slots // T):
            rows = fold == val[g * T]
            zz = z[rows][:, g * T:(g + 1) * T]
            if T == 1:
                exact[g] = int(((zz[:, 0] > 0).astype(np.int64) == y[rows]).sum())
                slack[g] = int((np.abs(zz[:, 0]) < 1e-5).sum())
            else:
                exact[g] = int((zz.argmax(axis=1) == y[rows]).sum())
                top = np.sort(zz, axis=1)
                slack[g] = int(((top[:, -1] - top[:, -2]) < 1e-5).sum())
        rc, got = _score_static(ds, fold_d, W, b, sel, val, T)
        assert rc == 0, _cabi.load().dml_last_error(None)
        assert (np.abs(got - exact) <= slack).all(), (got, exact, slack)        # against numpy
        assert (np.abs(got - want) <= slack).all(), (got, want, slack)          # against dml_score
        assert (np.abs(want - exact) <= slack).all(), (want, exact, slack)
        rc, again = _score_static(ds, fold_d, W, b, sel, val, T)      # same call on a dirty array
        assert rc == 0
        np.testing.assert_array_equal(again, got)                     # leftovers of the first call do not leak


def test_static_scratch_limit_is_enforced(solver):
    X, y = make_dataset("binary", 4096, 8, seed=1)
    ds = solver.upload_dataset(X, y, classifier=True)
    fold_d = solver.to_device(folds.stratified_labels(y, 2), torch.int8)
    n_slots = (1 << 20) // 32 + 1                  # 32 row tiles x n_slots just over the 1 Mi-entry array
    W = np.zeros((n_slots, 8), np.float32)
    rc, out = _score_static(ds, fold_d, W, np.zeros(n_slots, np.float32), np.zeros(n_slots, np.int32),
                            np.zeros(n_slots, np.int32), 1)
    assert rc == 4 and (out == -1).all()           # refused, nothing written
    assert b"fixed scratch" in _cabi.load().dml_last_error(None)
