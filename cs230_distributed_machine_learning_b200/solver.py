"""Device-level sweep API: torch tensors in, numpy results out, all arithmetic in
the sm_100a library behind include/dml_b200.h.

PyTorch is used for device memory, streams and the NCCL bootstrap only.  One
``SweepSolver`` per (process, GPU).
"""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict
from dataclasses import dataclass

import numpy as np
import torch

from . import _cabi


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(None)


@dataclass
class DeviceDataset:
    """A dataset resident in HBM: X fp32 row-major, labels, row masks."""
    X: torch.Tensor                 # (n, d) float32, C-contiguous
    y_cls: torch.Tensor | None      # (n,) int32 class index, classifiers
    y_reg: torch.Tensor | None      # (n,) float32, regressors
    classes: np.ndarray | None      # sorted unique labels (sklearn classes_)
    n: int
    d: int
    ready: object = None            # CUDA event of an upload still in flight on a copy stream
    copy_events: object = None      # (start, end) events of that upload, for the worker's statistics

    @property
    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in (self.X, self.y_cls, self.y_reg) if t is not None)


class SweepSolver:
    def __init__(self, device: int = 0):
        if not torch.cuda.is_available():
            raise RuntimeError("SweepSolver needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(self.device)
        self.stream = torch.cuda.current_stream(self.device)
        self.ctx = _cabi.Context(device, self.stream.cuda_stream)
        self.lib = self.ctx.lib
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        # label encodings of the datasets uploaded last, page-locked; bounded: an entry keeps its y alive (so that
        # id(y) cannot be recycled under the key) and its encoding pinned
        self._class_cache: "OrderedDict[tuple, tuple]" = OrderedDict()
        self._param_buf = torch.empty(1 << 16, dtype=torch.uint8, pin_memory=True)

    # ---- plumbing -------------------------------------------------------------------
    def to_device(self, a, dtype=None, pin: bool = False) -> torch.Tensor:
        if isinstance(a, torch.Tensor):
            t = a
            if dtype is not None and t.dtype != dtype:
                t = t.to(dtype)
            if t.device != self.device:
                self.h2d_bytes += t.numel() * t.element_size()
                t = t.to(self.device, non_blocking=True)
            return t.contiguous()
        arr = np.ascontiguousarray(a)
        t = torch.from_numpy(arr)
        if dtype is not None and t.dtype != dtype:
            t = t.to(dtype)
        if pin:
            t = t.pin_memory()
        self.h2d_bytes += t.numel() * t.element_size()
        return t.to(self.device, non_blocking=True)

    def params(self, *arrays):
        """Small per-candidate arrays handed to the library without a copy engine: they are
        written into one page-locked buffer, which the device reads in place (pinned host memory
        is device-addressable under UVA).  A host->device copy here would queue behind a dataset
        prefetch on the same DMA engine and hold the sweep back for the whole transfer."""
        arrays = [np.ascontiguousarray(a) for a in arrays]
        need = sum((a.nbytes + 255) // 256 * 256 for a in arrays)
        if need > self._param_buf.numel():
            self._param_buf = torch.empty(2 * need, dtype=torch.uint8, pin_memory=True)
        host = self._param_buf.numpy()
        base, off, ptrs = self._param_buf.data_ptr(), 0, []
        for a in arrays:
            host[off:off + a.nbytes] = a.view(np.uint8).reshape(-1)
            ptrs.append(C.c_void_p(base + off))
            off += (a.nbytes + 255) // 256 * 256
            self.h2d_bytes += a.nbytes
        return ptrs

    def to_host(self, t: torch.Tensor) -> np.ndarray:
        self.d2h_bytes += t.numel() * t.element_size()
        return t.cpu().numpy()

    def upload_replicated(self, X: np.ndarray, replica) -> torch.Tensor:
        """X for a GPU that is one of ``world`` replicas of the same dataset on one box (candidate blocks,
        one process per GPU): this rank copies rows [rank * per, (rank + 1) * per) over its own PCIe link and
        the replicas complete each other with one NCCL all-gather over NVLink, so the host side moves the
        dataset once instead of ``world`` times.  Collective: every rank of ``group`` must upload the same
        dataset in the same order.  ``replica`` = (process group, rank, world)."""
        from .dist import replicate_rows
        group, rank, world = replica
        n = X.shape[0]
        per = -(-n // world)
        self.h2d_bytes += max(0, min(n, (rank + 1) * per) - min(n, rank * per)) * X.shape[1] * 4
        return replicate_rows(X, self.device, group, rank, world)

    def upload_dataset(self, X, y, classifier: bool, stream=None, classes=None, replica=None) -> DeviceDataset:
        """X (n, d) any float dtype -> fp32 on device; y -> class indices or fp32.
        With ``stream`` the copies are queued there and ``ready`` holds the event to wait on.
        ``classes`` (sorted unique labels of the WHOLE dataset) is for row shards: a shard may miss a
        class, and every rank must number the classes alike."""
        if stream is not None:
            with torch.cuda.stream(stream):
                t0 = torch.cuda.Event(enable_timing=True)
                t0.record(stream)
                ds = self.upload_dataset(X, y, classifier, classes=classes, replica=replica)
                ds.ready = torch.cuda.Event(enable_timing=True)
                ds.ready.record(stream)
                ds.copy_events = (t0, ds.ready)
            for t in (ds.X, ds.y_cls, ds.y_reg):
                if t is not None:
                    t.record_stream(self.stream)
            return ds
        if replica is not None and isinstance(X, np.ndarray) and X.dtype == np.float32 and X.ndim == 2 and replica[2] > 1:
            Xd = self.upload_replicated(X, replica)
        else:
            Xd = self.to_device(X, torch.float32)
        if Xd.dim() != 2:
            raise ValueError("X must be 2-D")
        n, d = Xd.shape
        if isinstance(y, torch.Tensor):
            y = y.cpu().numpy()
        key = (id(y), classifier, None if classes is None else len(classes))
        if key not in self._class_cache:      # label encoding of a registered dataset is computed once,
            if classifier and classes is not None:
                classes = np.asarray(classes)
                idx = np.searchsorted(classes, np.asarray(y))
                enc = torch.from_numpy(idx.astype(np.int32)).pin_memory()
            elif classifier:                  # into page-locked memory so that its copy is asynchronous
                classes, idx = np.unique(np.asarray(y), return_inverse=True)
                enc = torch.from_numpy(idx.astype(np.int32)).pin_memory()
            else:
                classes, enc = None, torch.from_numpy(np.ascontiguousarray(y, dtype=np.float32)).pin_memory()
            self._class_cache[key] = (classes, enc, y)
            while len(self._class_cache) > 16:
                self._class_cache.popitem(last=False)
        else:
            self._class_cache.move_to_end(key)
        classes, enc, _ = self._class_cache[key]
        yd = self.to_device(enc)
        if classifier:
            return DeviceDataset(Xd, yd, None, classes, n, d)
        return DeviceDataset(Xd, None, yd, None, n, d)

    def init_comm(self, n_ranks: int, rank: int, unique_id: bytes):
        self.ctx.init_comm(n_ranks, rank, unique_id)

    # ---- sweeps ---------------------------------------------------------------------
    def logreg_sweep(self, ds: DeviceDataset, fold_id, n_folds: int, C_vals, tol, fit_intercept, max_iter,
                     hold_id=None, use_tensor_cores: int = -1, want_scores: bool = True, class_weight=None):
        """All candidates x (n_folds CV fits [+ hold-out fit]).  Returns dict of numpy arrays
        shaped [n_cand, fits_per_cand, ...].  ``class_weight``: (n_cand, fits_per_cand, n_classes) weights of
        each class for each fit (dml_logreg_sweep_weighted), or None."""
        n_cand = len(C_vals)
        n_classes = int(len(ds.classes))
        T = 1 if n_classes == 2 else n_classes
        fpc = n_folds + (1 if hold_id is not None else 0)
        F = n_cand * fpc
        dev = self.device
        fold_d = self.to_device(fold_id, torch.int8)
        hold_d = self.to_device(hold_id, torch.int8) if hold_id is not None else None
        C_p, tol_p, fi_p, mi_p = self.params(
            np.asarray(C_vals, dtype=np.float64), np.broadcast_to(np.asarray(tol, dtype=np.float64), (n_cand,)),
            np.broadcast_to(np.asarray(fit_intercept, dtype=np.uint8), (n_cand,)),
            np.broadcast_to(np.asarray(max_iter, dtype=np.int32), (n_cand,)))
        # every result lives in ONE device buffer (64-bit fields first), so a sweep ends with one
        # device->host copy instead of seven
        fields = [("correct", np.int64, (F,)), ("n_test", np.int64, (F,))] if want_scores else []
        fields += [("coef", np.float32, (F, T, ds.d)), ("intercept", np.float32, (F, T)),
                   ("n_iter", np.int32, (F,)), ("status", np.int32, (F,)), ("nfev", np.int32, (F,))]
        offs, total = {}, 0
        for name, dt, shape in fields:
            offs[name] = total
            total += (int(np.prod(shape)) * np.dtype(dt).itemsize + 15) // 16 * 16
        buf = torch.empty(total, dtype=torch.uint8, device=dev)
        base = buf.data_ptr()
        ptr = lambda name: C.c_void_p(base + offs[name]) if name in offs else C.c_void_p(None)
        rounds = C.c_int32(0)
        if class_weight is None:
            rc = self.lib.dml_logreg_sweep(
                self.ctx.handle, _ptr(ds.X), ds.n, ds.d, ds.X.stride(0), _ptr(ds.y_cls), n_classes,
                _ptr(fold_d), n_folds, _ptr(hold_d), C_p, tol_p, fi_p, mi_p, n_cand,
                ptr("coef"), ptr("intercept"), ptr("n_iter"), ptr("status"), ptr("nfev"), ptr("correct"), ptr("n_test"),
                C.byref(rounds), int(use_tensor_cores))
        else:
            cw = np.ascontiguousarray(class_weight, dtype=np.float64)
            if cw.shape != (n_cand, fpc, n_classes):
                raise ValueError(f"class_weight must have shape {(n_cand, fpc, n_classes)}, got {cw.shape}")
            cw_d = self.to_device(cw)      # (the pinned parameter buffer is in use by C_p .. mi_p)
            rc = self.lib.dml_logreg_sweep_weighted(
                self.ctx.handle, _ptr(ds.X), ds.n, ds.d, ds.X.stride(0), _ptr(ds.y_cls), n_classes,
                _ptr(fold_d), n_folds, _ptr(hold_d), C_p, tol_p, fi_p, mi_p, n_cand, _ptr(cw_d),
                ptr("coef"), ptr("intercept"), ptr("n_iter"), ptr("status"), ptr("nfev"), ptr("correct"), ptr("n_test"),
                C.byref(rounds), int(use_tensor_cores))
        self.ctx.check(rc, "dml_logreg_sweep")
        host = self.to_host(buf)
        out = dict(rounds=int(rounds.value), timing=self.ctx.last_timing())
        for name, dt, shape in fields:
            n_el = int(np.prod(shape))
            a = host[offs[name]:offs[name] + n_el * np.dtype(dt).itemsize].view(dt)
            out[name] = a.reshape((n_cand, fpc) + tuple(shape[1:]))
        return out

    def ridge_sweep(self, ds: DeviceDataset, fold_id, n_folds: int, alphas, fit_intercept, classifier: bool,
                    hold_id=None, use_tensor_cores: int = -1):
        n_cand = len(alphas)
        n_classes = int(len(ds.classes)) if classifier else 0
        T = (1 if n_classes == 2 else n_classes) if classifier else 1
        fpc = n_folds + (1 if hold_id is not None else 0)
        F = n_cand * fpc
        dev = self.device
        fold_d = self.to_device(fold_id, torch.int8)
        hold_d = self.to_device(hold_id, torch.int8) if hold_id is not None else None
        a_p, fi_p = self.params(np.asarray(alphas, dtype=np.float64),
                                np.broadcast_to(np.asarray(fit_intercept, dtype=np.uint8), (n_cand,)))
        coef = torch.empty((F, T, ds.d), dtype=torch.float32, device=dev)
        icpt = torch.empty((F, T), dtype=torch.float32, device=dev)
        n_test = torch.empty(F, dtype=torch.int64, device=dev)
        info = torch.empty(F, dtype=torch.int32, device=dev)
        correct = torch.empty(F, dtype=torch.int64, device=dev) if classifier else None
        sse = torch.empty(F, dtype=torch.float64, device=dev) if not classifier else None
        sy = torch.empty(F, dtype=torch.float64, device=dev) if not classifier else None
        syy = torch.empty(F, dtype=torch.float64, device=dev) if not classifier else None
        rc = self.lib.dml_ridge_sweep(
            self.ctx.handle, _ptr(ds.X), ds.n, ds.d, ds.X.stride(0), _ptr(ds.y_reg), _ptr(ds.y_cls),
            n_classes, 1 if classifier else 0, _ptr(fold_d), n_folds, _ptr(hold_d), a_p, fi_p, n_cand,
            _ptr(coef), _ptr(icpt), _ptr(correct), _ptr(n_test), _ptr(sse), _ptr(sy), _ptr(syy), _ptr(info),
            int(use_tensor_cores))
        self.ctx.check(rc, "dml_ridge_sweep")
        out = dict(coef=self.to_host(coef).reshape(n_cand, fpc, T, ds.d),
                   intercept=self.to_host(icpt).reshape(n_cand, fpc, T),
                   n_test=self.to_host(n_test).reshape(n_cand, fpc),
                   info=self.to_host(info).reshape(n_cand, fpc), timing=self.ctx.last_timing())
        if classifier:
            out["correct"] = self.to_host(correct).reshape(n_cand, fpc)
        else:
            out["sse"] = self.to_host(sse).reshape(n_cand, fpc)
            out["sy"] = self.to_host(sy).reshape(n_cand, fpc)
            out["syy"] = self.to_host(syy).reshape(n_cand, fpc)
        return out

    def logreg_eval(self, ds: DeviceDataset, fold_id, W, b, mask_sel, mask_val, hold_id=None,
                    use_tensor_cores: int = -1) -> np.ndarray:
        """Loss / gradient sums of given columns (dml_logreg_eval): (n_slots, d + 2) float64."""
        W_d = self.to_device(np.ascontiguousarray(W, dtype=np.float32))
        b_d = self.to_device(np.ascontiguousarray(b, dtype=np.float32))
        n_slots = W_d.shape[0]
        sel_d = self.to_device(np.asarray(mask_sel, dtype=np.int32))
        val_d = self.to_device(np.asarray(mask_val, dtype=np.int32))
        fold_d = self.to_device(fold_id, torch.int8)
        hold_d = self.to_device(hold_id, torch.int8) if hold_id is not None else None
        out = torch.empty((n_slots, ds.d + 2), dtype=torch.float64, device=self.device)
        rc = self.lib.dml_logreg_eval(self.ctx.handle, _ptr(ds.X), ds.n, ds.d, ds.X.stride(0), _ptr(ds.y_cls),
                                      _ptr(fold_d), _ptr(hold_d), _ptr(W_d), _ptr(b_d), _ptr(sel_d), _ptr(val_d),
                                      n_slots, _ptr(out), int(use_tensor_cores))
        self.ctx.check(rc, "dml_logreg_eval")
        return self.to_host(out)

    def score(self, ds: DeviceDataset, fold_id, W, b, mask_sel, mask_val, T: int, hold_id=None):
        """Counts / squared errors of given coefficients (dml_score)."""
        W_d = self.to_device(np.ascontiguousarray(W, dtype=np.float32))
        b_d = self.to_device(np.ascontiguousarray(b, dtype=np.float32))
        n_slots = W_d.shape[0]
        sel_d = self.to_device(np.asarray(mask_sel, dtype=np.int32))
        val_d = self.to_device(np.asarray(mask_val, dtype=np.int32))
        fold_d = self.to_device(fold_id, torch.int8)
        hold_d = self.to_device(hold_id, torch.int8) if hold_id is not None else None
        cls = ds.y_cls is not None
        correct = torch.empty(n_slots // T, dtype=torch.int64, device=self.device) if cls else None
        sse = torch.empty(n_slots, dtype=torch.float64, device=self.device) if not cls else None
        rc = self.lib.dml_score(self.ctx.handle, _ptr(ds.X), ds.n, ds.d, ds.X.stride(0), _ptr(ds.y_cls),
                                _ptr(ds.y_reg), _ptr(fold_d), _ptr(hold_d), _ptr(W_d), _ptr(b_d), _ptr(sel_d),
                                _ptr(val_d), n_slots, T, _ptr(correct), _ptr(sse))
        self.ctx.check(rc, "dml_score")
        return self.to_host(correct) if cls else self.to_host(sse)

    def score_all_rows(self, ds: DeviceDataset, fold_id, W, b, T: int):
        """Counts (classifier) or squared errors (regressor) of every column over ALL rows of the dataset:
        the columns select the hold label with value 0, and without a hold array every row carries 0."""
        n_slots = np.asarray(W).shape[0]
        return self.score(ds, fold_id, W, b, np.ones(n_slots, np.int32), np.zeros(n_slots, np.int32), T, hold_id=None)

    def close(self):
        self.ctx.close()
