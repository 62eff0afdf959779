"""ctypes binding of include/dml_b200.h.  No CPU fallback: a missing library or a
machine without a B200 raises."""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

_LIB = None

SYMBOLS = {
    "dml_version": (C.c_int, []),
    "dml_ctx_create": (C.c_int, [C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]),
    "dml_ctx_destroy": (None, [C.c_void_p]),
    "dml_last_error": (C.c_char_p, [C.c_void_p]),
    "dml_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "dml_ctx_init_comm": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "dml_logreg_sweep": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64,
                                   C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int]),
    "dml_logreg_sweep_weighted": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64,
                                            C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.POINTER(C.c_int32), C.c_int]),
    "dml_ridge_sweep": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64,
                                  C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                  C.c_void_p, C.c_int, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "dml_logreg_eval": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_int, C.c_void_p, C.c_int]),
    "dml_score": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64,
                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                            C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                            C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "dml_last_cluster": (C.c_int, [C.c_void_p]),
    "dml_last_comm": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_double)]),
    "dml_last_counters": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    "dml_score_static": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int64,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "dml_last_first_round": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_int64)]),
    "dml_last_timing": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float),
                                  C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
}


def lib_path() -> str:
    return _build.LIB


def load():
    """dlopen the in-tree library (it must have been built: __graft_entry__.build())."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} is missing: build it with `python -m cs230_distributed_machine_learning_b200.build` "
            "(nvcc, sm_100a). There is no CPU fallback for the sweep kernels.")
    lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the export is missing
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib


class DmlError(RuntimeError):
    pass


class Context:
    """One per (process, GPU).  Wraps dml_ctx_create / dml_ctx_destroy."""

    def __init__(self, device: int = 0, stream_ptr: int = 0):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.dml_ctx_create(int(device), C.c_void_p(stream_ptr or None), C.byref(h))
        if rc != 0:
            raise DmlError(f"dml_ctx_create failed ({rc}): {self.lib.dml_last_error(None).decode()}")
        self.handle = h
        self.device = device

    def check(self, rc: int, what: str):
        if rc != 0:
            raise DmlError(f"{what} failed ({rc}): {self.lib.dml_last_error(self.handle).decode()}")

    def init_comm(self, n_ranks: int, rank: int, unique_id: bytes):
        buf = C.create_string_buffer(unique_id, 128)
        self.check(self.lib.dml_ctx_init_comm(self.handle, n_ranks, rank, buf), "dml_ctx_init_comm")

    def last_timing(self):
        t, c = C.c_float(), C.c_float()
        cl, tl = C.c_int32(), C.c_int32()
        self.lib.dml_last_timing(self.handle, C.byref(t), C.byref(c), C.byref(cl), C.byref(tl))
        cr, rd = C.c_int64(), C.c_int32()
        self.lib.dml_last_counters(self.handle, C.byref(cr), C.byref(rd))
        fm, fc = C.c_float(), C.c_int64()
        self.lib.dml_last_first_round(self.handle, C.byref(fm), C.byref(fc))
        cm, cc, cb = C.c_float(), C.c_int32(), C.c_double()
        self.lib.dml_last_comm(self.handle, C.byref(cm), C.byref(cc), C.byref(cb))
        return dict(allreduce_ms=cm.value, allreduce_calls=cc.value, allreduce_bytes=cb.value,
                    total_ms=t.value, contraction_ms=c.value, contraction_launches=cl.value,
                    total_launches=tl.value, column_rounds=cr.value, rounds=rd.value,
                    first_round_ms=fm.value, first_round_columns=fc.value,
                    cluster=int(self.lib.dml_last_cluster(self.handle)))

    def close(self):
        if getattr(self, "handle", None):
            self.lib.dml_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def nccl_unique_id() -> bytes:
    lib = load()
    buf = C.create_string_buffer(128)
    rc = lib.dml_nccl_unique_id(buf)
    if rc != 0:
        raise DmlError(f"dml_nccl_unique_id failed ({rc}): {lib.dml_last_error(None).decode()}")
    return buf.raw
