"""Local / synthetic datasets for the sweep.

The reference pulls CSVs from Kaggle/S3 into ``/mnt/efs/datasets/<id>`` and
re-parses them for every subtask (aws-prod/worker/worker.py:390-433, target =
last column :428-429).  BASELINE.json replaces that with local or synthetic
arrays of the named shape: a dataset is registered once by name and handed to the
executor as (X float32/float64 C-contiguous, y).  The generators follow
SURVEY.md 8d (``numpy.random.default_rng(seed)``, features ~ N(0,1) so lbfgs
stays well-conditioned).
"""
from __future__ import annotations

import threading

import numpy as np

_REGISTRY: dict[str, tuple[np.ndarray, np.ndarray]] = {}
_VERSION: dict[str, int] = {}      # bumped by every registration: device copies and row labels are cached per version
_LOCK = threading.Lock()


def register_dataset(name: str, X, y) -> None:
    """Make (X, y) available to jobs as ``dataset_name`` (replaces /download_data)."""
    X = np.ascontiguousarray(X)
    if X.dtype not in (np.float32, np.float64):
        X = X.astype(np.float64)
    with _LOCK:
        _REGISTRY[name] = (X, np.asarray(y))
        _VERSION[name] = _VERSION.get(name, 0) + 1


def dataset_version(name: str) -> int:
    """How many times ``name`` has been registered.  The reference re-reads the CSV for every subtask
    (worker.py:425) and so always sees current data; the GPU executor keeps device copies across jobs
    and drops them when this number changes."""
    with _LOCK:
        return _VERSION.get(name, 0)


def get_dataset(name: str):
    with _LOCK:
        if name not in _REGISTRY:
            raise KeyError(f"dataset {name!r} is not registered")
        return _REGISTRY[name]


def has_dataset(name: str) -> bool:
    with _LOCK:
        return name in _REGISTRY


def load_csv_dataset(name: str, path: str) -> None:
    """Reference convention: every column but the last is a feature (worker.py:428-429)."""
    import pandas as pd
    df = pd.read_csv(path)
    register_dataset(name, df.iloc[:, :-1].to_numpy(dtype=np.float64), df.iloc[:, -1].to_numpy())


def make_dataset(kind: str, n: int, d: int, seed: int, chunk: int = 1 << 18):
    """kind in {"binary", "regression", "multiclassK"}; returns (X fp32, y).

    Row-chunked so 16M x 512 never needs a second full-size temporary; the byte
    stream depends only on (kind, n, d, seed, chunk).
    """
    rng = np.random.default_rng(seed)
    X = np.empty((n, d), dtype=np.float32)
    if kind.startswith("multiclass"):
        k = int(kind[len("multiclass"):] or 10)
        W = rng.standard_normal((d, k)) / np.sqrt(d)
        y = np.empty(n, dtype=np.int32)
    else:
        w = rng.standard_normal(d) / np.sqrt(d)
        y = np.empty(n, dtype=np.float32 if kind == "regression" else np.int32)
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        xc = rng.standard_normal((b - a, d), dtype=np.float32)
        X[a:b] = xc
        if kind.startswith("multiclass"):
            y[a:b] = np.argmax(xc.astype(np.float64) @ W + rng.gumbel(size=(b - a, k)), axis=1)
        else:
            t = xc.astype(np.float64) @ w + 0.5 * rng.standard_normal(b - a)
            y[a:b] = t if kind == "regression" else (t > 0)
    return X, y


CHUNK_ROWS = 1 << 18


def make_dataset_rows(kind: str, n: int, d: int, seed: int, lo: int = 0, hi: int | None = None, threads: int = 4):
    """Rows [lo, hi) of a synthetic dataset whose bytes depend only on (kind, n, d, seed): chunk c (2^18 rows)
    is drawn from ``default_rng([seed, c])`` and the planted weights from ``default_rng([seed, 2**31 - 1])``,
    so any row range can be produced on its own -- by the rank that owns the row shard (BASELINE.json config
    5: 16M x 512 is 32.8 GB, no rank ever holds all of it on the host) or by an oracle checking a slice.
    Same distributions as ``make_dataset``, different stream.  lo must be chunk-aligned."""
    from concurrent.futures import ThreadPoolExecutor
    hi = n if hi is None else hi
    if lo % CHUNK_ROWS:
        raise ValueError("lo must be a multiple of CHUNK_ROWS")
    multi = kind.startswith("multiclass")
    k = int(kind[len("multiclass"):] or 10) if multi else 0
    rw = np.random.default_rng([seed, 2 ** 31 - 1])
    W = rw.standard_normal((d, k)) / np.sqrt(d) if multi else rw.standard_normal(d) / np.sqrt(d)
    X = np.empty((hi - lo, d), dtype=np.float32)
    y = np.empty(hi - lo, dtype=np.float32 if kind == "regression" else np.int32)

    def fill(c):
        a, full = c * CHUNK_ROWS, min(n, (c + 1) * CHUNK_ROWS)     # the whole chunk is drawn (the stream position of
        b = min(hi, full)                                           # the noise depends on it), then cut to [lo, hi)
        rng = np.random.default_rng([seed, c])
        xc = rng.standard_normal((full - a, d), dtype=np.float32)
        X[a - lo:b - lo] = xc[:b - a]
        if multi:
            yc = np.argmax(xc.astype(np.float64) @ W + rng.gumbel(size=(full - a, k)), axis=1)
        else:
            t = xc.astype(np.float64) @ W + 0.5 * rng.standard_normal(full - a)
            yc = t if kind == "regression" else (t > 0)
        y[a - lo:b - lo] = yc[:b - a]
    chunks = range(lo // CHUNK_ROWS, -(-hi // CHUNK_ROWS))
    with ThreadPoolExecutor(max(1, threads)) as pool:
        list(pool.map(fill, chunks))
    return X, y
