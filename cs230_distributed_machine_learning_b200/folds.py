"""Row labels for the sweep kernels: CV fold ids and the hold-out split.

``cross_val_score(model, X, y, cv=int)`` (reference: aws-prod/worker/worker.py:326,341)
resolves an int ``cv`` to an un-shuffled ``StratifiedKFold`` for classifiers and
``KFold`` for regressors (sklearn/model_selection/_split.py:2697-2776).  The
kernels take the result as one int8 label per row -- the fold whose test set
holds the row -- instead of index lists, so no row of X is ever gathered.
"""
from __future__ import annotations

import numpy as np


def kfold_labels(n: int, n_folds: int) -> np.ndarray:
    """KFold without shuffle: contiguous blocks, the first ``n % n_folds`` one row
    longer (_split.py:517-531)."""
    if not 2 <= n_folds <= n:
        raise ValueError(f"cv={n_folds} is not usable with {n} samples")
    sizes = np.full(n_folds, n // n_folds, dtype=np.int64)
    sizes[: n % n_folds] += 1
    return np.repeat(np.arange(n_folds, dtype=np.int8), sizes)


def stratified_labels(y_idx: np.ndarray, n_folds: int) -> np.ndarray:
    """StratifiedKFold without shuffle (_split.py:774-842).

    Classes are numbered by first appearance; the per-fold quota of a class is the
    round-robin count over the sorted labels; within a class, rows keep their order
    and receive fold ids in blocks.
    """
    y_idx = np.asarray(y_idx)
    n = y_idx.shape[0]
    if y_idx.dtype.kind in "iu" and n and 0 <= int(y_idx.min()) and int(y_idx.max()) < (1 << 16):
        # small non-negative integers (the worker's label encoding): counting instead of sorting
        inv = y_idx.astype(np.intp, copy=False)
        span = int(y_idx.max()) + 1
        first = np.full(span, n, dtype=np.intp)
        first[inv[::-1]] = np.arange(n - 1, -1, -1, dtype=np.intp)     # duplicates: the last write = smallest index
        present = np.nonzero(first < n)[0]
        remap = np.zeros(span, dtype=np.intp)
        remap[present] = np.arange(present.shape[0])
        inv, first = remap[inv], first[present]
    else:
        _, first, inv = np.unique(y_idx, return_index=True, return_inverse=True)
    n_classes = first.shape[0]
    counts = np.bincount(inv, minlength=n_classes)
    if np.all(n_folds > counts):
        raise ValueError(f"n_splits={n_folds} cannot be greater than the number of members in each class.")
    rank_of_class = np.empty(n_classes, dtype=np.int64)
    rank_of_class[np.argsort(first, kind="stable")] = np.arange(n_classes)
    enc = rank_of_class[inv]                       # class id by order of appearance
    cnt = np.bincount(enc, minlength=n_classes)
    starts = np.concatenate(([0], np.cumsum(cnt)[:-1]))  # offset of each class in sorted order
    out = np.empty(n, dtype=np.int8)
    for k in range(n_classes):
        # rows of class k occupy positions [starts[k], starts[k]+cnt[k]) of the sorted
        # label vector; position p goes to fold p % n_folds, so the quota of fold f is
        # the number of such p with p % n_folds == f
        f = np.arange(n_folds)
        quota = cnt[k] // n_folds + (((f - starts[k]) % n_folds) < (cnt[k] % n_folds))
        out[enc == k] = np.repeat(np.arange(n_folds, dtype=np.int8), quota)
    return out


def holdout_labels(n: int, test_size, random_state) -> np.ndarray:
    """``train_test_split(X, y, test_size=..., random_state=...)`` (worker.py:302 with
    keyword arguments) as a 0/1 label per row: 1 = test row.

    ShuffleSplit draws ``rng.permutation(n)``; the first n_test entries are the test
    set (sklearn/model_selection/_split.py:2233-2263, 2365-2441).  Rows beyond
    n_test + n_train (possible when both sizes are given) are not modelled: the
    reference only passes test_size.
    """
    if isinstance(test_size, float) or (isinstance(test_size, (int, np.integer)) is False):
        n_test = int(np.ceil(float(test_size) * n))
    else:
        n_test = int(test_size)
    if not 0 < n_test < n:
        raise ValueError(f"test_size={test_size} leaves an empty train or test set for n={n}")
    if random_state is None:
        rng = np.random.mtrand._rand
    else:
        rng = np.random.RandomState(int(random_state))
    perm = rng.permutation(n)
    lab = np.zeros(n, dtype=np.int8)
    lab[perm[:n_test]] = 1
    return lab


def stratified_labels_slice(y_local: np.ndarray, before: np.ndarray, total: np.ndarray, first_seen: np.ndarray,
                            n_folds: int) -> np.ndarray:
    """``stratified_labels`` for a row shard, without the other shards' labels (row-sharded sweeps whose
    dataset never exists in one place, BASELINE.json config 5).

    y_local    class indices 0..K-1 of this shard's rows, in row order
    before[k]  members of class k in the rows that precede this shard
    total[k]   members of class k in the whole dataset
    first_seen[k]  global row index of the first member of class k (classes are numbered by first
               appearance, _split.py:774-842)
    The j-th member (0-based, in row order) of a class gets the fold whose block of that class's quota
    contains j -- exactly what the un-sharded routine assigns."""
    y_local = np.asarray(y_local).astype(np.intp, copy=False)
    K = len(total)
    order = np.empty(K, dtype=np.int64)
    order[np.argsort(np.asarray(first_seen), kind="stable")] = np.arange(K)   # class -> id by first appearance
    cnt = np.zeros(K, dtype=np.int64)
    cnt[order] = np.asarray(total, dtype=np.int64)                            # counts in first-appearance numbering
    starts = np.concatenate(([0], np.cumsum(cnt)[:-1]))
    out = np.empty(y_local.shape[0], dtype=np.int8)
    f = np.arange(n_folds)
    for k in range(K):
        e = order[k]
        quota = cnt[e] // n_folds + (((f - starts[e]) % n_folds) < (cnt[e] % n_folds))
        edges = np.cumsum(quota)
        rows = np.nonzero(y_local == k)[0]
        j = int(before[k]) + np.arange(rows.shape[0])
        out[rows] = np.searchsorted(edges, j, side="right").astype(np.int8)
    return out
