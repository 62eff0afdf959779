"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

numpy restatement of the un-shuffled CV splitters that ``cross_val_score(...,
cv=int)`` picks (reference call site aws-prod/worker/worker.py:326,341):
``StratifiedKFold`` for classifiers, ``KFold`` otherwise
(sklearn/model_selection/_split.py:2697-2776, 774-842, 517-531).
"""
import numpy as np


def kfold_ids(n, n_folds):
    """Contiguous blocks; the first n % f folds hold one extra row (_split.py:524-525)."""
    sizes = np.full(n_folds, n // n_folds, dtype=np.int64)
    sizes[: n % n_folds] += 1
    return np.repeat(np.arange(n_folds), sizes).astype(np.int8)


def stratified_fold_ids(y, n_folds):
    """Per-class blocks of fold ids in row order (_split.py:792-841)."""
    y = np.asarray(y)
    _, first, inv = np.unique(y, return_index=True, return_inverse=True)
    _, perm = np.unique(first, return_inverse=True)  # classes by first appearance
    enc = perm[inv]
    n_classes = len(first)
    order = np.sort(enc)
    alloc = np.asarray([np.bincount(order[i::n_folds], minlength=n_classes)
                        for i in range(n_folds)])
    out = np.empty(len(y), dtype=np.int8)
    for k in range(n_classes):
        out[enc == k] = np.arange(n_folds).repeat(alloc[:, k])
    return out
