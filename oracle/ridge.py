"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

numpy fp64 restatement of ``Ridge(solver="auto" -> "cholesky").fit`` and
``RidgeClassifier.fit`` for dense X with n_samples >= n_features
(sklearn/linear_model/_ridge.py:215-234, 746-761, 867-878, 919-1020,
1291-1348; centering sklearn/linear_model/_base.py:189-199; intercept
:318-335).  North-star models; the reference registry
(aws-prod/worker/worker.py:36-57) is extended with them.
"""
import numpy as np


def fit_ridge(X, Y, alpha, fit_intercept=True):
    """Y is (n,) or (n, T).  Returns coef (T, d), intercept (T,)."""
    X = np.asarray(X, dtype=np.float64)
    Y = np.asarray(Y, dtype=np.float64)
    Y2 = Y[:, None] if Y.ndim == 1 else Y
    if fit_intercept:
        xm, ym = X.mean(axis=0), Y2.mean(axis=0)
        Xc, Yc = X - xm, Y2 - ym
    else:
        xm, ym = np.zeros(X.shape[1]), np.zeros(Y2.shape[1])
        Xc, Yc = X, Y2
    A = Xc.T @ Xc
    A.flat[:: A.shape[0] + 1] += alpha
    L = np.linalg.cholesky(A)
    W = np.linalg.solve(L.T, np.linalg.solve(L, Xc.T @ Yc)).T  # (T, d)
    b = ym - W @ xm if fit_intercept else np.zeros(Y2.shape[1])
    return W, b


def classifier_targets(y_idx, n_classes):
    """LabelBinarizer(pos_label=1, neg_label=-1): binary -> one column for
    classes_[1]; multiclass -> one-vs-rest columns (_ridge.py:1337-1338)."""
    y_idx = np.asarray(y_idx)
    if n_classes == 2:
        return np.where(y_idx == 1, 1.0, -1.0)[:, None]
    return np.where(y_idx[:, None] == np.arange(n_classes)[None, :], 1.0, -1.0)
