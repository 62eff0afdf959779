"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference worker's fit-and-score body, restated call for call on the
installed scikit-learn: aws-prod/worker/worker.py:289-363 (``train_model``),
with the three repairs SURVEY.md 3.4 lists (keyword ``train_test_split``
arguments, optional hold-out for search jobs, ``cv`` taken from
``train_params``) and ``Ridge``/``RidgeClassifier`` added to the registry
(worker.py:36-57).  ``sweep`` runs it for every candidate and applies the
``GridSearchCV`` result rules (sklearn/model_selection/_search.py:1208-1294).

This is both the parity checker's ground truth for scores/counts and the CPU
baseline that bench.py times (``cpu_baseline.kind = "port"``: the reference's
own call sequence, on the third-party library it lands in).
"""
import time
import warnings

import numpy as np
from scipy.stats import rankdata
from sklearn.base import clone, is_classifier
from sklearn.exceptions import ConvergenceWarning
from sklearn.linear_model import LogisticRegression, Ridge, RidgeClassifier
from sklearn.metrics import accuracy_score, mean_squared_error, r2_score
from sklearn.model_selection import (KFold, StratifiedKFold, cross_val_score,
                                     train_test_split)

REGISTRY = {"LogisticRegression": LogisticRegression, "Ridge": Ridge,
            "RidgeClassifier": RidgeClassifier}


def make_model(model_type, parameters):
    """worker.py:436-455 ``get_model_instance`` without exec/eval."""
    if model_type not in REGISTRY:
        raise ValueError(f"Unsupported model: {model_type}")
    return REGISTRY[model_type](**parameters)


def train_model(task, X, y):
    """worker.py:289-363 for one subtask; X, y replace ``load_dataset``."""
    params = task.get("parameters", {})
    tp = task.get("train_params", {})
    model = make_model(task["model_type"], params)
    clf = is_classifier(model)
    cv = tp.get("cv", 5)
    out = {}
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", ConvergenceWarning)
        if "test_size" in tp:
            Xtr, Xte, ytr, yte = train_test_split(
                X, y, test_size=tp["test_size"], random_state=tp.get("random_state"))
            t0 = time.time()
            model.fit(Xtr, ytr)
            out["training_time"] = time.time() - t0
            pred = model.predict(Xte)
            if clf:
                out["accuracy"] = accuracy_score(yte, pred)
            else:
                out["r2_score"] = r2_score(yte, pred)
                out["mean_squared_error"] = mean_squared_error(yte, pred)
        scores = cross_val_score(model, X, y, cv=cv,
                                 scoring="accuracy" if clf else "r2")
    out["cv_scores"] = scores.tolist()
    out["mean_cv_score"] = scores.mean()
    out["parameters"] = params
    return out


def cv_fits(model_type, parameters, X, y, cv=5):
    """Per-fold fitted estimators + integer correct counts (classifiers) or
    (sse, sst) (regressors), using the same splitter cross_val_score picks."""
    model = make_model(model_type, parameters)
    clf = is_classifier(model)
    splitter = StratifiedKFold(cv) if clf else KFold(cv)
    fits = []
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", ConvergenceWarning)
        for tr, te in splitter.split(X, y):
            m = clone(model).fit(X[tr], y[tr])
            pred = m.predict(X[te])
            rec = dict(model=m, n_test=len(te))
            if clf:
                rec["correct"] = int((pred == y[te]).sum())
                rec["score"] = rec["correct"] / len(te)
            else:
                rec["sse"] = float(((y[te] - pred) ** 2).sum())
                rec["sst"] = float(((y[te] - y[te].mean()) ** 2).sum())
                rec["score"] = r2_score(y[te], pred)
            fits.append(rec)
    return fits


def search_rules(fold_scores):
    """mean / std / rank / best_index exactly as GridSearchCV forms them
    (_search.py:1228-1270, 862-874): np.average over folds, population std,
    rankdata(-mean, 'min') with NaN ranked worst, best = argmin(rank)."""
    a = np.asarray(fold_scores, dtype=np.float64)
    mean = np.average(a, axis=1)
    std = np.sqrt(np.average((a - mean[:, None]) ** 2, axis=1))
    if np.isnan(mean).all():
        rank = np.ones_like(mean, dtype=np.int32)
    else:
        m = np.nan_to_num(mean, nan=np.nanmin(mean) - 1)
        rank = rankdata(-m, method="min").astype(np.int32)
    return mean, std, rank, int(rank.argmin())


def sweep(model_type, base_params, candidates, X, y, cv=5):
    """All candidates x folds, serially, as the reference's workers do."""
    rows = []
    for cand in candidates:
        fits = cv_fits(model_type, {**base_params, **cand}, X, y, cv)
        rows.append(fits)
    scores = [[f["score"] for f in r] for r in rows]
    mean, std, rank, best = search_rules(scores)
    return dict(fits=rows, scores=np.asarray(scores), mean=mean, std=std, rank=rank,
                best_index=best, best_params=candidates[best], best_score=float(mean[best]))
