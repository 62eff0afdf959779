"""Worker: consumes (block) subtasks and runs the fit-and-score body.

Counterpart of aws-prod/worker/worker.py.  The loop (register, heartbeat every
5 s, consume only messages keyed to this worker, publish ``metrics`` then
``result``; worker.py:156-286) is kept; the body of ``train_model``
(worker.py:289-363: 1 hold-out fit + predict/score + 5 more fits inside
``cross_val_score``, per candidate, through scikit-learn) is replaced for
LogisticRegression(lbfgs) / Ridge / RidgeClassifier by ONE batched call into the
sm_100a library for every candidate and fold of the block.

Executors
  "b200"    the CUDA path.  Raises if the library or a B200 is missing -- never
            falls back to the CPU.
  "sklearn" the reference's own path (scikit-learn on host cores), kept for the
            model types the reference registry lists and the GPU path does not
            cover (worker.py:36-57) and for float64 data (BASELINE.json config 1
            "runs on CPU": sklearn computes float64 inputs in float64, the GPU
            kernels are fp32 / 3xTF32, and SURVEY.md 8c shows ill-conditioned
            float64 problems cannot be tracked at that precision).
``train_params["executor"]`` or $DML_EXECUTOR picks one; the default "auto" sends a
candidate to "b200" exactly when the model/solver is covered and X is float32.
"""
from __future__ import annotations

import inspect
import logging
import os
import pickle
import threading
import weakref
import time
import warnings

import numpy as np

from . import datasets, folds

HEARTBEAT_INTERVAL_S = 5.0   # worker.py:33
log = logging.getLogger("dml_b200.worker")
# scorers the GPU path forms from the sums its scoring pass produces (cv_params.scoring, core.py:123-130)
GPU_SCORINGS = {True: (None, "accuracy"), False: (None, "r2", "neg_mean_squared_error")}

MODEL_IMPORTS = {   # worker.py:36-57 plus the two north-star models
    "RandomForestClassifier": "sklearn.ensemble", "RandomForestRegressor": "sklearn.ensemble",
    "LogisticRegression": "sklearn.linear_model", "SVC": "sklearn.svm", "SVR": "sklearn.svm",
    "GradientBoostingClassifier": "sklearn.ensemble", "GradientBoostingRegressor": "sklearn.ensemble",
    "KNeighborsClassifier": "sklearn.neighbors", "KNeighborsRegressor": "sklearn.neighbors",
    "LinearRegression": "sklearn.linear_model", "Ridge": "sklearn.linear_model",
    "RidgeClassifier": "sklearn.linear_model",
}
GPU_MODELS = ("LogisticRegression", "Ridge", "RidgeClassifier", "LinearRegression")
REGRESSORS = ("Ridge", "LinearRegression")     # KFold + R^2 (sklearn check_cv / default scorer)


def get_model_instance(algorithm_name: str, parameters: dict):
    """worker.py:436-455 without exec/eval.  Keys the installed class does not know
    (e.g. ``multi_class`` sent by an older client, library_testing.ipynb:640) are dropped."""
    if algorithm_name not in MODEL_IMPORTS:
        raise ValueError(f"Unsupported model: {algorithm_name}")
    module = __import__(MODEL_IMPORTS[algorithm_name], fromlist=[algorithm_name])
    cls = getattr(module, algorithm_name)
    known = set(inspect.signature(cls.__init__).parameters) - {"self"}
    return cls(**{k: v for k, v in parameters.items() if k in known})


def gpu_eligible(model_type: str, params: dict) -> tuple[bool, str]:
    """Does the batched GPU solver reproduce this estimator configuration?"""
    p = params
    if model_type == "LogisticRegression":
        if p.get("solver", "lbfgs") != "lbfgs":
            return False, f"solver={p.get('solver')!r}"
        if p.get("penalty", "deprecated") not in ("deprecated", "l2"):
            return False, f"penalty={p.get('penalty')!r}"
        if p.get("l1_ratio", 0.0) not in (0, 0.0, None):
            return False, "l1_ratio != 0"
        cw = p.get("class_weight")
        if cw is not None and not (cw == "balanced" or isinstance(cw, dict)):
            return False, f"class_weight={cw!r}"
        if p.get("dual") or p.get("warm_start"):
            return False, "dual/warm_start"
        c = p.get("C", 1.0)
        if not (isinstance(c, (int, float)) and np.isfinite(c) and c > 0):
            return False, "C must be finite and positive"
        return True, ""
    if model_type in ("Ridge", "RidgeClassifier"):
        if p.get("solver", "auto") not in ("auto", "cholesky"):
            return False, f"solver={p.get('solver')!r}"
        if p.get("positive") or p.get("class_weight") is not None:
            return False, "positive/class_weight"
        a = p.get("alpha", 1.0)
        if not (isinstance(a, (int, float)) and np.isfinite(a) and a >= 0):
            return False, "alpha must be a finite scalar"
        return True, ""
    if model_type == "LinearRegression":      # ordinary least squares = the Ridge normal equations with alpha = 0
        if p.get("positive"):
            return False, "positive=True"
        return True, ""
    return False, "model type not covered by the GPU path"


def class_weight_table(params: list[dict], classes, y_idx, fold, hold, cv: int):
    """Weights of each class for every fit (candidate, k): what scikit-learn derives inside each fit from
    ``class_weight`` and the fit's own training labels (sklearn/utils/class_weight.py: "balanced" =
    n_train / (n_classes * bincount(y_train)); a dict maps class label -> weight, missing labels weigh 1).
    Returns (n_cand, cv + (hold is not None), n_classes) float64, or None when no candidate sets class_weight.
    Dict keys are matched by value or by their string form (a JSON round trip turns them into strings)."""
    if all(p.get("class_weight") is None for p in params):
        return None
    K = len(classes)
    total = np.bincount(y_idx, minlength=K).astype(np.float64)
    train_counts = [total - np.bincount(y_idx[fold == k], minlength=K) for k in range(cv)]
    if hold is not None:
        train_counts.append(total - np.bincount(y_idx[hold == 1], minlength=K))
    out = np.ones((len(params), len(train_counts), K), dtype=np.float64)
    for i, p in enumerate(params):
        cw = p.get("class_weight")
        if cw is None:
            continue
        if cw == "balanced":
            for k, cnt in enumerate(train_counts):
                present = cnt > 0
                out[i, k, present] = cnt.sum() / (present.sum() * cnt[present])
        else:
            lookup = {str(k): float(v) for k, v in cw.items()}
            unknown = set(lookup) - {str(c) for c in classes}
            if unknown:
                raise ValueError(f"class_weight names labels that are not in y: {sorted(unknown)}")
            out[i, :, :] = np.array([lookup.get(str(c), 1.0) for c in classes])[None, :]
    return out


def fitted_estimator(model_type: str, parameters: dict, coef, intercept, classes=None, n_iter=None):
    """A scikit-learn estimator carrying coefficients fitted on the GPU, so that the artefact
    the worker pickles (worker.py:351-356) loads and predicts like the reference's."""
    est = get_model_instance(model_type, parameters)
    coef = np.asarray(coef, dtype=np.float64)
    icpt = np.asarray(intercept, dtype=np.float64)
    if model_type in REGRESSORS:
        est.coef_ = coef.reshape(-1)
        est.intercept_ = float(icpt.reshape(-1)[0])
        est.n_features_in_ = est.coef_.shape[0]
    else:
        est.coef_ = coef.reshape(-1, coef.shape[-1])
        est.intercept_ = icpt.reshape(-1)
        est.n_features_in_ = est.coef_.shape[1]
        if model_type == "RidgeClassifier":      # older scikit-learn reads classes_ off the fitted binarizer
            from sklearn.preprocessing import LabelBinarizer
            est._label_binarizer = LabelBinarizer(pos_label=1, neg_label=-1).fit(np.asarray(classes))
        try:
            est.classes_ = np.asarray(classes)
        except AttributeError:                   # classes_ is a read-only property there
            pass
    if n_iter is not None and model_type == "LogisticRegression":
        est.n_iter_ = np.asarray([int(n_iter)], dtype=np.int32)
    return est


def save_model(model, model_dir: str, model_id: str) -> str:
    os.makedirs(model_dir, exist_ok=True)
    path = os.path.join(model_dir, f"{model_id}.pkl")
    with open(path, "wb") as fh:
        pickle.dump(model, fh)
    return path


def _executor_choice(task: dict) -> str:
    return str(task.get("train_params", {}).get("executor") or os.environ.get("DML_EXECUTOR", "auto")).lower()


# ---------------------------------------------------------------------------------------
# sklearn executor: the reference's body, one candidate at a time
# ---------------------------------------------------------------------------------------
def train_candidate_sklearn(model_type: str, parameters: dict, train_params: dict, X, y, subtask_id: str,
                            model_dir: str | None = None) -> dict:
    """worker.py:289-363 with the SURVEY.md 3.4 repairs (keyword train_test_split
    arguments, optional hold-out, cv from train_params)."""
    from sklearn.base import is_classifier
    from sklearn.exceptions import ConvergenceWarning
    from sklearn.metrics import accuracy_score, mean_squared_error, r2_score
    from sklearn.model_selection import cross_validate, train_test_split
    model = get_model_instance(model_type, parameters)
    clf = is_classifier(model)
    results = {}
    if train_params.get("refit_fit"):      # GridSearchCV's refit (_search.py:1146-1159): best params, all rows
        with warnings.catch_warnings():
            warnings.simplefilter("ignore", ConvergenceWarning)
            t0 = time.time()
            model.fit(X, y)
            results["training_time"] = time.time() - t0
        results.update(refit=True, model_id=f"{subtask_id}_model", parameters=parameters, executor="sklearn",
                       model_path=save_model(model, model_dir, f"{subtask_id}_model") if model_dir else None)
        if hasattr(model, "coef_"):
            results["coef"] = np.asarray(model.coef_).tolist()
            results["intercept"] = np.asarray(model.intercept_).tolist()
        return results
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", ConvergenceWarning)
        if "test_size" in train_params:
            Xtr, Xte, ytr, yte = train_test_split(X, y, test_size=train_params["test_size"],
                                                  random_state=train_params.get("random_state"))
            t0 = time.time()
            model.fit(Xtr, ytr)
            results["training_time"] = time.time() - t0
            pred = model.predict(Xte)
            if clf:
                results["accuracy"] = float(accuracy_score(yte, pred))
            else:
                results["r2_score"] = float(r2_score(yte, pred))
                results["mean_squared_error"] = float(mean_squared_error(yte, pred))
        t0 = time.time()
        # cross_val_score (worker.py:326,341) is cross_validate()["test_score"]; the fitted fold models come back too
        # so that the result can say how many L-BFGS iterations each fold took (master.IterationModel)
        cvr = cross_validate(model, X, y, cv=train_params.get("cv", 5),
                             scoring=train_params.get("scoring") or ("accuracy" if clf else "r2"),
                             error_score="raise", return_estimator=True,
                             return_train_score=bool(train_params.get("return_train_score")))
        scores = cvr["test_score"]
        if "train_score" in cvr:
            results["train_scores"] = [float(v) for v in cvr["train_score"]]
        results.setdefault("training_time", (time.time() - t0) / max(1, len(scores)))
        if model_type == "LogisticRegression":
            results["n_iter"] = [int(np.max(e.n_iter_)) for e in cvr["estimator"]]
    results["cv_scores"] = scores.tolist()
    results["mean_cv_score"] = float(scores.mean())
    results["model_id"] = f"{subtask_id}_model"
    results["model_path"] = None
    results["parameters"] = parameters
    results["executor"] = "sklearn"
    return results


# ---------------------------------------------------------------------------------------
# b200 executor: every eligible candidate x fold of the block in one library call
# ---------------------------------------------------------------------------------------
class GpuExecutor:
    """Owns the SweepSolver of one GPU and the datasets resident on it."""

    def __init__(self, gpu: int = 0):
        from .solver import SweepSolver   # raises without CUDA: no CPU fallback
        import torch
        self.gpu = gpu
        self.solver = SweepSolver(gpu)
        self.resident: dict[tuple, object] = {}       # (name, classifier[, shard, n_shards]); insertion order = LRU first
        self.versions: dict[str, int] = {}            # datasets.dataset_version() the cached copies belong to
        self.replica = None                           # (process group, rank, world): see SweepSolver.upload_replicated
        self.shard_solvers: dict[tuple[int, int], object] = {}   # (n_shards, shard) -> SweepSolver with an NCCL comm
        self.labels: dict[tuple, np.ndarray] = {}     # host-side row labels, computed once per dataset
        self.labels_dev: dict[tuple, object] = {}
        self.lock = threading.Lock()
        self.res_lock = threading.Lock()              # guards ``resident`` (prefetch runs beside a sweep)
        self.copy_stream = torch.cuda.Stream(self.solver.device)
        self.max_resident_bytes = int(0.7 * torch.cuda.get_device_properties(self.solver.device).total_memory)
        self.in_use: tuple | None = None
        self.hint: tuple | None = None                # dataset of the next queued task (prefetch_task)
        self.prefetched = 0
        self.stats = {"sweep_ms": [], "copy_ms": [], "host_ms": []}   # last jobs: GPU sweep, overlapped copy, wall

    def drop_resident(self):
        """Forget device copies (the next job pays the host->device copy again)."""
        with self.lock, self.res_lock:
            self.resident.clear()

    def _evict(self, incoming: int = 0, keep: tuple | None = None):
        """Least recently used first; never the dataset a sweep is reading, the one the
        running task is about to read (``keep``) or a prefetched copy nobody has used yet.
        The budget is soft: when nothing may go, the new copy is made anyway.  Call under res_lock."""
        total = sum(v.nbytes for v in self.resident.values()) + incoming
        for key in list(self.resident):
            if total <= self.max_resident_bytes:
                break
            if key == self.in_use or key == keep or self.resident[key].ready is not None:
                continue
            total -= self.resident.pop(key).nbytes

    def _check_version(self, name: str) -> None:
        """A name registered again is a new dataset: drop its device copies and row labels
        (the reference re-reads the CSV for every subtask, worker.py:425).  Call under res_lock."""
        v = datasets.dataset_version(name)
        if self.versions.get(name) == v:
            return
        for cache in (self.resident, self.labels, self.labels_dev):
            for key in [k for k in cache if k[0] == name]:
                del cache[key]
        self.versions[name] = v

    def prefetch(self, name: str, classifier: bool) -> bool:
        """Start the host->device copy of a dataset a queued task will need, on the copy
        stream, while the current sweep runs (the reference parses the CSV again inside
        every subtask, worker.py:425).  Returns True when a copy was queued."""
        key = (name, classifier)
        with self.res_lock:
            self._check_version(name)
            if key in self.resident:
                return False
            X, y = datasets.get_dataset(name)
            if X.dtype != np.float32:
                return False
            self._evict(incoming=X.nbytes)
            self._pin(X)
            self.resident[key] = self.solver.upload_dataset(X, y, classifier, stream=self.copy_stream, replica=self.replica)
            self.prefetched += 1
            return True

    def dataset(self, name: str, classifier: bool, shard: tuple[int, int] | None = None, solver=None):
        """Device copy of a registered dataset; ``shard`` = (index, count) keeps only that row slice
        (dist.row_slice) on this GPU, with the class numbering of the whole dataset."""
        key = (name, classifier) if shard is None else (name, classifier, shard[0], shard[1])
        solver = solver or self.solver
        with self.res_lock:
            self._check_version(name)
            if key not in self.resident:   # the reference re-parses the CSV per subtask (worker.py:425)
                X, y = datasets.get_dataset(name)
                if shard is not None:
                    from .dist import row_slice
                    lo, hi = row_slice(X.shape[0], shard[1], shard[0])
                    classes = np.unique(np.asarray(y)) if classifier else None
                    X, y = X[lo:hi], np.asarray(y)[lo:hi]
                    self._evict(incoming=X.nbytes)
                    self.resident[key] = solver.upload_dataset(X, y, classifier, classes=classes)
                else:
                    self._evict(incoming=X.nbytes)
                    self._pin(X)
                    self.resident[key] = solver.upload_dataset(X, y, classifier, replica=self.replica)
            ds = self.resident.pop(key)
            self.resident[key] = ds        # most recently used
            self.in_use = key
        if ds.ready is not None:           # uploaded by prefetch: order the sweep's stream after the copy
            solver.stream.wait_event(ds.ready)
            ds.ready = None
        return ds

    def shard_solver(self, n_shards: int, shard: int, exchange, group: str):
        """The SweepSolver of this GPU for row-sharded blocks: its own context, joined to an NCCL
        communicator of ``n_shards`` ranks (dml_ctx_init_comm).  Created on the first row-sharded block;
        ``exchange(key, payload)`` is the side channel that carries shard 0's 128-byte ncclUniqueId to the
        other workers (Scheduler.exchange here; Redis in the reference's deployment).  Collective: every
        worker of the shard set must get here, which the master guarantees by queueing a job's blocks to
        all workers in one step."""
        key = (n_shards, shard)
        if key not in self.shard_solvers:
            from . import _cabi
            from .solver import SweepSolver
            sv = SweepSolver(self.gpu)
            uid = exchange(f"nccl-uid/{group}/{n_shards}", _cabi.nccl_unique_id() if shard == 0 else None)
            sv.init_comm(n_shards, shard, uid)
            self.shard_solvers[key] = sv
        return self.shard_solvers[key]

    def _pin(self, X: np.ndarray) -> None:
        """Page-lock a registered float32 array in place (once per process: the executors of all GPUs share the
        host array), so every later upload is a pinned-memory DMA instead of a staged copy."""
        import torch
        if X.dtype != np.float32 or not X.flags.c_contiguous:
            return
        with _PIN_LOCK:
            if id(X) in _PINNED:
                return
            rt = torch.cuda.cudart()
            ptr = X.ctypes.data
            rc = rt.cudaHostRegister(ptr, X.nbytes, 0)
            if int(rc) == 0:
                # The registration must end before the memory does: a freed array's address range is handed out
                # again by the allocator, and a host buffer that overlaps a stale registration makes the next
                # cudaMemcpy on it fail with "invalid argument".  numpy runs weak-reference callbacks at the start
                # of an array's deallocation, before its buffer is released.
                _PINNED[id(X)] = ptr
                fin = weakref.finalize(X, _unpin, id(X), ptr)
                fin.atexit = False      # at interpreter exit the driver releases everything itself
            else:                       # e.g. already registered by someone else: not an error for us; do not
                _clear_cuda_error()     # leave it pending for the next CUDA call torch makes on this thread

    def row_labels(self, dataset_id: str, classifier: bool, cv: int, n: int, y_host, test_size=None, random_state=None,
                   shard: tuple[int, int] | None = None, solver=None):
        """Fold labels (and hold-out labels) of the WHOLE dataset, computed once on the host; the device
        copy holds this GPU's rows of them (all, or its row shard)."""
        key = (dataset_id, classifier, cv)
        if key not in self.labels:
            if classifier:
                _, y_idx = np.unique(np.asarray(y_host), return_inverse=True)
                self.labels[key] = folds.stratified_labels(y_idx, cv)
            else:
                self.labels[key] = folds.kfold_labels(n, cv)
        hold = None
        if test_size is not None:
            hkey = (dataset_id, "hold", test_size, random_state)
            if hkey not in self.labels:
                self.labels[hkey] = folds.holdout_labels(n, test_size, random_state)
            hold = self.labels[hkey]
        return self._on_device(key, shard, solver), (self._on_device(hkey, shard, solver) if hold is not None else None)

    def _on_device(self, key, shard: tuple[int, int] | None = None, solver=None):
        """Row labels go to the GPU once per dataset and stay (1 byte per row)."""
        import torch
        dkey = key if shard is None else key + ("shard",) + tuple(shard)
        if dkey not in self.labels_dev:
            lab = self.labels[key]
            if shard is not None:
                from .dist import row_slice
                lo, hi = row_slice(len(lab), shard[1], shard[0])
                lab = lab[lo:hi]
            self.labels_dev[dkey] = (solver or self.solver).to_device(lab, torch.int8)
        return self.labels_dev[dkey]

    def run(self, model_type: str, cands: list[dict], train_params: dict, dataset_id: str, model_dir: str,
            shard: tuple[int, int] | None = None, exchange=None) -> list[dict]:
        """cands: [{index, subtask_id, parameters}] -> result dict per candidate (A9 shape).

        ``shard`` = (index, count): this GPU holds rows dist.row_slice(n, count, index) and the sweep runs
        together with the other shards' workers, the library summing gradients / moments / counts over NCCL
        (SURVEY.md 8e-2); every shard returns the same, complete results."""
        classifier = model_type not in REGRESSORS
        with self.lock:
            X_host, y_host = datasets.get_dataset(dataset_id)
            n_total = X_host.shape[0]
            solver = self.solver
            if shard is not None:
                if exchange is None:
                    raise ValueError("a row-sharded block needs the scheduler's exchange channel")
                solver = self.shard_solver(shard[1], shard[0], exchange, group=str(train_params.get("shard_group", "default")))
            ds = self.dataset(dataset_id, classifier, shard, solver)
            hint, self.hint = self.hint, None
            if hint is not None and hint != (dataset_id, classifier) and shard is None:
                try:
                    self.prefetch(*hint)          # the next task's copy runs beside this sweep
                except Exception:
                    pass
            cv = train_params.get("cv", 5)
            if not isinstance(cv, int):
                raise ValueError("the GPU executor takes an integer cv (un-shuffled (Stratified)KFold)")
            refit = bool(train_params.get("refit_fit"))
            fold, hold = self.row_labels(dataset_id, classifier, cv, n_total, y_host,
                                         None if refit else train_params.get("test_size"),
                                         train_params.get("random_state"), shard, solver)
            if refit:        # the all-rows fit is the "hold-out" fit of an empty hold-out set
                zkey = (dataset_id, "all-rows")
                if zkey not in self.labels:
                    self.labels[zkey] = np.zeros(n_total, dtype=np.int8)
                hold = self._on_device(zkey, shard, solver)
            P = [c["parameters"] for c in cands]
            fi = [bool(p.get("fit_intercept", True)) for p in P]
            if refit and model_type == "LogisticRegression":
                cv = 0           # only the all-rows column; Ridge keeps its folds (one Gram pass either way)
            t0 = time.time()
            if model_type == "LogisticRegression":
                cw = None
                if any(p.get("class_weight") is not None for p in P):
                    ykey = (dataset_id, "y-index")
                    if ykey not in self.labels:
                        self.labels[ykey] = np.unique(np.asarray(y_host), return_inverse=True)[1].astype(np.int64)
                    hold_host = None
                    if refit:
                        hold_host = self.labels[(dataset_id, "all-rows")]
                    elif train_params.get("test_size") is not None:
                        hold_host = self.labels[(dataset_id, "hold", train_params.get("test_size"), train_params.get("random_state"))]
                    fold_host = self.labels[(dataset_id, classifier, train_params.get("cv", 5))]
                    cw = class_weight_table(P, ds.classes, self.labels[ykey], fold_host, hold_host, cv)
                out = solver.logreg_sweep(
                    ds, fold, cv, [float(p.get("C", 1.0)) for p in P], [float(p.get("tol", 1e-4)) for p in P], fi,
                    [int(p.get("max_iter", 100)) for p in P], hold_id=hold, want_scores=not refit, class_weight=cw)
            else:
                alphas = [0.0 if model_type == "LinearRegression" else float(p.get("alpha", 1.0)) for p in P]
                out = solver.ridge_sweep(ds, fold, cv, alphas, fi, classifier=classifier, hold_id=hold)
            train_all = None
            if train_params.get("return_train_score") and not refit and cv > 0:
                # counts / squared errors of every CV fit over ALL rows; minus the test rows = the training rows
                T = out["coef"].shape[2]
                W = out["coef"][:, :cv].reshape(-1, out["coef"].shape[3])
                b = out["intercept"][:, :cv].reshape(-1)
                train_all = solver.score_all_rows(ds, fold, W, b, T).reshape(len(cands), cv)
            wall = time.time() - t0
            if len(self.stats["sweep_ms"]) < 4096:
                self.stats["sweep_ms"].append(out["timing"]["total_ms"])
                self.stats["host_ms"].append(wall * 1e3)
                if ds.copy_events is not None:      # the copy ended before the sweep began (wait_event)
                    self.stats["copy_ms"].append(ds.copy_events[0].elapsed_time(ds.copy_events[1]))
                    ds.copy_events = None
            with self.res_lock:
                self.in_use = None
                self._evict()
        n_fits = len(cands) * (cv + (hold is not None))
        scoring = train_params.get("scoring")
        results = []
        for i, c in enumerate(cands):
            r = {}
            info = out.get("info")
            if info is not None and np.any(info[i] != 0):
                # the Cholesky factorisation met a non-positive or negligible pivot (dml_ridge_sweep, info_out):
                # scikit-learn answers such systems with lstsq / SVD; the caller re-runs the candidate there
                results.append({"rerun_on_sklearn": f"rank-deficient normal equations (info={info[i].tolist()})",
                                "parameters": c["parameters"]})
                continue
            if refit:
                r.update(refit=True, training_time=wall, coef=out["coef"][i][cv].tolist(),
                         intercept=out["intercept"][i][cv].tolist())
            else:
                if classifier:
                    sc = out["correct"][i] / out["n_test"][i]        # float64 count / n_test (A14)
                else:
                    sst = out["syy"][i] - out["sy"][i] ** 2 / out["n_test"][i]
                    sc = 1.0 - out["sse"][i] / sst
                    if scoring == "neg_mean_squared_error":
                        sc = -out["sse"][i] / out["n_test"][i]
                cvs = sc[:cv]
                if hold is not None:
                    if classifier:
                        r["accuracy"] = float(out["correct"][i][cv] / out["n_test"][i][cv])
                    else:
                        sst_h = out["syy"][i][cv] - out["sy"][i][cv] ** 2 / out["n_test"][i][cv]
                        r["r2_score"] = float(1.0 - out["sse"][i][cv] / sst_h)
                        r["mean_squared_error"] = float(out["sse"][i][cv] / out["n_test"][i][cv])
                r["cv_scores"] = [float(v) for v in cvs]
                r["mean_cv_score"] = float(np.mean(cvs))
                r["training_time"] = wall / n_fits
                if train_all is not None:
                    n_tr = n_total - out["n_test"][i][:cv]
                    if classifier:
                        tr = (train_all[i] - out["correct"][i][:cv]) / n_tr
                    else:
                        sse_tr = train_all[i] - out["sse"][i][:cv]
                        sy_tr = out["sy"][i][:cv].sum() - out["sy"][i][:cv]
                        syy_tr = out["syy"][i][:cv].sum() - out["syy"][i][:cv]
                        tr = (-sse_tr / n_tr) if scoring == "neg_mean_squared_error" else \
                            (1.0 - sse_tr / (syy_tr - sy_tr ** 2 / n_tr))
                    r["train_scores"] = [float(v) for v in tr]
            r["model_id"] = f"{c['subtask_id']}_model"
            r["model_path"] = None
            if hold is not None and model_dir and (shard is None or shard[0] == 0):   # worker.py:351-356: the fitted model is pickled
                n_it = out["n_iter"][i][cv] if model_type == "LogisticRegression" else None
                est = fitted_estimator(model_type, c["parameters"], out["coef"][i][cv], out["intercept"][i][cv],
                                       ds.classes if classifier else None, n_it)
                r["model_path"] = save_model(est, model_dir, r["model_id"])
            r["parameters"] = c["parameters"]
            r["executor"] = "b200"
            if shard is not None:
                r["row_shards"] = shard[1]
            if model_type == "LogisticRegression":
                r["n_iter"] = [int(v) for v in out["n_iter"][i]]
                r["fit_status"] = [int(v) for v in out["status"][i]]
            results.append(r)
        self.last_timing = out["timing"]
        return results


_PINNED: dict[int, int] = {}        # id() of registered arrays that are page-locked, process-wide -> base address
_PIN_LOCK = threading.RLock()       # re-entrant: a finalizer may run inside _pin when an allocation triggers a collection


def _unpin(key: int, ptr: int) -> None:
    """End the page-lock of an array that is being deallocated (weakref.finalize callback of GpuExecutor._pin)."""
    with _PIN_LOCK:
        if _PINNED.get(key) != ptr:
            return
        del _PINNED[key]
    try:
        import torch
        rt = torch.cuda.cudart()
        if int(rt.cudaHostUnregister(ptr)) != 0:
            _clear_cuda_error()
    except Exception:                   # noqa: BLE001 -- never raise from a finalizer
        pass


_CUDART = None


def _clear_cuda_error() -> None:
    """cudaGetLastError() of the CUDA runtime torch uses (torch.cuda.cudart() does not export it; the library's own
    runtime is a separate, statically linked instance with its own per-thread error state)."""
    global _CUDART
    try:
        if _CUDART is None:
            import ctypes
            with open("/proc/self/maps") as f:
                paths = {ln.split()[-1] for ln in f if "libcudart" in ln}
            _CUDART = [ctypes.CDLL(p) for p in sorted(paths)]
        for lib in _CUDART:
            lib.cudaGetLastError()
    except Exception:                   # noqa: BLE001
        pass
_EXECUTORS: dict[int, GpuExecutor] = {}
_EXEC_LOCK = threading.Lock()


def gpu_executor(gpu: int = 0) -> GpuExecutor:
    with _EXEC_LOCK:
        if gpu not in _EXECUTORS:
            _EXECUTORS[gpu] = GpuExecutor(gpu)
        return _EXECUTORS[gpu]


# ---------------------------------------------------------------------------------------
# the reference entry points
# ---------------------------------------------------------------------------------------
def _as_block(task: dict) -> dict:
    """A legacy one-candidate subtask is a block with one candidate and all folds."""
    if "candidates" in task:
        return task
    t = dict(task)
    t["candidates"] = [{"index": 0, "subtask_id": task["subtask_id"], "parameters": task.get("parameters", {})}]
    return t


def train_model(task: dict, gpu: int = 0, model_dir: str = "./models", exchange=None):
    """worker.py:289.  Returns the reference's result dict for a legacy subtask, or a
    list of {subtask_id, status, result|error} for a block subtask.

    A block with ``partition == "rows"`` and ``n_shards > 1`` is one of n_shards identical blocks, one per
    GPU: this worker keeps row shard ``block["shard"]`` and runs the sweep together with the others
    (``exchange`` = the scheduler's side channel for the NCCL id).  Every shard computes the complete
    results; only shard 0 reports them, the others report an empty list."""
    block = _as_block(task)
    model_type = block["model_type"]
    tp = block.get("train_params", {})
    X, y = datasets.get_dataset(block["dataset_id"])
    choice = _executor_choice(block)
    if choice not in ("auto", "b200", "sklearn"):
        raise ValueError(f"unknown executor {choice!r}")
    shard = None
    if block.get("partition") == "rows" and int(block.get("n_shards", 1)) > 1:
        shard = (int(block.get("shard", block.get("gpu", 0))), int(block["n_shards"]))
        if choice == "sklearn":
            raise ValueError("row shards exist on the b200 executor only")
    classifier = model_type not in REGRESSORS
    scoring_ok = tp.get("scoring") in GPU_SCORINGS[classifier]
    on_gpu, on_cpu = [], []
    for c in block["candidates"]:
        ok, why = gpu_eligible(model_type, c["parameters"])
        if ok and not scoring_ok:
            ok, why = False, f"scoring={tp.get('scoring')!r}"
        if ok and choice == "auto" and X.dtype != np.float32:
            ok, why = False, f"X is {X.dtype}: float64 problems are computed in float64 by scikit-learn"
        if (choice == "b200" or shard is not None) and not ok:
            raise ValueError(f"executor 'b200' cannot run {model_type}({why})")
        if choice == "sklearn" or not ok:
            if choice == "auto":       # never silent: say which candidate leaves the GPU path and why
                log.warning("candidate %s (%s) runs on the scikit-learn executor: %s", c["subtask_id"], model_type, why)
            c = dict(c, why=why if choice == "auto" else "executor='sklearn'")
            on_cpu.append(c)
        else:
            on_gpu.append(c)
    entries = {}
    if on_gpu:
        got = gpu_executor(gpu).run(model_type, on_gpu, tp, block["dataset_id"], model_dir, shard=shard, exchange=exchange)
        for c, r in zip(on_gpu, got):
            if "rerun_on_sklearn" in r:
                if shard is not None:
                    raise ValueError(f"row-sharded block: {r['rerun_on_sklearn']}")
                log.warning("candidate %s (%s) re-runs on the scikit-learn executor: %s", c["subtask_id"], model_type,
                            r["rerun_on_sklearn"])
                on_cpu.append(dict(c, why=r["rerun_on_sklearn"], rerouted=True))
            else:
                entries[c["subtask_id"]] = {"subtask_id": c["subtask_id"], "status": "completed", "result": r}
    for c in on_cpu:
        try:
            r = train_candidate_sklearn(model_type, c["parameters"], tp, X, y, c["subtask_id"], model_dir)
            if c.get("rerouted"):
                r["rerouted"] = c["why"]
            elif choice == "auto":
                r["executor_reason"] = c["why"]
            entries[c["subtask_id"]] = {"subtask_id": c["subtask_id"], "status": "completed", "result": r}
        except Exception as e:   # worker.py:264-286: a failing fit is reported, not raised
            if "candidates" not in task:
                raise
            entries[c["subtask_id"]] = {"subtask_id": c["subtask_id"], "status": "failed", "error": str(e)}
    ordered = [entries[c["subtask_id"]] for c in block["candidates"]]
    if "candidates" not in task:
        return ordered[0]["result"]
    if shard is not None and shard[0] != 0:
        return []                   # shard 0 reports for the set
    return ordered


def _gpu_block_key(task) -> tuple | None:
    """(dataset_id, classifier) when the b200 executor would take at least one candidate of the task."""
    if not isinstance(task, dict) or "model_type" not in task or "dataset_id" not in task:
        return None
    try:
        block = _as_block(task)
        choice = _executor_choice(block)
        if choice == "sklearn":
            return None
        if block.get("partition") == "rows" and int(block.get("n_shards", 1)) > 1:
            return None   # a row-sharded block uploads its own slice; nothing to copy ahead
        if not any(gpu_eligible(block["model_type"], c["parameters"])[0] for c in block["candidates"]):
            return None
        if choice == "auto" and datasets.get_dataset(block["dataset_id"])[0].dtype != np.float32:
            return None
        return (block["dataset_id"], block["model_type"] not in REGRESSORS)
    except Exception:
        return None       # a look-ahead never fails a task; the task itself reports the error


def prefetch_task(task: dict, gpu: int = 0, current: dict | None = None) -> bool:
    """Look one message ahead in the worker's queue: arrange for that task's dataset to be
    copied to the GPU while ``current`` runs.  Only acts when this GPU already has an
    executor (a b200 block ran before).  When ``current`` is itself a b200 block the copy
    starts right after it has acquired its own dataset (GpuExecutor.run), so the two never
    compete for the residency budget; otherwise it starts at once."""
    ex = _EXECUTORS.get(gpu)
    key = _gpu_block_key(task) if ex is not None else None
    if key is None:
        return False
    if _gpu_block_key(current) is None:
        try:
            return ex.prefetch(*key)
        except Exception:
            return False
    ex.hint = key
    return True


def process_task(task: dict, gpu: int = 0, exchange=None):
    """worker.py:366-387."""
    return train_model(task, gpu=gpu, exchange=exchange)


class Worker(threading.Thread):
    """worker.py:156-286 against the in-process scheduler."""

    def __init__(self, scheduler, gpu: int | None = None, mem_capacity_mb: int = 180_000, slot: int | None = None):
        """``gpu`` = CUDA device this worker drives; ``slot`` = the block index it serves
        (block["gpu"], master.create_blocks) -- they differ when a process owns one GPU of many."""
        super().__init__(daemon=True)
        self.scheduler = scheduler
        self.gpu = gpu
        self.worker_id = scheduler.subscribe(mem_capacity_mb, host="localhost",
                                             gpu=gpu if slot is None else slot)["worker_id"]
        self.queue = scheduler.get_queue(self.worker_id)
        self.stop_event = threading.Event()
        self.hb = threading.Thread(target=self._heartbeat, daemon=True)

    def _heartbeat(self):
        while not self.stop_event.wait(HEARTBEAT_INTERVAL_S):
            self.scheduler.heartbeat(self.worker_id)

    def run(self):
        self.hb.start()
        while not self.stop_event.is_set():
            try:
                key, task = self.queue.get(timeout=0.1)
            except Exception:
                continue
            if task is None:
                break
            if key != self.worker_id:      # worker.py:185: skip messages keyed to another worker
                continue
            with self.queue.mutex:         # look one message ahead: its dataset copy overlaps this task
                nxt = self.queue.queue[0] if self.queue.queue else None
            if nxt is not None and nxt[0] == self.worker_id and nxt[1] is not None:
                prefetch_task(nxt[1], gpu=self.gpu or 0, current=task)
            self.handle(task)

    def handle(self, task: dict):
        received = started = time.time()
        try:
            import psutil
            cpu0, mem0 = psutil.cpu_percent(None), psutil.virtual_memory().percent
        except Exception:
            psutil, cpu0, mem0 = None, 0.0, 0.0
        try:
            result = process_task(task, gpu=self.gpu or 0, exchange=getattr(self.scheduler, "exchange", None))
            message = {"subtask_id": task["subtask_id"], "job_id": task["job_id"], "session_id": task.get("session_id"),
                       "status": "completed", "result": result}
        except Exception as e:
            err = str(e)
            log.exception("subtask %s of job %s failed", task.get("subtask_id"), task.get("job_id"))
            if "candidates" in task:
                result = [{"subtask_id": c["subtask_id"], "status": "failed", "error": err} for c in task["candidates"]]
                message = {"subtask_id": task["subtask_id"], "job_id": task["job_id"],
                           "session_id": task.get("session_id"), "status": "failed", "error": err, "result": result}
            else:
                message = {"subtask_id": task["subtask_id"], "job_id": task["job_id"],
                           "session_id": task.get("session_id"), "status": "failed", "error": err}
        finished = time.time()
        n_fits = len(task.get("candidates", [0])) * (len(task.get("folds", [])) or 1)
        self.scheduler.publish_metrics({
            "worker_id": self.worker_id, "subtask_id": task["subtask_id"], "status": "DONE",
            "received_at": received, "started_at": started, "finished_at": finished,
            "cpu_percent_avg": psutil.cpu_percent(None) if psutil else cpu0,
            "mem_percent_avg": psutil.virtual_memory().percent if psutil else mem0,
            "algo": task.get("model_type"), "gpu_id": self.gpu, "fits": n_fits,
            "fits_per_sec": n_fits / max(finished - started, 1e-9)})
        self.scheduler.publish_result(self.worker_id, message)

    def shutdown(self):
        self.stop_event.set()
        self.queue.put((self.worker_id, None))
        self.scheduler.unsubscribe(self.worker_id)
