"""In-process master: job intake, splitter, result aggregation.

Mirrors aws-prod/master/{master.py,task_handler.py,redis_util.py} of the reference
for the training path only: the Flask routes become methods, Redis becomes a dict,
the Kafka topics become the scheduler's in-process queues.  Message dicts keep the
reference's keys so that a maintainer can put the real transports back
(INTEGRATION.md).

What changes, per BASELINE.json north_star:
  * splitter: the one-subtask-per-candidate list (task_handler.py:186-236) is
    regrouped into per-GPU candidate x fold *block* subtasks (``create_blocks``);
  * aggregation: besides the reference's {results, best_result, completion_time}
    (task_handler.py:102-114, 254-263) the job result carries cv_results_ /
    best_params_ / best_score_ / best_index_ formed by GridSearchCV's rules
    (sklearn/model_selection/_search.py:1208-1294, 862-874).
"""
from __future__ import annotations

import copy
import threading
import time
import uuid
from datetime import datetime

import numpy as np
from scipy.stats import rankdata
from sklearn.model_selection import ParameterGrid, ParameterSampler

from . import datasets


# --------------------------------------------------------------------------------------
# splitter
# --------------------------------------------------------------------------------------
def create_subtasks(job_config: dict) -> list[dict]:
    """Reference semantics of task_handler.py:156-252, key for key.

    GridSearchCV -> ParameterGrid order (sorted keys, last key fastest);
    RandomizedSearchCV -> ParameterSampler(..., random_state=cv_params.random_state or 42)
    (the client drops the estimator's own random_state, core.py:117-130);
    plain estimator -> one subtask whose ``parameters`` are the hyperparameters.
    """
    job_id = job_config["job_id"]
    common = {"job_id": job_id, "session_id": job_config["session_id"], "dataset_id": job_config["dataset_id"]}
    details = job_config["model_details"]
    train_params = job_config.get("train_params", {})
    if "search_type" not in details:
        return [{"subtask_id": f"{job_id}-subtask-1", **common, "model_type": details["model_type"],
                 "parameters": details["hyperparameters"], "train_params": train_params}]
    hyper = details["hyperparameters"]
    base = hyper["base_estimator_params"]
    cv_params = hyper.get("cv_params", {})
    search = hyper.get("search_params", {})
    cv = cv_params.get("cv", 5)
    if details["search_type"] == "GridSearchCV":
        combos = list(ParameterGrid(search.get("param_grid", {})))
    elif details["search_type"] == "RandomizedSearchCV":
        combos = list(ParameterSampler(search.get("param_distributions", {}), search.get("n_iter", 10),
                                       random_state=cv_params.get("random_state", 42)))
    else:
        combos = []
    return [{"subtask_id": f"{job_id}-subtask-{i + 1}", **common, "model_type": details["model_type"],
             "parameters": {**base, **params}, "train_params": {"cv": cv, **train_params}}
            for i, params in enumerate(combos)]


def create_blocks(subtasks: list[dict], n_gpus: int = 1, partition: str = "candidates",
                  cost: list[float] | None = None, extra_train_params: dict | None = None) -> list[dict]:
    """Regroup per-candidate subtasks into per-GPU block subtasks.

    A block keeps the subtask keys and adds ``candidates`` [{index, subtask_id,
    parameters}], ``folds``, ``gpu`` and ``partition``.  "candidates": GPU g gets
    candidates g, g+G, g+2G, ... (round-robin spreads slow large-C columns);
    "rows": every GPU gets all candidates and row shard ``shard`` of ``n_shards``
    (dist.row_slice); the workers run the sweep together and the library sums
    gradients / moments / counts over NCCL (worker.train_model).  A block with a
    single candidate and all folds is the legacy subtask.  ``extra_train_params``
    (the search's ``scoring`` / ``return_train_score``, which the reference client
    serialises under cv_params and nobody reads, core.py:123-130) is merged into the
    block's ``train_params``; the per-candidate subtasks stay as the reference makes them.

    ``cost`` (one predicted number of batched rounds per candidate, ``IterationModel``)
    replaces the round-robin deal by longest-first dealing: candidates sorted by cost,
    each to the GPU with the least work so far.  A GPU's sweep time follows the sum of
    its columns' rounds, and a randomized search enumerates candidates in draw order,
    where round-robin leaves that sum to chance.
    """
    if not subtasks:
        return []
    n_gpus = max(1, min(n_gpus, len(subtasks))) if partition == "candidates" else max(1, n_gpus)
    first = subtasks[0]
    cv = first.get("train_params", {}).get("cv", 5)
    blocks = []
    deal = deal_by_cost(cost, n_gpus) if (cost is not None and partition == "candidates") else None
    for g in range(n_gpus):
        if partition != "candidates":
            own = list(range(len(subtasks)))
        else:
            own = deal[g] if deal is not None else list(range(len(subtasks)))[g::n_gpus]
        mine = [subtasks[i] for i in own]
        blocks.append({
            "subtask_id": f"{first['job_id']}-block-{g + 1}",
            "job_id": first["job_id"], "session_id": first["session_id"], "dataset_id": first["dataset_id"],
            "model_type": first["model_type"],
            "train_params": {**copy.deepcopy(first.get("train_params", {})), **(extra_train_params or {}),
                             **({"shard_group": f"{n_gpus}gpus"} if partition == "rows" else {})},
            "candidates": [{"index": i, "subtask_id": s["subtask_id"], "parameters": s["parameters"]}
                           for i, s in zip(own, mine)],
            "folds": list(range(cv)) if isinstance(cv, int) else cv,
            "gpu": g, "partition": partition, "n_shards": n_gpus if partition == "rows" else 1,
            "shard": g if partition == "rows" else 0,
        })
    return blocks


def deal_by_cost(cost: list[float], n_gpus: int) -> list[list[int]]:
    """Longest-processing-time dealing: indices by falling cost (ties: enumeration order), each to
    the least-loaded GPU (ties: lowest GPU).  Every GPU gets at least one candidate when there are
    enough; each GPU's list comes back in enumeration order."""
    order = sorted(range(len(cost)), key=lambda i: (-float(cost[i]), i))
    load = [0.0] * n_gpus
    own: list[list[int]] = [[] for _ in range(n_gpus)]
    for i in order:
        g = min(range(n_gpus), key=lambda k: (-1.0 if not own[k] else load[k], k))
        own[g].append(i)
        load[g] += max(float(cost[i]), 0.0)
    return [sorted(o) for o in own]


class IterationModel:
    """Predicted batched rounds of one candidate: the scheduler's runtime estimate for this path.

    The reference predicts a subtask's runtime with a GBRT over dataset and model features
    (scheduler_service.py:40-84) and places by ``load/speed + est/speed`` (:173-191).  On the GPU a
    block's time follows the sum over its columns of L-BFGS rounds (every round streams X once for
    the columns still live), so the quantity to predict is ``n_iter_`` per candidate.  Ridge-family
    candidates are direct solves (cost 1).  ``n_iter_`` is not monotone in C (scikit-learn on 50 000 x 128
    synthetic rows, C = 1e-6 .. 1e2: 7, 5, 3, 6, 7, 7, 7, 7, 7 iterations at tol 1e-4 and about one more
    at tol 1e-5), so nothing is assumed about C before a job has run: the prior is flat in C (equal costs
    deal out like round-robin) and only counts the decades of tol below 1e-4.  Finished jobs report
    ``n_iter`` per fold (worker result dict): per (model, dataset) the model keeps the mean rounds seen
    at each (log10 C, log10 tol) and interpolates in log10 C among the observations with the nearest tol,
    holding the edge values outside the observed range.
    """

    def __init__(self):
        self.seen: dict[tuple[str, str], dict[tuple[float, float], tuple[float, int]]] = {}

    @staticmethod
    def _key(params: dict) -> tuple[float, float]:
        c = float(params.get("C", 1.0) or 1.0)
        tol = float(params.get("tol", 1e-4) or 1e-4)
        return (float(np.log10(max(c, 1e-300))), float(np.log10(max(tol, 1e-300))))

    @staticmethod
    def prior(lc: float, lt: float, max_iter: float) -> float:
        return float(min(max_iter, 6.0 + max(0.0, -4.0 - lt)))

    def predict(self, model_type: str, dataset_id: str, params: dict) -> float:
        if model_type != "LogisticRegression":
            return 1.0
        lc, lt = self._key(params)
        max_iter = float(params.get("max_iter", 100) or 100)
        obs = self.seen.get((model_type, dataset_id))
        if not obs:
            return self.prior(lc, lt, max_iter)
        tol_near = min({k[1] for k in obs}, key=lambda t: abs(t - lt))
        pts = sorted((k[0], v[0]) for k, v in obs.items() if k[1] == tol_near)
        est = float(np.interp(lc, [p[0] for p in pts], [p[1] for p in pts]))
        return float(min(max_iter, max(1.0, est)))       # np.interp holds the edge values outside the range

    def observe(self, model_type: str, dataset_id: str, params: dict, n_iter) -> None:
        if model_type != "LogisticRegression" or n_iter is None:
            return
        it = np.asarray(n_iter, dtype=np.float64).ravel()
        it = it[np.isfinite(it)]
        if it.size == 0:
            return
        tab = self.seen.setdefault((model_type, dataset_id), {})
        k = self._key(params)
        mean, cnt = tab.get(k, (0.0, 0))
        tab[k] = ((mean * cnt + float(it.mean())) / (cnt + 1), cnt + 1)


# --------------------------------------------------------------------------------------
# aggregation
# --------------------------------------------------------------------------------------
def aggregate_results(results: list[dict]):
    """Best result = highest mean_cv_score; NaN (failed fits) ranks last; ties go to
    the lowest candidate index.  The reference sorts by mean_cv_score descending and
    takes the first (task_handler.py:254-263) but feeds it Redis KEYS order
    (:76), which makes ties nondeterministic; sklearn's rule is used instead."""
    if not results:
        return None
    scores = np.array([_nan_if_none(r.get("mean_cv_score")) for r in results], dtype=np.float64)
    if np.isnan(scores).all():
        return results[0]
    return results[int(np.nanargmax(scores))]  # nanargmax returns the first maximum


def _nan_if_none(v):
    return np.nan if v is None else float(v)


def build_cv_results(candidates: list[dict], fold_scores, train_scores=None) -> dict:
    """cv_results_-style dict + best_* (sklearn _search.py:1208-1294, 862-874).  ``train_scores``
    (return_train_score=True) adds split{k}_train_score / mean_train_score / std_train_score."""
    a = np.asarray(fold_scores, dtype=np.float64)
    n_cand, n_folds = a.shape
    mean = np.average(a, axis=1)
    std = np.sqrt(np.average((a - mean[:, None]) ** 2, axis=1))
    if np.isnan(mean).all():
        rank = np.ones(n_cand, dtype=np.int32)
    else:
        filled = np.nan_to_num(mean, nan=np.nanmin(mean) - 1.0)
        rank = rankdata(-filled, method="min").astype(np.int32)
    best = int(rank.argmin())
    res = {f"split{k}_test_score": a[:, k].tolist() for k in range(n_folds)}
    res.update(mean_test_score=mean.tolist(), std_test_score=std.tolist(), rank_test_score=rank.tolist(),
               params=candidates)
    if train_scores is not None:
        t = np.asarray(train_scores, dtype=np.float64)
        tm = np.average(t, axis=1)
        res.update({f"split{k}_train_score": t[:, k].tolist() for k in range(t.shape[1])})
        res.update(mean_train_score=tm.tolist(),
                   std_train_score=np.sqrt(np.average((t - tm[:, None]) ** 2, axis=1)).tolist())
    keys = sorted({k for c in candidates for k in c})
    for k in keys:
        res[f"param_{k}"] = [c.get(k) for c in candidates]
    return {"cv_results_": res, "best_index_": best, "best_params_": candidates[best],
            "best_score_": float(mean[best])}


# --------------------------------------------------------------------------------------
# master service
# --------------------------------------------------------------------------------------
class Master:
    """Session/job state + the routes the client uses (master.py:27-62, 115-268)."""

    def __init__(self, scheduler, n_gpus: int | None = None, partition: str = "candidates"):
        self.scheduler = scheduler
        self.n_gpus = n_gpus
        self.partition = partition
        self.sessions: set[str] = set()
        self.jobs: dict[tuple[str, str], dict] = {}
        self.lock = threading.Lock()
        self.submit_lock = threading.Lock()
        self.iterations = IterationModel()      # rounds per candidate, learnt from finished jobs
        scheduler.attach_master(self)

    # /create_session
    def create_session(self) -> dict:
        sid = str(uuid.uuid4())
        with self.lock:
            self.sessions.add(sid)
        return {"session_id": sid}

    # /check_data and /download_data (local arrays / CSV only)
    def check_data(self, session_id: str, dataset_name: str) -> dict:
        return {"exists": datasets.has_dataset(dataset_name), "dataset_name": dataset_name}

    # /train and /train_status
    def train(self, session_id: str, payload: dict) -> dict:
        with self.lock:
            if session_id not in self.sessions:
                return {"status": "error", "message": "Invalid or expired session_id"}
        if not datasets.has_dataset(payload["dataset_id"]):
            return {"status": "error", "message": f"Dataset '{payload['dataset_id']}' not found"}
        subtasks = create_subtasks(payload)
        if not subtasks:
            return {"status": "error", "message": "No subtasks created"}
        details = payload["model_details"]
        search = details.get("hyperparameters", {}).get("search_params", {}) if "search_type" in details else {}
        base = details.get("hyperparameters", {}).get("base_estimator_params", {}) if "search_type" in details else {}
        job = {"job_id": payload["job_id"], "session_id": session_id, "subtasks": subtasks,
               "results": {}, "failed": {}, "status": "pending", "created_at": time.time(),
               "search_type": details.get("search_type"), "base_params": base, "search_params": search,
               "metrics": [], "done": threading.Event(),
               # GridSearchCV's refit step (_search.py:1146-1159).  The reference never refits (its
               # client serialises cv_params.refit and nobody reads it), so it is opt-in here.
               "refit": bool(payload.get("train_params", {}).get("refit")) and "search_type" in details}
        with self.lock:
            self.jobs[(session_id, payload["job_id"])] = job
        n_gpus = self.n_gpus or max(1, self.scheduler.n_workers())
        with self.lock:
            cost = [self.iterations.predict(st["model_type"], st["dataset_id"], st["parameters"]) for st in subtasks]
        cvp = details.get("hyperparameters", {}).get("cv_params", {}) if "search_type" in details else {}
        extra = {}
        if cvp.get("scoring") is not None:
            extra["scoring"] = cvp["scoring"]
        if cvp.get("return_train_score"):
            extra["return_train_score"] = True
        partition = payload.get("train_params", {}).get("partition", self.partition)
        blocks = create_blocks(subtasks, n_gpus=n_gpus, partition=partition, cost=cost, extra_train_params=extra)
        # one step for all blocks of a job: the workers of a row-sharded job meet in collectives, so every
        # worker's queue must hold the jobs' blocks in the same order
        with self.submit_lock:
            for block in blocks:
                self.scheduler.submit(block)       # topic "tasks"
        return {"status": "Training started", "job_id": payload["job_id"], "total_subtasks": len(subtasks)}

    # topic "result" consumer (task_handler.py:18-123)
    def on_result(self, message: dict) -> None:
        key = (message.get("session_id"), message.get("job_id"))
        with self.lock:
            job = self.jobs.get(key) or next((j for (s, jid), j in self.jobs.items() if jid == message.get("job_id")), None)
            if job is None or job["status"] == "completed":
                return          # at-least-once delivery: a requeued block may report a finished job again
            entries = message["result"] if isinstance(message.get("result"), list) else [message]
            if message.get("subtask_id") == f"{job['job_id']}-refit":
                if not entries or "pending_result" not in job:
                    return
                e = entries[0]
                ok = e.get("status", message.get("status")) == "completed"
                self._complete(job, job.pop("pending_result"), e.get("result") if ok else None,
                               None if ok else e.get("error", message.get("error", "failed")))
                return
            if job.get("aggregated"):
                return          # all candidates are in and the refit is under way
            known = {st["subtask_id"] for st in job["subtasks"]}
            fresh = 0
            for e in entries:
                sid = e.get("subtask_id", message.get("subtask_id"))
                if sid not in known or sid in job["results"]:
                    continue    # not a candidate of this job (a block id), or already recorded (another shard / a rerun)
                fresh += 1
                if e.get("status", message.get("status")) == "completed":
                    job["results"][sid] = e.get("result", e)
                    job["failed"].pop(sid, None)
                else:
                    job["failed"][sid] = e.get("error", message.get("error", "failed"))
            if isinstance(message.get("result"), list) and message.get("status") == "failed" and not entries:
                for sid in known - set(job["results"]):
                    job["failed"][sid] = message.get("error", "failed")
                    fresh += 1
            total = len(job["subtasks"])
            finished = len(job["results"]) + len(job["failed"])
            job["status"] = str(len(job["results"]) / total * 100.0)
            refit_task = None
            if fresh and finished == total:
                job["aggregated"] = True
                refit_task = self._finish(job)
        if refit_task is not None:
            self.scheduler.submit(refit_task)

    def on_metrics(self, record: dict) -> None:
        with self.lock:
            for (s, jid), job in self.jobs.items():
                if record.get("subtask_id", "").startswith(jid):
                    job["metrics"].append(record)

    def _finish(self, job: dict) -> dict | None:
        """All candidates are in: aggregate.  Returns the refit block to submit, if one is due."""
        ordered = []
        for st in job["subtasks"]:
            r = job["results"].get(st["subtask_id"])
            if r is None:  # failed candidate: error_score = nan row, like GridSearchCV (_search.py:1251-1265)
                cv = st.get("train_params", {}).get("cv", 5)
                r = {"cv_scores": [float("nan")] * (cv if isinstance(cv, int) else 5),
                     "mean_cv_score": float("nan"), "parameters": st["parameters"],
                     "error": job["failed"].get(st["subtask_id"])}
            else:
                self.iterations.observe(st["model_type"], st["dataset_id"], st["parameters"], r.get("n_iter"))
            ordered.append(r)
        out = {"results": ordered, "best_result": aggregate_results(ordered), "completion_time": time.time()}
        if job["search_type"]:
            keys = _searched(job)
            cands = [{k: st["parameters"][k] for k in keys if k in st["parameters"]} for st in job["subtasks"]]
            nf = max(len(r.get("cv_scores", [])) for r in ordered)
            scores = [list(r.get("cv_scores", [])) + [float("nan")] * (nf - len(r.get("cv_scores", []))) for r in ordered]
            train = None
            if all("train_scores" in r or "error" in r for r in ordered) and any("train_scores" in r for r in ordered):
                train = [list(r.get("train_scores", [])) + [float("nan")] * (nf - len(r.get("train_scores", []))) for r in ordered]
            out.update(build_cv_results(cands, scores, train))
        best = out.get("best_index_")
        if job["refit"] and best is not None and np.isfinite(out["best_score_"]):
            st = job["subtasks"][best]       # best candidate again, all rows, one more block subtask
            rid = f"{job['job_id']}-refit"
            tp = {k: v for k, v in st.get("train_params", {}).items() if k != "test_size"}
            job["pending_result"] = out
            return {"subtask_id": rid, "job_id": job["job_id"], "session_id": job["session_id"],
                    "dataset_id": st["dataset_id"], "model_type": st["model_type"],
                    "train_params": {**tp, "refit_fit": True},
                    "candidates": [{"index": best, "subtask_id": rid, "parameters": st["parameters"]}],
                    "folds": [], "gpu": 0, "partition": "candidates", "n_shards": 1}
        self._complete(job, out)
        return None

    def _complete(self, job: dict, out: dict, refit_result: dict | None = None, refit_error: str | None = None) -> None:
        if refit_result is not None:
            out["best_estimator_"] = {k: refit_result.get(k) for k in
                                      ("model_id", "model_path", "coef", "intercept", "parameters", "executor")}
            out["refit_time_"] = refit_result.get("training_time")
        elif refit_error is not None:
            out["refit_error"] = refit_error
        out["completion_time"] = time.time()
        job["job_result"] = out
        job["status"] = "completed"
        job["done"].set()

    # /check_status (master.py:115-167)
    def check_status(self, session_id: str, job_id: str) -> dict:
        with self.lock:
            job = self.jobs.get((session_id, job_id))
            if job is None:
                return {"status": "error", "message": f"job {job_id} not found"}
            total = len(job["subtasks"])
            pending = total - len(job["results"]) - len(job["failed"])
            resp = {"session_id": session_id, "tasks_pending": pending, "job_id": job_id,
                    "job_status": job["status"], "total_subtasks": total}
            if job["status"] == "completed":
                resp["job_result"] = job["job_result"]
                if total > 1:
                    resp["best_result"] = job["job_result"]["best_result"]
            return resp

    # /metrics (master.py:294-340): one resource record per finished block
    def metrics(self, session_id: str, job_id: str) -> list:
        with self.lock:
            job = self.jobs.get((session_id, job_id))
            return list(job["metrics"]) if job else []

    def wait(self, session_id: str, job_id: str, timeout: float | None = None) -> bool:
        job = self.jobs.get((session_id, job_id))
        return bool(job and job["done"].wait(timeout))


def _searched(job) -> list[str]:
    sp = job.get("search_params") or {}
    grid = sp.get("param_grid") or sp.get("param_distributions") or {}
    if isinstance(grid, list):
        return sorted({k for g in grid for k in g})
    return sorted(grid)
