"""``MLTaskManager``: the reference client API over the in-process master.

Same constructor, methods, payloads and return shapes as
DistributedLibrary/src/distributed_ml/core.py:15-213.  The reference speaks HTTP to
a Flask master; here ``url`` selects a backend:
  * ``"local://"`` (or ``"local://<n_gpus>"``)  -- master, scheduler and one worker
    thread per GPU inside this process (the drop-in for the train path);
  * a ``Master`` instance                       -- an already-wired backend (tests).
Two reference defects are not copied: ``train`` talks to the
non-streaming /train route, and ``check_job_status`` returns the /check_status
envelope that ``_wait_for_completion`` (core.py:180-199) and demo_results.py:12-19
actually expect.
"""
from __future__ import annotations

import json
import math
import time
import uuid
from datetime import datetime

import numpy as np
from sklearn.model_selection import GridSearchCV, RandomizedSearchCV

from . import datasets
from .master import Master
from .scheduler import Scheduler
from .worker import Worker


def _json_serializer(obj):
    """core.py:60-72: numpy scalars -> python, arrays -> lists, anything else -> str."""
    if isinstance(obj, np.integer):
        return int(obj)
    if isinstance(obj, np.floating):
        return float(obj)
    if isinstance(obj, np.ndarray):
        return obj.tolist()
    if isinstance(obj, np.bool_):
        return bool(obj)
    return str(obj)


def _clean_nan(obj):
    """core.py:74-80: NaN / Inf -> None so the payload is strict JSON."""
    if isinstance(obj, dict):
        return {k: _clean_nan(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return [_clean_nan(v) for v in obj]
    if isinstance(obj, float) and (math.isnan(obj) or math.isinf(obj)):
        return None
    return obj


def wire(payload):
    """What the payload looks like after the client's JSON round trip (core.py:51-52)."""
    return json.loads(json.dumps(_clean_nan(json.loads(json.dumps(payload, default=_json_serializer)))))


class LocalBackend:
    """Master + scheduler + worker threads in this process."""

    def __init__(self, n_gpus: int | None = None, partition: str = "candidates", gpus: list[int] | None = None,
                 monitor_interval: float | None = 15.0):
        if gpus is not None:
            n_gpus = len(gpus)
        if n_gpus is None:
            try:
                import torch
                n_gpus = max(1, torch.cuda.device_count())
            except Exception:
                n_gpus = 1
        self.scheduler = Scheduler()
        self.master = Master(self.scheduler, n_gpus=n_gpus, partition=partition)
        self.workers = [Worker(self.scheduler, gpu=g, slot=i) for i, g in enumerate(gpus or range(n_gpus))]
        for w in self.workers:
            w.start()
        if monitor_interval:       # scheduler_service.py:205-225: dead-worker check every 15 s
            self.scheduler.start_monitor(monitor_interval)

    def shutdown(self):
        self.scheduler.stop_monitor()
        for w in self.workers:
            w.shutdown()


class MLTaskManager:
    def __init__(self, url="local://"):
        """core.py:15-27: stores the endpoint and opens a session."""
        self.result = None
        self.backend = None
        if isinstance(url, Master):
            self.master = url
        elif isinstance(url, LocalBackend):
            self.backend, self.master = url, url.master
        elif isinstance(url, str) and url.startswith("local://"):
            tail = url[len("local://"):]
            self.backend = LocalBackend(int(tail) if tail.isdigit() else None)
            self.master = self.backend.master
        else:
            raise ValueError("this build serves the training path in-process: use url='local://' "
                             "(see INTEGRATION.md for putting the Flask/Kafka transports back)")
        self.api_url = url if isinstance(url, str) else "local://"
        self.session_id = self._create_session()
        self.job_id = None

    # -- transport (core.py:29-58): errors become {"status": "error", ...}, never exceptions
    def _api_request(self, endpoint: str, method: str, data=None):
        try:
            parts = endpoint.strip("/").split("/")
            route = parts[0]
            if route == "create_session":
                return self.master.create_session()
            if route in ("train", "train_status"):
                return self.master.train(parts[1], wire(data))
            if route == "check_status":
                return self.master.check_status(parts[1], parts[2])
            if route == "metrics":
                return self.master.metrics(parts[1], parts[2])
            if route == "check_data":
                return self.master.check_data(parts[1], (data or {}).get("dataset_name"))
            return {"status": "error", "message": f"route {route!r} is outside the training path"}
        except Exception as e:
            return {"status": "error", "message": str(e)}

    def _create_session(self):
        return self._api_request("create_session", "post").get("session_id")

    # -- datasets: Kaggle/S3 download is replaced by local arrays (BASELINE.json north_star)
    def check_data(self, dataset_name):
        return self._api_request(f"check_data/{self.session_id}", "get", data={"dataset_name": dataset_name})

    def download_data(self, data_type, data_link, data_name):
        """core.py:82-94.  ``data_type`` "local": ``data_link`` is a CSV path whose last
        column is the target (worker.py:428-429); remote sources are out of scope."""
        if data_type != "local":
            return {"status": "error", "message": "only data_type='local' is available offline"}
        try:
            datasets.load_csv_dataset(data_name, data_link)
            return {"status": "success", "dataset_name": data_name}
        except Exception as e:
            return {"status": "error", "message": str(e)}

    def register_data(self, data_name, X, y):
        """Local/synthetic arrays of the named shape."""
        datasets.register_dataset(data_name, X, y)
        return {"status": "success", "dataset_name": data_name}

    def preprocess(self, dataset_name, yaml):
        return {"status": "error", "message": "preprocessing is outside the training path (SURVEY.md section 2 row 9)"}

    # -- core.py:96-150
    def _extract_model_details(self, estimator):
        if isinstance(estimator, (GridSearchCV, RandomizedSearchCV)):
            base = estimator.estimator
            is_grid = isinstance(estimator, GridSearchCV)
            search = ({"param_grid": estimator.param_grid} if is_grid else
                      {"param_distributions": estimator.param_distributions, "n_iter": estimator.n_iter})
            cv_params = {"cv": estimator.cv, "scoring": estimator.scoring, "refit": estimator.refit,
                         "verbose": estimator.verbose, "error_score": estimator.error_score,
                         "return_train_score": estimator.return_train_score}
            return {"model_type": type(base).__name__,
                    "search_type": "GridSearchCV" if is_grid else "RandomizedSearchCV",
                    "hyperparameters": {
                        "base_estimator_params": {k.split("__")[-1]: v for k, v in base.get_params().items()},
                        "search_params": search, "cv_params": cv_params}}
        return {"model_type": type(estimator).__name__, "hyperparameters": dict(estimator.get_params())}

    # -- core.py:152-174
    def train(self, estimator, dataset_name, train_params=None, wait_for_completion=False):
        """Submits a training job."""
        self.job_id = str(uuid.uuid4())
        if train_params is None:
            train_params = {}
        if "test_size" not in train_params and not isinstance(estimator, (GridSearchCV, RandomizedSearchCV)):
            train_params["test_size"] = 0.2
        payload = {"job_id": self.job_id, "session_id": self.session_id, "dataset_id": dataset_name,
                   "model_details": self._extract_model_details(estimator), "train_params": train_params,
                   "timestamp": datetime.now().isoformat()}
        api_response = self._api_request(f"train/{self.session_id}", "post", data=payload)
        if wait_for_completion and api_response.get("status") != "error":
            return self._wait_for_completion(self.job_id)
        return api_response

    # -- core.py:176-178 (route fixed, see module docstring)
    def check_job_status(self, job_id):
        return self._api_request(f"check_status/{self.session_id}/{job_id}", "get")

    # -- core.py:180-199
    def _wait_for_completion(self, job_id, polling_interval=0.01, timeout=3600):
        start = time.time()
        while True:
            if time.time() - start > timeout:
                return {"status": "error", "message": "Timeout exceeded"}
            self.master.wait(self.session_id, job_id, polling_interval)
            status = self.check_job_status(job_id)
            if status.get("status") == "error" or status.get("job_status") in ("completed", "failed"):
                self.result = status
                return status

    def download_best_model(self, job_id, model_path, model_id):
        """core.py:201-206: artefacts are local files; returns the path when it exists."""
        import os
        if model_path and os.path.exists(model_path):
            return {"status": "success", "model_path": model_path, "model_id": model_id}
        return {"status": "error", "message": "model artefact not found"}

    def close(self):
        if self.backend is not None:
            self.backend.shutdown()
