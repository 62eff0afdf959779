"""Scheduler shim: static GPU placement behind the reference's worker protocol.

The reference scheduler (aws-prod/scheduler/{scheduler.py,scheduler_service.py})
registers workers over HTTP (/subscribe, /heartbeat, /unsubscribe), predicts a
runtime per subtask with a GBRT and re-keys each message to the worker that
minimises load/speed + est/speed.  On one box with 1-8 identical B200s the
placement problem is static: a block subtask already
names its GPU, so this shim keeps the registration / heartbeat / keyed-delivery
contract (and the dead-worker requeue, scheduler_service.py:205-247) and drops
the predictor.  Kafka topics "tasks"/"train"/"result"/"metrics" become in-process
queues and callbacks carrying the same dicts.
"""
from __future__ import annotations

import queue
import threading
import time
import uuid

HEARTBEAT_TIMEOUT_S = 10.0   # scheduler_service.py:214


def _task_fits(task: dict) -> int:
    """Work of a subtask in candidate-fold fits (a legacy subtask is one candidate, all folds)."""
    cv = task.get("train_params", {}).get("cv", 5)
    folds = len(task["folds"]) if task.get("folds") else (cv if isinstance(cv, int) else 5)
    return max(1, len(task.get("candidates", [0])) * max(1, folds))


class Scheduler:
    def __init__(self):
        self.lock = threading.Lock()
        self.workers: dict[str, dict] = {}
        self.master = None
        self.backlog: list[dict] = []
        self.mailbox: dict[str, bytes] = {}
        self.mail_cv = threading.Condition()
        self.monitor = None
        self.monitor_stop = threading.Event()

    def attach_master(self, master) -> None:
        self.master = master

    # POST /subscribe {mem_capacity_mb, host} -> {status, worker_id}   (scheduler.py:105-117)
    def subscribe(self, mem_capacity_mb: int = 0, host: str = "localhost", gpu: int | None = None) -> dict:
        wid = f"worker-{uuid.uuid4().hex[:8]}"
        with self.lock:
            self.workers[wid] = {"queue": queue.Queue(), "last_seen": time.time(), "gpu": gpu,
                                 "mem_capacity_mb": mem_capacity_mb, "host": host, "inflight": {}, "done": 0}
            backlog, self.backlog = self.backlog, []
        for task in backlog:
            self.submit(task)
        return {"status": "subscribed", "worker_id": wid}

    # POST /heartbeat {worker_id}
    def heartbeat(self, worker_id: str) -> dict:
        with self.lock:
            w = self.workers.get(worker_id)
            if w is None:
                return {"status": "unknown_worker"}
            w["last_seen"] = time.time()
        return {"status": "ok"}

    # POST /unsubscribe {worker_id}: hand its queue to the survivors (scheduler.py:120-139)
    def unsubscribe(self, worker_id: str) -> dict:
        with self.lock:
            w = self.workers.pop(worker_id, None)
        if w is None:
            return {"status": "unknown_worker"}
        self._requeue(w)
        return {"status": "unsubscribed"}

    def n_workers(self) -> int:
        with self.lock:
            return len(self.workers)

    def get_queue(self, worker_id: str) -> queue.Queue:
        return self.workers[worker_id]["queue"]

    # topic "tasks" -> topic "train" keyed by worker id (scheduler_service.py:249-293)
    def submit(self, task: dict) -> str | None:
        with self.lock:
            if not self.workers:
                self.backlog.append(task)
                return None
            wid = self._select_worker(task)
            w = self.workers[wid]
            w["inflight"][task["subtask_id"]] = task
            w["queue"].put((wid, task))
        return wid

    def _select_worker(self, task: dict) -> str:
        gpu = task.get("gpu")
        ids = sorted(self.workers)
        for wid in ids:                       # the block names its GPU
            if gpu is not None and self.workers[wid]["gpu"] == gpu:
                return wid
        # scheduler_service.py:173-191: minimise load/speed + est/speed; "speed" is the fits/s this worker
        # reported on its last blocks (publish_metrics), work is counted in candidate-fold fits
        est = _task_fits(task)

        def cost(w):
            rec = self.workers[w]
            load = sum(_task_fits(t) for t in rec["inflight"].values())
            return ((load + est) / rec.get("speed", 1.0), w)
        return min(ids, key=cost)

    # topics "result" / "metrics"
    def publish_result(self, worker_id: str, message: dict) -> None:
        with self.lock:
            w = self.workers.get(worker_id)
            if w is not None:
                w["inflight"].pop(message.get("subtask_id"), None)
                w["done"] += 1
        if self.master is not None:
            self.master.on_result(message)

    def publish_metrics(self, record: dict) -> None:
        fps = record.get("fits_per_sec")
        if fps:                               # speed factor of the placement rule: moving average of measured fits/s
            with self.lock:
                w = self.workers.get(record.get("worker_id"))
                if w is not None:
                    w["speed"] = 0.5 * w["speed"] + 0.5 * float(fps) if "speed" in w else float(fps)
        if self.master is not None:
            self.master.on_metrics(record)

    # monitor (scheduler_service.py:205-247)
    def reap_dead_workers(self, now: float | None = None) -> list[str]:
        now = time.time() if now is None else now
        with self.lock:
            dead = [wid for wid, w in self.workers.items() if now - w["last_seen"] > HEARTBEAT_TIMEOUT_S]
            gone = [self.workers.pop(wid) for wid in dead]
        for w in gone:
            self._requeue(w)
        return dead

    def _requeue(self, w: dict) -> None:
        tasks = list(w["inflight"].values())
        while True:
            try:
                _, t = w["queue"].get_nowait()
            except queue.Empty:
                break
            if t is not None and t["subtask_id"] not in w["inflight"]:
                tasks.append(t)
        for t in tasks:
            t = dict(t)
            t.pop("gpu", None)               # any surviving GPU may take it
            self.submit(t)

    # side channel between the workers of one job (Redis in the reference's deployment): the worker of row
    # shard 0 posts the 128-byte ncclUniqueId of a new communicator, the other shards' workers wait for it
    def exchange(self, key: str, payload: bytes | None = None, timeout: float = 120.0) -> bytes:
        with self.mail_cv:
            if payload is not None:
                self.mailbox[key] = payload
                self.mail_cv.notify_all()
                return payload
            if not self.mail_cv.wait_for(lambda: key in self.mailbox, timeout):
                raise TimeoutError(f"nothing was posted under {key!r} within {timeout} s")
            return self.mailbox[key]

    # monitor loop (scheduler_service.py:205-225): every `interval` seconds drop workers whose last
    # heartbeat is older than HEARTBEAT_TIMEOUT_S and hand their tasks to the survivors
    def start_monitor(self, interval: float = 15.0) -> None:
        if self.monitor is not None:
            return
        self.monitor_stop.clear()

        def loop():
            while not self.monitor_stop.wait(interval):
                try:
                    self.reap_dead_workers()
                except Exception:
                    pass
        self.monitor = threading.Thread(target=loop, daemon=True, name="scheduler-monitor")
        self.monitor.start()

    def stop_monitor(self) -> None:
        self.monitor_stop.set()
        self.monitor = None

    def workers_view(self) -> dict:      # GET /workers
        with self.lock:
            return {wid: {"gpu": w["gpu"], "inflight": len(w["inflight"]), "done": w["done"],
                          "last_seen": w["last_seen"], "speed": w.get("speed")} for wid, w in self.workers.items()}
