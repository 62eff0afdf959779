"""Host-side logic: splitter, blocks, aggregation, scheduler, client API, C-ABI exports."""
import ctypes
import json
import os
import re
import time

import numpy as np
import pytest
from sklearn.linear_model import LogisticRegression, RidgeClassifier
from sklearn.model_selection import GridSearchCV, ParameterSampler, RandomizedSearchCV

import cs230_distributed_machine_learning_b200 as pkg
from cs230_distributed_machine_learning_b200 import worker as wk
from cs230_distributed_machine_learning_b200 import _cabi, build, folds, master, scheduler, worker
from cs230_distributed_machine_learning_b200.client import MLTaskManager, wire

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_create_subtasks_matches_notebook_fixture(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "notebook_subtasks.json")))
    got = master.create_subtasks(g["job_config"])
    assert json.loads(json.dumps(got)) == json.loads(json.dumps(g["subtasks"]))
    assert [s["parameters"]["C"] for s in got] == [0.1, 0.1, 1.0, 1.0, 10.0, 10.0, 100, 100]
    assert [s["parameters"]["solver"] for s in got] == ["liblinear", "lbfgs"] * 4


def test_create_subtasks_randomized_matches_reference(golden_dir):
    g = json.load(open(os.path.join(golden_dir, "notebook_subtasks.json")))
    got = master.create_subtasks(g["random_job_config"])
    assert json.loads(json.dumps(got)) == json.loads(json.dumps(g["random_subtasks"]))
    # sampling without replacement from an all-list grid, seed 42 (task_handler.py:213)
    assert len({json.dumps(s["parameters"], sort_keys=True) for s in got}) == 7


def test_single_estimator_subtask():
    cfg = {"job_id": "j", "session_id": "s", "dataset_id": "d", "train_params": {"test_size": 0.2},
           "model_details": {"model_type": "Ridge", "hyperparameters": {"alpha": 2.0}}}
    (st,) = master.create_subtasks(cfg)
    assert st["subtask_id"] == "j-subtask-1" and st["parameters"] == {"alpha": 2.0}
    assert st["train_params"] == {"test_size": 0.2}


@pytest.mark.parametrize("n_gpus", [1, 2, 4, 8])
def test_blocks_cover_every_candidate_once(golden_dir, n_gpus):
    g = json.load(open(os.path.join(golden_dir, "notebook_subtasks.json")))
    subtasks = master.create_subtasks(g["job_config"])
    blocks = master.create_blocks(subtasks, n_gpus=n_gpus)
    seen = sorted(c["index"] for b in blocks for c in b["candidates"])
    assert seen == list(range(8))
    assert [b["gpu"] for b in blocks] == list(range(min(n_gpus, 8)))
    for b in blocks:
        assert b["folds"] == [0, 1, 2, 3, 4] and b["partition"] == "candidates"
        for c in b["candidates"]:
            assert c["parameters"] == subtasks[c["index"]]["parameters"]
    rows = master.create_blocks(subtasks, n_gpus=n_gpus, partition="rows")
    assert all(len(b["candidates"]) == 8 and b["n_shards"] == n_gpus for b in rows)


def test_aggregate_ties_and_nan():
    res = [{"mean_cv_score": float("nan"), "i": 0}, {"mean_cv_score": 0.9, "i": 1},
           {"mean_cv_score": 0.98, "i": 2}, {"mean_cv_score": 0.98, "i": 3}, {"mean_cv_score": None, "i": 4}]
    assert master.aggregate_results(res)["i"] == 2          # lowest index among exact ties
    assert master.aggregate_results(list(reversed(res)))["i"] == 3
    cv = master.build_cv_results([{"C": c} for c in range(5)],
                                 [[np.nan] * 2, [0.9, 0.9], [0.98, 0.98], [0.98, 0.98], [np.nan] * 2])
    assert cv["cv_results_"]["rank_test_score"] == [4, 3, 1, 1, 4]
    assert cv["best_index_"] == 2 and cv["best_params_"] == {"C": 2}


def test_build_cv_results_equals_gridsearchcv():
    from sklearn.datasets import load_iris
    X, y = load_iris(return_X_y=True)
    grid = {"alpha": [0.01, 0.1, 1.0, 10.0]}
    gs = GridSearchCV(RidgeClassifier(), grid, cv=5).fit(X, y)
    scores = np.stack([gs.cv_results_[f"split{k}_test_score"] for k in range(5)], axis=1)
    mine = master.build_cv_results([{"alpha": a} for a in grid["alpha"]], scores)
    assert mine["cv_results_"]["rank_test_score"] == gs.cv_results_["rank_test_score"].tolist()
    np.testing.assert_array_equal(mine["cv_results_"]["mean_test_score"], gs.cv_results_["mean_test_score"])
    np.testing.assert_allclose(mine["cv_results_"]["std_test_score"], gs.cv_results_["std_test_score"], rtol=1e-12)
    assert mine["best_index_"] == gs.best_index_ and mine["best_score_"] == gs.best_score_


def test_wire_format_cleans_numpy_and_nan():
    p = wire({"a": np.float64(1.5), "b": np.arange(3), "c": float("nan"), "d": {"e": np.int64(2)}, "f": LogisticRegression})
    assert p["a"] == 1.5 and p["b"] == [0, 1, 2] and p["c"] is None and p["d"]["e"] == 2 and isinstance(p["f"], str)


def test_client_config1_iris_known_answers():
    """BASELINE.json config 1 through MLTaskManager -> master -> scheduler -> worker (sklearn executor)."""
    from sklearn.datasets import load_iris
    X, y = load_iris(return_X_y=True)
    pkg.register_dataset("iris", X, y)
    m = MLTaskManager("local://1")
    try:
        gs = GridSearchCV(LogisticRegression(), {"C": [0.1, 1.0, 10.0, 100], "solver": ["liblinear", "lbfgs"]}, cv=5)
        first = m.train(gs, "iris", train_params={"random_state": 42})
        assert first["status"] == "Training started" and first["total_subtasks"] == 8
        r = m._wait_for_completion(m.job_id)
        assert r["job_status"] == "completed" and r["tasks_pending"] == 0 and r["total_subtasks"] == 8
        jr = r["job_result"]
        assert jr["best_params_"] == {"C": 10.0, "solver": "lbfgs"} and jr["best_index_"] == 5
        assert jr["best_score_"] == 0.9800000000000001
        assert jr["cv_results_"]["rank_test_score"] == [5, 4, 5, 3, 5, 1, 5, 1]
        assert r["best_result"]["parameters"]["C"] == 10.0           # demo_results.py:12-19 shape
        assert len(jr["results"]) == 8 and "training_time" in jr["results"][1]
        assert np.isnan(jr["results"][0]["mean_cv_score"])
        assert m.check_job_status("nope")["status"] == "error"       # errors are dicts, never raises
        assert m.train(gs, "missing-dataset")["status"] == "error"
        assert len(m._api_request(f"metrics/{m.session_id}/{m.job_id}", "get")) == 0  # failed submit: no job
    finally:
        m.close()


def test_refit_returns_best_estimator_artefact(tmp_path, monkeypatch):
    """train_params {"refit": True}: after the search the best candidate is fitted on all rows
    (GridSearchCV.best_estimator_) and pickled the way the reference pickles its models."""
    import pickle
    from sklearn.datasets import load_iris
    monkeypatch.chdir(tmp_path)
    X, y = load_iris(return_X_y=True)
    pkg.register_dataset("iris", X, y)
    m = MLTaskManager("local://1")
    try:
        grid = {"C": [0.1, 1.0, 10.0]}
        r = m.train(GridSearchCV(LogisticRegression(max_iter=500), grid, cv=5), "iris",
                    train_params={"refit": True}, wait_for_completion=True)
        assert r["job_status"] == "completed", r
        jr = r["job_result"]
        gs = GridSearchCV(LogisticRegression(max_iter=500), grid, cv=5).fit(X, y)
        assert jr["best_params_"] == gs.best_params_ and len(jr["results"]) == 3
        be = jr["best_estimator_"]
        assert be["parameters"]["C"] == gs.best_params_["C"] and jr["refit_time_"] >= 0
        np.testing.assert_allclose(be["coef"], gs.best_estimator_.coef_, rtol=1e-12)
        got = m.download_best_model(m.job_id, be["model_path"], be["model_id"])
        assert got["status"] == "success"
        with open(got["model_path"], "rb") as fh:
            est = pickle.load(fh)
        assert (est.predict(X) == gs.best_estimator_.predict(X)).all()
        # without the flag nothing is refitted (the reference never does)
        r2 = m.train(GridSearchCV(LogisticRegression(max_iter=500), grid, cv=5), "iris", wait_for_completion=True)
        assert "best_estimator_" not in r2["job_result"]
    finally:
        m.close()


def test_fitted_estimator_predicts_like_sklearn():
    """The artefact built from GPU coefficients is a working scikit-learn estimator."""
    from sklearn.linear_model import Ridge
    rng = np.random.default_rng(0)
    X = rng.normal(size=(200, 6))
    yb = (X[:, 0] + 0.3 * rng.normal(size=200) > 0).astype(int) * 2 + 3      # labels {3, 5}
    for name, ref in (("LogisticRegression", LogisticRegression(C=0.5).fit(X, yb)),
                      ("RidgeClassifier", RidgeClassifier(alpha=2.0).fit(X, yb))):
        est = wk.fitted_estimator(name, ref.get_params(), ref.coef_, ref.intercept_, ref.classes_, 7)
        assert (est.predict(X) == ref.predict(X)).all() and list(est.classes_) == [3, 5]
    ym = rng.integers(0, 4, size=200)
    ref = LogisticRegression().fit(X, ym)
    est = wk.fitted_estimator("LogisticRegression", ref.get_params(), ref.coef_, ref.intercept_, ref.classes_, 9)
    np.testing.assert_allclose(est.predict_proba(X), ref.predict_proba(X), rtol=1e-12)
    yr = X @ rng.normal(size=6) + 1.0
    ref = Ridge(alpha=0.3).fit(X, yr)
    est = wk.fitted_estimator("Ridge", ref.get_params(), ref.coef_, ref.intercept_)
    np.testing.assert_allclose(est.predict(X), ref.predict(X), rtol=1e-12)


def test_client_single_estimator_defaults_test_size():
    from sklearn.datasets import load_iris
    X, y = load_iris(return_X_y=True)
    pkg.register_dataset("iris", X, y)
    m = MLTaskManager("local://1")
    try:
        r = m.train(RidgeClassifier(alpha=1.0), "iris", train_params={"random_state": 42}, wait_for_completion=True)
        res = r["job_result"]["results"][0]
        assert r["total_subtasks"] == 1 and "best_result" not in r
        assert "accuracy" in res and len(res["cv_scores"]) == 5
        assert res["mean_cv_score"] == pytest.approx(0.82, abs=1e-12)   # alpha=1 -> 0.82
        rec = m._api_request(f"metrics/{m.session_id}/{m.job_id}", "get")
        assert rec and {"worker_id", "subtask_id", "status", "received_at", "started_at", "finished_at",
                        "cpu_percent_avg", "mem_percent_avg", "algo"} <= set(rec[0])
    finally:
        m.close()


def test_scheduler_contract_and_requeue():
    s = scheduler.Scheduler()
    got = []

    class M:
        def on_result(self, msg): got.append(msg)
        def on_metrics(self, rec): pass
    s.attach_master(M())
    assert s.submit({"subtask_id": "t0", "gpu": 0}) is None            # no worker yet: backlog
    w0 = s.subscribe(1000, "h", gpu=0)["worker_id"]
    w1 = s.subscribe(1000, "h", gpu=1)["worker_id"]
    assert s.get_queue(w0).qsize() == 1                                # backlog delivered on subscribe
    assert s.submit({"subtask_id": "t1", "gpu": 1}) == w1              # a block names its GPU
    assert s.heartbeat(w0)["status"] == "ok" and s.heartbeat("x")["status"] == "unknown_worker"
    # worker 1 goes silent for > 10 s: its task is handed to the survivor
    s.workers[w1]["last_seen"] -= 11
    assert s.reap_dead_workers() == [w1]
    assert s.get_queue(w0).qsize() == 2
    s.publish_result(w0, {"subtask_id": "t0", "job_id": "j", "status": "completed", "result": {}})
    assert got and s.workers_view()[w0]["done"] == 1
    assert s.unsubscribe(w0)["status"] == "unsubscribed" and s.n_workers() == 0


def test_scheduler_places_by_measured_speed():
    """scheduler_service.py:173-191: a task without a GPU goes to the worker with the least
    (queued fits + its own fits) / measured fits per second."""
    s = scheduler.Scheduler()

    class M:
        def on_result(self, msg): pass
        def on_metrics(self, rec): pass
    s.attach_master(M())
    fast = s.subscribe(1000, "h", gpu=0)["worker_id"]
    slow = s.subscribe(1000, "h", gpu=1)["worker_id"]
    s.publish_metrics({"worker_id": fast, "subtask_id": "x", "fits_per_sec": 20000.0})
    s.publish_metrics({"worker_id": slow, "subtask_id": "y", "fits_per_sec": 5.0})
    view = s.workers_view()
    assert view[fast]["speed"] == 20000.0 and view[slow]["speed"] == 5.0
    blk = lambda i: {"subtask_id": f"b{i}", "candidates": [{}] * 64, "folds": [0, 1, 2, 3, 4]}
    assert [s.submit(blk(i)) for i in range(4)] == [fast] * 4          # 4 x 320 fits queued is still cheaper there
    s.publish_metrics({"worker_id": fast, "subtask_id": "x", "fits_per_sec": 10000.0})
    assert s.workers_view()[fast]["speed"] == 15000.0                   # moving average
    # equal speeds: the shorter queue wins
    s.workers[slow]["speed"] = 15000.0
    assert s.submit(blk(9)) == slow


@pytest.mark.parametrize("kind", ["binary", "strings", "many", "sorted"])
def test_fold_labels_equal_sklearn_splitters(kind):
    """The int8 row labels the kernels take = the un-shuffled (Stratified)KFold of cross_val_score(cv=int)."""
    from sklearn.model_selection import KFold, StratifiedKFold
    from cs230_distributed_machine_learning_b200 import folds as pf
    rng = np.random.default_rng(3)
    n = 1237
    y = {"binary": rng.integers(0, 2, n), "strings": rng.choice(np.array(["b", "a", "c"]), n, p=[0.6, 0.3, 0.1]),
         "many": rng.integers(0, 37, n) * 3 - 5, "sorted": np.sort(rng.integers(0, 4, n))}[kind]
    for nf in (2, 5, 7):
        want = np.empty(n, dtype=np.int8)
        for i, (_, te) in enumerate(StratifiedKFold(nf).split(y, y)):
            want[te] = i
        _, idx = np.unique(y, return_inverse=True)
        assert (pf.stratified_labels(idx, nf) == want).all()
        assert (pf.stratified_labels(y, nf) == want).all()
        for i, (_, te) in enumerate(KFold(nf).split(y)):
            want[te] = i
        assert (pf.kfold_labels(n, nf) == want).all()


def test_gpu_eligibility_rules():
    ok = worker.gpu_eligible
    assert ok("LogisticRegression", {"C": 1.0, "solver": "lbfgs"})[0]
    assert not ok("LogisticRegression", {"solver": "liblinear"})[0]
    assert ok("LogisticRegression", {"class_weight": "balanced"})[0] and ok("LogisticRegression", {"class_weight": {"0": 2.0}})[0]
    assert not ok("LogisticRegression", {"class_weight": [1.0, 2.0]})[0]
    assert not ok("RidgeClassifier", {"class_weight": "balanced"})[0]
    assert not ok("LogisticRegression", {"l1_ratio": 0.5})[0]
    assert ok("Ridge", {"alpha": 1.0})[0] and not ok("Ridge", {"solver": "sag"})[0]
    assert ok("RidgeClassifier", {})[0] and not ok("RandomForestClassifier", {})[0]
    with pytest.raises(ValueError):
        worker.get_model_instance("XGBClassifier", {})
    # an older client's extra key is dropped, not fatal (library_testing.ipynb:640)
    assert worker.get_model_instance("LogisticRegression", {"C": 2.0, "multi_class": "auto"}).C == 2.0


def test_b200_executor_never_falls_back_to_cpu():
    """float32 data + covered model => the CUDA path, which must raise when there is no GPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present: the loud-failure branch is not reachable")
    X = np.random.default_rng(0).standard_normal((200, 4)).astype(np.float32)
    y = (X[:, 0] > 0).astype(np.int32)
    pkg.register_dataset("f32", X, y)
    task = {"subtask_id": "s-1", "job_id": "j", "dataset_id": "f32", "model_type": "LogisticRegression",
            "parameters": {"C": 1.0}, "train_params": {"cv": 5}}
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        worker.train_model(task)
    task["train_params"]["executor"] = "sklearn"                      # explicit request for the reference path
    assert len(worker.train_model(task)["cv_scores"]) == 5


def test_holdout_labels_match_train_test_split():
    from sklearn.model_selection import train_test_split
    for n, ts, rs in [(150, 0.25, 42), (1001, 0.2, 7), (64, 0.33, 0)]:
        _, te = train_test_split(np.arange(n), test_size=ts, random_state=rs)
        assert set(np.nonzero(folds.holdout_labels(n, ts, rs))[0]) == set(te)


def test_cabi_library_exports_every_declared_symbol():
    """The .so builds for sm_100a, loads without a GPU, and exports what include/dml_b200.h declares."""
    path = build.build()
    lib = ctypes.CDLL(path)
    header = open(os.path.join(ROOT, "include", "dml_b200.h")).read()
    declared = set(re.findall(r"\b(dml_[a-z0-9_]+)\s*\(", header))
    assert declared >= {"dml_version", "dml_ctx_create", "dml_ctx_destroy", "dml_last_error", "dml_logreg_sweep",
                        "dml_ridge_sweep", "dml_score", "dml_ctx_init_comm", "dml_nccl_unique_id", "dml_last_timing",
                        "dml_last_counters", "dml_logreg_eval"}
    for name in declared:
        assert hasattr(lib, name), name
    assert set(_cabi.SYMBOLS) == declared
    assert _cabi.load().dml_version() >= 100
    import torch
    if not torch.cuda.is_available():
        h = ctypes.c_void_p()
        assert _cabi.load().dml_ctx_create(0, None, ctypes.byref(h)) != 0   # loud, not a crash
        assert b"CUDA" in _cabi.load().dml_last_error(None) or b"device" in _cabi.load().dml_last_error(None)


# ---- dataset prefetch / residency (host logic only; the copies themselves are GPU tests) ----
class _FakeDs:
    def __init__(self, nbytes, ready=None):
        self.nbytes, self.ready = nbytes, ready


def _bare_executor(budget):
    import threading
    ex = object.__new__(wk.GpuExecutor)
    ex.resident, ex.in_use, ex.max_resident_bytes = {}, None, budget
    ex.res_lock = threading.Lock()
    return ex


def test_resident_eviction_order():
    ex = _bare_executor(budget=250)
    ex.resident = {("a", True): _FakeDs(100), ("b", True): _FakeDs(100), ("c", True): _FakeDs(100, ready=object())}
    ex.in_use = ("a", True)
    ex._evict()                       # a is being read, c is a prefetched copy nobody used: b goes
    assert list(ex.resident) == [("a", True), ("c", True)]
    ex.in_use = None
    ex._evict(incoming=100, keep=("a", True))   # nothing may go: the budget is soft
    assert list(ex.resident) == [("a", True), ("c", True)]
    ex.resident[("c", True)].ready = None
    ex._evict(incoming=100)           # least recently used first
    assert list(ex.resident) == [("c", True)]


def test_prefetch_task_routes_to_the_executor(monkeypatch):
    calls = []

    class Fake:
        hint = None

        def prefetch(self, name, classifier):
            calls.append((name, classifier))
            return True

    fake = Fake()
    monkeypatch.setitem(wk._EXECUTORS, 0, fake)
    X = np.zeros((8, 2), dtype=np.float32)
    pkg.register_dataset("pf-now", X, np.arange(8) % 2)
    pkg.register_dataset("pf-next", X, np.arange(8) % 2)
    pkg.register_dataset("pf-f64", X.astype(np.float64), np.arange(8) % 2)
    block = {"model_type": "LogisticRegression", "dataset_id": "pf-next", "train_params": {"cv": 3},
             "candidates": [{"index": 0, "subtask_id": "s0", "parameters": {"C": 1.0}}]}
    # nothing on the GPU right now (no current task, or a scikit-learn one): the copy starts at once
    assert wk.prefetch_task(block, gpu=0)
    assert wk.prefetch_task(block, gpu=0, current=dict(block, model_type="RandomForestClassifier"))
    assert calls == [("pf-next", True), ("pf-next", True)] and fake.hint is None
    # the running task is a b200 block: leave a hint, GpuExecutor.run copies once it holds its own dataset
    cur = dict(block, model_type="Ridge", dataset_id="pf-now")
    assert wk.prefetch_task(block, gpu=0, current=cur)
    assert fake.hint == ("pf-next", True) and len(calls) == 2
    # tasks the b200 executor would not take are not prefetched
    assert not wk.prefetch_task(dict(block, model_type="RandomForestClassifier"), gpu=0)
    assert not wk.prefetch_task(dict(block, train_params={"executor": "sklearn"}), gpu=0)
    assert not wk.prefetch_task(dict(block, dataset_id="pf-f64"), gpu=0)      # auto keeps float64 on scikit-learn
    assert not wk.prefetch_task(dict(block, dataset_id="missing"), gpu=0)
    assert not wk.prefetch_task(block, gpu=1)          # no executor on that GPU yet
    assert len(calls) == 2


def test_blocks_dealt_by_predicted_rounds():
    """a randomized search enumerates candidates in draw order; dealing by the rounds a
    finished job of the same (model, dataset) reported balances the GPUs' column-rounds where round-robin
    leaves them to chance."""
    rng = np.random.default_rng(0)
    grid = np.logspace(-6, 2, 9)
    seen_iters = dict(zip(grid.tolist(), [7, 5, 3, 6, 7, 7, 7, 7, 7]))      # scikit-learn, 50 000 x 128 synthetic rows
    Cs = rng.permutation(np.logspace(-6, 2, 64))
    subtasks = [{"subtask_id": f"j-subtask-{i + 1}", "job_id": "j", "session_id": "s", "dataset_id": "d",
                 "model_type": "LogisticRegression", "parameters": {"C": float(c)}, "train_params": {"cv": 5}}
                for i, c in enumerate(Cs)]
    model = master.IterationModel()
    flat = [model.predict("LogisticRegression", "d", st["parameters"]) for st in subtasks]
    assert len(set(flat)) == 1 and flat[0] >= 1.0            # nothing known yet: equal costs ...
    assert sorted(len(b["candidates"]) for b in master.create_blocks(subtasks, n_gpus=4, cost=flat)) == [16] * 4   # ... equal blocks
    assert model.predict("LogisticRegression", "d", {"C": 1.0, "tol": 1e-6}) > flat[0]   # tighter tol, more rounds
    for c, it in seen_iters.items():
        model.observe("LogisticRegression", "d", {"C": c}, [it] * 5)
    cost = [model.predict("LogisticRegression", "d", st["parameters"]) for st in subtasks]
    assert 3.0 <= min(cost) < 3.5 and max(cost) == 7.0      # interpolated in log10 C between the observed points
    for n_gpus in (2, 4, 8):
        dealt = master.create_blocks(subtasks, n_gpus=n_gpus, cost=cost)
        robin = master.create_blocks(subtasks, n_gpus=n_gpus)
        assert sorted(c["index"] for b in dealt for c in b["candidates"]) == list(range(64))
        for b in dealt:
            idx = [c["index"] for c in b["candidates"]]
            assert idx == sorted(idx) and all(c["parameters"] == subtasks[c["index"]]["parameters"] for c in b["candidates"])
        load = lambda blocks: [sum(cost[c["index"]] for c in b["candidates"]) for b in blocks]
        assert max(load(dealt)) <= max(load(robin)) + 1e-9
        assert max(load(dealt)) - min(load(dealt)) <= max(cost)            # LPT bound
    # Ridge-family candidates are direct solves: equal cost, equal block sizes
    assert model.predict("Ridge", "d", {"alpha": 3.0}) == 1.0
    sizes = sorted(len(o) for o in master.deal_by_cost([1.0] * 10, 4))
    assert sizes == [2, 2, 3, 3]


def test_iteration_model_learns_from_results():
    model = master.IterationModel()
    before = model.predict("LogisticRegression", "d", {"C": 1.0})
    for c, it in ((1e-4, [2, 2, 3]), (1e-2, [4, 4, 4]), (1.0, [9, 9, 9]), (1e2, [12, 12, 12])):
        model.observe("LogisticRegression", "d", {"C": c}, it)
    assert model.predict("LogisticRegression", "d", {"C": 1.0}) == 9.0 != before
    assert model.predict("LogisticRegression", "d", {"C": 0.1}) == pytest.approx(6.5)        # interpolated in log10 C
    assert model.predict("LogisticRegression", "d", {"C": 1e3}) == 12.0                      # edge value held
    assert model.predict("LogisticRegression", "d", {"C": 1e3, "max_iter": 10}) == 10.0
    assert model.predict("LogisticRegression", "other", {"C": 1.0}) == before                # per dataset
    model.observe("LogisticRegression", "d", {"C": 1.0}, [11, 11, 11])
    assert model.predict("LogisticRegression", "d", {"C": 1.0}) == 10.0                      # running mean


def test_master_learns_rounds_from_a_finished_job():
    """After one job on a dataset the splitter deals the next job's candidates by the rounds the first one
    reported (results carry n_iter on both executors)."""
    from sklearn.datasets import load_iris
    X, y = load_iris(return_X_y=True)
    pkg.register_dataset("iris-rounds", X, y)
    m = MLTaskManager("local://1")
    try:
        model = m.master.iterations
        flat = model.predict("LogisticRegression", "iris-rounds", {"C": 0.01})
        assert model.predict("LogisticRegression", "iris-rounds", {"C": 100.0}) == flat
        gs = GridSearchCV(LogisticRegression(max_iter=500), {"C": [0.01, 1.0, 100.0]}, cv=5)
        m.train(gs, "iris-rounds")
        r = m._wait_for_completion(m.job_id)
        assert r["job_status"] == "completed"
        want = [float(np.mean(res["n_iter"])) for res in r["job_result"]["results"]]
        got = [model.predict("LogisticRegression", "iris-rounds", {"C": c, "max_iter": 500}) for c in (0.01, 1.0, 100.0)]
        assert got == want and want[0] < want[2]                     # weak penalty on iris: many more iterations
        assert model.predict("LogisticRegression", "iris", {"C": 100.0}) == flat          # other datasets: still the prior
    finally:
        m.close()


# ---------------------------------------------------------------------------------------------------
# round 2: row-shard blocks, at-least-once delivery, live failure detection, scoring / train scores
# ---------------------------------------------------------------------------------------------------
def _subtasks(n, cv=5):
    return [{"subtask_id": f"j-subtask-{i + 1}", "job_id": "j", "session_id": "s", "dataset_id": "d",
             "model_type": "LogisticRegression", "parameters": {"C": 10.0 ** i}, "train_params": {"cv": cv}}
            for i in range(n)]


def test_row_partition_blocks_carry_shard_and_all_candidates():
    blocks = master.create_blocks(_subtasks(5), n_gpus=4, partition="rows",
                                  extra_train_params={"scoring": "accuracy", "return_train_score": True})
    assert [b["gpu"] for b in blocks] == [0, 1, 2, 3] and [b["shard"] for b in blocks] == [0, 1, 2, 3]
    assert all(b["n_shards"] == 4 and b["partition"] == "rows" and len(b["candidates"]) == 5 for b in blocks)
    assert all(b["train_params"]["scoring"] == "accuracy" and b["train_params"]["return_train_score"] for b in blocks)
    assert len({b["train_params"]["shard_group"] for b in blocks}) == 1
    cand = master.create_blocks(_subtasks(5), n_gpus=4)
    assert all(b["n_shards"] == 1 and b["shard"] == 0 and "shard_group" not in b["train_params"] for b in cand)
    assert sorted(c["index"] for b in cand for c in b["candidates"]) == list(range(5))


def test_master_accepts_one_result_per_shard_set_and_ignores_reruns():
    """Row shards: every block names all candidates, shard 0 reports them, the others report nothing.  A block
    that runs twice (requeue of a worker that was slow, not dead) must not finish the job twice."""
    s = scheduler.Scheduler()
    m = master.Master(s, n_gpus=2, partition="rows")
    sid = m.create_session()["session_id"]
    pkg.register_dataset("rows-host", np.zeros((8, 2), np.float32), np.arange(8) % 2)
    est = {"model_type": "LogisticRegression", "search_type": "GridSearchCV",
           "hyperparameters": {"base_estimator_params": {}, "search_params": {"param_grid": {"C": [0.1, 1.0, 10.0]}},
                               "cv_params": {"cv": 2, "refit": True}}}
    r = m.train(sid, {"job_id": "jr", "session_id": sid, "dataset_id": "rows-host", "model_details": est,
                      "train_params": {"refit": True}})
    assert r["status"] == "Training started" and len(s.backlog) == 2      # no worker yet: both shard blocks wait
    blocks = list(s.backlog)
    assert [b["shard"] for b in blocks] == [0, 1]
    entries = [{"subtask_id": c["subtask_id"], "status": "completed",
                "result": {"cv_scores": [0.5 + 0.1 * i, 0.5], "mean_cv_score": 0.5 + 0.05 * i, "parameters": c["parameters"],
                           "n_iter": [3, 3]}} for i, c in enumerate(blocks[0]["candidates"])]
    finished = []
    orig = m._finish
    m._finish = lambda job: (finished.append(1), orig(job))[1]
    m.on_result({"subtask_id": blocks[1]["subtask_id"], "job_id": "jr", "session_id": sid, "status": "completed", "result": []})
    assert m.check_status(sid, "jr")["job_status"] in ("pending", "0.0")
    m.on_result({"subtask_id": blocks[0]["subtask_id"], "job_id": "jr", "session_id": sid, "status": "completed", "result": entries})
    m.on_result({"subtask_id": blocks[0]["subtask_id"], "job_id": "jr", "session_id": sid, "status": "completed", "result": entries})
    m.on_result({"subtask_id": blocks[1]["subtask_id"], "job_id": "jr", "session_id": sid, "status": "completed", "result": []})
    assert finished == [1]                                                 # aggregated exactly once
    refits = [t for t in s.backlog if t["subtask_id"] == "jr-refit"]
    assert len(refits) == 1 and refits[0]["candidates"][0]["parameters"]["C"] == 10.0
    done = {"subtask_id": "jr-refit", "job_id": "jr", "session_id": sid, "status": "completed",
            "result": [{"subtask_id": "jr-refit", "status": "completed", "result": {"coef": [[1.0, 2.0]], "intercept": [0.0]}}]}
    m.on_result(done)
    m.on_result(done)                                                      # a second refit report: no KeyError, no change
    st = m.check_status(sid, "jr")
    assert st["job_status"] == "completed" and st["job_result"]["best_params_"] == {"C": 10.0}


def test_scheduler_exchange_is_a_rendezvous():
    import threading
    s = scheduler.Scheduler()
    got = []
    t = threading.Thread(target=lambda: got.append(s.exchange("k")))
    t.start()
    s.exchange("k", b"\x01" * 128)
    t.join(5)
    assert got == [b"\x01" * 128]
    with pytest.raises(TimeoutError):
        s.exchange("never", timeout=0.05)


def test_monitor_thread_requeues_the_blocks_of_a_worker_that_stops_heartbeating(monkeypatch):
    """scheduler_service.py:205-247 live: a worker dies mid-job (no unsubscribe, no heartbeat); the monitor loop of
    the LocalBackend notices after HEARTBEAT_TIMEOUT_S and hands its block to the surviving worker."""
    import time
    from cs230_distributed_machine_learning_b200.client import LocalBackend, MLTaskManager
    from sklearn.linear_model import LogisticRegression
    from sklearn.model_selection import GridSearchCV
    monkeypatch.setattr(scheduler, "HEARTBEAT_TIMEOUT_S", 0.3)
    monkeypatch.setattr(worker, "HEARTBEAT_INTERVAL_S", 0.05)
    rng = np.random.default_rng(5)
    X = rng.standard_normal((300, 4)); y = (X[:, 0] + 0.3 * rng.standard_normal(300) > 0).astype(int)
    pkg.register_dataset("mon-f64", X, y)                                  # float64: the scikit-learn executor (CPU test)
    backend = LocalBackend(n_gpus=2, monitor_interval=0.1)
    victim = backend.workers[1]
    victim.handle = lambda task: time.sleep(3600)                          # takes its block and never answers
    victim._heartbeat = lambda: None
    orig_hb = backend.scheduler.heartbeat
    backend.scheduler.heartbeat = lambda wid: {"status": "dropped"} if wid == victim.worker_id else orig_hb(wid)
    mgr = MLTaskManager(backend)
    try:
        r = mgr.train(GridSearchCV(LogisticRegression(), {"C": [0.1, 1.0, 10.0, 100.0]}, cv=3), "mon-f64",
                      wait_for_completion=False)
        assert r["status"] == "Training started"
        st = mgr._wait_for_completion(mgr.job_id, timeout=30)
        assert st.get("job_status") == "completed", st
        assert len(st["job_result"]["results"]) == 4 and victim.worker_id not in backend.scheduler.workers_view()
        want = GridSearchCV(LogisticRegression(), {"C": [0.1, 1.0, 10.0, 100.0]}, cv=3).fit(X, y)
        assert st["job_result"]["best_params_"] == want.best_params_
    finally:
        backend.scheduler.stop_monitor()
        for w in backend.workers:
            if w is not victim:
                w.shutdown()


def test_scoring_string_and_train_scores_through_the_api_equal_gridsearchcv():
    """cv_params.scoring / return_train_score (core.py:123-130) reach the executor; here the scikit-learn one."""
    from sklearn.linear_model import Ridge
    from sklearn.model_selection import GridSearchCV
    rng = np.random.default_rng(6)
    X = rng.standard_normal((400, 5)); y = X @ rng.standard_normal(5) + 0.2 * rng.standard_normal(400)
    pkg.register_dataset("score-f64", X, y)
    m = pkg.MLTaskManager("local://1")
    try:
        est = GridSearchCV(Ridge(), {"alpha": [0.1, 10.0, 1000.0]}, cv=4, scoring="neg_mean_squared_error",
                           return_train_score=True)
        r = m.train(est, "score-f64", wait_for_completion=True)
        assert r["job_status"] == "completed", r
        res, want = r["job_result"]["cv_results_"], est.fit(X, y).cv_results_
        for k in ("mean_test_score", "mean_train_score", "split2_train_score", "std_train_score"):
            np.testing.assert_allclose(res[k], want[k], rtol=1e-9)
        assert r["job_result"]["best_params_"] == est.best_params_
    finally:
        m.close()


def test_dataset_version_changes_on_reregistration():
    from cs230_distributed_machine_learning_b200 import datasets
    v0 = datasets.dataset_version("ver-x")
    pkg.register_dataset("ver-x", np.zeros((4, 2), np.float32), np.zeros(4))
    pkg.register_dataset("ver-x", np.ones((4, 2), np.float32), np.zeros(4))
    assert datasets.dataset_version("ver-x") == v0 + 2


def test_stratified_labels_of_a_row_shard_equal_the_unsharded_labels():
    rng = np.random.default_rng(0)
    for K, n, nf in [(2, 10007, 10), (3, 5000, 5), (4, 999, 7)]:
        y = rng.integers(0, K, n)
        if K == 3:
            y[:50] = 2                                        # first-appearance order differs from value order
        full = folds.stratified_labels(y, nf)
        total = np.bincount(y, minlength=K)
        first = np.array([np.nonzero(y == k)[0][0] for k in range(K)])
        cuts = [0, n // 3, n // 3 * 2 + 5, n]
        got = np.concatenate([folds.stratified_labels_slice(y[a:b], np.bincount(y[:a], minlength=K), total, first, nf)
                              for a, b in zip(cuts, cuts[1:])])
        assert (got == full).all()


def test_chunk_seeded_generator_gives_any_row_range_on_its_own():
    from cs230_distributed_machine_learning_b200.datasets import CHUNK_ROWS, make_dataset_rows
    n = 2 * CHUNK_ROWS + 1000
    X, y = make_dataset_rows("binary", n, 8, 5)
    X2, y2 = make_dataset_rows("binary", n, 8, 5, lo=CHUNK_ROWS, hi=CHUNK_ROWS + 77)
    assert (X2 == X[CHUNK_ROWS:CHUNK_ROWS + 77]).all() and (y2 == y[CHUNK_ROWS:CHUNK_ROWS + 77]).all()
    assert abs(float(y.mean()) - 0.5) < 0.01 and abs(float(X.std()) - 1.0) < 0.01


def test_class_weight_table_equals_sklearn_compute_class_weight():
    from sklearn.utils.class_weight import compute_class_weight
    rng = np.random.default_rng(0)
    y = rng.choice([0, 1, 2], 500, p=[.6, .3, .1])
    fold = folds.stratified_labels(y, 5)
    hold = folds.holdout_labels(500, .2, 1)
    t = worker.class_weight_table([{"class_weight": "balanced"}, {"class_weight": {"0": 2.0, "2": 0.5}}, {}],
                                  np.array([0, 1, 2]), y, fold, hold, 5)
    assert t.shape == (3, 6, 3)
    for k in range(5):
        np.testing.assert_allclose(t[0, k], compute_class_weight("balanced", classes=np.array([0, 1, 2]), y=y[fold != k]))
    np.testing.assert_allclose(t[0, 5], compute_class_weight("balanced", classes=np.array([0, 1, 2]), y=y[hold != 1]))
    np.testing.assert_allclose(t[1, 3], [2, 1, .5])
    assert (t[2] == 1).all()
    assert worker.class_weight_table([{}, {}], np.array([0, 1]), y % 2, fold, None, 5) is None
    with pytest.raises(ValueError):
        worker.class_weight_table([{"class_weight": {"7": 1.0}}], np.array([0, 1, 2]), y, fold, None, 5)


def test_unpin_bookkeeping_without_a_gpu():
    """worker._unpin is a finalizer: it must drop the entry it was made for, leave a newer entry of a recycled id()
    alone, and never raise (here: no CUDA runtime to call)."""
    from cs230_distributed_machine_learning_b200 import worker as W
    W._PINNED[123456789] = 4096
    W._unpin(123456789, 8192)                 # the id was recycled by another array at another address: not ours
    assert W._PINNED.get(123456789) == 4096
    W._unpin(123456789, 4096)
    assert 123456789 not in W._PINNED
    W._unpin(123456789, 4096)                 # a second call is a no-op
    W._clear_cuda_error()                     # nothing loaded here: must not raise
