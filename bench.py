"""Headline benchmark: candidate-fold fits/sec of the (candidate x CV-fold) sweep.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config 2|3|3c|4|5]

Default workload = BASELINE.json configs[1] ("config 2"): GridSearchCV(LogisticRegression lbfgs, 64 C values =
logspace(-6, 2, 64), cv=5) on synthetic binary 1M x 256 fp32 (seed 2).  A "step" is one whole
sweep: all 64*5 = 320 fits trained and scored.  X (1.024 GB) is larger than the 126 MB L2, so no flush is
needed between steps.  --config 3 / 3c / 4 / 5 measure the other BASELINE.json configs the same way (Ridge /
RidgeClassifier 128 alphas on 4M x 512; 10-class multinomial RandomizedSearchCV, 512 candidates, 2M x 512;
256 C x cv 10 on 16M x 512) with their own roofline and cpu_baseline; their lines are kept under profiles/.

N > 1 (torchrun, one rank per GPU).  The headline line is weak scaling by candidate blocks: the grid grows to
64*N values of C over the same range, rank r owns candidates r, r+N, ...; X is replicated, there is no
data-path collective, rank 0 gathers the scores.  The same line carries
  "strong": config 2 itself (64 C in total) over the N GPUs, by candidate blocks and by row shards,
  "rows"  : BASELINE.json config 5 (16M x 512, 256 C x cv 10) row-sharded over the N GPUs with an NCCL
            all-reduce of [G | sum r | sum loss] per round, plus "dist_parity" (a small row-sharded sweep
            equals the one-GPU sweep),
  "api_all_gpus": one job through MLTaskManager -> master.create_blocks -> one worker per GPU in ONE
            process, by candidate blocks and by row shards.

One JSON line is printed by rank 0 (see README / the driver contract).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# stdout carries exactly one JSON line.  Native libraries print there too (NCCL's version banner), so the
# process keeps a private copy of the original stdout for that line and points fd 1 at stderr.
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
_RESULT_FD = None


def _claim_stdout():
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC, UNIT = "candidate_fold_fits_per_sec", "fits/s"

# BASELINE.json configs (shapes, seeds, candidate lists)
CONFIGS = {
    "2": dict(kind="binary", n=1_000_000, d=256, model="LogisticRegression", n_cand=64, cv=5, seed=2),
    "3": dict(kind="regression", n=4_000_000, d=512, model="Ridge", n_cand=128, cv=5, seed=3),
    "3c": dict(kind="binary", n=4_000_000, d=512, model="RidgeClassifier", n_cand=128, cv=5, seed=3),
    "4": dict(kind="multiclass10", n=2_000_000, d=512, model="LogisticRegression", n_cand=512, cv=5, seed=4),
    "5": dict(kind="binary", n=16_000_000, d=512, model="LogisticRegression", n_cand=256, cv=10, seed=5),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="2", choices=sorted(CONFIGS))
    ap.add_argument("--rows", type=int, default=None)       # smaller sizes are for debugging only
    ap.add_argument("--features", type=int, default=None)
    ap.add_argument("--candidates", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="N > 1: skip the strong / rows / api_all_gpus records")
    ap.add_argument("--rows-config-rows", type=int, default=None, help="N > 1: rows of the config-5 record (default 16M)")
    ap.add_argument("--tc", type=int, default=-1, help="use_tensor_cores flag passed to the library")
    ap.add_argument("--api-probe", type=int, default=0, help=argparse.SUPPRESS)
    a = ap.parse_args()
    cfg = dict(CONFIGS[a.config])
    cfg["n"] = a.rows or cfg["n"]; cfg["d"] = a.features or cfg["d"]; cfg["n_cand"] = a.candidates or cfg["n_cand"]
    a.cfg = cfg
    a.full_size = (cfg == CONFIGS[a.config])
    return a


def candidates_of(cfg: dict, n_cand: int):
    """The candidate list of a config, in enumeration order."""
    if cfg["model"] in ("Ridge", "RidgeClassifier"):
        return [{"alpha": float(v)} for v in np.logspace(-3, 5, n_cand)]
    if cfg["kind"].startswith("multiclass"):
        # param_distributions are lists: a 64 x 8 x 2 grid sampled without replacement, sampler seed 42
        # (task_handler.py:213; sklearn _search.py:319-335); candidate index = draw order
        from sklearn.model_selection import ParameterSampler
        dists = {"C": [float(v) for v in np.logspace(-5, 2, 64)], "tol": [float(v) for v in np.logspace(-5, -3, 8)],
                 "fit_intercept": [True, False]}
        return list(ParameterSampler(dists, n_cand, random_state=42))
    return [{"C": float(v)} for v in np.logspace(-6, 2, n_cand)]


def workload_name(cfg: dict, n_cand_total: int):
    if cfg["model"] in ("Ridge", "RidgeClassifier"):
        return (f"GridSearchCV({cfg['model']}, {n_cand_total} alphas logspace(-3,5), cv={cfg['cv']}) on synthetic "
                f"{cfg['kind']} {cfg['n']}x{cfg['d']} fp32 (seed {cfg['seed']})")
    if cfg["kind"].startswith("multiclass"):
        return (f"RandomizedSearchCV(LogisticRegression multinomial 10-class, {n_cand_total} candidates over "
                f"C/tol/fit_intercept, sampler seed 42, cv={cfg['cv']}) on synthetic {cfg['n']}x{cfg['d']} fp32 (seed {cfg['seed']})")
    return (f"GridSearchCV(LogisticRegression lbfgs, {n_cand_total} C values logspace(-6,2), cv={cfg['cv']}) on synthetic "
            f"binary {cfg['n']}x{cfg['d']} fp32 (seed {cfg['seed']})")
# ------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock and throttle reasons DURING the timed region: NVML every 20 ms (the nvidia-smi query of
    B200_PROFILING.md, one process per sample, is the fallback when the NVML binding is missing)."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown"}

    def __init__(self, gpu_index: int):
        super().__init__(daemon=True)
        self.gpu_index, self.stop_ev = gpu_index, threading.Event()
        self.sm, self.mx, self.reasons = [], [], set()
        self.watts, self.limit_w = [], None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        except Exception:
            self.nv = None

    def sample(self):
        if self.nv is not None:
            nv = self.nv
            self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
            self.mx.append(float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)))
            try:
                bits = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            except Exception:
                bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for bit, name in self.REASONS.items():
                if bits & bit:
                    self.reasons.add(name)
            try:                                     # board power next to the clocks: the sweep runs at the cap
                self.watts.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                if self.limit_w is None:
                    self.limit_w = nv.nvmlDeviceGetEnforcedPowerLimit(self.h) / 1000.0
            except Exception:
                pass
            return
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                              str(self.gpu_index)], capture_output=True, text=True, timeout=5).stdout
        r = [c.strip() for c in out.strip().split(",")]
        self.sm.append(float(r[0])); self.mx.append(float(r[1]))
        for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:]):
            if v.lower().startswith("active"):
                self.reasons.add(name)

    def run(self):
        while not self.stop_ev.is_set():
            try:
                self.sample()
            except Exception:
                pass
            self.stop_ev.wait(0.02 if self.nv is not None else 0.2)

    def summary(self):
        self.stop_ev.set()
        self.join(timeout=6)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None,
                "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm),
                "power_w": float(np.median(self.watts)) if self.watts else None, "power_limit_w": self.limit_w,
                "source": "nvml" if self.nv is not None else "nvidia-smi"}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def measure_tf32_peak(torch):
    """cuBLAS TF32 GEMM 8192^3 the way MEASURED_PEAKS.json measures bf16: best of 10 (burst, for a kernel
    timed alone) and back to back for ~2 s (sustained, under the power cap: the figure that applies to
    kernels timed inside a long step).  Returns (burst, sustained) in TFLOP/s."""
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(8192, 8192, device="cuda"); b = torch.randn(8192, 8192, device="cuda")
        for _ in range(3):
            a @ b
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); a @ b; e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        flops = 2 * 8192 ** 3
        reps = max(50, int(2000.0 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            a @ b
        e1.record(); torch.cuda.synchronize()
        sustained = flops * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12
        return flops / (best * 1e-3) / 1e12, sustained
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old



def reset_sampler(s):
    if s is not None:
        s.sm.clear(); s.mx.clear(); s.reasons.clear(); s.watts.clear()


# ------------------------------------------------------------------------------------------
# parity of a config-2 sweep against the cached full-size scikit-learn (float64) answers
def search_best_index(scores):
    """GridSearchCV's rule (sklearn _search.py:1228-1270, 862-874): rankdata(-mean, 'min').argmin()."""
    from scipy.stats import rankdata
    mean = np.average(np.asarray(scores, dtype=np.float64), axis=1)
    return int(rankdata(-mean, method="min").argmin())


def parity_report(out: dict, X, y, fold, cv: int, golden_path: str, eps: float = 2e-5):
    """north_star's parity gates for one sweep result against tests/golden/cfg2_full.npz (scikit-learn 1.9.0 on
    the same values as float64, produced by tests/golden/make_cfg_full.py): nothing that differs is hidden."""
    if not os.path.exists(golden_path):
        return {"unavailable": f"{os.path.relpath(golden_path, ROOT)} missing"}
    z = np.load(golden_path)
    n_cand = out["n_iter"].shape[0]
    if z["n_iter"].shape != out["n_iter"].shape:
        return {"unavailable": "golden file is for another candidate list"}
    same_iter = out["n_iter"] == z["n_iter"]
    coef_err = np.zeros((n_cand, cv)); diffs = np.zeros((n_cand, cv), np.int64); ties = np.zeros((n_cand, cv), np.int64)
    for k in range(cv):
        rows = fold == k
        Z = X[rows].astype(np.float64) @ z["coef"][:, k, :].T + z["intercept"][:, k][None, :]
        ties[:, k] = (np.abs(Z) < eps * np.maximum(1.0, np.abs(Z).max(axis=0))[None, :]).sum(axis=0)
        coef_err[:, k] = np.abs(out["coef"][:, k, 0] - z["coef"][:, k]).max(axis=1) / np.abs(z["coef"][:, k]).max(axis=1)
        diffs[:, k] = np.abs(out["correct"][:, k] - z["correct"][:, k])
    scores = out["correct"] / out["n_test"]
    best = search_best_index(scores)
    return {"against": os.path.relpath(golden_path, ROOT) + " (scikit-learn float64, full size)",
            "fits_checked": int(same_iter.size), "n_iter_equal": int(same_iter.sum()),
            "nfev_equal": int((out["nfev"] == z["nfev"]).sum()),
            "max_coef_relerr": float(coef_err[same_iter].max()) if same_iter.any() else None,
            "count_diffs": int((diffs > 0).sum()), "max_count_diff": int(diffs.max()),
            "count_diffs_beyond_ties": int((diffs > ties).sum()),
            "rows_excused_as_ties": int(np.minimum(diffs, ties).sum()), "rows_near_ties": int(ties.sum()),
            "n_test_equal": bool((out["n_test"] == z["n_test"]).all()),
            "max_mean_score_diff": float(np.abs(np.average(scores, axis=1) - z["mean_test_score"]).max()),
            "best_index": best, "best_index_oracle": int(z["best_index"]), "best_index_equal": bool(best == int(z["best_index"]))}


GOLDEN_CFG2 = os.path.join(ROOT, "tests", "golden", "cfg2_full.npz")
GOLDEN_CFG3 = os.path.join(ROOT, "tests", "golden", "cfg3_full.npz")


def parity_report_ridge(full: dict, golden_path: str):
    """Config 3 against tests/golden/cfg3_full.npz: the float64 Cholesky solution at full size for 8 of the 128
    alphas x 5 folds (tests/golden/make_cfg3_full.py; pinned there by one scikit-learn Ridge fit)."""
    if not os.path.exists(golden_path):
        return {"unavailable": f"{os.path.relpath(golden_path, ROOT)} missing"}
    z = np.load(golden_path)
    idx = json.loads(str(z["meta"]))["alpha_index"]
    if max(idx) >= full["coef"].shape[0]:
        return {"unavailable": "golden file is for another candidate list"}
    coef = full["coef"][idx][:, :, 0, :]
    err = np.abs(coef - z["coef"]).max(axis=2) / np.abs(z["coef"]).max(axis=2)
    r2 = 1.0 - full["sse"][idx] / (full["syy"][idx] - full["sy"][idx] ** 2 / full["n_test"][idx])
    return {"against": os.path.relpath(golden_path, ROOT) + " (float64 Cholesky at full size, pinned by scikit-learn Ridge)",
            "fits_checked": int(err.size), "max_coef_relerr": float(err.max()),
            "max_intercept_err": float(np.abs(full["intercept"][idx][:, :, 0] - z["intercept"]).max()),
            "max_sse_relerr": float(np.abs(full["sse"][idx] / z["sse"] - 1.0).max()),
            "max_r2_diff": float(np.abs(r2 - z["r2"]).max()),
            "n_test_equal": bool((full["n_test"][idx] == z["n_test"][None, :]).all()),
            "best_of_checked_equal": bool(int(np.argmax(r2.mean(axis=1))) == int(np.argmax(z["r2"].mean(axis=1))))}


GOLDEN_CFG4 = os.path.join(ROOT, "tests", "golden", "cfg4_full.npz")


def parity_report_multi(out: dict, X, fold, cv: int, golden_path: str, pos=None, eps: float = 2e-5):
    """Config 4 against tests/golden/cfg4_full.npz: scikit-learn (float64) at full size on 8 of the 512 sampled
    candidates x 5 folds (tests/golden/make_cfg4_full.py).  `pos` = where those candidates sit in `out`
    (default: their index in the 512-candidate draw order).  A count may differ only on rows whose two largest
    decision values are within eps of each other (an arg-max tie at fp32 accuracy)."""
    if not os.path.exists(golden_path):
        return {"unavailable": f"{os.path.relpath(golden_path, ROOT)} missing"}
    z = np.load(golden_path)
    pos = [int(i) for i in z["index"]] if pos is None else list(pos)
    if max(pos) >= out["n_iter"].shape[0]:
        return {"unavailable": "golden file is for another candidate list"}
    m = len(pos)
    same_iter = out["n_iter"][pos] == z["n_iter"]
    coef_err = np.zeros((m, cv)); icpt_err = np.zeros((m, cv))
    diffs = np.zeros((m, cv), np.int64); ties = np.zeros((m, cv), np.int64)
    for k in range(cv):
        Xk = X[fold == k].astype(np.float64)
        for i in range(m):
            W, b = z["coef"][i, k], z["intercept"][i, k]
            Z = Xk @ W.T + b[None, :]                                       # scikit-learn's decision values
            top2 = np.partition(Z, Z.shape[1] - 2, axis=1)[:, -2:]
            ties[i, k] = int(((top2[:, 1] - top2[:, 0]) < eps * max(1.0, float(np.abs(Z).max()))).sum())
            coef_err[i, k] = np.abs(out["coef"][pos[i], k] - W).max() / np.abs(W).max()
            icpt_err[i, k] = np.abs(out["intercept"][pos[i], k] - b).max() / max(1.0, float(np.abs(W).max()))
        diffs[:, k] = np.abs(out["correct"][pos][:, k] - z["correct"][:, k])
    return {"against": os.path.relpath(golden_path, ROOT) + " (scikit-learn float64, full size, 8 of the 512 candidates)",
            "fits_checked": int(same_iter.size), "n_iter_equal": int(same_iter.sum()),
            "nfev_equal": int((out["nfev"][pos] == z["nfev"]).sum()),
            "max_coef_relerr": float(coef_err[same_iter].max()) if same_iter.any() else None,
            "max_intercept_err": float(icpt_err[same_iter].max()) if same_iter.any() else None,
            "count_diffs": int((diffs > 0).sum()), "max_count_diff": int(diffs.max()),
            "count_diffs_beyond_ties": int((diffs > ties).sum()),
            "rows_excused_as_ties": int(np.minimum(diffs, ties).sum()), "rows_near_ties": int(ties.sum()),
            "n_test_equal": bool((out["n_test"][pos] == z["n_test"]).all()),
            "max_mean_score_diff": float(np.abs((out["correct"][pos] / out["n_test"][pos]).mean(axis=1)
                                                - (z["correct"] / z["n_test"]).mean(axis=1)).max()),
            "best_of_checked_equal": bool(int(np.argmax((out["correct"][pos] / out["n_test"][pos]).mean(axis=1)))
                                          == int(np.argmax((z["correct"] / z["n_test"]).mean(axis=1))))}


GOLDEN_CFG5 = os.path.join(ROOT, "tests", "golden", "cfg5_full.npz")


def _guarded(fn, *args):
    """A parity record must never cost the measured line: a fault in the comparison is reported in its place."""
    try:
        return fn(*args)
    except Exception as e:                                    # noqa: BLE001
        return {"unavailable": f"{type(e).__name__}: {e}"}


def parity_report_cached_ties(out: dict, golden_path: str, pos=None):
    """Config 5 against tests/golden/cfg5_full.npz: the oracle's L-BFGS restatement evaluated in float64 at full
    size (16M x 512) for 8 of the 256 C values x 10 folds, pinned against scikit-learn on a slice
    (tests/golden/make_cfg5_full.py).  The dataset is never whole on the host here, so the number of test rows
    within 2e-5 of a decision tie comes from the cache instead of being recomputed."""
    if not os.path.exists(golden_path):
        return {"unavailable": f"{os.path.relpath(golden_path, ROOT)} missing"}
    z = np.load(golden_path)
    pos = [int(i) for i in z["index"]] if pos is None else list(pos)
    if max(pos) >= out["n_iter"].shape[0] or out["n_iter"].shape[1] != z["n_iter"].shape[1]:
        return {"unavailable": "golden file is for another candidate list"}
    same_iter = out["n_iter"][pos] == z["n_iter"]
    wmax = np.abs(z["coef"]).max(axis=2)
    coef_err = np.abs(out["coef"][pos][:, :, 0, :] - z["coef"]).max(axis=2) / wmax
    icpt_err = np.abs(out["intercept"][pos][:, :, 0] - z["intercept"]) / np.maximum(1.0, wmax)
    diffs = np.abs(out["correct"][pos] - z["correct"]); ties = z["ties"]
    sc, so = (out["correct"][pos] / out["n_test"][pos]).mean(axis=1), (z["correct"] / z["n_test"]).mean(axis=1)
    return {"against": os.path.relpath(golden_path, ROOT) + " (float64 L-BFGS restatement at full size, pinned by scikit-learn; 8 of the 256 C values)",
            "fits_checked": int(same_iter.size), "n_iter_equal": int(same_iter.sum()),
            "nfev_equal": int((out["nfev"][pos] == z["nfev"]).sum()),
            "max_coef_relerr": float(coef_err[same_iter].max()) if same_iter.any() else None,
            "max_intercept_err": float(icpt_err[same_iter].max()) if same_iter.any() else None,
            "count_diffs": int((diffs > 0).sum()), "max_count_diff": int(diffs.max()),
            "count_diffs_beyond_ties": int((diffs > ties).sum()),
            "rows_excused_as_ties": int(np.minimum(diffs, ties).sum()), "rows_near_ties": int(ties.sum()),
            "n_test_equal": bool((out["n_test"][pos] == z["n_test"]).all()),
            "max_mean_score_diff": float(np.abs(sc - so).max()),
            "best_of_checked_equal": bool(int(np.argmax(sc)) == int(np.argmax(so)))}


def golden_nfev_sum():
    """Function evaluations of all 320 fits of config 2 as scikit-learn makes them (for roofline.frac_step)."""
    if not os.path.exists(GOLDEN_CFG2):
        return None
    return int(np.load(GOLDEN_CFG2)["nfev"].sum())


# ------------------------------------------------------------------------------------------
# CPU legs: the reference's own path on the box's host cores (oracle.reference_path = the worker's call
# sequence, aws-prod/worker/worker.py:315-326).  Only these functions touch oracle/.
def _cv_fits_limited(model, params, X, y, cv, threads):
    from threadpoolctl import threadpool_limits
    from oracle import reference_path as rp
    with threadpool_limits(limits=threads):
        return [(f["score"], f.get("correct"), f["n_test"]) for f in rp.cv_fits(model, params, X, y, cv)]


def cpu_worker_path(model, cands, X, y, cv, workers, threads_per_worker):
    """C1 of BASELINE.md: per candidate a serial cross_val_score (worker.py:326), candidates spread over
    `workers` processes (the reference runs 4 worker containers, docker-compose.yml:133-199), each with
    `threads_per_worker` BLAS threads.  Returns (seconds, per-candidate fold records)."""
    from joblib import Parallel, delayed
    t0 = time.perf_counter()
    if workers <= 1 or len(cands) == 1:
        res = [_cv_fits_limited(model, c, X, y, cv, threads_per_worker) for c in cands]
    else:
        res = Parallel(n_jobs=min(workers, len(cands)))(
            delayed(_cv_fits_limited)(model, c, X, y, cv, threads_per_worker) for c in cands)
    return time.perf_counter() - t0, res


def cpu_gridsearch_njobs(model, cands, X, y, cv):
    """C2 of BASELINE.md: scikit-learn's own GridSearchCV(..., n_jobs=-1) as in results1.py:71,98 (one loky
    worker per core; joblib limits each worker's BLAS to cores // n_jobs = 1 thread)."""
    from sklearn.model_selection import GridSearchCV
    from oracle import reference_path as rp
    keys = sorted(cands[0])
    grid = [{k: [c[k]] for k in keys} for c in cands]          # the sampled candidates, not their cross product
    gs = GridSearchCV(rp.make_model(model, {}), grid, cv=cv, n_jobs=-1, refit=False)
    t0 = time.perf_counter()
    gs.fit(X, y)
    return time.perf_counter() - t0, gs


def host_data(cfg, rows=None):
    """(X fp32, y) of a config on the host.  Config 5 uses the chunk-seeded generator (any row range on its own)."""
    from cs230_distributed_machine_learning_b200 import make_dataset
    from cs230_distributed_machine_learning_b200.datasets import make_dataset_rows
    if cfg["seed"] == 5:
        lo, hi = rows if rows else (0, cfg["n"])
        return make_dataset_rows(cfg["kind"], cfg["n"], cfg["d"], cfg["seed"], lo, hi, threads=max(2, (os.cpu_count() or 8) // 2))
    return make_dataset(cfg["kind"], cfg["n"], cfg["d"], seed=cfg["seed"])


def run_reference(a):
    """--impl reference: scikit-learn on the box's host cores, a bounded sample of the workload per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = a.cfg
    cores = os.cpu_count() or 1
    # the reference gets every host thread: torchrun exports OMP_NUM_THREADS=1 to its ranks, which would pin
    # this process's BLAS (already loaded) and the joblib workers' to one thread
    os.environ.pop("OMP_NUM_THREADS", None)
    from threadpoolctl import threadpool_limits
    threadpool_limits(limits=cores)
    n_sub = cfg["n"] if cfg["n"] <= 4_000_000 else 1_000_000     # config 5: a 1M-row slice, fits/s scaled by n (stated)
    X, y = host_data(cfg, rows=(0, n_sub) if cfg["seed"] == 5 else None)
    if cfg["model"] == "Ridge":
        y = y.astype(np.float32)
    cands_all = candidates_of(cfg, cfg["n_cand"] * a.gpus)
    model, cv = cfg["model"], cfg["cv"]
    # calibrate on one candidate (cv fits, BLAS on all cores), then size each leg of a step to ~3 s
    t1, _ = cpu_worker_path(model, cands_all[len(cands_all) // 2:len(cands_all) // 2 + 1], X, y, cv, 1, cores)
    workers = max(1, min(cores // 4, 16))
    per_step = max(1, min(len(cands_all), int(3.0 / max(t1, 1e-3) * max(1, workers // 2))))
    sample = [cands_all[i] for i in np.linspace(0, len(cands_all) - 1, per_step).astype(int)]
    w_eff = workers if per_step > 1 else 1
    thr = max(1, cores // w_eff)
    for _ in range(max(1, a.warmup) if a.warmup else 0):
        cpu_worker_path(model, sample[:1], X, y, cv, 1, cores)
    t_c1 = t_c2 = 0.0
    for _ in range(a.steps):
        t_c1 += cpu_worker_path(model, sample, X, y, cv, w_eff, thr)[0]
        t_c2 += cpu_gridsearch_njobs(model, sample, X, y, cv)[0]
    fits = per_step * cv * a.steps
    scale = n_sub / cfg["n"]                                   # cost of a fit is linear in n
    v1, v2 = fits / t_c1 * scale, fits / t_c2 * scale
    v, t = (v1, t_c1) if v1 >= v2 else (v2, t_c2)
    import sklearn
    sample_txt = (f"{per_step} of {len(cands_all)} candidates x {cv} folds per step at n={n_sub}"
                  + (f" of {cfg['n']} rows (fits/s scaled by {scale:.4f}: a fit's cost is linear in n)" if scale != 1 else " (full n)")
                  + f"; value = the faster of C1 worker path ({w_eff} processes x {thr} BLAS threads: {v1:.3f} fits/s) and "
                    f"C2 GridSearchCV(n_jobs=-1) ({v2:.3f} fits/s); float32 X; scikit-learn {sklearn.__version__}")
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": t / a.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(cfg, len(cands_all)), "l2": "inputs larger than L2"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample_txt,
                         "c1_worker_path_fits_per_s": v1, "c2_gridsearch_njobs_fits_per_s": v2},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0})


# ------------------------------------------------------------------------------------------
# GPU arm
class Env:
    """Rank bookkeeping + the small collectives the bench needs."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise RuntimeError("bench.py --impl b200 needs a B200; there is no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        try:      # run on the CPUs next to this rank's GPU, so that the host arrays it creates (first touch) and
            import pynvml      # pins sit on the GPU's NUMA node: a copy across sockets runs at 30 instead of 52 GB/s
            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(self.local))
        except Exception:
            pass
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_ms(self, v):
        from cs230_distributed_machine_learning_b200 import dist as ddist
        return ddist.max_over_ranks(float(v), device=self.dev)


def device_dataset(cfg, solver, rows=None):
    """Config data resident on this rank's GPU: (DeviceDataset, host y of ALL rows it holds, host X or None).
    Config 5 is generated and uploaded 4M rows at a time (32.8 GB never sits on the host at once)."""
    import torch
    from cs230_distributed_machine_learning_b200.datasets import CHUNK_ROWS, make_dataset_rows
    from cs230_distributed_machine_learning_b200.solver import DeviceDataset
    classifier = cfg["model"] != "Ridge"
    if cfg["seed"] != 5:
        X, y = host_data(cfg)
        if not classifier:
            y = y.astype(np.float32)
        return solver.upload_dataset(X, y, classifier=classifier), y, X
    lo, hi = rows if rows else (0, cfg["n"])
    return device_dataset_rows(cfg, solver, lo, hi)


def make_step(cfg, solver, ds, fold_d, cands, tc):
    """One whole sweep of `cands` (all folds) on this rank."""
    n = len(cands)
    if cfg["model"] == "LogisticRegression":
        Cs = [c["C"] for c in cands]
        tol = [c.get("tol", 1e-4) for c in cands]; fi = [c.get("fit_intercept", True) for c in cands]; mi = [100] * n
        return lambda: solver.logreg_sweep(ds, fold_d, cfg["cv"], Cs, tol, fi, mi, use_tensor_cores=tc)
    al = [c["alpha"] for c in cands]
    clf = cfg["model"] == "RidgeClassifier"
    return lambda: solver.ridge_sweep(ds, fold_d, cfg["cv"], al, [True] * n, classifier=clf, use_tensor_cores=tc)


def timed_steps(env, step, steps, warmup, sampler=None):
    """W untimed steps, then exactly K steps between barriers; device time, max over ranks.  Returns
    (ms total, last result, summed timing counters)."""
    torch = env.torch
    out = None
    for _ in range(warmup):
        out = step()
    reset_sampler(sampler)
    env.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    acc = {}
    for _ in range(steps):
        out = step()
        for k, v in out["timing"].items():
            acc[k] = acc.get(k, 0) + v
    e1.record()
    env.barrier()
    return env.max_ms(e0.elapsed_time(e1)), out, acc


def scores_of(cfg, out):
    if cfg["model"] == "Ridge":
        sst = out["syy"] - out["sy"] ** 2 / out["n_test"]
        return 1.0 - out["sse"] / sst
    return out["correct"] / out["n_test"]


def traffic_record(key):
    """DRAM bytes per launch of the dominant kernels, from the committed ncu --set full capture."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    return json.load(open(p)).get(key)


def roofline_logreg(cfg, acc, ms_total, steps, fits_local, T, p_tf32, p_burst, peaks, src, oracle_nfev=None):
    """Tensor roofline of the forward + gradient contraction pair (rank 0's kernels)."""
    n, d, cv = cfg["n"], cfg["d"], cfg["cv"]
    n_tr, n_te, p3 = n * (cv - 1) / cv, n / cv, p_tf32 / 3.0
    col_rounds, first_cols = acc["column_rounds"], acc["first_round_columns"]
    contraction_ms, first_ms = acc["contraction_ms"], acc["first_round_ms"]
    # a column-round = one pass of the training rows through one column, forward + gradient = 4 n_tr (d+1) flops
    flops = (col_rounds - first_cols) * 4.0 * n_tr * (d + 1)
    sec = (contraction_ms - first_ms) * 1e-3
    ach = flops / sec / 1e12 if sec > 0 else None
    # whole step, SURVEY.md 8d: sum over fits of nfev * 4 n_tr (d+1) T + scoring 2 n_te d T, over the driver-timed step
    if oracle_nfev is not None:
        evals, evals_src = oracle_nfev * steps, "scikit-learn's nfev per fit (tests/golden/cfg2_full.npz)"
    else:
        evals, evals_src = col_rounds / T, "the GPU's own evaluated column-rounds (round 0 counted once per row mask)"
    flops_step = evals * 4.0 * n_tr * (d + 1) * T + steps * fits_local * 2.0 * n_te * d * T
    ach_step = flops_step / (ms_total * 1e-3) / 1e12
    rounds = acc["rounds"] / steps
    full_rounds = rounds - (1 if first_cols else 0)
    tr = traffic_record(f"cfg{cfg['seed']}_fwd_grad_pair") if (n, d) == (1_000_000, 256) else None
    return {"bound": "tensor", "achieved": ach, "peak": p3, "unit": "TFLOP/s", "frac": (ach / p3) if ach else None,
            "frac_step": ach_step / p3, "achieved_step": ach_step,
            "frac_step_note": f"all algorithmic flops of the step (round 0 and the scoring pass included; evaluations = {evals_src}) "
                              "over the driver-timed step, against the same peak",
            "kernel": "forward+gradient contraction pair per batched round (frac: their CUDA-event time only, round 0 excluded)",
            "traffic": tr["bytes"] if tr else None,
            "traffic_algorithmic": float(n) * d * 4,
            "traffic_note": (f"measured DRAM bytes of one forward + one gradient launch ({tr['source']}) against the algorithmic "
                             f"X-once {n * d * 4 / 1e9:.3f} GB per round: {tr['bytes'] / (n * d * 4):.2f}x -- the residuals R (n x K fp32) "
                             "cross HBM once each way and X is streamed by both kernels; DESIGN.md has the on-chip budget "
                             "that rules the single-kernel form out") if tr else "no ncu capture for this shape",
            "peak_source": f"cuBLAS TF32 8192^3 measured in this run, sustained over ~2 s = {p_tf32:.1f} TFLOP/s "
                           f"(burst {p_burst:.1f}); /3 for 3xTF32; MEASURED_PEAKS.json has no TF32 entry",
            "burst_frac": (ach / (p_burst / 3.0)) if ach else None, "naive_tf32_frac": (ach / p_tf32) if ach else None,
            "bf16_sustained_frac": (ach / (peaks.get("bf16_tflops_sustained", 0) / 6.0)) if ach and peaks.get("bf16_tflops_sustained") else None,
            "hbm_achieved_gbs": (full_rounds * steps * n * d * 4.0) / sec / 1e9 if sec > 0 else None,
            "hbm_peak_gbs": peaks.get("hbm_gbs"), "hbm_peak_source": src,
            "kernel_ms_per_step": contraction_ms / steps, "kernel_share_of_step": contraction_ms / ms_total,
            "round0": {"ms_per_step": first_ms / steps, "columns": first_cols // max(1, steps),
                       "note": "w = 0: evaluated once per row mask and copied to every candidate; not in `achieved`, in `frac_step`"},
            "rounds_per_step": rounds}


def roofline_ridge(cfg, acc, ms_total, steps, n_alpha, p_tf32, p_burst, peaks, src):
    n, d, cv, T = cfg["n"], cfg["d"], cfg["cv"], 1
    p3 = p_tf32 / 3.0
    # moments once for all folds; the kernel exploits symmetry (upper block triangle), so the d^2 term is halved
    # as SURVEY.md 8d prescribes: n d^2 for X^T X + 2 n d T for X^T Y
    gram = float(n) * d * d + 2.0 * n * d * T
    solves = cv * n_alpha * (d ** 3 / 3.0 + 2.0 * d * d * T)
    score = 2.0 * n * d * n_alpha * T       # every row is a test row of exactly one fold, for every alpha
    flops_step = gram + solves + score
    ach_step = flops_step * steps / (ms_total * 1e-3) / 1e12
    sec = acc["contraction_ms"] * 1e-3
    ach = (gram + score) * steps / sec / 1e12 if sec > 0 else None
    return {"bound": "tensor", "achieved": ach, "peak": p3, "unit": "TFLOP/s", "frac": (ach / p3) if ach else None,
            "frac_step": ach_step / p3, "achieved_step": ach_step,
            "kernel": "Gram (tc_gram_kernel) + scoring pass; frac: their CUDA-event time only; frac_step: Gram + fp64 solves + "
                      "scoring over the whole step",
            "flops_note": "n d^2 + 2 n d T for the moments (symmetry is exploited -- the kernel multiplies the upper block "
                          "triangle -- so the d^2 term of SURVEY.md 8d is halved) + cv * n_alpha * (d^3/3 + 2 d^2 T) fp64 "
                          "solve flops + 2 n d n_alpha T scoring",
            "traffic": None, "traffic_algorithmic": 2.0 * n * d * 4,
            "peak_source": f"cuBLAS TF32 8192^3 sustained = {p_tf32:.1f} TFLOP/s (burst {p_burst:.1f}); /3 for 3xTF32",
            "hbm_achieved_gbs": 2.0 * n * d * 4 * steps / sec / 1e9 if sec > 0 else None, "hbm_peak_gbs": peaks.get("hbm_gbs"),
            "hbm_peak_source": src, "kernel_ms_per_step": acc["contraction_ms"] / steps,
            "kernel_share_of_step": acc["contraction_ms"] / ms_total,
            "non_contraction_share_of_step": 1.0 - acc["contraction_ms"] / ms_total}


def search_estimator(cfg, cands):
    """The scikit-learn search object a user hands to MLTaskManager.train for these candidates."""
    from sklearn.linear_model import LogisticRegression, Ridge, RidgeClassifier
    from sklearn.model_selection import GridSearchCV
    base = {"LogisticRegression": LogisticRegression, "Ridge": Ridge, "RidgeClassifier": RidgeClassifier}[cfg["model"]]()
    keys = sorted(cands[0])
    if len(keys) == 1:
        return GridSearchCV(base, {keys[0]: [c[keys[0]] for c in cands]}, cv=cfg["cv"])
    # several searched parameters (config 4: the RandomizedSearchCV draw): one single-point grid per candidate,
    # which ParameterGrid enumerates in list order
    return GridSearchCV(base, [{k: [c[k]] for k in keys} for c in cands], cv=cfg["cv"])


def run_e2e(env, a, cfg, X, y, cands, n_cand_total):
    """The user's call with HOST buffers: a stream of MLTaskManager.train jobs on alternating registered
    datasets, every job paying the host->device copy of its X and y inside the timed region."""
    torch, dist = env.torch, env.dist
    from cs230_distributed_machine_learning_b200 import register_dataset
    from cs230_distributed_machine_learning_b200 import worker as wk
    from cs230_distributed_machine_learning_b200.client import LocalBackend, MLTaskManager
    names = [f"bench-cfg{a.config}-a", f"bench-cfg{a.config}-b"]
    for nm in names:
        register_dataset(nm, X, y)
    backend = LocalBackend(gpus=[env.local], monitor_interval=None)
    mgr = MLTaskManager(backend)
    est = search_estimator(cfg, cands)
    ex = wk.gpu_executor(env.local)
    if env.world > 1:
        # the ranks hold replicas of the same dataset: each copies 1/N of the rows over its own PCIe link and one
        # NCCL all-gather over NVLink completes the copies (SweepSolver.upload_replicated)
        ex.replica = (dist.new_group(list(range(env.world))), env.rank, env.world)
    # the worker may keep 1.5 datasets: the one a job reads plus the next one arriving, so
    # EVERY job pays the full host->device copy of its X and y inside the timed region
    ex.max_resident_bytes = int(1.5 * X.nbytes)
    cv = cfg["cv"]
    for nm in names:                                                 # warm-up (module import, arena, labels)
        r = mgr.train(est, nm, wait_for_completion=True)
        assert r.get("job_status") == "completed", r
    k_ser = max(1, min(a.steps, 4))
    env.barrier()
    t0 = time.perf_counter()
    for i in range(k_ser):                                           # one job at a time: copy, then sweep
        r = mgr.train(est, names[i % 2], wait_for_completion=True)
        assert r["job_status"] == "completed"
    torch.cuda.synchronize()
    t_ser = env.max_ms(time.perf_counter() - t0)
    k_e2e = max(2, a.steps if X.nbytes < (2 << 30) else min(a.steps, 4))
    env.barrier()
    t0 = time.perf_counter()
    h2d0, d2h0, pf0 = ex.solver.h2d_bytes, ex.solver.d2h_bytes, ex.prefetched
    for v in ex.stats.values():
        v.clear()
    jobs = []
    for i in range(k_e2e):                                           # the queue holds the next job: its copy
        resp = mgr.train(est, names[i % 2], wait_for_completion=False)   # runs beside the current sweep
        assert resp.get("status") == "Training started", resp
        jobs.append(mgr.job_id)
    for j in jobs:
        r = mgr._wait_for_completion(j)
        assert r["job_status"] == "completed"
    torch.cuda.synchronize()
    t_e2e = env.max_ms(time.perf_counter() - t0)
    e2e = {"value": n_cand_total * cv * k_e2e / t_e2e, "unit": UNIT,
           "h2d_bytes_per_step": (ex.solver.h2d_bytes - h2d0) // k_e2e,
           "d2h_bytes_per_step": (ex.solver.d2h_bytes - d2h0) // k_e2e,
           "api": "MLTaskManager.train(GridSearchCV(...)) x steps, then _wait_for_completion per job",
           "steps": k_e2e, "ms_per_step": t_e2e / k_e2e * 1e3,
           "overlap": "the worker copies job i+1's dataset on a copy stream while job i's sweep runs; "
                      f"{ex.prefetched - pf0} of {k_e2e} copies were started that way",
           "one_job_at_a_time": {"value": n_cand_total * cv * k_ser / t_ser, "steps": k_ser,
                                 "ms_per_step": t_ser / k_ser * 1e3},
           "per_job_ms": {k: (float(np.mean(v)) if v else None) for k, v in ex.stats.items()},
           "best_params_": r["job_result"]["best_params_"]}
    if env.world > 1:
        e2e["upload"] = (f"per rank: rows [r n/{env.world}, (r+1) n/{env.world}) host->device over its own link, replicas completed "
                         "by one NCCL all-gather over NVLink; h2d_bytes_per_step is per rank")
    mgr.close()
    return e2e


# ------------------------------------------------------------------------------------------
# N > 1 records that ride on the headline line
def rows_solver(env):
    """A second context on this GPU, joined to an NCCL communicator of all ranks (row-sharded sweeps)."""
    from cs230_distributed_machine_learning_b200 import _cabi
    from cs230_distributed_machine_learning_b200 import dist as ddist
    from cs230_distributed_machine_learning_b200.solver import SweepSolver
    sv = SweepSolver(env.local)
    uid = ddist.broadcast_bytes(_cabi.nccl_unique_id() if env.rank == 0 else None, 128, device=env.dev)
    sv.init_comm(env.world, env.rank, uid)
    return sv


def strong_record(env, a, cfg, X, y, fold, solver, ds, fold_d, sv_rows):
    """Config 2 itself (64 candidates in total) over the N GPUs: candidate blocks (64 / N candidates per GPU,
    X replicated, no collective) and row shards (all 64 candidates per GPU on n / N rows, one all-reduce per round)."""
    torch = env.torch
    from cs230_distributed_machine_learning_b200 import dist as ddist
    cands = candidates_of(cfg, cfg["n_cand"])
    F = cfg["n_cand"] * cfg["cv"]
    k = max(2, min(a.steps, 10))
    rec = {"workload": workload_name(cfg, cfg["n_cand"]), "steps": k}
    # -- candidate blocks
    mine = ddist.candidate_indices(len(cands), env.world, env.rank)
    step = make_step(cfg, solver, ds, fold_d, [cands[i] for i in mine], a.tc)
    ms, out, acc = timed_steps(env, step, k, 3)
    keys = ["correct", "n_test", "n_iter", "nfev", "coef", "intercept"]
    full = ddist.gather_candidates({q: out[q] for q in keys}, len(cands), env.world, env.rank, keys, device=env.dev)
    rec["candidates"] = {"fits_per_s": F * k / (ms * 1e-3), "ms_per_step": ms / k, "columns_per_gpu": len(mine) * cfg["cv"],
                         "rounds": acc["rounds"] / k, "kernel_share_of_step": acc["contraction_ms"] / ms}
    if env.rank == 0 and a.full_size and a.config == "2":
        rec["candidates"]["parity"] = parity_report(full, X, y, fold, cfg["cv"], GOLDEN_CFG2)
    # -- row shards
    lo, hi = ddist.row_slice(cfg["n"], env.world, env.rank)
    ds_r = sv_rows.upload_dataset(X[lo:hi], y[lo:hi], classifier=True, classes=np.unique(y))
    fold_r = sv_rows.to_device(fold[lo:hi], torch.int8)
    step = make_step(cfg, sv_rows, ds_r, fold_r, cands, a.tc)
    ms, out, acc = timed_steps(env, step, k, 3)
    rec["rows"] = {"fits_per_s": F * k / (ms * 1e-3), "ms_per_step": ms / k, "columns_per_gpu": F, "rows_per_gpu": hi - lo,
                   "rounds": acc["rounds"] / k, "kernel_share_of_step": acc["contraction_ms"] / ms,
                   "allreduce_ms_per_step": acc["allreduce_ms"] / k, "allreduce_calls_per_step": acc["allreduce_calls"] / k,
                   "allreduce_us_per_call": 1e3 * acc["allreduce_ms"] / max(1, acc["allreduce_calls"]),
                   "allreduce_bytes_per_step": acc["allreduce_bytes"] / k}
    if env.rank == 0 and a.full_size and a.config == "2":
        rec["rows"]["parity"] = parity_report(out, X, y, fold, cfg["cv"], GOLDEN_CFG2)
    del ds_r, fold_r
    best = max(("candidates", "rows"), key=lambda m: rec[m]["fits_per_s"])
    rec["partition"] = best
    rec["fits_per_s"], rec["ms_per_step"] = rec[best]["fits_per_s"], rec[best]["ms_per_step"]
    rec["limiter"] = ("candidate blocks leave 320 / N columns per GPU: below the HBM/tensor ridge (about 40 columns) a round "
                      "costs one pass over X whatever its width, and the per-round optimiser and reduction kernels (tens of "
                      "microseconds) plus the host gap between sweeps stop shrinking; row shards keep 320 columns per GPU and "
                      "pay one all-reduce of the fp64 [G | sum r | sum loss] record per round")
    return rec


def dist_parity_record(env, solver, sv_rows):
    """A small row-sharded sweep must equal the one-GPU sweep: counts and n_iter identical, coefficients to
    rounding (different summation order across shards).  The driver's GPU-test box has one GPU and skips
    tests/test_gpu_dist.py, so the check rides on the scaling run."""
    torch = env.torch
    from cs230_distributed_machine_learning_b200 import dist as ddist
    from cs230_distributed_machine_learning_b200 import folds, make_dataset
    n, d, cv = 40_000, 64, 5
    X, y = make_dataset("binary", n, d, seed=77)
    fold = folds.stratified_labels(y, cv)
    Cs = [float(c) for c in np.logspace(-4, 1, 6)]
    args = ([1e-4] * 6, [True] * 6, [100] * 6)
    want = solver.logreg_sweep(solver.upload_dataset(X, y, classifier=True), fold, cv, Cs, *args)
    lo, hi = ddist.row_slice(n, env.world, env.rank)
    ds = sv_rows.upload_dataset(X[lo:hi], y[lo:hi], classifier=True, classes=np.unique(y))
    got = sv_rows.logreg_sweep(ds, fold[lo:hi], cv, Cs, *args)
    ok = dict(counts_equal=bool((got["correct"] == want["correct"]).all()), n_test_equal=bool((got["n_test"] == want["n_test"]).all()),
              n_iter_equal=bool((got["n_iter"] == want["n_iter"]).all()),
              coef_relerr=float(np.abs(got["coef"] - want["coef"]).max() / np.abs(want["coef"]).max()))
    t = torch.tensor(got["coef"].ravel()[:256].astype(np.float64), device=env.dev)
    tmax, tmin = t.clone(), t.clone()
    env.dist.all_reduce(tmax, op=env.dist.ReduceOp.MAX); env.dist.all_reduce(tmin, op=env.dist.ReduceOp.MIN)
    ok["ranks_identical"] = bool((tmax == tmin).all().item())
    flags = torch.tensor([int(ok["counts_equal"] and ok["n_iter_equal"] and ok["n_test_equal"])], device=env.dev)
    env.dist.all_reduce(flags, op=env.dist.ReduceOp.MIN)
    ok["all_ranks_agree_with_one_gpu"] = bool(flags.item())
    ok["shape"] = f"binary {n}x{d}, 6 C x cv {cv}, {env.world} row shards"
    return ok


def rows_config5_record(env, a, sv_rows):
    """BASELINE.json config 5 (16M x 512, 256 C x cv 10) row-sharded over the N GPUs: every rank generates and
    holds only its rows; per round one ncclAllReduce of the fp64 [G | sum r | sum loss] record."""
    torch = env.torch
    from cs230_distributed_machine_learning_b200 import dist as ddist
    from cs230_distributed_machine_learning_b200 import folds
    cfg = dict(CONFIGS["5"])
    cfg["n"] = a.rows_config_rows or cfg["n"]
    lo, hi = ddist.row_slice(cfg["n"], env.world, env.rank)
    t0 = time.perf_counter()
    ds, y_loc, _ = device_dataset_rows(cfg, sv_rows, lo, hi)
    gen_s = time.perf_counter() - t0
    # StratifiedKFold labels of a shard from class counts alone (folds.stratified_labels_slice)
    cnt = torch.tensor(np.bincount(y_loc, minlength=2), device=env.dev, dtype=torch.int64)
    allc = [torch.zeros_like(cnt) for _ in range(env.world)]
    env.dist.all_gather(allc, cnt)
    allc = np.stack([c.cpu().numpy() for c in allc])
    first_loc = np.array([lo + int(np.argmax(y_loc == k)) if (y_loc == k).any() else cfg["n"] for k in range(2)])
    ft = torch.tensor(first_loc, device=env.dev, dtype=torch.int64)
    env.dist.all_reduce(ft, op=env.dist.ReduceOp.MIN)
    fold = folds.stratified_labels_slice(y_loc, allc[:env.rank].sum(axis=0), allc.sum(axis=0), ft.cpu().numpy(), cfg["cv"])
    fold_d = sv_rows.to_device(fold, torch.int8)
    cands = candidates_of(cfg, cfg["n_cand"])
    step = make_step(cfg, sv_rows, ds, fold_d, cands, a.tc)
    k = 2
    ms, out, acc = timed_steps(env, step, k, 1)
    F = cfg["n_cand"] * cfg["cv"]
    n_tr = cfg["n"] * (cfg["cv"] - 1) / cfg["cv"]
    flops = acc["column_rounds"] * 4.0 * n_tr * (cfg["d"] + 1) / env.world     # this rank's share of every column-round
    scores = out["correct"] / out["n_test"]
    return {"workload": workload_name(cfg, cfg["n_cand"]) + f", row-sharded over {env.world} GPUs",
            "fits_per_s": F * k / (ms * 1e-3), "ms_per_step": ms / k, "steps": k, "warmup": 1, "rows_per_gpu": hi - lo,
            "rounds": acc["rounds"] / k, "kernel_share_of_step": acc["contraction_ms"] / ms,
            "algorithmic_tflops_per_gpu": flops / (acc["contraction_ms"] * 1e-3) / 1e12,
            "allreduce_ms_per_step": acc["allreduce_ms"] / k, "allreduce_calls_per_step": acc["allreduce_calls"] / k,
            "allreduce_us_per_call": 1e3 * acc["allreduce_ms"] / max(1, acc["allreduce_calls"]),
            "allreduce_bytes_per_step": acc["allreduce_bytes"] / k,
            "n_test_total": int(out["n_test"][0].sum()), "status_ok": bool((out["status"] == 0).all()),
            "mean_n_iter": float(out["n_iter"].mean()), "best_C": cands[search_best_index(scores)]["C"],
            "parity": _guarded(parity_report_cached_ties, out, GOLDEN_CFG5) if cfg == CONFIGS["5"] else None,
            "data": f"chunk-seeded generator (datasets.make_dataset_rows, seed 5), {gen_s:.1f} s per rank to generate and upload its rows"}


def device_dataset_rows(cfg, solver, lo, hi):
    """Rows [lo, hi) of a chunk-seeded config on this rank's GPU (lo need not be chunk-aligned)."""
    import torch
    from cs230_distributed_machine_learning_b200.datasets import CHUNK_ROWS, make_dataset_rows
    from cs230_distributed_machine_learning_b200.solver import DeviceDataset
    Xd = torch.empty((hi - lo, cfg["d"]), dtype=torch.float32, device=solver.device)
    ys = []
    step = 8 * CHUNK_ROWS
    a0 = lo // CHUNK_ROWS * CHUNK_ROWS
    threads = max(2, (os.cpu_count() or 8) // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))))
    while a0 < hi:
        b0 = min(hi, a0 + step)
        Xc, yc = make_dataset_rows(cfg["kind"], cfg["n"], cfg["d"], cfg["seed"], a0, b0, threads=threads)
        s0 = max(lo, a0)
        Xd[s0 - lo:b0 - lo].copy_(torch.from_numpy(Xc[s0 - a0:]))
        ys.append(yc[s0 - a0:])
        a0 = b0
        del Xc
    y = np.concatenate(ys)
    yd = torch.from_numpy(y.astype(np.int32)).to(solver.device)
    return DeviceDataset(Xd, yd, None, np.array([0, 1]), hi - lo, cfg["d"]), y, None


def api_probe(n_gpus: int):
    """One job through MLTaskManager -> master.create_blocks -> one Worker thread per GPU, all in THIS process
    (the reference's master + N workers on one box), by candidate blocks and by row shards.  Run by rank 0 in a
    fresh process (bench.py --api-probe N) so that a fault here cannot take the benchmark line with it."""
    import torch
    from sklearn.linear_model import LogisticRegression
    from sklearn.model_selection import GridSearchCV
    from cs230_distributed_machine_learning_b200 import make_dataset, register_dataset
    from cs230_distributed_machine_learning_b200.client import LocalBackend, MLTaskManager
    n_gpus = min(n_gpus, torch.cuda.device_count())
    X, y = make_dataset("binary", 400_000, 256, seed=2)
    register_dataset("api-probe", X, y)
    est = GridSearchCV(LogisticRegression(), {"C": [float(c) for c in np.logspace(-6, 2, 64)]}, cv=5)
    rec = {"dataset": "binary 400000x256 fp32 (seed 2), 64 C x cv 5", "gpus": n_gpus}
    means = {}
    for label, gpus, partition in (("one_gpu", [0], "candidates"), ("candidate_blocks", list(range(n_gpus)), "candidates"),
                                   ("row_shards", list(range(n_gpus)), "rows")):
        backend = LocalBackend(gpus=gpus, partition=partition, monitor_interval=None)
        mgr = MLTaskManager(backend)
        try:
            r = mgr.train(est, "api-probe", wait_for_completion=True)      # first job: uploads, NCCL set-up
            assert r.get("job_status") == "completed", r
            t0 = time.perf_counter()
            r = mgr.train(est, "api-probe", wait_for_completion=True)
            dt = time.perf_counter() - t0
            assert r.get("job_status") == "completed", r
            jr = r["job_result"]
            means[label] = np.asarray(jr["cv_results_"]["mean_test_score"])
            recs = mgr._api_request(f"metrics/{mgr.session_id}/{mgr.job_id}", "get")
            blocks = {m["subtask_id"] for m in recs}
            from cs230_distributed_machine_learning_b200 import worker as wk
            block_ms = [round((m["finished_at"] - m["started_at"]) * 1e3, 2) for m in recs]
            sweep_ms = {g: round(ex.stats["sweep_ms"][-1], 2) for g, ex in wk._EXECUTORS.items() if ex.stats["sweep_ms"]}
            host_ms = {g: round(ex.stats["host_ms"][-1], 2) for g, ex in wk._EXECUTORS.items() if ex.stats["host_ms"]}
            rec[label] = {"ms_per_job_resident": dt * 1e3, "fits_per_s": 320 / dt, "blocks": len(blocks),
                          "block_wall_ms": block_ms, "last_sweep_gpu_ms": sweep_ms, "last_sweep_host_ms": host_ms,
                          "best_params_": jr["best_params_"], "executors": sorted({x.get("executor", "failed: " + str(x.get("error"))[:200]) for x in jr["results"]}),
                          "row_shards": sorted({x.get("row_shards", 1) for x in jr["results"]})}
        finally:
            mgr.close()
    rec["candidate_blocks"]["max_mean_test_score_diff_vs_one_gpu"] = float(np.abs(means["candidate_blocks"] - means["one_gpu"]).max())
    rec["candidate_blocks"]["best_params_equal_one_gpu"] = rec["candidate_blocks"]["best_params_"] == rec["one_gpu"]["best_params_"]
    rec["row_shards"]["max_mean_test_score_diff_vs_one_gpu"] = float(np.abs(means["row_shards"] - means["one_gpu"]).max())
    rec["row_shards"]["best_params_equal_one_gpu"] = rec["row_shards"]["best_params_"] == rec["one_gpu"]["best_params_"]
    emit(rec)


def run_api_probe_subprocess(n_gpus: int, timeout: float = 240.0):
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE",
                                                              "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_RUN_ID",
                                                              "GROUP_RANK", "ROLE_RANK", "ROLE_WORLD_SIZE", "OMP_NUM_THREADS")}
    try:
        res = subprocess.run([sys.executable, os.path.abspath(__file__), "--api-probe", str(n_gpus)], capture_output=True,
                             text=True, timeout=timeout, env=env, cwd=ROOT)
        lines = [l for l in res.stdout.splitlines() if l.startswith("{")]
        if res.returncode == 0 and lines:
            return json.loads(lines[-1])
        return {"error": (res.stderr or res.stdout)[-600:]}
    except subprocess.TimeoutExpired:
        return {"error": f"timed out after {timeout} s"}


def run_b200(a):
    env = Env()
    torch, dist = env.torch, env.dist
    from cs230_distributed_machine_learning_b200 import dist as ddist
    from cs230_distributed_machine_learning_b200 import folds
    from cs230_distributed_machine_learning_b200.solver import SweepSolver
    cfg, rank, world = a.cfg, env.rank, env.world
    cv, model = cfg["cv"], cfg["model"]
    classifier = model != "Ridge"
    T = int(cfg["kind"][len("multiclass"):]) if cfg["kind"].startswith("multiclass") else 1
    n_cand_total = cfg["n_cand"] * world
    cands_all = candidates_of(cfg, n_cand_total)
    mine = ddist.candidate_indices(n_cand_total, world, rank)
    cands = [cands_all[i] for i in mine]

    sampler = ClockSampler(env.local) if rank == 0 else None
    if sampler:
        sampler.start()
    solver = SweepSolver(env.local)
    ds, y, X = device_dataset(cfg, solver)          # identical bytes on every rank
    fold = folds.stratified_labels(y, cv) if classifier else folds.kfold_labels(len(y), cv)
    fold_d = solver.to_device(fold, torch.int8)
    # The cuBLAS TF32 peak is measured right before the sweeps (about 3 s at the power cap): it is the roofline's
    # denominator and it brings the board to its sustained clocks before any sweep is timed, so that the W
    # warm-up steps below are the warm-up and nothing else (a process on a freshly booted box otherwise runs
    # its first second of sweeps above the sustained clock).
    p_burst, p_tf32 = measure_tf32_peak(torch)
    step = make_step(cfg, solver, ds, fold_d, cands, a.tc)
    n_warm = max(3, a.warmup)
    ms, out, acc = timed_steps(env, step, a.steps, n_warm, sampler)
    clocks = sampler.summary() if sampler else None
    fits_step = n_cand_total * cv
    value = fits_step * a.steps / (ms * 1e-3)

    # scores of all candidates on rank 0 (not timed)
    keys = ["correct", "n_test", "n_iter", "nfev", "coef", "intercept"] if model == "LogisticRegression" else \
        (["correct", "n_test"] if classifier else ["sse", "sy", "syy", "n_test", "coef", "intercept"])
    full = ddist.gather_candidates({k: out[k] for k in keys}, n_cand_total, world, rank, keys, device=env.dev)

    # ---- e2e: the user's call, host buffers, H2D/D2H inside the timed region ----
    e2e = None
    ds = fold_d = step = None
    solver.close()
    del solver
    torch.cuda.empty_cache()
    if not a.no_e2e and X is not None:
        e2e = run_e2e(env, a, cfg, X, y, cands, n_cand_total)

    line = None
    if rank == 0:
        peaks, src = measured_peaks()
        if model == "LogisticRegression":
            nfev = golden_nfev_sum() if (a.config == "2" and a.full_size and world == 1) else None
            roof = roofline_logreg(cfg, acc, ms, a.steps, len(cands) * cv, T, p_tf32, p_burst, peaks, src, nfev)
            scores = full["correct"] / full["n_test"]
        else:
            roof = roofline_ridge(cfg, acc, ms, a.steps, len(cands), p_tf32, p_burst, peaks, src)
            scores = scores_of(cfg, full)
        best = search_best_index(scores)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": n_warm, "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 (3xTF32-equivalent accuracy, fp64 optimiser state)" if model == "LogisticRegression"
            else "f32 (3xTF32 Gram, fp64 solves)",
            "data": "synthetic",
            "config": {"workload": workload_name(cfg, n_cand_total), "partition": f"candidate blocks x{world}",
                       "l2": f"inputs ({cfg['n'] * cfg['d'] * 4 / 1e9:.2f} GB) larger than the 126 MB L2; no flush",
                       "fits_per_step": fits_step, "best_params_": cands_all[best], "best_index_": best,
                       "x_tile_multicast": out["timing"].get("cluster") == 2,
                       "preheat": "cuBLAS TF32 peak measurement (about 3 s) runs right before the warm-up steps"},
            "roofline": roof, "clocks": clocks, "gpu_launches": int(acc["total_launches"]),
            "contraction_launches": int(acc["contraction_launches"]),
        }
        if model == "LogisticRegression":
            line["config"]["mean_n_iter"] = float(np.mean(full["n_iter"]))
            if a.config == "2" and a.full_size and world == 1:
                line["parity"] = parity_report(full, X, y, fold, cv, GOLDEN_CFG2)
            if a.config == "4" and a.full_size and world == 1:
                line["parity"] = _guarded(parity_report_multi, full, X, fold, cv, GOLDEN_CFG4)
            if a.config == "5" and a.full_size and world == 1:
                line["parity"] = _guarded(parity_report_cached_ties, full, GOLDEN_CFG5)
        if a.config == "3" and a.full_size and world == 1:
            line["parity"] = parity_report_ridge(full, GOLDEN_CFG3)
        line["e2e"] = e2e if e2e else {"value": None, "unit": UNIT, "h2d_bytes_per_step": None, "d2h_bytes_per_step": None,
                                       "note": "not measured for this config (the dataset is generated per row range, never whole on the host)"
                                       if X is None else "skipped (--no-e2e)"}
        if not a.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline_record(cfg, cands_all, X, y)

    # ---- N > 1: strong scaling of config 2, config 5 row-sharded, the API over all GPUs of the box ----
    if world > 1 and not a.no_extras and a.config == "2":
        extras = {}

        def give_up():      # a hung collective must not take the measured line with it
            if rank == 0:
                line.update(extras)
                line["extras_error"] = "the multi-GPU records did not finish within 420 s"
                emit(line)
            os._exit(0)
        deadline = threading.Timer(420.0, give_up)
        deadline.daemon = True
        deadline.start()
        solver = SweepSolver(env.local)
        ds = solver.upload_dataset(X, y, classifier=True)
        fold_d = solver.to_device(fold, torch.int8)
        sv_rows = rows_solver(env)
        extras["strong"] = strong_record(env, a, cfg, X, y, fold, solver, ds, fold_d, sv_rows)
        extras["dist_parity"] = dist_parity_record(env, solver, sv_rows)
        ds = fold_d = None
        solver.close()
        del solver
        torch.cuda.empty_cache()
        try:
            extras["rows"] = rows_config5_record(env, a, sv_rows)
        except torch.cuda.OutOfMemoryError as e:
            extras["rows"] = {"error": str(e)[:200]}
        extras["rows"]["dist_parity"] = extras["dist_parity"]
        sv_rows.close()
        del sv_rows
        torch.cuda.empty_cache()
        env.barrier()
        # The other ranks wait on the host (a key in the rendezvous store), not in an NCCL barrier: a pending NCCL
        # kernel spins on their GPUs and the probe's sweeps would share the SMs with it (measured: 31 ms per job
        # instead of 9).
        store = dist.distributed_c10d._get_default_store()
        if rank == 0:
            extras["api_all_gpus"] = run_api_probe_subprocess(world)
            store.set("dml-bench-api-probe-done", "1")
        else:
            store.wait(["dml-bench-api-probe-done"])
        env.barrier()
        deadline.cancel()
        if rank == 0:
            line.update(extras)
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline_record(cfg, cands_all, X, y):
    """The oracle (the worker's call sequence on scikit-learn) on the box's host cores, a bounded sample."""
    cores = os.cpu_count() or 1
    model, cv = cfg["model"], cfg["cv"]
    scale = 1.0
    if X is None:                                  # config 5: a 1M-row slice, a fit's cost is linear in n
        X, y = host_data(cfg, rows=(0, 1_000_000))
        scale = len(y) / cfg["n"]
    pick = [len(cands_all) // 4, 3 * len(cands_all) // 4] if model == "LogisticRegression" and scale == 1.0 and \
        cfg["n"] * cfg["d"] <= 300_000_000 else [len(cands_all) // 2]
    sample = [cands_all[i] for i in pick]
    t, _ = cpu_worker_path(model, sample, X, y, cv, 1, cores)
    return {"value": len(sample) * cv / t * scale, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{len(sample)} of {len(cands_all)} candidates x {cv} folds at n={len(y)}"
                      + (f" of {cfg['n']} rows, fits/s scaled by {scale:.4f} (a fit's cost is linear in n)" if scale != 1.0 else " (full n)")
                      + ", serial fits as in worker.py:326, BLAS on all cores, float32 X"}


if __name__ == "__main__":
    a = parse()
    _claim_stdout()
    if a.api_probe:
        api_probe(a.api_probe)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
