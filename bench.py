"""Headline benchmark: candidate-fold fits/sec of the (candidate x CV-fold) sweep.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Workload (BASELINE.json configs[1]): GridSearchCV(LogisticRegression lbfgs, 64 C
values = logspace(-6, 2, 64), cv=5) on synthetic binary 1M x 256 fp32 (seed 2,
SURVEY.md 8d).  A "step" is one whole sweep: all 64*5 = 320 fits trained and scored.
X (1.024 GB) is larger than the 126 MB L2, so no flush is needed between steps.
N > 1 (torchrun, one rank per GPU): weak scaling by candidate blocks -- the grid
grows to 64*N values of C over the same range, rank r owns candidates r, r+N, ...;
X is replicated, there is no data-path collective, rank 0 gathers the scores.

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

N_ROWS, N_FEAT, N_CAND, CV, SEED = 1_000_000, 256, 64, 5, 2
METRIC, UNIT = "candidate_fold_fits_per_sec", "fits/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=N_ROWS)       # smaller sizes are for debugging only
    ap.add_argument("--features", type=int, default=N_FEAT)
    ap.add_argument("--candidates", type=int, default=N_CAND)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--tc", type=int, default=-1, help="use_tensor_cores flag passed to the library")
    return ap.parse_args()


def workload_name(a, n_cand_total):
    return (f"GridSearchCV(LogisticRegression lbfgs, {n_cand_total} C values logspace(-6,2), cv={CV}) on synthetic "
            f"binary {a.rows}x{a.features} fp32 (seed {SEED})")


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


# ------------------------------------------------------------------------------------------
def cpu_reference_step(X, y, Cs, n_jobs):
    """The reference's CPU implementation of the path on host cores: the worker's call
    sequence (oracle.reference_path, aws-prod/worker/worker.py:315-326) for the given
    candidates.  n_jobs > 1 spreads candidates over processes (one sklearn worker each,
    like the reference's 4 worker containers, docker-compose.yml:133-199)."""
    from joblib import Parallel, delayed
    from oracle import reference_path as rp
    t0 = time.perf_counter()
    if n_jobs == 1 or len(Cs) == 1:
        for C in Cs:
            rp.cv_fits("LogisticRegression", {"C": float(C)}, X, y, CV)
    else:
        Parallel(n_jobs=min(n_jobs, len(Cs)))(
            delayed(rp.cv_fits)("LogisticRegression", {"C": float(C)}, X, y, CV) for C in Cs)
    return time.perf_counter() - t0


def run_reference(a):
    """--impl reference: sklearn on the box's host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from cs230_distributed_machine_learning_b200 import make_dataset
    cores = os.cpu_count() or 1
    # the reference gets every host thread: torchrun exports OMP_NUM_THREADS=1 to its ranks, which would pin
    # this process's BLAS (already loaded) and the joblib workers' to one thread
    os.environ.pop("OMP_NUM_THREADS", None)
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=cores)
    except Exception:
        pass
    X, y = make_dataset("binary", a.rows, a.features, seed=SEED)
    Cs_all = np.logspace(-6, 2, a.candidates * a.gpus)
    # calibrate on one candidate (5 fits, BLAS on all cores), then size the step to ~8 s
    t1 = cpu_reference_step(X, y, Cs_all[len(Cs_all) // 2:len(Cs_all) // 2 + 1], 1)
    workers = max(1, min(cores // 4, 16))          # processes, each with its BLAS threads
    per_step = max(1, min(len(Cs_all), int(8.0 / max(t1, 1e-3) * max(1, workers // 2))))
    sample = Cs_all[np.linspace(0, len(Cs_all) - 1, per_step).astype(int)]
    for _ in range(a.warmup):
        cpu_reference_step(X, y, sample[:1], 1)
    t = 0.0
    for _ in range(a.steps):
        t += cpu_reference_step(X, y, sample, workers if per_step > 1 else 1)
    fits = per_step * CV * a.steps
    v = fits / t
    sample_txt = (f"{per_step} of {len(Cs_all)} candidates x {CV} folds per step at full n={a.rows}, "
                  f"{workers if per_step > 1 else 1} process(es), BLAS threads left at default")
    emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": t / a.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(a, len(Cs_all)), "l2": "inputs larger than L2"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample_txt},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0})


# ------------------------------------------------------------------------------------------
def run_b200(a):
    import torch
    import torch.distributed as dist
    from cs230_distributed_machine_learning_b200 import dist as ddist
    from cs230_distributed_machine_learning_b200 import folds, make_dataset, register_dataset
    from cs230_distributed_machine_learning_b200.solver import SweepSolver

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a B200; there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    try:      # run on the CPUs next to this rank's GPU, so that the host arrays it creates (first touch) and
        import pynvml      # pins sit on the GPU's NUMA node: a copy across sockets runs at 30 instead of 52 GB/s
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
    except Exception:
        pass
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n_cand_total = a.candidates * world
    Cs_all = np.logspace(-6, 2, n_cand_total)
    mine = ddist.candidate_indices(n_cand_total, world, rank)
    Cs = Cs_all[mine]

    X, y = make_dataset("binary", a.rows, a.features, seed=SEED)   # identical bytes on every rank
    fold = folds.stratified_labels(y, CV)
    solver = SweepSolver(local_rank)
    ds = solver.upload_dataset(X, y, classifier=True)
    fold_d = solver.to_device(fold, torch.int8)
    n = len(Cs)
    args = dict(tol=[1e-4] * n, fit_intercept=[True] * n, max_iter=[100] * n)

    def step():
        return solver.logreg_sweep(ds, fold_d, CV, Cs, args["tol"], args["fit_intercept"], args["max_iter"],
                                   use_tensor_cores=a.tc)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # Warm-up: the requested steps (at least 3), then more untimed steps until 1.5 s have passed.  On a box
    # that has just booted the first process launches kernels slowly for a while (measured: 20.4 ms per
    # step with the kernels taking 12.7 ms of it, against 14.1 ms for the same command run again), and the
    # sweep settles at the power cap only after about a second.  The count that was run is what is reported.
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    n_warm, t_warm = 0, time.perf_counter()
    while n_warm < max(3, a.warmup) or (time.perf_counter() - t_warm < 1.5 and n_warm < 200):
        out = step()
        n_warm += 1
    if world > 1:       # every rank runs the same number of warm-up steps' worth of time; align before timing
        dist.barrier()
    if sampler:
        sampler.sm.clear(); sampler.mx.clear(); sampler.reasons.clear(); sampler.watts.clear()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    contraction_ms = launches = contraction_launches = 0
    col_rounds = first_ms = first_cols = 0
    for _ in range(a.steps):
        out = step()
        contraction_ms += out["timing"]["contraction_ms"]
        launches += out["timing"]["total_launches"]
        contraction_launches += out["timing"]["contraction_launches"]
        col_rounds += out["timing"]["column_rounds"]
        first_ms += out["timing"]["first_round_ms"]
        first_cols += out["timing"]["first_round_columns"]
    e1.record()
    barrier()
    ms = ddist.max_over_ranks(e0.elapsed_time(e1), device=dev)
    clocks = sampler.summary() if sampler else None
    fits_total = n_cand_total * CV * a.steps
    value = fits_total / (ms * 1e-3)

    # gather scores for the record (not timed): mean CV score and winner
    full = ddist.gather_candidates({"correct": out["correct"], "n_test": out["n_test"], "n_iter": out["n_iter"]},
                                   n_cand_total, world, rank, ["correct", "n_test", "n_iter"], device=dev)

    # ---- e2e: the user's call, host buffers, H2D/D2H inside the timed region ----
    e2e = None
    if not a.no_e2e:
        from sklearn.linear_model import LogisticRegression
        from sklearn.model_selection import GridSearchCV
        from cs230_distributed_machine_learning_b200 import worker as wk
        from cs230_distributed_machine_learning_b200.client import LocalBackend, MLTaskManager
        names = ["bench-cfg2-a", "bench-cfg2-b"]     # a stream of jobs on alternating datasets
        for nm in names:
            register_dataset(nm, X, y)
        ds = None
        solver.close()
        del solver
        torch.cuda.empty_cache()
        backend = LocalBackend(gpus=[local_rank])
        mgr = MLTaskManager(backend)
        est = GridSearchCV(LogisticRegression(), {"C": [float(c) for c in Cs]}, cv=CV)
        ex = wk.gpu_executor(local_rank)
        # the worker may keep 1.5 datasets: the one a job reads plus the next one arriving, so
        # EVERY job pays the full host->device copy of its X and y inside the timed region
        ex.max_resident_bytes = int(1.5 * X.nbytes)
        for nm in names:                                                 # warm-up (module import, arena, labels)
            r = mgr.train(est, nm, wait_for_completion=True)
            assert r.get("job_status") == "completed", r
        k_ser = max(1, min(a.steps, 4))
        barrier()
        t0 = time.perf_counter()
        for i in range(k_ser):                                           # one job at a time: copy, then sweep
            r = mgr.train(est, names[i % 2], wait_for_completion=True)
            assert r["job_status"] == "completed"
        torch.cuda.synchronize()
        t_ser = ddist.max_over_ranks(time.perf_counter() - t0, device=dev)
        k_e2e = max(2, a.steps)
        barrier()
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
        t_e2e = ddist.max_over_ranks(time.perf_counter() - t0, device=dev)
        e2e = {"value": n_cand_total * CV * k_e2e / t_e2e, "unit": UNIT,
               "h2d_bytes_per_step": (ex.solver.h2d_bytes - h2d0) // k_e2e,
               "d2h_bytes_per_step": (ex.solver.d2h_bytes - d2h0) // k_e2e,
               "api": "MLTaskManager.train(GridSearchCV(...)) x steps, then _wait_for_completion per job",
               "steps": k_e2e, "ms_per_step": t_e2e / k_e2e * 1e3,
               "overlap": "the worker copies job i+1's dataset on a copy stream while job i's sweep runs; "
                          f"{ex.prefetched - pf0} of {k_e2e} copies were started that way",
               "one_job_at_a_time": {"value": n_cand_total * CV * k_ser / t_ser, "steps": k_ser,
                                     "ms_per_step": t_ser / k_ser * 1e3},
               "per_job_ms": {k: (float(np.mean(v)) if v else None) for k, v in ex.stats.items()},
               "best_params_": r["job_result"]["best_params_"]}
        mgr.close()

    if rank == 0:
        peaks, src = measured_peaks()
        p_burst, p_tf32 = measure_tf32_peak(torch)   # the sweep is a long step: sustained figure
        p3 = p_tf32 / 3.0
        n_tr = a.rows * (CV - 1) / CV
        # algorithmic flops: every column-round is one pass of the training rows through one
        # column, forward + gradient = 4 * n_tr * (d + 1) flops
        # Round 0 (w = 0) runs once per row mask (5 columns, HBM-bound) and is left out of the
        # tensor roofline: flops and time below are those of the full-width rounds.
        flops = (col_rounds - first_cols) * 4.0 * n_tr * (a.features + 1)
        sec = (contraction_ms - first_ms) * 1e-3
        ach = flops / sec / 1e12 if sec > 0 else None
        rounds = out["timing"]["rounds"]
        full_rounds = rounds - (1 if first_cols else 0)
        hbm_ach = (full_rounds * a.steps * a.rows * a.features * 4) / sec / 1e9 if sec > 0 else None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": n_warm, "warmup_requested": a.warmup, "ms_per_step": ms / a.steps, "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 (3xTF32-equivalent accuracy, fp64 optimiser state)",
            "data": "synthetic",
            "config": {"workload": workload_name(a, n_cand_total), "partition": f"candidate blocks x{world}",
                       "l2": "inputs (1.02 GB) larger than the 126 MB L2; no flush", "fits_per_step": n_cand_total * CV,
                       "rounds_per_step": rounds, "mean_n_iter": float(np.mean(full["n_iter"])),
                       "best_C": float(Cs_all[int(np.argmax(np.mean(full["correct"] / full["n_test"], axis=1)))])},
            "roofline": {"bound": "tensor", "achieved": ach, "peak": p3, "unit": "TFLOP/s",
                         "frac": (ach / p3) if ach else None,
                         # DRAM bytes of one forward + one gradient launch (ncu --set full, dram__bytes_read.sum +
                         # dram__bytes_write.sum, profiles/r01_ncu_full_n1_fwd_grad2.csv): only known for config 2
                         "traffic": 4.20e9 if (a.rows, a.features, a.candidates) == (1_000_000, 256, 64) else None,
                         "traffic_note": "bytes per full-width round (forward 1.03 GB read + 0.98 GB written, gradient "
                                         "2.15 GB read + 0.04 GB written); algorithmic: X once per kernel (2 x 1.02 GB) + "
                                         "fp32 residuals of the training rows once each way (2 x 1.02 GB for 320 columns, "
                                         "4 of 5 folds) = 4.10 GB",
                         "kernel": "forward+gradient contraction pair per batched round",
                         "peak_source": f"cuBLAS TF32 8192^3 measured in this run, sustained over ~2 s = {p_tf32:.1f} TFLOP/s "
                                        f"(burst {p_burst:.1f}); /3 for 3xTF32; MEASURED_PEAKS.json has no TF32 entry",
                         "burst_frac": (ach / (p_burst / 3.0)) if ach else None,
                         "naive_tf32_frac": (ach / p_tf32) if ach else None,
                         "hbm_achieved_gbs": hbm_ach, "hbm_peak_gbs": peaks.get("hbm_gbs"), "hbm_peak_source": src,
                         "kernel_ms_per_step": contraction_ms / a.steps,
                         "round0": {"ms_per_step": first_ms / a.steps, "columns": first_cols // max(1, a.steps),
                                    "note": "w = 0: evaluated once per row mask and copied to every candidate; "
                                            "not counted in achieved"},
                         "kernel_share_of_step": contraction_ms / ms},
            "clocks": clocks, "gpu_launches": int(launches), "contraction_launches": int(contraction_launches),
        }
        if e2e:
            line["e2e"] = e2e
        if not a.no_cpu_baseline and world == 1:
            cores = os.cpu_count() or 1
            sample = Cs_all[[len(Cs_all) // 4, 3 * len(Cs_all) // 4]]
            t = cpu_reference_step(X, y, sample, 1)
            line["cpu_baseline"] = {"value": len(sample) * CV / t, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{len(sample)} of {len(Cs_all)} candidates x {CV} folds at full "
                                              f"n={a.rows}, serial fits as in worker.py:326, BLAS on all cores"}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    _claim_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
