"""Run under torchrun with N >= 2 GPUs (tests/test_gpu_dist.py launches it):
row-sharded sweeps over NCCL must reproduce the single-GPU result."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cs230_distributed_machine_learning_b200 import _cabi, folds, make_dataset  # noqa: E402
from cs230_distributed_machine_learning_b200 import dist as ddist  # noqa: E402
from cs230_distributed_machine_learning_b200.solver import SweepSolver  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    out = {}
    n, d, cv = 40_000, 64, 5
    X, y = make_dataset("binary", n, d, seed=77)
    fold = folds.stratified_labels(y, cv)
    Cs = np.logspace(-4, 1, 6)
    solver = SweepSolver(local)
    # ---- single-GPU answers on every rank (no comm yet) ----
    ds_full = solver.upload_dataset(X, y, classifier=True)
    ref = {tc: solver.logreg_sweep(ds_full, fold, cv, Cs, [1e-4] * 6, [True] * 6, [100] * 6, use_tensor_cores=tc)
           for tc in (0, 1)}
    Xr, yr = make_dataset("regression", n, d, seed=78)
    kf = folds.kfold_labels(n, cv)
    dr_full = solver.upload_dataset(Xr, yr, classifier=False)
    ref_r = solver.ridge_sweep(dr_full, kf, cv, [0.1, 10.0], [True, True], classifier=False)
    # ---- row-sharded ----
    uid = ddist.broadcast_bytes(_cabi.nccl_unique_id() if rank == 0 else None, 128, device=dev)
    solver.init_comm(world, rank, uid)
    lo, hi = ddist.row_slice(n, world, rank)
    ds = solver.upload_dataset(X[lo:hi], y[lo:hi], classifier=True)
    ds.classes = ds_full.classes
    for tc in (0, 1):
        got = solver.logreg_sweep(ds, fold[lo:hi], cv, Cs, [1e-4] * 6, [True] * 6, [100] * 6, use_tensor_cores=tc)
        want = ref[tc]
        out[f"logreg_tc{tc}"] = dict(
            counts_equal=bool((got["correct"] == want["correct"]).all()),
            n_test_equal=bool((got["n_test"] == want["n_test"]).all()),
            n_iter_equal=bool((got["n_iter"] == want["n_iter"]).all()),
            coef_relerr=float(np.abs(got["coef"] - want["coef"]).max() / np.abs(want["coef"]).max()))
    dr = solver.upload_dataset(Xr[lo:hi], yr[lo:hi], classifier=False)
    got = solver.ridge_sweep(dr, kf[lo:hi], cv, [0.1, 10.0], [True, True], classifier=False)
    out["ridge"] = dict(coef_relerr=float(np.abs(got["coef"] - ref_r["coef"]).max() / np.abs(ref_r["coef"]).max()),
                        sse_relerr=float(np.abs(got["sse"] / ref_r["sse"] - 1).max()),
                        n_test_equal=bool((got["n_test"] == ref_r["n_test"]).all()))
    # every rank must hold the same answer
    t = torch.tensor(got["coef"].ravel()[:64].astype(np.float64), device=dev)
    tmax, tmin = t.clone(), t.clone()
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX); dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
    out["ranks_identical"] = bool((tmax == tmin).all().item())
    if rank == 0:
        print("DIST_CHECK " + json.dumps(out), flush=True)
    solver.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
