"""One process per GPU: how a sweep is split over ranks and put back together.

The reference's only parallelism is one candidate per worker process over Kafka
(aws-prod/master/task_handler.py:190-207).  Here:
  * "candidates": rank r takes candidates r, r+W, r+2W, ...; X is replicated; no
    collective during the solve, one gather of per-candidate results at the end;
  * "rows": every rank takes all candidates and the row slice
    [r*n/W, (r+1)*n/W); the library all-reduces gradients / moments / counts over
    NCCL each round (dml_ctx_init_comm), so every rank ends with identical results.
torch.distributed is the plumbing: NCCL on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def candidate_indices(n_cand: int, world: int, rank: int) -> np.ndarray:
    return np.arange(rank, n_cand, world)


def row_slice(n: int, world: int, rank: int, align: int = 128) -> tuple[int, int]:
    """Contiguous, ``align``-row aligned shards (the kernels tile rows by 128)."""
    per = -(-n // world)
    per = -(-per // align) * align
    lo = min(n, rank * per)
    return lo, min(n, lo + per)


def gather_candidates(local: dict, n_cand: int, world: int, rank: int, keys, device=None) -> dict | None:
    """Per-candidate arrays computed on each rank (round-robin ownership) -> full arrays
    on rank 0.  Arrays have the candidate axis first."""
    out = {}
    for k in keys:
        a = np.ascontiguousarray(local[k])
        shape = (n_cand,) + a.shape[1:]
        full = torch.zeros(shape, dtype=torch.from_numpy(a).dtype, device=device or "cpu")
        idx = candidate_indices(n_cand, world, rank)
        if len(idx):
            full[torch.as_tensor(idx, device=full.device)] = torch.from_numpy(a).to(full.device)
        if world > 1:
            dist.reduce(full, dst=0, op=dist.ReduceOp.SUM)   # disjoint ownership: sum == gather
        out[k] = full.cpu().numpy()
    return out if rank == 0 else None


def broadcast_bytes(payload: bytes | None, n: int, src: int = 0, device=None) -> bytes:
    """Ship a small blob (the 128-byte ncclUniqueId) from ``src`` to every rank."""
    t = torch.zeros(n, dtype=torch.uint8, device=device or "cpu")
    if dist.get_rank() == src:
        t.copy_(torch.frombuffer(bytearray(payload), dtype=torch.uint8))
    dist.broadcast(t, src=src)
    return bytes(t.cpu().numpy().tobytes())


def max_over_ranks(value: float, device=None) -> float:
    t = torch.tensor([value], dtype=torch.float64, device=device or "cpu")
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def replicate_rows(X: np.ndarray, device, group, rank: int, world: int) -> torch.Tensor:
    """A full copy of ``X`` on ``device`` of every rank, moving each row over the host link once in total:
    rank r copies rows [r * per, (r + 1) * per) (per = ceil(n / world)) from its host array and one
    all-gather completes the replicas (NCCL over NVLink on the GPU box, gloo in the CPU tests).
    Collective over ``group``; every rank must hold the same ``X``."""
    n, d = X.shape
    per = -(-n // world)
    full = torch.empty((per * world, d), dtype=torch.float32, device=device)
    lo, hi = min(n, rank * per), min(n, (rank + 1) * per)
    stage = torch.zeros((per, d), dtype=torch.float32, device=device)
    if hi > lo:
        stage[:hi - lo].copy_(torch.from_numpy(X[lo:hi]), non_blocking=True)
    dist.all_gather_into_tensor(full.view(-1), stage.view(-1), group=group)
    return full[:n]
