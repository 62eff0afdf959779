"""N > 1 host logic on CPU: world_size-2 gloo processes partition candidates and gather."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

from cs230_distributed_machine_learning_b200 import dist as ddist


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_cand = 7
    idx = ddist.candidate_indices(n_cand, world, rank)
    local = {"correct": (idx[:, None] * 10 + np.arange(5)[None, :]).astype(np.int64),
             "coef": np.tile(idx[:, None, None].astype(np.float32), (1, 5, 3))}
    full = ddist.gather_candidates(local, n_cand, world, rank, ["correct", "coef"])
    blob = ddist.broadcast_bytes(bytes(range(128)) if rank == 0 else None, 128)
    mx = ddist.max_over_ranks(float(rank + 1))
    # replicas completing each other: every rank copies its own row slice, one all-gather does the rest
    X = np.random.default_rng(3).standard_normal((1001, 7)).astype(np.float32)
    rep = ddist.replicate_rows(X, "cpu", dist.group.WORLD, rank, world)
    assert rep.shape == (1001, 7) and (rep.numpy() == X).all()
    if rank == 0:
        q.put((full["correct"].tolist(), full["coef"][:, 0, 0].tolist(), list(blob) == list(range(128)), mx))
    else:
        q.put((None, None, list(blob) == list(range(128)), mx))
    dist.destroy_process_group()


def test_candidate_partition_and_gather_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    full = next(g for g in got if g[0] is not None)
    assert full[0] == [[c * 10 + k for k in range(5)] for c in range(7)]
    assert full[1] == [float(c) for c in range(7)]
    assert all(g[2] for g in got) and all(g[3] == 2.0 for g in got)


def test_row_slices_tile_the_rows():
    for n, w in [(1_000_000, 8), (1000, 3), (16_000_000, 4), (130, 2)]:
        cuts = [ddist.row_slice(n, w, r) for r in range(w)]
        assert cuts[0][0] == 0 and cuts[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
        assert all(lo % 128 == 0 for lo, _ in cuts)
