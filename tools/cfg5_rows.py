"""Manual timing (not a pytest file): BASELINE.json config 5 (binary, 256 C x 10 folds, 16M x 512) ROW-SHARDED
over the GPUs of one box.  torchrun --nproc-per-node N tools/cfg5_rows.py [rows]
Every rank generates its own row slice on its GPU (device data of the named shape), the ranks share one NCCL
communicator inside the library (dml_ctx_init_comm) and each calls the same sweep on its slice."""
import os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, ".")
from cs230_distributed_machine_learning_b200 import _cabi, folds
from cs230_distributed_machine_learning_b200 import dist as ddist
from cs230_distributed_machine_learning_b200.solver import SweepSolver, DeviceDataset

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n, d, nc, cv = (int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000), 512, 256, 10
lo, hi = ddist.row_slice(n, world, rank)
gw = torch.Generator(device="cuda").manual_seed(5)
w = torch.randn(d, device="cuda", generator=gw) / d ** 0.5              # same on every rank
g = torch.Generator(device="cuda").manual_seed(1000 + rank)
X = torch.randn(hi - lo, d, device="cuda", generator=g)
y = ((X @ w + 0.5 * torch.randn(hi - lo, device="cuda", generator=g)) > 0).to(torch.int32).contiguous()
parts = [torch.empty(ddist.row_slice(n, world, r)[1] - ddist.row_slice(n, world, r)[0], dtype=torch.int32, device=dev)
         for r in range(world)]
dist.all_gather(parts, y)
fold = folds.stratified_labels(torch.cat(parts).cpu().numpy(), cv)[lo:hi]      # folds of the whole data set
del parts
s = SweepSolver(local)
uid = ddist.broadcast_bytes(_cabi.nccl_unique_id() if rank == 0 else None, 128, device=dev)
s.init_comm(world, rank, uid)
ds = DeviceDataset(X, y, None, np.array([0, 1]), hi - lo, d)
Cs = np.logspace(-6, 2, nc)
for rep in range(2):
    dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
    out = s.logreg_sweep(ds, fold, cv, Cs, [1e-4] * nc, [True] * nc, [100] * nc)
    torch.cuda.synchronize(); dt = time.perf_counter() - t0
    t = torch.tensor([dt], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        acc = (out["correct"] / out["n_test"]).mean(1)
        print(f"cfg5 rows x{world}: n={n} ({hi - lo} per rank) {t.item():.3f} s ({nc * cv / t.item():.0f} fits/s), rounds {out['rounds']}, "
              f"n_iter mean {out['n_iter'].mean():.2f}, n_test sum {int(out['n_test'][0].sum())}, best acc {acc.max():.4f}", flush=True)
s.close()
dist.destroy_process_group()
