"""Manual probe: cudaHostRegister'ed numpy memory -> device, host-side call time and throughput, per rank."""
import os, time, numpy as np, torch, torch.distributed as dist
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
X = np.random.default_rng(0).standard_normal((1_000_000, 256), dtype=np.float32)
rc = torch.cuda.cudart().cudaHostRegister(X.ctypes.data, X.nbytes, 0)
t = torch.from_numpy(X)
print(f"rank {rank}: register rc={int(rc)} is_pinned={t.is_pinned()} OMP={os.environ.get('OMP_NUM_THREADS')}", flush=True)
s = torch.cuda.Stream()
for rep in range(3):
    if world > 1: dist.barrier()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    with torch.cuda.stream(s):
        d = t.to("cuda", non_blocking=True)
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"rank {rank} rep {rep}: call {1e3*(t1-t0):.1f} ms, total {1e3*(t2-t0):.1f} ms = {X.nbytes/(t2-t0)/1e9:.1f} GB/s", flush=True)
    del d
