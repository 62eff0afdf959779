"""Manual probe (not a pytest file): host->device bandwidth of each rank alone and of all ranks at once.
torchrun --nproc-per-node N tools/h2d_probe.py"""
import os, time, torch, torch.distributed as dist
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("gloo")
h = torch.empty(1 << 28, dtype=torch.float32).pin_memory()      # 1 GiB
d = torch.empty_like(h, device="cuda")
def run(tag):
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(5):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"rank {rank} {tag}: {5 * h.numel() * 4 / dt / 1e9:.1f} GB/s", flush=True)
run("warm")
for r in range(world):
    if world > 1: dist.barrier()
    if r == rank: run("alone")
if world > 1:
    dist.barrier(); run("together")
