import os, time, torch, torch.distributed as dist
local = int(os.environ.get("LOCAL_RANK", 0)); torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
try:
    import pynvml; pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(local)
    words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
    cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
    print(rank, "nvml cpus", min(cpus), max(cpus), len(cpus), "allowed", len(os.sched_getaffinity(0)), "of", os.cpu_count(), flush=True)
    if cpus & os.sched_getaffinity(0): os.sched_setaffinity(0, cpus & os.sched_getaffinity(0))
except Exception as e:
    print(rank, "nvml failed", e, flush=True)
n = 170_000_000
hin, hout = torch.empty(n, dtype=torch.float64).pin_memory(), torch.empty(n, dtype=torch.float64).pin_memory()
hin.fill_(1.0); hout.fill_(0.0)
d1, d2 = torch.empty(n, dtype=torch.float64, device="cuda"), torch.ones(n, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(up, down, reps=4):
    torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s1): d1.copy_(hin, non_blocking=True)
        if down:
            with torch.cuda.stream(s2): hout.copy_(d2, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / reps
    return n * 8 / dt / 1e9
for name, u, d in (("H2D only", 1, 0), ("D2H only", 0, 1), ("both", 1, 1)):
    g = run(u, d); t = torch.tensor([g], device="cuda"); dist.all_reduce(t)
    if rank == 0: print(f"{world} ranks: {name}: rank0 {g:.1f} GB/s per direction, sum over ranks {float(t):.1f}", flush=True)
dist.destroy_process_group()
