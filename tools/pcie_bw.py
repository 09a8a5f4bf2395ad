"""Concurrent pinned-memory copy bandwidth of every rank (torchrun): H2D alone, D2H alone, both at once.
Explains the N=8 e2e figure: the host pipeline needs 512 MiB each way per 59 ms step and GPU (9.1 GB/s)."""
import os, time, json
import torch, torch.distributed as dist
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
n = 512 << 20
h_in = torch.empty(n, dtype=torch.uint8).pin_memory(); h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_a = torch.empty(n, dtype=torch.uint8, device=dev); d_b = torch.empty(n, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(h2d, d2h, reps=4):
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1): d_a.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2): h_out.copy_(d_b, non_blocking=True)
    torch.cuda.synchronize()
    return n * reps / (time.perf_counter() - t0) / 1e9
res = {}
for name, a, b in (("h2d", 1, 0), ("d2h", 0, 1), ("both_each_way", 1, 1)):
    run(a, b, 1)
    v = run(a, b)
    if world > 1:
        t = torch.tensor([v], dtype=torch.float64, device=dev); ts = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(ts, t); res[name] = [round(float(x.item()), 1) for x in ts]
    else:
        res[name] = [round(v, 1)]
if rank == 0:
    print(json.dumps({"what": "GB/s per rank, all ranks copying 512 MiB pinned buffers at once", "world": world, **res}))
if world > 1: dist.destroy_process_group()
