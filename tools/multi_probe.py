#!/usr/bin/env python
"""Where does the gather time of lifeb200_multi_soup_survey go?  Runs under torchrun (one process per GPU) or alone
(all GPUs in this process): several survey calls of growing size, printing wall time, the device-timed gather and
the number of hit records; then the same with gather mode 1 (direct downloads) in single-process mode."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lifeapi_b200 import HIT_DTYPE, MultiEngine  # noqa: E402

SEED = 0x4C696665415049
rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
if world > 1:
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    uid = [MultiEngine.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    g = MultiEngine(device=int(os.environ["LOCAL_RANK"]), rank=rank, world=world, unique_id=uid[0])
    bar = dist.barrier
else:
    g = MultiEngine()
    bar = lambda: None  # noqa: E731
info = g.info()
hits = np.zeros(1 << 22, dtype=HIT_DTYPE)
for mode in ((0, 1) if info["n_local"] == info["world"] and info["world"] > 1 else (0,)):
    g.set_gather(mode)
    for n, gens in ((1 << 14, 16), (1 << 18, 16), (1 << 20, 16), (1 << 22, 16), (1 << 22, 16), (1 << 24, 64), (1 << 24, 64), (1 << 24, 4096)):
        bar()
        t0 = time.perf_counter()
        h, ps, hx = g.soup_survey(0, n, SEED, 0, gens, len(hits), hits=hits)
        dt = time.perf_counter() - t0
        if rank == 0:
            print(f"world {info['world']} local {info['n_local']} nccl {info['nccl_version']} mode {mode} soups {n} gens {gens}: "
                  f"{dt * 1e3:9.3f} ms wall, gather {g.last_gather_ms():8.3f} ms, {len(h)} hits ({len(h) * 16 / 1e6:.2f} MB)", flush=True)
g.close()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
