#!/usr/bin/env python
"""lifeb200_step_batch_host on the BASELINE batch (1 Mi soups x 1024 generations) from three kinds of host memory:
lifeb200_host_alloc (pinned), a caller-owned numpy array page-locked with lifeb200_host_register, plain pageable."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lifeapi_b200 import Engine  # noqa: E402

SEED = 0x4C696665415049
eng = Engine(0)
n, gens = 1 << 20, 1024
d = eng.malloc(n * 512)
eng.soup_dev(d, 0, n, SEED, 0)
pin_in, pin_out = eng.host_alloc((n, 64)), eng.host_alloc((n, 64))
eng.d2h(pin_in, d)
eng.sync()
own_in, own_out = pin_in.copy(), np.empty((n, 64), dtype=np.uint64)
own_out[:] = 0  # touch the pages


def timed(src, dst, reps=3):
    eng.step_batch_host(src, gens, out=dst)
    t0 = time.perf_counter()
    for _ in range(reps):
        eng.step_batch_host(src, gens, out=dst)
    return (time.perf_counter() - t0) / reps * 1e3


ms_pin = timed(pin_in, pin_out)
ms_page = timed(own_in, own_out)
t0 = time.perf_counter()
eng.host_register(own_in)
eng.host_register(own_out)
ms_reg = (time.perf_counter() - t0) * 1e3
ms_own = timed(own_in, own_out)
eng.host_unregister(own_in)
eng.host_unregister(own_out)
assert (own_out == pin_out).all()
cells = n * 4096.0 * gens
for name, ms in (("lifeb200_host_alloc (pinned)", ms_pin), ("pageable numpy", ms_page), ("numpy + lifeb200_host_register", ms_own)):
    print(f"{name:34s} {ms:8.2f} ms per step = {cells / ms / 1e9:8.1f} Tcell/s")
print(f"registering 2 x 512 MiB took {ms_reg:.1f} ms (once)")
