#!/usr/bin/env python
"""Measured crossovers of the round-2 layout-8 kernels against the lane-group kernels they replace:
LifeWeld::Step (weld_step_kernel_x vs weld_step_kernel<8>) and the explicit-pattern match (step_match_kernel_x vs
step_match_kernel<8>), by generations per launch.  Prints one line per (kernel, generations)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench_extras import D4, GLIDER, SEED, STILL_LIFES, _pattern, _time  # noqa: E402
from lifeapi_b200 import Engine  # noqa: E402

eng = Engine(0)
n = 1 << 17
d = eng.malloc(n * 2048)
eng.soup_dev(d, 0, 4 * n, SEED, 0)
recs = np.empty((n, 4, 64), dtype=np.uint64)
eng.d2h(recs, d)
eng.sync()
recs[:, 1] &= recs[:, 2] & recs[:, 3]
recs[:, 2] &= recs[:, 3]
eng.h2d(d, recs)
eng.sync()
for gens in (4, 8, 12, 16, 20, 24, 32, 48, 64, 128):
    ms = {}
    for name, layout in (("layout8", 8), ("lanes", 1)):
        eng.set_step_layout(layout)
        ms[name] = _time(eng, lambda: eng._check(eng.lib.lifeb200_weld_step_batch_dev(eng._h, d, n, gens)), reps=5)
    eng.set_step_layout(0)
    print(f"weld_step  gens {gens:4d}: layout 8 {ms['layout8']:8.4f} ms  lane groups {ms['lanes']:8.4f} ms  ratio {ms['lanes'] / ms['layout8']:.3f}", flush=True)
eng.free(d)

names = list(STILL_LIFES)
cats = np.stack([_pattern(STILL_LIFES[names[i % 8]], 28, 28, D4[(i // 8) * 2]) for i in range(32)])
n = 1 << 18
rng = np.random.default_rng(4)
c_idx = rng.integers(0, 32, size=n)
gl = np.stack([_pattern(GLIDER, int(x), int(y), D4[int(t)]) for x, y, t in zip(rng.integers(0, 64, 512), rng.integers(0, 64, 512), rng.integers(0, 8, 512))])
pats = np.ascontiguousarray(cats[c_idx] | gl[rng.integers(0, 512, size=n)])
tg = np.zeros((32, 2, 64), dtype=np.uint64)
for c in range(32):
    z = np.zeros(64, dtype=np.uint64)
    for dx in (-1, 0, 1):
        col = np.roll(cats[c], dx)
        z |= col | (col << np.uint64(1)) | (col >> np.uint64(63)) | (col >> np.uint64(1)) | (col << np.uint64(63))
    tg[c, 0], tg[c, 1] = cats[c], z & ~cats[c]
d_p, d_t, d_i, d_h, d_n = eng.malloc(pats.nbytes), eng.malloc(tg.nbytes), eng.malloc(n * 4), eng.malloc(n * 16), eng.malloc(8)
eng.h2d(d_p, pats)
eng.h2d(d_t, tg)
eng.h2d(d_i, c_idx.astype(np.uint32))
eng.sync()
for gens in (8, 16, 24, 32, 48, 64, 128, 256):
    ms = {}
    for name, layout in (("layout8", 8), ("lanes", 2)):
        eng.set_step_layout(layout)
        ms[name] = _time(eng, lambda: eng.step_match_dev(d_p, d_t, d_i, n, 0, gens, 8, d_h, n, d_n), reps=5)
    eng.set_step_layout(0)
    print(f"step_match gens {gens:4d}: layout 8 {ms['layout8']:8.4f} ms  lane groups {ms['lanes']:8.4f} ms  ratio {ms['lanes'] / ms['layout8']:.3f}", flush=True)
eng.close()
