#!/usr/bin/env python
"""One launch each of the secondary kernels at their bench shapes, for `ncu --set full -k regex:...` captures:
config 4 enumeration (enumerate_kernel_x), LifeWeld::Step x 64 (weld_step_kernel*), the explicit-pattern fused match
(step_match_kernel*), config 3 three-valued step (unknown_step_kernel*).  Usage: python tools/ncu_targets.py [names...]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench_extras import D4, GLIDER, SEED, STILL_LIFES, _pattern  # noqa: E402
from lifeapi_b200 import Engine  # noqa: E402


def main():
    which = set(sys.argv[1:]) or {"enumerate", "weld", "match", "unknown"}
    eng = Engine(0)
    names = list(STILL_LIFES)
    cats = np.stack([_pattern(STILL_LIFES[names[i % 8]], 28, 28, D4[(i // 8) * 2]) for i in range(32)])
    g0 = np.stack([_pattern(GLIDER, 8, 8, tf) for tf in D4])
    phases = [g0]
    for _ in range(3):
        phases.append(eng.step_batch_host(phases[-1], 1))
    gliders = np.ascontiguousarray(np.stack(phases, axis=1).reshape(32, 64))
    if "enumerate" in which:
        hits, skipped = eng.enumerate_collisions(cats, gliders, 0, 32 * 32 * 4096, 256, 8, 1 << 20)
        print("enumerate", len(hits), skipped)
    if "weld" in which:
        n = 1 << 17
        d = eng.malloc(n * 2048)
        eng.soup_dev(d, 0, 4 * n, SEED, 0)
        recs = np.empty((n, 4, 64), dtype=np.uint64)
        eng.d2h(recs, d)
        eng.sync()
        recs[:, 1] &= recs[:, 2] & recs[:, 3]
        recs[:, 2] &= recs[:, 3]
        eng.h2d(d, recs)
        eng._check(eng.lib.lifeb200_weld_step_batch_dev(eng._h, d, n, 64))
        eng.sync()
        eng.free(d)
        print("weld ok")
    if "match" in which:
        # 256 Ki explicit placements of catalyst 0..31 x glider variants, one target per catalyst
        n = 1 << 18
        rng = np.random.default_rng(1)
        ids = rng.integers(0, 32 * 32 * 4096, size=n)
        d_p = eng.malloc(n * 512)
        eng.soup_dev(d_p, 0, n, SEED, 1)  # contents do not matter for the profile: sparse-ish soups against still-life targets
        tg = np.zeros((32, 2, 64), dtype=np.uint64)
        for c in range(32):
            z = np.zeros(64, dtype=np.uint64)
            for dx in (-1, 0, 1):
                col = np.roll(cats[c], dx)
                z |= col | (col << np.uint64(1)) | (col >> np.uint64(63)) | (col >> np.uint64(1)) | (col << np.uint64(63))
            tg[c, 0], tg[c, 1] = cats[c], z & ~cats[c]
        d_t, d_i, d_h, d_n = eng.malloc(tg.nbytes), eng.malloc(n * 4), eng.malloc(n * 16), eng.malloc(8)
        eng.h2d(d_t, tg)
        eng.h2d(d_i, (ids >> 17).astype(np.uint32))
        eng.step_match_dev(d_p, d_t, d_i, n, 0, 256, 8, d_h, n, d_n)
        eng.sync()
        print("match ok")
    if "unknown" in which:
        n = 1 << 18
        d_s, d_u, d_a, d_b = (eng.malloc(n * 512) for _ in range(4))
        eng.soup_dev(d_s, 0, n, SEED, 0)
        host = np.empty((n, 64), dtype=np.uint64)
        eng.d2h(host, d_s)
        eng.sync()
        box = np.zeros(64, dtype=np.uint64)
        box[16:48] = np.uint64(0x0000FFFFFFFF0000)
        unk = (np.roll(host, 1, axis=0) & np.roll(host, 2, axis=0) & np.roll(host, 3, axis=0)) & box
        st = (host & box) & ~unk
        eng.h2d(d_s, st)
        eng.h2d(d_u, unk)
        eng.unknown_step_batch_dev(d_s, d_u, d_a, d_b, n, 16)
        eng.sync()
        print("unknown ok")
    eng.close()


if __name__ == "__main__":
    main()
