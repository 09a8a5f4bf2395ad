"""Where do the 4-patterns-per-lane-group kernels (layout 8) overtake the lane-group ones (layout 2) for the soup survey
and the collision enumeration?  Forced layouts over short generation counts."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench_extras as B
from lifeapi_b200 import Engine
eng = Engine(0)
n = 1 << 19
for gens in (8, 16, 24, 32, 48, 64, 96, 128):
    row = []
    for layout in (2, 8):
        eng.set_step_layout(layout)
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter(); eng.soup_survey(0, n, B.SEED, 0, gens, n); best = min(best, time.perf_counter() - t0)
        row.append(f"layout {layout}: {best*1e3:8.3f} ms")
    print(f"survey    G={gens:4d}  " + "   ".join(row))
names = list(B.STILL_LIFES)
cats = np.stack([B._pattern(B.STILL_LIFES[names[i % 8]], 28, 28, B.D4[(i // 8) * 2]) for i in range(32)])
g0 = np.stack([B._pattern(B.GLIDER, 8, 8, tf) for tf in B.D4])
phases = [g0]
eng.set_step_layout(0)
for _ in range(3):
    phases.append(eng.step_batch_host(phases[-1], 1))
gliders = np.ascontiguousarray(np.stack(phases, axis=1).reshape(32, 64))
total = 32 * 32 * 4096 // 4
for gens in (8, 16, 24, 32, 48, 64, 96):
    row = []
    for layout in (2, 8):
        eng.set_step_layout(layout)
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter(); eng.enumerate_collisions(cats, gliders, 0, total, gens, 8, 1 << 20); best = min(best, time.perf_counter() - t0)
        row.append(f"layout {layout}: {best*1e3:8.3f} ms")
    print(f"enumerate G={gens:4d}  " + "   ".join(row))
eng.close()
