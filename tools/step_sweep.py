"""LifeState::Step(n) throughput over generations per launch (1 Mi resident soups)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from lifeapi_b200 import Engine
from bench_extras import _time, SEED
eng = Engine(0)
n = 1 << 20
d_in, d_out = eng.malloc(n * 512), eng.malloc(n * 512)
eng.soup_dev(d_in, 0, n, SEED, 0)
for g in (1, 2, 4, 8, 16, 32, 48, 64, 96, 128, 256, 1024):
    ms = _time(eng, lambda: eng.step_batch_dev(d_in, d_out, n, g), reps=5)
    print(f"G={g:5d}  {ms:9.4f} ms  {n*4096.0*g/(ms*1e-3)/1e12:7.2f} Tcell/s  {2*n*512/(ms*1e-3)/1e9:7.0f} GB/s")
eng.close()
# layout crossover: forced register layouts (1 = reference, 2 = even | odd rows, 8 = row classes) over short runs
eng = Engine(0)
d_in, d_out = eng.malloc(n * 512), eng.malloc(n * 512)
eng.soup_dev(d_in, 0, n, SEED, 0)
for g in (8, 16, 24, 32, 40, 48, 64, 96, 128):
    row = []
    for layout in (1, 2, 8):
        eng.set_step_layout(layout)
        ms = _time(eng, lambda: eng.step_batch_dev(d_in, d_out, n, g), reps=5)
        row.append(f"layout {layout}: {ms:8.4f} ms")
    print(f"G={g:5d}  " + "   ".join(row))
eng.close()
