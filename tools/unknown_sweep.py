"""Three-valued step (config 3 shape) over lanes-per-pattern x register layout x generations x path.
    python tools/unknown_sweep.py > gpurun_out/unknown_sweep.txt   (on the GPU box)"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from lifeapi_b200 import Engine  # noqa: E402

SEED = 0x4C696665415049


def main():
    eng = Engine(0)
    n = 1 << 18
    d_s, d_u, d_t, d_o = (eng.malloc(n * 512) for _ in range(4))
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
    eng.sync()
    p_int, _ = eng.measure_int_peak()
    print(f"int peak {p_int / 1e12:.2f} T lane-instr/s")
    for path in (0, 1):
        eng.set_three_valued_path(path)
        for layout in ((1, 2) if path == 0 else (1,)):
            eng.set_step_layout(layout)
            for lpp in (4, 8, 16, 32):
                eng.set_lanes_per_pattern(lpp)
                row = []
                for gens in (1, 16, 64, 256):
                    eng.unknown_step_batch_dev(d_s, d_u, d_t, d_o, n, gens)
                    eng.sync()
                    eng.timer_start()
                    for _ in range(3):
                        eng.unknown_step_batch_dev(d_s, d_u, d_t, d_o, n, gens)
                    ms = eng.timer_stop() / 3
                    row.append(f"g{gens}: {ms:8.3f} ms {n * 4096.0 * gens / ms / 1e9:8.1f} Tcell/s")
                print(f"path {path} layout {layout} lpp {lpp:2d}  " + "  ".join(row), flush=True)
    eng.close()


if __name__ == "__main__":
    main()
