set -x
python -m pytest tests/test_next_gpu.py -m gpu -x -q 2>&1 | tail -3
python - <<'PY' 2>&1 | tee gpurun_out/r01o_interaction.txt
import numpy as np
from lifeapi_b200 import Engine
from bench_extras import _time, SEED
eng = Engine(0)
for n in (1 << 18, 1 << 20):
    d_s, d_p = eng.malloc(n * 512), eng.malloc(n * 2048)
    eng.soup_dev(d_s, 0, n, SEED, 0)
    for lpp in (0, 8, 16, 32):
        eng.set_lanes_per_pattern(lpp)
        ms = _time(eng, lambda: eng._check(eng.lib.lifeb200_interaction_counts_dev(eng._h, d_s, d_p, n)), reps=10)
        print(f"interaction_counts n={n} lpp={lpp or 'warp-persistent'}: {ms*1e3:.1f} us  {n*2560/(ms*1e-3)/1e9:.0f} GB/s")
    eng.set_lanes_per_pattern(0)
    eng.free(d_s); eng.free(d_p)
eng.close()
PY
