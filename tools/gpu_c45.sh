python -m pytest tests/test_match_gpu.py -m gpu -x -q 2>&1 | tail -2
python - <<'PY' 2>&1 | tee gpurun_out/r01v_config45.txt
import json
from lifeapi_b200 import Engine
import bench_extras
eng = Engine(0)
for _ in range(2):
    r4 = bench_extras.config4_enumeration(eng, 18.5e12); r4.pop('cpu_baseline', None)
    print(json.dumps(r4))
    print(json.dumps(bench_extras.config5_survey(eng, 18.5e12)))
eng.close()
PY
