python -m pytest tests/test_locate_gpu.py -m gpu -x -q 2>&1 | tail -2
python - <<'PY' 2>&1 | tee gpurun_out/r01t_locate.txt
import json
from lifeapi_b200 import Engine
import bench_extras
eng = Engine(0)
r = bench_extras.locate_rows(eng, 18.5e12)
print(json.dumps({k: {kk: vv for kk, vv in v.items() if kk != 'cpu_baseline'} for k, v in r.items()}))
eng.close()
PY
