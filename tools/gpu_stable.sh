set -x
python -m pytest tests/test_circuits_gpu.py -m gpu -x -q 2>&1 | tail -3
python - <<'PY' 2>&1 | tee gpurun_out/r01p_stable.txt
import json
from lifeapi_b200 import Engine
import bench_extras
eng = Engine(0)
print(json.dumps(bench_extras.stable_propagate(eng, 18.5e12)))
eng.close()
PY
