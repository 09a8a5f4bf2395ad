"""config 4 (collision enumeration) through enumerate_kernel (layout 2) and enumerate_kernel_x (default)."""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench_extras as B  # noqa: E402
from lifeapi_b200 import Engine  # noqa: E402

e = Engine(0)
p, _ = e.measure_int_peak()
for lay in (2, 0):
    e.set_step_layout(lay)
    r = B.config4_enumeration(e, p)
    r.pop("cpu_baseline", None)
    print("layout", lay, json.dumps(r))
e.close()
