"""config 4 through the default path only (for ncu captures of enumerate_kernel_x)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench_extras as B  # noqa: E402
from lifeapi_b200 import Engine  # noqa: E402

e = Engine(0)
B._cpu_ref = lambda: None
print(B.config4_enumeration(e, 1.0)["s"])
e.close()
