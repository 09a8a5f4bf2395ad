#!/usr/bin/env python
"""Runs named rows of bench_extras once each, for `ncu --set full -k regex:... -c N` captures of the side kernels at
their bench shapes (the numbers printed under ncu are not bench values).
Usage: python tools/ncu_extras.py [config3 stable locate next config5 ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_extras as bx  # noqa: E402
from lifeapi_b200 import Engine  # noqa: E402

ROWS = {"config3": bx.config3_unknown_step, "config4": bx.config4_enumeration, "match": bx.step_match_explicit,
        "config5": bx.config5_survey, "stable": bx.stable_propagate, "locate": bx.locate_rows, "next": bx.next_rows}


def main():
    bx._cpu_ref = lambda: None  # no CPU legs under the profiler
    eng = Engine(0)
    try:
        for name in sys.argv[1:] or list(ROWS):
            out = ROWS[name](eng, 18.5e12)
            print(name, "ok", list(out) if isinstance(out, dict) else "")
    finally:
        eng.close()


if __name__ == "__main__":
    main()
