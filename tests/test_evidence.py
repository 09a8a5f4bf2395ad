"""CPU tests that tie the numbers bench.py reads from profiles/ to the library that is actually shipped: the SASS
histogram is regenerated from lifeapi_b200/_lib/liblifeapi_b200.so (cuobjdump, no GPU) and must agree with the
committed one; no kernel may spill; the committed ncu capture the DRAM-traffic figure comes from must parse."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def fresh_mix(tmp_path_factory):
    from lifeapi_b200.build import build
    build()
    out = str(tmp_path_factory.mktemp("sass") / "mix")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sass_mix.py"), "--out", out], check=True, capture_output=True)
    return json.load(open(out + ".json"))["kernels"]


def test_committed_sass_histogram_matches_the_shipped_library(fresh_mix):
    committed = json.load(open(os.path.join(ROOT, "profiles", "r02_sass_mix.json")))["kernels"]
    for name in ("step_kernel_x<8>", "survey_kernel_x<8>", "enumerate_kernel_x", "step_match_kernel_x", "weld_step_kernel_x"):
        assert name in fresh_mix, name
        assert fresh_mix[name]["generation_loop"]["pipes"] == committed[name]["generation_loop"]["pipes"], name
    g = fresh_mix["step_kernel_x<8>"]["generation_loop"]
    assert g["ops"]["LOP3"] == 480 and g["ops"]["SHF"] == 16 and g["ops"]["SHFL"] == 32 and g["instructions"] == 531
    assert fresh_mix["step_kernel_x<8>"]["alu_instr_per_32_cells"] == 7.75
    assert fresh_mix["weld_step_kernel_x"]["alu_instr_per_32_cells"] == 12.25


def test_no_kernel_spills(fresh_mix):
    spilling = {k: v["local_memory_instructions"] for k, v in fresh_mix.items() if v["local_memory_instructions"]}
    assert not spilling, spilling


def test_bench_reads_its_roofline_inputs_from_profiles():
    sys.path.insert(0, ROOT)
    import bench
    per32, src = bench.executed_instr_per_32_cells("step_kernel_x<8>")
    assert per32 == 7.75 and src["file"] == "profiles/r02_sass_mix.json" and src["loop_instructions"] == 531
    traffic, rel = bench.ncu_dram_traffic()
    assert rel == "profiles/r02v_step_kernel_x8_raw.csv"  # the capture of the shipped tree
    assert 0.9 * 2 * (1 << 29) < traffic < 1.1 * 2 * (1 << 29)  # about 512 MiB in + 512 MiB out
