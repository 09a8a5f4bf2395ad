"""The C++ drop-in header (include/LifeAPI_b200.hpp): compiles against the C ABI on CPU; on the GPU
box the reference-style test binary (tests/cpp/step_alt_test.cpp) must pass."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "_build", "step_alt_test")


def build_binary():
    from lifeapi_b200.build import build
    from oracle.loader import build_port
    so = build()
    port = build_port()
    os.makedirs(os.path.dirname(BIN), exist_ok=True)
    src = os.path.join(ROOT, "tests", "cpp", "step_alt_test.cpp")
    if not os.path.exists(BIN) or os.path.getmtime(BIN) < max(os.path.getmtime(p) for p in (
            src, so, port, os.path.join(ROOT, "include", "LifeAPI_b200.hpp"))):
        subprocess.run(["g++", "-std=c++20", "-O2", "-Wall", "-Wextra", "-I" + os.path.join(ROOT, "include"), src, "-o", BIN,
                        so, port, "-Wl,-rpath," + os.path.dirname(so), "-Wl,-rpath," + os.path.dirname(port)], check=True)
    return BIN


def test_header_compiles_and_fails_loudly_without_gpu():
    import torch
    exe = build_binary()
    if torch.cuda.is_available():
        pytest.skip("GPU present: the gpu test runs the binary")
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode != 0 and "lifeb200_create" in r.stderr  # abort, never a silent CPU path


@pytest.mark.gpu
def test_reference_style_cpp_test_passes_on_gpu():
    r = subprocess.run([build_binary()], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "all passed" in r.stdout
