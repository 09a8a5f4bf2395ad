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


MULTI_BIN = os.path.join(ROOT, "tests", "cpp", "_build", "multi_survey")


def build_multi_survey():
    from lifeapi_b200.build import build
    so = build()
    src = os.path.join(ROOT, "tests", "cpp", "multi_survey.cpp")
    os.makedirs(os.path.dirname(MULTI_BIN), exist_ok=True)
    if not os.path.exists(MULTI_BIN) or os.path.getmtime(MULTI_BIN) < max(os.path.getmtime(p) for p in (
            src, so, os.path.join(ROOT, "include", "LifeAPI_b200.hpp"))):
        subprocess.run(["g++", "-std=c++20", "-O2", "-Wall", "-Wextra", "-I" + os.path.join(ROOT, "include"), src, "-o", MULTI_BIN,
                        so, "-Wl,-rpath," + os.path.dirname(so)], check=True)
    return MULTI_BIN


def test_multi_gpu_cpp_program_builds():
    assert os.path.exists(build_multi_survey())


@pytest.mark.gpu
def test_multi_gpu_cpp_program_matches_single_gpu():
    """A C++ program (no Python in the data path) surveys soups on every visible GPU through lifeb200_multi_* and gets
    the byte-identical hit list a single context returns; on a one-GPU box the group has one rank."""
    r = subprocess.run([build_multi_survey(), str(200003), "300"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "identical" in r.stdout and "all passed" in r.stdout


REF_TEST = "/root/reference/tests/StepAltTest.cpp"
REF_BIN = os.path.join(ROOT, "tests", "cpp", "_build", "reference_StepAltTest")


def build_reference_test():
    """The reference's OWN tests/StepAltTest.cpp, body verbatim: the file is read where it lies and piped to g++
    with only its `#include "../LifeAPI.hpp"` line pointed at the drop-in header (nothing is copied into the repo);
    <gtest/gtest.h> resolves to tests/cpp/gtest_shim.  Built only where /root/reference exists; the binary travels
    to the GPU box in tests/cpp/_build/ like the other prebuilt checkers."""
    from lifeapi_b200.build import build
    so = build()
    if not os.path.exists(REF_TEST):
        return REF_BIN if os.path.exists(REF_BIN) else None
    os.makedirs(os.path.dirname(REF_BIN), exist_ok=True)
    hdr = os.path.join(ROOT, "include", "LifeAPI_b200.hpp")
    shim = os.path.join(ROOT, "tests", "cpp", "gtest_shim")
    if not os.path.exists(REF_BIN) or os.path.getmtime(REF_BIN) < max(os.path.getmtime(p) for p in (
            REF_TEST, so, hdr, os.path.join(shim, "gtest", "gtest.h"))):
        text = open(REF_TEST).read()
        assert '#include "../LifeAPI.hpp"' in text
        text = text.replace('#include "../LifeAPI.hpp"', '#include "LifeAPI_b200.hpp"')
        subprocess.run(["g++", "-std=c++20", "-O2", "-Wall", "-Wextra", "-I" + os.path.join(ROOT, "include"),
                        "-I" + shim, "-o", REF_BIN, "-x", "c++", "-", "-x", "none", so, "-Wl,-rpath," + os.path.dirname(so)], input=text, text=True, check=True)
    return REF_BIN


def test_reference_own_step_alt_test_compiles_against_the_drop_in_header():
    if not os.path.exists(REF_TEST):
        pytest.skip("needs /root/reference (the binary is prebuilt for the GPU box)")
    import torch
    exe = build_reference_test()
    assert exe and os.path.exists(exe)
    if not torch.cuda.is_available():
        r = subprocess.run([exe], capture_output=True, text=True)
        assert r.returncode != 0 and "lifeb200_create" in r.stderr  # no GPU: abort, never a silent CPU path


@pytest.mark.gpu
def test_reference_own_step_alt_test_passes_on_gpu():
    """tests/StepAltTest.cpp:5-13 verbatim: 10000 RandomState soups, Step() == StepAlt() -- here two different GPU
    circuits (the step kernel's pair-shared select against the interaction kernel's two-full-adder form)."""
    exe = build_reference_test()
    if exe is None:
        pytest.skip("tests/cpp/_build/reference_StepAltTest was not prebuilt (needs /root/reference at build time)")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "all passed" in r.stdout and "StepAltTest.Random" in r.stdout


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
