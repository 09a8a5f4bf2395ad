"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol the public
header declares, the Python stub covers all of them, and compute calls FAIL LOUDLY without a GPU
(no CPU fallback anywhere in the product path)."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from lifeapi_b200.build import build
    from lifeapi_b200.capi import load_library
    build()
    return load_library()


def test_header_symbols_exported(lib):
    from lifeapi_b200.capi import SIGNATURES, declared_functions
    names = declared_functions()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/lifeapi_b200.h but not exported"
    assert set(names) == set(SIGNATURES), set(names) ^ set(SIGNATURES)


def test_abi_version_and_struct_layout(lib):
    from lifeapi_b200.capi import HIT_DTYPE, Hit
    assert lib.lifeb200_abi_version() >= 1
    assert C.sizeof(Hit) == 16 == HIT_DTYPE.itemsize


def test_product_does_not_touch_the_oracle():
    """Nothing under lifeapi_b200/ or include/ may import, link or name oracle/."""
    bad = []
    for base in ("lifeapi_b200", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                    text = open(os.path.join(dp, f), errors="ignore").read()
                    if re.search(r"(from|import)\s+oracle|oracle/|liblifeoracle|libliferef", text):
                        if f == "__init__.py" and "imports oracle/" in text:
                            continue
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_library_has_sm100a_code_only():
    so = os.path.join(ROOT, "lifeapi_b200", "_lib", "liblifeapi_b200.so")
    out = subprocess.run(["cuobjdump", "-lelf", so], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_no_cpu_fallback_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu tests")
    from lifeapi_b200 import E_CUDA, E_NODEVICE, Engine, LifeB200Error
    with pytest.raises(LifeB200Error) as ei:
        Engine(0)
    assert ei.value.code in (E_NODEVICE, E_CUDA)
    from lifeapi_b200 import MultiEngine
    with pytest.raises(LifeB200Error) as ei:
        MultiEngine([0])  # the multi-GPU group has no fallback either
    assert ei.value.code in (E_NODEVICE, E_CUDA)
    # null-context compute call is an argument error, never a silent success
    rc = lib.lifeb200_step_batch_dev(None, None, None, 1, 1)
    assert rc != 0 and lib.lifeb200_last_error()
