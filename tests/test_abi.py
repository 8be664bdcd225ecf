"""CPU-only checks of the C ABI: the shared library loads, exports every symbol include/paffy_b200.h declares,
and refuses to work without a CUDA device (there is no CPU fallback to fall into)."""
import ctypes as C
import os
import re

import pytest

import paffy_b200
from paffy_b200 import lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "paffy_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pafgpu_\w+)\s*\(", text)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(L.ABI_SYMBOLS)


def test_library_exports_every_declared_symbol():
    lib = paffy_b200.load()
    for name in declared_symbols():
        assert hasattr(lib, name), f"libpaffy_b200.so does not export {name}"
    assert lib.pafgpu_abi_version() == 2


def test_reference_named_entry_points_are_exported():
    lib = paffy_b200.load()
    for name in L.REFERENCE_SYMBOLS:
        assert hasattr(lib, name), name
    text = open(os.path.join(ROOT, "include", "paf_b200_compat.h")).read() + open(os.path.join(ROOT, "include", "paffy_b200.h")).read()
    for name in L.REFERENCE_SYMBOLS:
        assert re.search(r"\b" + name + r"\s*\(", text), f"{name} is not declared in include/"


def test_result_struct_layout_matches_header():
    # 5 x int64 + 4 x float, padded to 8
    assert C.sizeof(L.Result) == 56 and C.sizeof(L.Batch) == 80
    assert C.sizeof(L.FileOpts) == 40 and C.sizeof(L.FileReport) == 160


def test_no_cpu_fallback():
    lib = paffy_b200.load()
    if lib.pafgpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(paffy_b200.PafGpuError) as e:
        paffy_b200.Context(0)
    assert e.value.code == -1 and "no CPU fallback" in str(e.value)


def test_product_does_not_touch_the_oracle():
    """Nothing under paffy_b200/ may import, link or execute oracle/ (the oracle is the checker only)."""
    pkg = os.path.join(ROOT, "paffy_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".c")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in text.lower().replace("the oracle", "").replace("oracle's", ""), f
