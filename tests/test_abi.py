"""CPU-side checks of the drop-in boundary: libtitan_b200.so loads and exports every symbol include/titan_b200.h
declares, and the product path fails loudly (no CPU fallback) when no CUDA device is present."""
import ctypes
import os
import re

import pytest

import titan_b200 as tb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "titan_b200.h")).read()
    return sorted(set(re.findall(r"^TB_API[^;(]*?\b(tb_[a-z0-9_]+)\(", text, flags=re.M)))


def test_header_declares_what_the_binding_lists():
    assert declared_symbols() == sorted(tb.EXPORTS)


def test_library_loads_and_exports_every_declared_symbol():
    L = ctypes.CDLL(tb.LIB_PATH)
    for name in declared_symbols():
        assert hasattr(L, name), name
    assert L.tb_abi_version() == 1


def test_no_cpu_fallback_without_a_device():
    if tb.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(tb.TitanError) as ei:
        tb.Engine(0)
    assert ei.value.status == 2 and "no CPU path" in str(ei.value)


def test_product_package_does_not_touch_the_oracle():
    # the oracle is test infrastructure: nothing under titan_b200/ may import, link or call it
    for dirpath, _, files in os.walk(os.path.join(ROOT, "titan_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "titan_oracle" not in text and "import oracle" not in text and "from oracle" not in text, f
