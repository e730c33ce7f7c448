"""CPU: the C-ABI library loads and exports every symbol include/capital_b200.h declares; without a GPU
the product path fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "capital_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cap_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported():
    from capital_b200 import _lib
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 40
    for name in names:
        assert hasattr(lib, name), "missing export " + name
    assert set(_lib.EXPORTS) <= set(names)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from capital_b200 import _lib
    lib = _lib.load()
    h = ctypes.c_void_p()
    assert lib.cap_create(ctypes.byref(h), 0) != 0
    assert b"no CUDA device" in lib.cap_last_error()
    from capital_b200.engine import Engine
    with pytest.raises(_lib.CapitalError):
        Engine(0)


def test_product_does_not_import_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "capital_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                assert "capital_oracle" not in open(os.path.join(dirpath, f)).read(), f
