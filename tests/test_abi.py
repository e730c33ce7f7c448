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


def test_every_declared_symbol_has_a_ctypes_signature():
    from capital_b200 import _lib
    assert set(declared_symbols()) == set(_lib.EXPORTS)


def test_header_is_plain_c_and_links_from_c(tmp_path):
    """The boundary is a C ABI: the header must compile as C99 on its own, and a C program must link against the
    library and call it (here: the host-only layout helpers of cap_potrf_1d, which need no GPU)."""
    import shutil
    import subprocess
    if not shutil.which("gcc"):
        pytest.skip("no gcc")
    from capital_b200 import _lib
    from capital_b200.build import nccl_dirs
    src = tmp_path / "abi.c"
    src.write_text('#include "capital_b200.h"\n#include <stdio.h>\n'
                   'int main(void) {\n'
                   '  int64_t cols = 0; int blocks = 0, owner = -1, local = -1;\n'
                   '  if (cap_potrf_1d_layout(32768, 1024, 8, 3, &cols, &blocks) != CAP_OK) return 1;\n'
                   '  if (cap_potrf_1d_block(12, 8, &owner, &local) != CAP_OK) return 2;\n'
                   '  if (cap_potrf_1d_layout(-1, 1024, 8, 3, &cols, &blocks) != CAP_ERR_ARG) return 3;\n'
                   '  printf("%lld %d %d %d %s\\n", (long long)cols, blocks, owner, local, cap_last_error());\n'
                   '  return 0;\n}\n')
    exe = tmp_path / "abi"
    libdir = os.path.dirname(_lib.LIB_PATH)
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                    "-L" + libdir, "-l:libcapital_b200.so", "-Wl,-rpath," + libdir, "-Wl,-rpath," + nccl_dirs()[1],
                    "-Wl,-rpath,/usr/local/cuda/lib64", "-Wl,--allow-shlib-undefined"], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0, out.stdout + out.stderr
    cols, blocks, owner, local = out.stdout.split()[:4]
    assert (int(cols), int(blocks), int(owner), int(local)) == (4096, 4, 3, 1)      # block 12: round 1 (odd) -> rank 7 - 4 = 3
    assert "cap_potrf_1d_layout" in out.stdout
