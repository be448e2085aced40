"""CPU: the C-ABI shared library loads and exports every symbol include/ballista_b200.h declares; without a CUDA
device the product path fails loudly (no CPU fallback).  No compute calls are made here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(ROOT, "include", "ballista_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported():
    from ballista_b200 import _capi
    lib = _capi.lib()
    names = header_functions()
    assert len(names) >= 40
    for n in names:
        assert getattr(lib, n) is not None, n
    # and the ctypes binding covers the whole header (no unbound entry point)
    assert set(names) == set(lib._b2_symbols), set(names) ^ set(lib._b2_symbols)


def test_version_and_row_counts_need_no_gpu():
    from ballista_b200 import _capi
    lib = _capi.lib()
    assert lib.b2_version().decode().startswith("ballista_b200")
    assert lib.b2_tpch_num_rows(b"lineitem", C.c_double(100.0)) == 600_000_000
    assert lib.b2_tpch_num_rows(b"orders", C.c_double(1.0)) == 1_500_000
    assert lib.b2_tpch_num_rows(b"customer", C.c_double(10.0)) == 1_500_000


def test_no_cpu_fallback():
    """Without a GPU the context cannot be created and every operator is unreachable: B2_ERR_CUDA, not a CPU path."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import ballista_b200 as b2
    with pytest.raises(b2.BallistaError) as e:
        b2.Context(0)
    assert e.value.code == -5


def test_product_does_not_import_the_oracle():
    """oracle/ is test infrastructure: nothing under ballista_b200/ may reference it."""
    pkg = os.path.join(ROOT, "ballista_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".sh")):
                src = open(os.path.join(dirpath, f)).read()
                bad = re.findall(r"^\s*(?:from|import)\s+oracle|liboracle|#include\s+\"[^\"]*oracle|oracle_py|dlopen\([^)]*oracle",
                                 src, flags=re.M)
                assert not bad, (os.path.join(dirpath, f), bad)


def test_header_is_plain_c(tmp_path):
    """include/ballista_b200.h is the drop-in boundary: it must compile as C99 (what bindgen / cgo-style tools consume) and
    as C++11, with no torch / CUDA type in any signature."""
    import subprocess
    src = tmp_path / "hdr.c"
    src.write_text('#include "ballista_b200.h"\nint main(void) { return b2_version() != 0 ? 0 : 1; }\n')
    inc = os.path.join(ROOT, "include")
    for cmd in (["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-I", inc, str(src)],
                ["g++", "-std=c++11", "-Wall", "-Werror", "-fsyntax-only", "-x", "c++", "-I", inc, str(src)]):
        p = subprocess.run(cmd, capture_output=True, text=True)
        assert p.returncode == 0, p.stderr
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(inc, "ballista_b200.h")).read(), flags=re.S)  # declarations only
    assert "torch" not in text and "cudaStream_t" not in text and "#include <cuda" not in text
