"""CPU: the C-ABI library loads and exports every symbol include/gvss.h declares; nothing is computed here."""
import os
import re
import ctypes
import pytest
from duckdb_vss_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    text = open(os.path.join(ROOT, "include", "gvss.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gvss_[a-z_0-9]+)\s*\(", text)))


def test_header_declares_what_the_binding_binds():
    assert header_functions() == sorted(capi.SIGNATURES)


def test_library_exports_every_declared_symbol():
    L = ctypes.CDLL(capi.LIB_PATH)
    for name in header_functions():
        assert hasattr(L, name), name


def test_every_declaration_cites_the_reference():
    text = open(os.path.join(ROOT, "include", "gvss.h")).read()
    for needle in ("hnsw_index.cpp", "index_dense.hpp", "index.hpp", "hnsw_index_physical_create.cpp", "array_kernels.hpp"):
        assert needle in text


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    cfg = capi.Config(4, b"e", 0, 0, 0, 0, -1)
    h = ctypes.c_void_p()
    rc = capi.lib().gvss_create(ctypes.byref(cfg), ctypes.byref(h))
    assert rc != 0 and b"no CUDA device" in capi.lib().gvss_last_error()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "duckdb_vss_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("hnsw_oracle", "oracle") or "import oracle" not in src, f
                assert "from oracle" not in src and "import oracle" not in src and "hnsw_oracle" not in src, f
