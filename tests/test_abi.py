"""CPU-only checks of the drop-in boundary: the C-ABI library builds, loads and exports every symbol include/qdsp.h
declares; the ctypes table covers all of them; without a GPU the product path fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib_path():
    from quadrants_b200.build import build_library

    return build_library()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "qdsp.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(qdsp_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(lib_path):
    lib = ctypes.CDLL(lib_path)
    syms = declared_symbols()
    assert len(syms) >= 35
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/qdsp.h but not exported by libqdsp.so"


def test_ctypes_table_matches_header(lib_path):
    from quadrants_b200 import _lib

    assert sorted(_lib.SIGNATURES) == declared_symbols()
    lib = _lib.load()
    assert lib.qdsp_abi_version() == 1


def test_only_the_abi_is_exported(lib_path):
    import subprocess

    out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True).stdout
    names = [l.split()[-1] for l in out.splitlines() if " T " in l]
    assert names and all(n.startswith("qdsp_") for n in names), names


def test_sm100a_only(lib_path):
    import subprocess

    out = subprocess.run(["cuobjdump", "-lelf", lib_path], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_pure_functions_without_gpu(lib_path):
    from quadrants_b200 import _lib

    lib = _lib.load()
    assert lib.qdsp_scan_tmp_bytes(1 << 20) >= 8 * (1 << 20) // 16384  # one 8-byte descriptor per 16384-element tile
    n = 1 << 20
    assert lib.qdsp_sort_tmp_bytes(n, _lib.KEY_I32, 4) >= 8 * n
    assert lib.qdsp_sort_tmp_bytes(n, _lib.KEY_U64, 8) >= 16 * n
    assert lib.qdsp_launch_count() == 0 or lib.qdsp_launch_count() > 0


def test_fails_loudly_without_gpu(lib_path):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import quadrants_b200 as qd
    from quadrants_b200 import _lib

    with pytest.raises(qd.QuadrantsRuntimeError):
        qd.init(arch=qd.cuda)
    with pytest.raises(qd.QuadrantsRuntimeError):
        qd.init(arch=qd.cpu)  # no other backend either
    lib = _lib.load()
    assert lib.qdsp_init(0) != 0 and b"CUDA" in lib.qdsp_last_error() or b"cuda" in lib.qdsp_last_error()
    h = ctypes.c_void_p()
    st = lib.qdsp_builder_create(4, 4, 10, 0, b"col_major", ctypes.byref(h))
    assert st != 0 and not h.value
    with pytest.raises(qd.QuadrantsRuntimeError):
        qd.linalg.SparseMatrixBuilder(4, 4, max_num_triplets=10)


def test_product_does_not_touch_the_oracle():
    """The product package must not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "quadrants_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "qd_oracle" not in text and "eigen_ref" not in text, f


def test_product_does_not_call_vendor_sparse_libraries(lib_path):
    """cuSPARSE / cuBLAS are the reference's path (the kernel to beat, tools/bench_cusparse.py); the product neither
    links nor loads them."""
    import subprocess

    needed = subprocess.run(["readelf", "-d", lib_path], capture_output=True, text=True).stdout
    assert "cusparse" not in needed and "cublas" not in needed
    pkg = os.path.join(ROOT, "quadrants_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")) and f != "build.py":
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"#include\s*<(cusparse|cublas)", text), f
                assert "libcusparse" not in text and "libcublas" not in text, f
