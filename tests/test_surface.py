"""The drop-in boundary, checked name by name: every method the reference binds with nanobind for the
sparse path (tests/golden/nanobind_surface.json, extracted from quadrants/python/export_lang.cpp by
tests/golden/make_surface.py) exists on the object of the same name in quadrants_b200._core; the reference's Python
classes and functions of qd.linalg / qd.algorithms exist with the same parameter names and defaults.  CPU only."""
import inspect
import json
import os

import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden", "nanobind_surface.json")

# what is mirrored, and what is deliberately not
MIRRORED = {"SparseMatrixBuilder": "SparseMatrixBuilder", "SparseMatrix": "CuSparseMatrix",
            "CuSparseMatrix": "CuSparseMatrix", "CGf": "CGf", "CGd": "CGd", "CUCG": "CUCG",
            "CuSparseSolver": "CuSparseSolver", "SparseSolver": "CuSparseSolver"}
MIRRORED_FUNCTIONS = {"make_cucg_solver", "make_float_cg_solver", "make_double_cg_solver", "make_sparse_solver",
                      "make_cusparse_solver"}
NOT_MIRRORED_PREFIXES = ()
NOT_MIRRORED_FUNCTIONS = set()


@pytest.fixture(scope="module")
def surface():
    return json.load(open(GOLD))


def test_golden_matches_reference_when_present(surface):
    if not os.path.exists("/root/reference/quadrants/python/export_lang.cpp"):
        pytest.skip("reference tree absent")
    import subprocess
    import sys
    import tempfile

    script = os.path.join(os.path.dirname(GOLD), "make_surface.py")
    src = open(script).read().replace('os.path.join(os.path.dirname(os.path.abspath(__file__)), "nanobind_surface.json")',
                                      "os.environ['SURFACE_OUT']")
    with tempfile.TemporaryDirectory() as d:
        out = os.path.join(d, "s.json")
        subprocess.run([sys.executable, "-c", src], check=True, env={**os.environ, "SURFACE_OUT": out},
                       capture_output=True)
        assert json.load(open(out)) == surface, "tests/golden/nanobind_surface.json is stale"


def test_core_objects_carry_every_bound_method(surface):
    from quadrants_b200 import _core

    for ref_cls, methods in surface["classes"].items():
        if ref_cls.startswith(NOT_MIRRORED_PREFIXES):
            continue
        if ref_cls.startswith("EigenSparseSolver"):  # the twelve CPU instantiations: one device class stands for them
            cls = _core.CuSparseSolver
        elif ref_cls.endswith("_EigenSparseMatrix"):  # the CPU twins: one device class stands for all four
            cls = _core.CuSparseMatrix
            methods = [m for m in methods if m != "mat_vec_mul"] + ["mat_vec_mul_host"]
        else:
            cls = getattr(_core, MIRRORED[ref_cls])
        for m in methods:
            assert callable(getattr(cls, m, None)), f"{cls.__name__}.{m} missing (reference: {ref_cls}.{m})"
    for op in ("__add__", "__sub__", "__mul__", "__rmul__"):
        assert hasattr(_core.CuSparseMatrix, op)
    from quadrants_b200 import lang

    for m in surface["program_methods"]:
        assert callable(getattr(lang.Program, m, None)), f"Program.{m} missing"
    for f in surface["functions"]:
        if f in NOT_MIRRORED_FUNCTIONS:
            assert not hasattr(_core, f)
        else:
            assert f in MIRRORED_FUNCTIONS and callable(getattr(_core, f))


def test_python_layer_signatures():
    import quadrants_b200 as qd

    def params(f):
        return [(p.name, p.default) for p in inspect.signature(f).parameters.values() if p.name != "self"]

    # python/quadrants/linalg/sparse_matrix.py:219-236, sparse_cg.py:12-24, sparse_solver.py:24
    assert params(qd.linalg.SparseMatrixBuilder.__init__) == [
        ("num_rows", None), ("num_cols", None), ("max_num_triplets", 0), ("dtype", qd.f32),
        ("storage_format", "col_major")]
    p = params(qd.linalg.SparseCG.__init__)
    assert [n for n, _ in p][:5] == ["A", "b", "x0", "max_iter", "atol"]
    assert dict(p)["x0"] is None and dict(p)["max_iter"] == 50 and dict(p)["atol"] == 1e-6
    assert params(qd.linalg.SparseSolver.__init__) == [("dtype", qd.f32), ("solver_type", "LLT"), ("ordering", "AMD")]
    # python/quadrants/algorithms: names and argument lists
    A = qd.algorithms
    assert [n for n, _ in params(A.parallel_sort)] == ["keys", "values"]
    for op in ("add", "min", "max"):
        assert [n for n, _ in params(getattr(A, f"reduce_{op}"))] == ["arr", "out", "scratch", "n", "dtype", "log256_max_n"]
        assert [n for n, _ in params(getattr(A, f"exclusive_scan_{op}"))] == ["arr", "out", "scratch", "n", "dtype",
                                                                             "log256_max_n"]
    assert [n for n, _ in params(A.select)] == ["arr", "flags", "out", "num_out", "scratch", "n", "log256_max_n"]
    assert [n for n, _ in params(A.reduce_by_key_add)] == ["keys_in", "values_in", "keys_out", "values_out", "num_runs",
                                                           "scratch", "n", "value_dtype", "log256_max_n"]
    assert [n for n, _ in params(A.sort)] == ["keys", "tmp_keys", "values", "tmp_values", "scratch", "n", "key_dtype",
                                              "has_values", "end_bit", "log256_max_n"]
    assert sorted(A.__all__) == sorted([
        "PrefixSumExecutor", "exclusive_scan_add", "exclusive_scan_max", "exclusive_scan_min",
        "exclusive_scan_scratch_slots", "parallel_sort", "reduce_add", "reduce_by_key_add",
        "reduce_by_key_scratch_slots", "reduce_max", "reduce_min", "reduce_scratch_slots", "select",
        "select_scratch_slots", "sort", "sort_scratch_slots"])  # python/quadrants/algorithms/__init__.py:26-43


def test_sparse_solver_has_no_hidden_fallback():
    import quadrants_b200 as qd
    from quadrants_b200 import _core

    with pytest.raises(qd.QuadrantsRuntimeError):
        qd.linalg.SparseSolver(solver_type="QR")  # argument validation is the reference's
    for src in (inspect.getsource(qd.linalg.sparse_solver), inspect.getsource(_core.CuSparseSolver)):
        assert "SparseCG(" not in src and "scipy" not in src and "linalg.solve" not in src, "no hidden fallback"
