"""ctypes binding of the C ABI in include/qdsp.h (quadrants_b200/lib/libqdsp.so).

There is no CPU fallback and no alternative backend: if the CUDA library is missing or fails to load, importing the
compute entry points raises.  (The reference loads cuSPARSE/cuBLAS with dlopen and QD_ERRORs when absent,
quadrants/rhi/cuda/cuda_driver.cpp:172-258.)
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libqdsp.so")

QDSP_OK = 0
F32, F64 = 0, 1
KEY_U32, KEY_I32, KEY_F32, KEY_U64, KEY_I64, KEY_F64 = 0, 1, 2, 3, 4, 5
ELEM_I32, ELEM_U32, ELEM_F32, ELEM_I64, ELEM_U64, ELEM_F64 = 0, 1, 2, 3, 4, 5  # qdsp_elem_type
OP_ADD, OP_MIN, OP_MAX = 0, 1, 2  # qdsp_reduce_op
ALGO_REDUCE, ALGO_SCAN, ALGO_SELECT, ALGO_REDUCE_BY_KEY, ALGO_SORT = 0, 1, 2, 3, 4  # qdsp_algo
CG_EIGEN_JACOBI, CG_CUCG = 0, 1
SOLVER_TYPES = {"LLT": 0, "LDLT": 1, "LU": 2}
SOLVER_ORDERINGS = {"AMD": 0, "COLAMD": 1}

_vp, _i, _i64, _sz, _d = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_double
_P = C.POINTER

# name -> (restype, argtypes); must list every symbol include/qdsp.h declares (tests/test_abi.py checks this)
SIGNATURES = {
    "qdsp_last_error": (C.c_char_p, []),
    "qdsp_abi_version": (_i, []),
    "qdsp_init": (_i, [_i]),
    "qdsp_device_info": (_i, [_P(_i), _P(_i), _P(_i), _P(_i), _P(_sz), _P(_sz)]),
    "qdsp_launch_count": (_i64, []),
    "qdsp_release_scratch": (_i, []),
    "qdsp_scratch_info": (_i, [_P(_sz), _P(_sz)]),
    "qdsp_builder_create": (_i, [_i, _i, _i64, _i, C.c_char_p, _P(_vp)]),
    "qdsp_builder_buffer_ptr": (_i, [_vp, _P(_vp)]),
    "qdsp_builder_count": (_i, [_vp, _vp, _P(_i64)]),
    "qdsp_builder_fill_from_host": (_i, [_vp, _vp, _i64, _vp]),
    "qdsp_builder_clear": (_i, [_vp, _vp]),
    "qdsp_builder_build": (_i, [_vp, _vp, _P(_vp)]),
    "qdsp_builder_print_triplets": (_i, [_vp]),
    "qdsp_builder_destroy": (_i, [_vp]),
    "qdsp_csr_from_device_coo": (_i, [_i, _i, _i64, _vp, _vp, _vp, _i, _vp, _P(_vp)]),
    "qdsp_csr_from_device_csr": (_i, [_i, _i, _i64, _vp, _vp, _vp, _i, _i, _vp, _P(_vp)]),
    "qdsp_csr_info": (_i, [_vp, _P(_i), _P(_i), _P(_i64), _P(_i)]),
    "qdsp_csr_export": (_i, [_vp, _P(_vp), _P(_vp), _P(_vp)]),
    "qdsp_csr_spmv": (_i, [_vp, _vp, _vp, _vp]),
    "qdsp_csr_spmv_config": (_i, [_vp, _i]),
    "qdsp_csr_spmv_info": (_i, [_vp, _vp, _P(_i), _P(_i), _P(_i)]),
    "qdsp_spmv_plan_host": (_i, [_vp, _i, _vp, _i64, _P(_i64), _P(_i), _P(_i)]),
    "qdsp_csr_spmv_host": (_i, [_vp, _vp, _vp, _vp]),
    "qdsp_csr_get_element": (_i, [_vp, _i, _i, _P(_d)]),
    "qdsp_csr_add": (_i, [_vp, _vp, _d, _d, _vp, _P(_vp)]),
    "qdsp_csr_scale": (_i, [_vp, _d, _vp, _P(_vp)]),
    "qdsp_csr_transpose": (_i, [_vp, _vp, _P(_vp)]),
    "qdsp_csr_matmul": (_i, [_vp, _vp, _vp, _P(_vp)]),
    "qdsp_csr_mmwrite": (_i, [_vp, C.c_char_p]),
    "qdsp_csr_destroy": (_i, [_vp]),
    "qdsp_cg_create": (_i, [_vp, _i, _d, _i, _i, _P(_vp)]),
    "qdsp_cg_solve": (_i, [_vp, _vp, _vp, _vp, _P(_i), _P(_d), _P(_i)]),
    "qdsp_cg_solve_host": (_i, [_vp, _vp, _vp, _vp, _P(_i), _P(_d), _P(_i)]),
    "qdsp_cg_run_fixed": (_i, [_vp, _vp, _vp, _i, _vp]),
    "qdsp_cg_destroy": (_i, [_vp]),
    "qdsp_dist_plan_host": (_i, [_i, _i, _vp, _vp, _i64, _i, _vp, _i64, _P(_i64), _vp]),
    "qdsp_dist_create": (_i, [_i, _i, _P(_vp)]),
    "qdsp_dist_destroy": (_i, [_vp]),
    "qdsp_dist_localize": (_i, [_vp, _vp, _vp, _vp, _P(_i), _P(_i), _P(_i), _vp]),
    "qdsp_dist_ghost_ids": (_i, [_vp, _P(_vp)]),
    "qdsp_dist_set_send_list": (_i, [_vp, _i, _vp, _i, _i, _i]),
    "qdsp_dist_comm_handle": (_i, [_vp, _vp]),
    "qdsp_dist_attach": (_i, [_vp, _vp]),
    "qdsp_dist_nccl_unique_id": (_i, [_vp]),
    "qdsp_dist_nccl_init": (_i, [_vp, _vp]),
    "qdsp_dist_spmv": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "qdsp_cg_create_dist": (_i, [_vp, _vp, _i, _d, _i, _i, _P(_vp)]),
    "qdsp_sort_tmp_bytes": (_sz, [_i64, _i, _i]),
    "qdsp_sort_pairs": (_i, [_vp, _vp, _i64, _i, _i, _vp, _sz, _vp]),
    "qdsp_scan_tmp_bytes": (_sz, [_i64]),
    "qdsp_inclusive_scan_i32": (_i, [_vp, _i64, _vp, _sz, _vp]),
    "qdsp_scan_static_max_n": (_i64, []),
    "qdsp_inclusive_scan_i32_static": (_i, [_vp, _i64, _vp]),
    "qdsp_algo_scratch_bytes": (_sz, [_i, _i64, _i]),
    "qdsp_reduce": (_i, [_vp, _vp, _i64, _vp, _i, _i, _vp, _sz, _vp]),
    "qdsp_exclusive_scan": (_i, [_vp, _vp, _i64, _vp, _i, _i, _vp, _sz, _vp]),
    "qdsp_select": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _i, _vp, _sz, _vp]),
    "qdsp_reduce_by_key_add": (_i, [_vp, _vp, _vp, _vp, _vp, _i64, _vp, _i, _i, _vp, _sz, _vp]),
    "qdsp_sort": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _i, _i, _i, _vp, _sz, _vp]),
    "qdsp_builder_append": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp]),
    "qdsp_builder_append_elements": (_i, [_vp, _vp, _vp, _i64, _i, _vp]),
    "qdsp_vec_dot2": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _vp, _P(_d)]),
    "qdsp_vec_lincomb": (_i, [_vp, _i64, _i, _i, _P(_vp), _P(_d), _vp]),
    "qdsp_vec_cg_step": (_i, [_vp, _vp, _vp, _vp, _d, _i64, _i, _vp, _P(_d)]),
    "qdsp_solver_create": (_i, [_i, _i, _i, _P(_vp)]),
    "qdsp_solver_analyze_pattern": (_i, [_vp, _vp, _vp]),
    "qdsp_solver_factorize": (_i, [_vp, _vp, _vp, _P(_i), _P(_i)]),
    "qdsp_solver_solve": (_i, [_vp, _vp, _vp, _vp]),
    "qdsp_solver_solve_host": (_i, [_vp, _vp, _vp, _vp]),
    "qdsp_solver_info": (_i, [_vp, _P(_i)]),
    "qdsp_solver_stats": (_i, [_vp, _P(_i64), _P(_i), _P(_d)]),
    "qdsp_solver_destroy": (_i, [_vp]),
    "qdsp_solver_analyze_host": (_i, [_i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _i64, _P(_i64), _P(_i)]),
    "qdsp_solver_lu_row_matching_host": (_i, [_i, _vp, _vp, _vp, _vp, _P(_i)]),
}

_lib = None


class QdspError(RuntimeError):
    """The reference throws std::string -> Python RuntimeError (quadrants/python/py_exception_translator.cpp:15-24)."""

    def __init__(self, status, message):
        super().__init__(message)
        self.status = status


def load():
    """Loads libqdsp.so (building nothing: use quadrants_b200.build / __graft_entry__.build()).  Raises if absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA library is the only implementation of this package "
            "(no CPU fallback).  Build it with `python -m quadrants_b200.build`.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.qdsp_abi_version() != 1:
        raise ImportError("libqdsp.so ABI version mismatch")
    _lib = lib
    return lib


def check(status: int):
    if status != QDSP_OK:
        raise QdspError(status, load().qdsp_last_error().decode(errors="replace"))


def launch_count() -> int:
    return int(load().qdsp_launch_count())
