"""ctypes binding of libeigsol_b200.so (C ABI: include/eigsol_b200.h).

This is the Python-side stub INTEGRATION.md shows: plain pointers and sizes, no torch types.  The
library is built in-tree by ``__graft_entry__.build()`` / ``csrc/Makefile``; importing this module
fails loudly if it is missing -- there is no fallback implementation of the path.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libeigsol_b200.so")

OK, ERR_INVALID, ERR_NOT_SQUARE, ERR_ZERO_SIZE, ERR_DEVICE, ERR_STATE, ERR_IO = range(7)
F64, C128 = 0, 1
KERNEL_AUTO, KERNEL_CSR_SCALAR, KERNEL_CSR_VECTOR, KERNEL_CSR_STREAM, KERNEL_SELL, KERNEL_GEMV, KERNEL_CSR_WINDOW = range(7)
KERNEL_NAMES = {0: "auto", 1: "csr_scalar", 2: "csr_vector", 3: "csr_stream", 4: "sell", 5: "gemv", 6: "csr_window"}

# every symbol include/eigsol_b200.h declares (tests check the library exports all of them)
EXPORTS = [
    "eigsol_b200_last_error", "eigsol_b200_device_count", "eigsol_b200_version",
    "eigsol_b200_power_csc", "eigsol_b200_power_dense", "eigsol_b200_power_csc_keep", "eigsol_b200_power_dense_keep",
    "eigsol_b200_matrix_create_csc", "eigsol_b200_matrix_create_csr_rows", "eigsol_b200_matrix_create_dense",
    "eigsol_b200_matrix_destroy", "eigsol_b200_matrix_get_info", "eigsol_b200_matrix_set_kernel",
    "eigsol_b200_power_run", "eigsol_b200_power_begin", "eigsol_b200_power_step", "eigsol_b200_power_end",
    "eigsol_b200_matvec", "eigsol_b200_rayleigh", "eigsol_b200_block_power_run", "eigsol_b200_matrix_launch_count",
    "eigsol_b200_matrix_peer_stats", "eigsol_b200_hbm_probe", "eigsol_b200_matrices_created",
    "eigsol_b200_read_matrix_text", "eigsol_b200_read_matrix_bin", "eigsol_b200_write_matrix_bin",
    "eigsol_b200_file_matrix_release",
    "eigsol_b200_comm_unique_id", "eigsol_b200_comm_create", "eigsol_b200_comm_destroy",
    "eigsol_b200_comm_max_float",
    "eigsol_b200_csc_to_csr", "eigsol_b200_partition_rows", "eigsol_b200_halo_plan",
    "eigsol_b200_choose_kernel", "eigsol_b200_choose_kernel_csr", "eigsol_b200_random_unit",
]


class FileMatrix(C.Structure):
    """eigsol_b200_file_matrix (include/eigsol_b200.h)."""
    _fields_ = [
        ("dtype", C.c_int32), ("is_dense", C.c_int32), ("rows", C.c_int64), ("cols", C.c_int64), ("nnz", C.c_int64),
        ("outer", C.POINTER(C.c_int32)), ("inner", C.POINTER(C.c_int32)), ("values", C.c_void_p), ("owner", C.c_void_p),
    ]


class MatrixInfo(C.Structure):
    _fields_ = [
        ("dtype", C.c_int32), ("is_dense", C.c_int32), ("n", C.c_int64), ("nnz", C.c_int64),
        ("local_rows", C.c_int64), ("local_nnz", C.c_int64), ("halo", C.c_int64),
        ("kernel", C.c_int32), ("lanes_per_row", C.c_int32), ("grid", C.c_int32), ("block", C.c_int32),
        ("row_len_mean", C.c_double), ("row_len_max", C.c_double), ("row_len_cv", C.c_double),
        ("alg_bytes", C.c_int64), ("stored_bytes", C.c_int64),
        ("nranks", C.c_int32), ("rank", C.c_int32), ("kernels_per_trip", C.c_int32), ("exchange", C.c_int32),
    ]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ }
        d["kernel_name"] = KERNEL_NAMES.get(self.kernel, "?")
        return d


class EigsolB200Error(RuntimeError):
    """Mirrors the std::runtime_error the reference throws (power_method.hpp:54,57,138)."""

    def __init__(self, status, message):
        super().__init__(message)
        self.status = status


_lib = None


def lib():
    """The loaded library; raises if it has not been built (no CPU fallback on this path)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C pcsc_eigenvalue_solver_project_b200/csrc`; this path has no CPU fallback")
    L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, i32p, i64p, dp, ip, fp = C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p
    i, i64, d = C.c_int, C.c_int64, C.c_double
    sig = {
        "eigsol_b200_last_error": (C.c_char_p, []),
        "eigsol_b200_device_count": (i, []),
        "eigsol_b200_version": (C.c_char_p, []),
        "eigsol_b200_power_csc": (i, [i, i64, i64, i32p, i32p, vp, vp, i, d, vp, dp, ip, ip]),
        "eigsol_b200_power_dense": (i, [i, i64, i64, vp, i64, vp, i, d, vp, dp, ip, ip]),
        "eigsol_b200_power_csc_keep": (i, [i, i64, i64, i32p, i32p, vp, vp, i, d, vp, dp, ip, ip, C.POINTER(vp)]),
        "eigsol_b200_power_dense_keep": (i, [i, i64, i64, vp, i64, vp, i, d, vp, dp, ip, ip, C.POINTER(vp)]),
        "eigsol_b200_matrix_create_csc": (i, [i, i64, i64, i32p, i32p, vp, vp, C.POINTER(vp)]),
        "eigsol_b200_matrix_create_csr_rows": (i, [i, i64, i64, i64, i32p, i32p, vp, vp, C.POINTER(vp)]),
        "eigsol_b200_matrix_create_dense": (i, [i, i64, i64, vp, i64, vp, C.POINTER(vp)]),
        "eigsol_b200_matrix_destroy": (i, [vp]),
        "eigsol_b200_matrix_get_info": (i, [vp, C.POINTER(MatrixInfo)]),
        "eigsol_b200_matrix_set_kernel": (i, [vp, i, i]),
        "eigsol_b200_power_run": (i, [vp, vp, i, d, vp, dp, ip, ip]),
        "eigsol_b200_power_begin": (i, [vp, vp, i, d]),
        "eigsol_b200_power_step": (i, [vp, i, fp]),
        "eigsol_b200_power_end": (i, [vp, vp, dp, ip, ip]),
        "eigsol_b200_matvec": (i, [vp, vp, vp]),
        "eigsol_b200_rayleigh": (i, [vp, vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "eigsol_b200_block_power_run": (i, [vp, i, i, vp, i, C.c_double, vp, dp, ip, ip, ip, fp]),
        "eigsol_b200_matrix_launch_count": (i64, [vp]),
        "eigsol_b200_matrix_peer_stats": (i, [vp, vp]),
        "eigsol_b200_matrices_created": (i64, []),
        "eigsol_b200_hbm_probe": (i, [i64, i, i, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "eigsol_b200_read_matrix_text": (i, [C.c_char_p, i, C.POINTER(FileMatrix)]),
        "eigsol_b200_read_matrix_bin": (i, [C.c_char_p, C.POINTER(FileMatrix)]),
        "eigsol_b200_write_matrix_bin": (i, [C.c_char_p, i, i, i64, i64, i64, vp, vp, vp]),
        "eigsol_b200_file_matrix_release": (i, [C.POINTER(FileMatrix)]),
        "eigsol_b200_comm_unique_id": (i, [vp]),
        "eigsol_b200_comm_create": (i, [i, i, vp, i, C.POINTER(vp)]),
        "eigsol_b200_comm_destroy": (i, [vp]),
        "eigsol_b200_comm_max_float": (i, [vp, fp]),
        "eigsol_b200_csc_to_csr": (i, [i, i64, i64, i32p, i32p, vp, i32p, i32p, vp]),
        "eigsol_b200_partition_rows": (i, [i64, i32p, i, i64p]),
        "eigsol_b200_halo_plan": (i, [i64, i32p, i32p, i, i64p, i, i64p, i32p, i64p]),
        "eigsol_b200_choose_kernel": (i, [i, i64, i32p, ip, ip]),
        "eigsol_b200_choose_kernel_csr": (i, [i, i64, i32p, i32p, ip, ip]),
        "eigsol_b200_random_unit": (i, [i, i64, i, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)  # AttributeError here == the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def last_error() -> str:
    return lib().eigsol_b200_last_error().decode("utf-8", "replace")


def check(status: int):
    if status == OK:
        return
    msg = last_error() or f"eigsol_b200 status {status}"
    if status == ERR_INVALID:
        raise ValueError(msg)
    raise EigsolB200Error(status, msg)


def ptr(a):
    """Raw host pointer of a C-contiguous numpy array (or None)."""
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)


def dtype_code(np_dtype) -> int:
    dt = np.dtype(np_dtype)
    if dt == np.float64:
        return F64
    if dt == np.complex128:
        return C128
    raise TypeError(f"unsupported scalar type {dt}: the B200 path handles float64 and complex128")


def np_dtype(code: int):
    return np.complex128 if code == C128 else np.float64
