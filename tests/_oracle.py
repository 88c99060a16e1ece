"""ctypes doorways onto the CHECKERS (test infrastructure only):

  oracle/libeigsol_oracle.so      plain-C restatement of the reference path (oracle/eigsol_oracle.c)
  oracle/_ref/libeigsol_ref.so    the UNMODIFIED reference compiled from /root/reference (oracle/ref_shim.cpp)

Nothing in the product package imports this module.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libeigsol_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libeigsol_ref.so")

_vp, _i, _i64, _d = C.c_void_p, C.c_int, C.c_int64, C.c_double


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def is_complex(dt):
    return 1 if np.dtype(dt) == np.complex128 else 0


class Oracle:
    def __init__(self):
        L = C.CDLL(ORACLE_SO)
        L.oracle_power_csc.restype = _i
        L.oracle_power_csc.argtypes = [_i, _i64, _i64, _vp, _vp, _vp, _vp, _i, _d, _vp, _vp, _vp, _vp]
        L.oracle_power_dense.restype = _i
        L.oracle_power_dense.argtypes = [_i, _i64, _i64, _vp, _i64, _vp, _i, _d, _vp, _vp, _vp, _vp]
        L.oracle_random_unit.restype = None
        L.oracle_random_unit.argtypes = [_i, _i64, _vp]
        L.oracle_srand.restype = None
        L.oracle_srand.argtypes = [C.c_uint]
        L.oracle_squared_norm.restype = _d
        L.oracle_squared_norm.argtypes = [_i64, _vp]
        L.oracle_dotc.restype = None
        L.oracle_dotc.argtypes = [_i, _i64, _vp, _vp, _vp]
        L.oracle_is_close_relative.restype = _i
        L.oracle_is_close_relative.argtypes = [_vp, _vp, _d]
        for name in ("oracle_spmv_csc_f64", "oracle_spmv_csc_c128"):
            f = getattr(L, name)
            f.restype = None
            f.argtypes = [_i64, _i64, _vp, _vp, _vp, _vp, _vp]
        for name in ("oracle_gemv_colmajor_f64", "oracle_gemv_colmajor_c128"):
            f = getattr(L, name)
            f.restype = None
            f.argtypes = [_i64, _i64, _vp, _i64, _vp, _vp]
        self.L = L

    def random_unit(self, n, dtype, seed=None):
        if seed is not None:
            self.L.oracle_srand(seed)
        x = np.empty(n, dtype=dtype)
        self.L.oracle_random_unit(is_complex(dtype), n, _p(x))
        return x

    def spmv_csc(self, shape, outer, inner, vals, x):
        y = np.empty(shape[0], dtype=vals.dtype)
        f = self.L.oracle_spmv_csc_c128 if is_complex(vals.dtype) else self.L.oracle_spmv_csc_f64
        f(shape[0], shape[1], _p(outer), _p(inner), _p(vals), _p(x), _p(y))
        return y

    def gemv(self, a, x):
        a = np.asfortranarray(a)
        y = np.empty(a.shape[0], dtype=a.dtype)
        f = self.L.oracle_gemv_colmajor_c128 if is_complex(a.dtype) else self.L.oracle_gemv_colmajor_f64
        f(a.shape[0], a.shape[1], _p(a), a.shape[0], _p(x), _p(y))
        return y

    def _result(self, st, x, lam, it, cv, cplx):
        eig = complex(lam[0], lam[1]) if cplx else float(lam[0])
        return st, eig, x, int(it.value), bool(cv.value)

    def power_csc(self, shape, outer, inner, vals, x0, max_it, tol):
        cplx = is_complex(vals.dtype)
        x = np.empty(max(shape[0], 1), dtype=vals.dtype)
        lam = np.zeros(2)
        it, cv = C.c_int(0), C.c_int(0)
        st = self.L.oracle_power_csc(cplx, shape[0], shape[1], _p(outer), _p(inner), _p(vals), _p(x0), max_it, tol,
                                     _p(x), _p(lam), C.byref(it), C.byref(cv))
        return self._result(st, x[:shape[0]], lam, it, cv, cplx)

    def power_dense(self, a, x0, max_it, tol):
        a = np.asfortranarray(a)
        cplx = is_complex(a.dtype)
        x = np.empty(max(a.shape[0], 1), dtype=a.dtype)
        lam = np.zeros(2)
        it, cv = C.c_int(0), C.c_int(0)
        st = self.L.oracle_power_dense(cplx, a.shape[0], a.shape[1], _p(a), max(a.shape[0], 1), _p(x0), max_it, tol,
                                       _p(x), _p(lam), C.byref(it), C.byref(cv))
        return self._result(st, x[:a.shape[0]], lam, it, cv, cplx)

    def is_close_relative(self, a, b, tol):
        aa = np.array([complex(a).real, complex(a).imag])
        bb = np.array([complex(b).real, complex(b).imag])
        return bool(self.L.oracle_is_close_relative(_p(aa), _p(bb), tol))


class Reference:
    """The compiled, unmodified reference (present when oracle/_ref was built and shipped)."""

    def __init__(self):
        L = C.CDLL(REF_SO)
        L.ref_last_error.restype = C.c_char_p
        L.ref_matrix_create_csc.restype = _vp
        L.ref_matrix_create_csc.argtypes = [_i, _i64, _i64, _vp, _vp, _vp]
        L.ref_matrix_create_dense.restype = _vp
        L.ref_matrix_create_dense.argtypes = [_i, _i64, _i64, _vp, _i64]
        L.ref_matrix_destroy.restype = None
        L.ref_matrix_destroy.argtypes = [_vp]
        L.ref_power_run.restype = _i
        L.ref_power_run.argtypes = [_vp, _i, _i, _i, _d, _vp, _vp, _vp, _vp, _vp]
        L.ref_random_unit.restype = None
        L.ref_random_unit.argtypes = [_i, _i64, _i, _vp]
        L.ref_matvec.restype = _i
        L.ref_matvec.argtypes = [_vp, _vp, _vp]
        L.ref_is_close_relative.restype = _i
        L.ref_is_close_relative.argtypes = [_i, _vp, _vp, _d]
        if hasattr(L, "ref_read_matrix_file"):
            L.ref_read_matrix_file.restype = _i
            L.ref_read_matrix_file.argtypes = [C.c_char_p, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
        self.L = L

    def last_error(self):
        return self.L.ref_last_error().decode()

    def create_csc(self, shape, outer, inner, vals):
        return self.L.ref_matrix_create_csc(is_complex(vals.dtype), shape[0], shape[1], _p(outer), _p(inner), _p(vals))

    def create_dense(self, a):
        a = np.asfortranarray(a)
        return self.L.ref_matrix_create_dense(is_complex(a.dtype), a.shape[0], a.shape[1], _p(a), max(a.shape[0], 1))

    def destroy(self, h):
        self.L.ref_matrix_destroy(h)

    def power(self, h, dtype, n, seed, max_it, tol, want_x=True):
        """Returns (status, eigenvalue, x, iterations, converged, seconds)."""
        cplx = is_complex(dtype)
        x = np.empty(max(n, 1), dtype=dtype) if want_x else None
        lam = np.zeros(2)
        it, cv, sec = C.c_int(0), C.c_int(0), C.c_double(0.0)
        st = self.L.ref_power_run(h, cplx, seed, max_it, tol, _p(x), _p(lam), C.byref(it), C.byref(cv), C.byref(sec))
        eig = complex(lam[0], lam[1]) if cplx else float(lam[0])
        return st, eig, (x[:n] if want_x else None), int(it.value), bool(cv.value), float(sec.value)

    def random_unit(self, n, dtype, seed):
        x = np.empty(n, dtype=dtype)
        self.L.ref_random_unit(is_complex(dtype), n, seed, _p(x))
        return x

    def matvec(self, h, x, n):
        y = np.empty(n, dtype=x.dtype)
        st = self.L.ref_matvec(h, _p(x), _p(y))
        assert st == 0, self.last_error()
        return y

    def read_matrix_file(self, path, dtype):
        """EigSol::readMatrixFromFile<Scalar>(path) of the compiled reference: returns
        (is_dense, rows, cols, outer, inner, values, seconds) or raises RuntimeError with its message."""
        cplx = is_complex(dtype)
        isd, rows, cols, nnz, sec = C.c_int(0), C.c_int64(0), C.c_int64(0), C.c_int64(0), C.c_double(0.0)
        args = (os.fsencode(path), cplx, C.byref(isd), C.byref(rows), C.byref(cols), C.byref(nnz))
        if self.L.ref_read_matrix_file(*args, None, None, None, C.byref(sec)) != 0:
            raise RuntimeError(self.last_error())
        vals = np.empty(max(nnz.value, 1), dtype=np.complex128 if cplx else np.float64)
        outer = np.empty(cols.value + 1, dtype=np.int32)
        inner = np.empty(max(nnz.value, 1), dtype=np.int32)
        if self.L.ref_read_matrix_file(*args, _p(outer), _p(inner), _p(vals), None) != 0:
            raise RuntimeError(self.last_error())
        n = nnz.value
        return bool(isd.value), rows.value, cols.value, outer, inner[:n], vals[:n], sec.value

    def is_close_relative(self, dtype, a, b, tol):
        aa = np.array([complex(a).real, complex(a).imag])
        bb = np.array([complex(b).real, complex(b).imag])
        return bool(self.L.ref_is_close_relative(is_complex(dtype), _p(aa), _p(bb), tol))


def have_reference():
    return os.path.exists(REF_SO)


def phase_aligned_error(x, xref):
    """min over phase of ||x - e^{i phi} xref||_2: e^{i phi} = (xref^H x)/|xref^H x|."""
    ip = np.vdot(xref, x)
    ph = ip / abs(ip) if abs(ip) > 0 else 1.0
    return float(np.linalg.norm(x - ph * xref))
