"""pcsc_eigenvalue_solver_project_b200 -- the B200-native power-iteration path of EigSol.

Python mirror of the reference's interface for this ONE path (names, argument meaning and error
behaviour follow the reference; citations are relative to the reference checkout):

    Matrix          src/matrix/matrix.hpp:36-246      type-erased dense / sparse owner
    SolverOptions   src/option/solver_option.hpp:14-20
    EigenResult     src/result/eigen_result.hpp:23-52
    powerMethod     src/power_method/power_method.hpp:135-148

Underneath, everything goes through the C ABI of ``libeigsol_b200.so`` (``include/eigsol_b200.h``)
into hand-written sm_100a kernels.  There is no CPU fallback: if the library is not built, or no
CUDA device is present, the calls raise.  The C++ drop-in headers with the reference's exact
signatures live in ``eigsol/src``.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import os

import numpy as np

from . import _capi
from ._capi import (C128, F64, KERNEL_AUTO, KERNEL_CSR_SCALAR, KERNEL_CSR_STREAM, KERNEL_CSR_VECTOR, KERNEL_CSR_WINDOW, KERNEL_GEMV,
                    KERNEL_NAMES, KERNEL_SELL, EigsolB200Error)

__all__ = ["Matrix", "SolverOptions", "EigenResult", "powerMethod", "power_method", "DeviceMatrix", "Comm",
           "BlockEigenResult", "blockPowerMethod", "random_block",
           "random_unit", "srand", "device_count", "hbm_probe", "matrices_created", "EigsolB200Error", "readMatrixFromFile",
           "writeMatrixBinary", "readMatrixBinary"]


# --------------------------------------------------------------------------------------------------
# src/option/solver_option.hpp:14-20 and src/result/eigen_result.hpp:23-52
# --------------------------------------------------------------------------------------------------
@dataclass
class SolverOptions:
    maxIterations: int = 1000
    tolerance: float = 1e-10


@dataclass
class EigenResult:
    eigenvalue: complex | float = 0.0
    eigenvector: np.ndarray = field(default_factory=lambda: np.zeros(0))
    iterations: int = 0
    converged: bool = False


@dataclass
class BlockEigenResult:
    """Result of the block power iteration: the block's Ritz values ordered by modulus (the first ``nev`` passed
    the convergence test), unit-norm Ritz vectors as columns, trips, flags."""
    eigenvalues: np.ndarray = field(default_factory=lambda: np.zeros(0, dtype=np.complex128))
    eigenvectors: np.ndarray = field(default_factory=lambda: np.zeros((0, 0)))
    iterations: int = 0
    converged: bool = False
    breakdown: bool = False
    device_ms: float = 0.0


def device_count() -> int:
    return int(_capi.lib().eigsol_b200_device_count())


def matrices_created() -> int:
    """Matrix handles (= uploads of a whole matrix) this process has created so far."""
    return int(_capi.lib().eigsol_b200_matrices_created())


def hbm_probe(plane_bytes: int = 128 << 20, planes: int = 12, reps: int = 20) -> tuple[float, float]:
    """(best, mean) GB/s of a read-dominated stream: ``planes`` read planes summed into one written plane
    (12 = the 92/8 read/write mix of an fp64 CSR matvec).  Measurement only; bench.py's ``peak_read_mix``."""
    best, mean = C.c_double(0.0), C.c_double(0.0)
    _capi.check(_capi.lib().eigsol_b200_hbm_probe(int(plane_bytes), int(planes), int(reps), C.byref(best), C.byref(mean)))
    return float(best.value), float(mean.value)


_libc = C.CDLL(None)


def srand(seed: int) -> None:
    """std::srand(seed): the reference draws x0 from the process-wide libc rand() stream
    (eigen/Eigen/src/Core/RandomImpl.h:46-51); seed it the way a caller of the reference would."""
    _libc.srand(C.c_uint(seed))


def random_unit(n: int, dtype=np.float64, seed: int = -1) -> np.ndarray:
    """The start vector of power_method.hpp:62 (Vector<Scalar>::Random(n), normalize())."""
    code = _capi.dtype_code(dtype)
    x = np.empty(n, dtype=_capi.np_dtype(code))
    _capi.check(_capi.lib().eigsol_b200_random_unit(code, n, seed, _capi.ptr(x)))
    return x


# --------------------------------------------------------------------------------------------------
# src/matrix/matrix.hpp:36-246
# --------------------------------------------------------------------------------------------------
class Matrix:
    """Type-erased owner of one dense (column-major) or sparse (CSC, int32) matrix.

    Like the reference class it always holds a value, is neither copyable nor movable, and reports
    ``isDense()``, ``scalar_type()``, ``type()``; ``cast(kind, dtype)`` returns the stored arrays or
    raises ``TypeError`` (the reference's ``std::bad_cast``, matrix.hpp:235-240).
    """

    def __init__(self, m, rows: int | None = None, cols: int | None = None):
        if rows is not None or cols is not None:
            # Matrix(const std::vector<Scalar>&, rows, cols): flat ROW-major data (matrix.hpp:109-116)
            data = np.asarray(m)
            if data.ndim != 1 or rows is None or cols is None or rows * cols != data.size:
                raise RuntimeError("Matrix: size mismatch in vector constructor")
            m = data.reshape(rows, cols)
        if hasattr(m, "tocsc"):  # any scipy.sparse matrix: normalised to CSC / int32 (matrix.hpp:198-204)
            s = m.tocsc()
            s.sort_indices()
            _capi.dtype_code(s.dtype)
            self._dense = False
            self._shape = tuple(int(v) for v in s.shape)
            self._outer = np.ascontiguousarray(s.indptr, dtype=np.int32)
            self._inner = np.ascontiguousarray(s.indices, dtype=np.int32)
            self._values = np.ascontiguousarray(s.data)
            self._dtype = np.dtype(s.dtype)
        elif isinstance(m, tuple) and len(m) == 4:  # (shape, outer, inner, values): raw CSC arrays
            shape, outer, inner, values = m
            values = np.ascontiguousarray(values)
            _capi.dtype_code(values.dtype)
            self._dense = False
            self._shape = (int(shape[0]), int(shape[1]))
            self._outer = np.ascontiguousarray(outer, dtype=np.int32)
            self._inner = np.ascontiguousarray(inner, dtype=np.int32)
            self._values = values
            self._dtype = values.dtype
            if self._outer.size != self._shape[1] + 1:
                raise ValueError("CSC outer index must have cols + 1 entries")
        else:
            a = np.asarray(m)
            if a.ndim != 2:
                raise TypeError("Matrix needs a 2-D dense array, a scipy.sparse matrix or CSC arrays")
            _capi.dtype_code(a.dtype)
            self._dense = True
            self._shape = (int(a.shape[0]), int(a.shape[1]))
            self._a = np.asfortranarray(a).copy(order="F")  # stored by value, column-major (matrix.hpp:39-40)
            self._dtype = a.dtype

    # Device-resident copy (the C++ drop-in does the same in eigsol/src/matrix/matrix.hpp): a Matrix
    # can neither move nor be copied, so it can carry its uploaded copy; powerMethod(M) uploads once per Matrix.
    # cast() hands out the stored (mutable) arrays: from then on the host copy may change behind our back, so the
    # device copy is dropped and never cached again -- every later solve uploads what the host holds at that moment.
    _dm = None
    _cacheable = True

    def _drop_device_copy(self):
        if self._dm is not None:
            self._dm.close()
            self._dm = None

    def __del__(self):
        try:
            self._drop_device_copy()
        except Exception:
            pass

    def __copy__(self):
        raise TypeError("Matrix is non-copyable by design")

    def __deepcopy__(self, memo):
        raise TypeError("Matrix is non-copyable by design")

    def isDense(self) -> bool:
        return self._dense

    def scalar_type(self):
        return self._dtype.type

    def type(self):
        return ("dense" if self._dense else "sparse", self._dtype.type)

    def rows(self) -> int:
        return self._shape[0]

    def cols(self) -> int:
        return self._shape[1]

    def cast(self, kind: str, dtype):
        if (kind, np.dtype(dtype).type) != self.type():
            raise TypeError("bad cast")  # std::bad_cast
        self._cacheable = False
        self._drop_device_copy()
        if self._dense:
            return self._a
        return self._outer, self._inner, self._values


def powerMethod(M: Matrix, opts: SolverOptions | None = None, scalar=None, x0: np.ndarray | None = None) -> EigenResult:
    """EigSol::powerMethod<Scalar>(M, opts) (power_method.hpp:135-148) on the B200.

    ``scalar`` is the template argument; it defaults to the stored scalar type.  Checks, their order
    and their messages follow the reference: scalar mismatch, then square, then zero size; only then
    is x0 drawn from libc rand() (so a raising call consumes no random numbers).
    """
    opts = opts or SolverOptions()
    scalar = M.scalar_type() if scalar is None else np.dtype(scalar).type
    if M.scalar_type() != scalar:
        raise RuntimeError("powerMethod: scalar type mismatch")
    n, ncols = M.rows(), M.cols()
    if n != ncols:
        raise RuntimeError("powerMethod: matrix must be square")
    if n == 0:
        raise RuntimeError("powerMethod: matrix has zero size")
    code = _capi.dtype_code(scalar)
    if x0 is None:
        x0 = random_unit(n, scalar)
    x0 = np.ascontiguousarray(x0, dtype=_capi.np_dtype(code))
    if x0.shape != (n,):
        raise ValueError("x0 must have length n")
    if M._cacheable and M._dm is not None:
        return M._dm.power(x0, opts)  # the Matrix is resident already: no upload
    x = np.empty(n, dtype=_capi.np_dtype(code))
    lam = np.zeros(2)
    iters, conv = C.c_int(0), C.c_int(0)
    L = _capi.lib()
    # first solve on this Matrix: upload + solve in one call; the device copy is kept with the Matrix unless mutable
    # access to its arrays was ever handed out (cast)
    keep = C.c_void_p()
    keep_ref = C.byref(keep) if M._cacheable else None
    if M.isDense():
        a = M._a
        st = L.eigsol_b200_power_dense_keep(code, n, ncols, _capi.ptr(a), a.strides[1] // a.itemsize if n > 1 else n,
                                            _capi.ptr(x0), int(opts.maxIterations), float(opts.tolerance), _capi.ptr(x),
                                            _capi.ptr(lam), C.byref(iters), C.byref(conv), keep_ref)
    else:
        st = L.eigsol_b200_power_csc_keep(code, n, ncols, _capi.ptr(M._outer), _capi.ptr(M._inner), _capi.ptr(M._values),
                                          _capi.ptr(x0), int(opts.maxIterations), float(opts.tolerance), _capi.ptr(x),
                                          _capi.ptr(lam), C.byref(iters), C.byref(conv), keep_ref)
    _capi.check(st)
    if keep.value:
        M._dm = DeviceMatrix(keep, code, n, None)
    eig = complex(lam[0], lam[1]) if code == C128 else float(lam[0])
    return EigenResult(eig, x, int(iters.value), bool(conv.value))


power_method = powerMethod


def random_block(n: int, k: int, dtype=np.float64, seed: int = 1) -> np.ndarray:
    """n x k start block: column j is the reference's start vector (power_method.hpp:62) drawn after srand(seed + j)."""
    return np.asfortranarray(np.stack([random_unit(n, dtype, seed + j) for j in range(k)], axis=1))


def blockPowerMethod(M: Matrix, block: int = 4, nev: int | None = None, opts: SolverOptions | None = None,
                     x0: np.ndarray | None = None) -> BlockEigenResult:
    """Block (simultaneous) power iteration on the B200: ``block`` vectors per pass over the matrix, Rayleigh-Ritz
    on the block (not in the reference API).  Same checks and messages as powerMethod
    (power_method.hpp:52-58); with ``block = 1`` it returns what powerMethod returns."""
    n, ncols = M.rows(), M.cols()
    if n != ncols:
        raise RuntimeError("powerMethod: matrix must be square")
    if n == 0:
        raise RuntimeError("powerMethod: matrix has zero size")
    if x0 is None:
        x0 = random_block(n, block, M.scalar_type())
    dm = DeviceMatrix.from_matrix(M)
    try:
        return dm.block_power(x0, nev, opts)
    finally:
        dm.close()


# --------------------------------------------------------------------------------------------------
# Handle API: device-resident matrix (upload once, iterate many) and multi-GPU ranks.
# --------------------------------------------------------------------------------------------------
def _matrix_from_file_struct(fm) -> "Matrix":
    """Copies the arrays of an eigsol_b200_file_matrix into a Matrix (which stores by value, matrix.hpp:187-204)."""
    dt = np.complex128 if fm.dtype else np.float64
    rows, cols, nnz = int(fm.rows), int(fm.cols), int(fm.nnz)
    vals = np.ctypeslib.as_array(C.cast(fm.values, C.POINTER(C.c_double)), shape=(nnz * (2 if fm.dtype else 1),))
    vals = vals.view(dt) if fm.dtype else vals
    if fm.is_dense:
        return Matrix(np.array(vals.reshape((cols, rows)).T, order="F", copy=True))
    outer = np.ctypeslib.as_array(fm.outer, shape=(cols + 1,)).copy()
    inner = np.ctypeslib.as_array(fm.inner, shape=(max(nnz, 1),))[:nnz].copy()
    return Matrix(((rows, cols), outer, inner, vals.copy()))


def readMatrixFromFile(filename: str, dtype=np.float64) -> "Matrix":
    """``EigSol::readMatrixFromFile<Scalar>(filename)`` (src/reader/file_matrix_reader.hpp:170-200): the reference's
    text format (README.md:87-171) parsed by all host threads.  ``dtype`` plays the template argument: complex
    reads two reals per value.  Raises ``RuntimeError`` with the reference's messages."""
    fm = _capi.FileMatrix()
    st = _capi.lib().eigsol_b200_read_matrix_text(os.fsencode(filename), _capi.dtype_code(np.dtype(dtype)), C.byref(fm))
    if st == _capi.ERR_IO:
        raise RuntimeError(_capi.last_error())
    _capi.check(st)
    try:
        return _matrix_from_file_struct(fm)
    finally:
        _capi.lib().eigsol_b200_file_matrix_release(C.byref(fm))


def writeMatrixBinary(m: "Matrix", filename: str) -> None:
    """Stores a Matrix in the mappable EIGSOLB2 container (64-byte header + 64-byte aligned raw arrays)."""
    L = _capi.lib()
    code = _capi.dtype_code(m._dtype)
    if m.isDense():
        a = m._a
        _capi.check(L.eigsol_b200_write_matrix_bin(os.fsencode(filename), code, 1, a.shape[0], a.shape[1], a.size,
                                                    None, None, _capi.ptr(a)))
    else:
        _capi.check(L.eigsol_b200_write_matrix_bin(os.fsencode(filename), code, 0, m._shape[0], m._shape[1],
                                                    int(m._outer[-1]), _capi.ptr(m._outer), _capi.ptr(m._inner),
                                                    _capi.ptr(m._values)))


def readMatrixBinary(filename: str) -> "Matrix":
    """Maps an EIGSOLB2 container (no parsing) and wraps its arrays in a Matrix."""
    fm = _capi.FileMatrix()
    st = _capi.lib().eigsol_b200_read_matrix_bin(os.fsencode(filename), C.byref(fm))
    if st == _capi.ERR_IO:
        raise RuntimeError(_capi.last_error())
    _capi.check(st)
    try:
        return _matrix_from_file_struct(fm)
    finally:
        _capi.lib().eigsol_b200_file_matrix_release(C.byref(fm))


class Comm:
    """One rank of a multi-GPU group (one process per GPU).  ``unique_id()`` on rank 0, broadcast the
    128 bytes with whatever the host program has (torch.distributed, MPI, a file), then construct."""

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        _capi.check(_capi.lib().eigsol_b200_comm_unique_id(buf))
        return buf.raw

    def __init__(self, rank: int, nranks: int, unique_id: bytes | None, device: int):
        self.rank, self.nranks, self.device = rank, nranks, device
        self._h = C.c_void_p()
        idbuf = C.create_string_buffer(unique_id, 128) if unique_id is not None else None
        _capi.check(_capi.lib().eigsol_b200_comm_create(rank, nranks, idbuf, device, C.byref(self._h)))

    def max_float(self, v: float) -> float:
        f = C.c_float(v)
        _capi.check(_capi.lib().eigsol_b200_comm_max_float(self._h, C.byref(f)))
        return float(f.value)

    def close(self):
        if self._h:
            _capi.lib().eigsol_b200_comm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DeviceMatrix:
    """A matrix resident in HBM plus its iteration state (eigsol_b200_matrix)."""

    def __init__(self, handle, dtype_code: int, n: int, comm: Comm | None):
        self._h, self._code, self.n, self._comm = handle, dtype_code, n, comm

    # ---- constructors ----
    @classmethod
    def from_matrix(cls, M: Matrix, comm: Comm | None = None):
        code = _capi.dtype_code(M.scalar_type())
        h = C.c_void_p()
        L = _capi.lib()
        ch = comm._h if comm else None
        if M.isDense():
            a = M._a
            lda = a.strides[1] // a.itemsize if M.rows() > 1 else M.rows()
            _capi.check(L.eigsol_b200_matrix_create_dense(code, M.rows(), M.cols(), _capi.ptr(a), lda, ch, C.byref(h)))
        else:
            _capi.check(L.eigsol_b200_matrix_create_csc(code, M.rows(), M.cols(), _capi.ptr(M._outer), _capi.ptr(M._inner),
                                                        _capi.ptr(M._values), ch, C.byref(h)))
        return cls(h, code, M.rows(), comm)

    @classmethod
    def from_csr_rows(cls, n: int, row_begin: int, rowptr, colidx, values, comm: Comm | None = None):
        values = np.ascontiguousarray(values)
        code = _capi.dtype_code(values.dtype)
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int32)
        colidx = np.ascontiguousarray(colidx, dtype=np.int32)
        h = C.c_void_p()
        _capi.check(_capi.lib().eigsol_b200_matrix_create_csr_rows(code, n, row_begin, rowptr.size - 1, _capi.ptr(rowptr),
                                                                   _capi.ptr(colidx), _capi.ptr(values),
                                                                   comm._h if comm else None, C.byref(h)))
        return cls(h, code, n, comm)

    # ---- queries / knobs ----
    def info(self) -> dict:
        inf = _capi.MatrixInfo()
        _capi.check(_capi.lib().eigsol_b200_matrix_get_info(self._h, C.byref(inf)))
        return inf.as_dict()

    def set_kernel(self, kernel: int, lanes: int = 0):
        _capi.check(_capi.lib().eigsol_b200_matrix_set_kernel(self._h, kernel, lanes))

    def launch_count(self) -> int:
        return int(_capi.lib().eigsol_b200_matrix_launch_count(self._h))

    def peer_stats(self) -> dict:
        """Where this rank's multi-GPU trips waited since the last call (needs EIGSOL_B200_PEER_STATS=1 at creation)."""
        out = np.zeros(8, dtype=np.uint64)
        _capi.check(_capi.lib().eigsol_b200_matrix_peer_stats(self._h, _capi.ptr(out)))
        return {"control_wait_ns": int(out[0]), "control_launches": int(out[1]), "halo_wait_ns": int(out[2]),
                "halo_waiting_warps": int(out[3]), "halo_wait_max_ns": int(out[4]), "halo_warps": int(out[5])}

    @property
    def dtype(self):
        return _capi.np_dtype(self._code)

    def _vec(self, x):
        x = np.ascontiguousarray(x, dtype=self.dtype)
        if x.shape != (self.n,):
            raise ValueError("vector must have length n")
        return x

    # ---- the path ----
    def power(self, x0, opts: SolverOptions | None = None) -> EigenResult:
        opts = opts or SolverOptions()
        x0 = self._vec(x0)
        x = np.empty(self.n, dtype=self.dtype)
        lam = np.zeros(2)
        iters, conv = C.c_int(0), C.c_int(0)
        _capi.check(_capi.lib().eigsol_b200_power_run(self._h, _capi.ptr(x0), int(opts.maxIterations),
                                                      float(opts.tolerance), _capi.ptr(x), _capi.ptr(lam),
                                                      C.byref(iters), C.byref(conv)))
        eig = complex(lam[0], lam[1]) if self._code == C128 else float(lam[0])
        return EigenResult(eig, x, int(iters.value), bool(conv.value))

    def begin(self, x0, max_iterations: int = 2**31 - 2, tolerance: float = -1.0):
        _capi.check(_capi.lib().eigsol_b200_power_begin(self._h, _capi.ptr(self._vec(x0)), max_iterations, tolerance))

    def step(self, trips: int) -> float:
        ms = C.c_float(0.0)
        _capi.check(_capi.lib().eigsol_b200_power_step(self._h, trips, C.byref(ms)))
        return float(ms.value)

    def end(self, want_vector: bool = True) -> EigenResult:
        x = np.empty(self.n, dtype=self.dtype) if want_vector else None
        lam = np.zeros(2)
        iters, conv = C.c_int(0), C.c_int(0)
        _capi.check(_capi.lib().eigsol_b200_power_end(self._h, _capi.ptr(x), _capi.ptr(lam), C.byref(iters),
                                                      C.byref(conv)))
        eig = complex(lam[0], lam[1]) if self._code == C128 else float(lam[0])
        return EigenResult(eig, x, int(iters.value), bool(conv.value))

    def matvec(self, x) -> np.ndarray:
        y = np.empty(self.n, dtype=self.dtype)
        _capi.check(_capi.lib().eigsol_b200_matvec(self._h, _capi.ptr(self._vec(x)), _capi.ptr(y)))
        return y

    def rayleigh(self, x):
        """``(x^H A x, ||A x||^2)`` from one fused launch: the ``x.dot(A * x)`` step of the reference's shifted
        inverse method (src/power_method/shifted_inverse_power_solver.hpp:62)."""
        dot = (C.c_double * 2)()
        nrm = C.c_double()
        _capi.check(_capi.lib().eigsol_b200_rayleigh(self._h, _capi.ptr(self._vec(x)), dot, C.byref(nrm)))
        d = complex(dot[0], dot[1])
        return (d if np.issubdtype(self.dtype, np.complexfloating) else d.real), float(nrm.value)

    def block_power(self, x0, nev: int | None = None, opts: SolverOptions | None = None, want_vectors: bool = True) -> BlockEigenResult:
        """Block power iteration (eigsol_b200_block_power_run): ``x0`` is the n x k start block, k in {1, 2, 4, 8}."""
        opts = opts or SolverOptions()
        x0 = np.asfortranarray(np.asarray(x0, dtype=self.dtype).reshape(self.n, -1))
        k = x0.shape[1]
        nev = k if nev is None else int(nev)
        x = np.empty((self.n, k), dtype=self.dtype, order="F") if want_vectors else None
        lam = np.zeros(2 * k)
        iters, conv, brk, ms = C.c_int(0), C.c_int(0), C.c_int(0), C.c_float(0)
        _capi.check(_capi.lib().eigsol_b200_block_power_run(
            self._h, k, nev, _capi.ptr(x0), int(opts.maxIterations), float(opts.tolerance),
            _capi.ptr(x) if want_vectors else None, _capi.ptr(lam), C.byref(iters), C.byref(conv), C.byref(brk), C.byref(ms)))
        return BlockEigenResult(lam.view(np.complex128).copy(), x if want_vectors else np.zeros((0, 0)), int(iters.value),
                                bool(conv.value), bool(brk.value), float(ms.value))

    def close(self):
        if self._h:
            _capi.lib().eigsol_b200_matrix_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
