"""Synthetic matrices with analytically known spectra (BASELINE.json configs 2-5).

numpy only.  Sparse matrices are returned as CSR triplets (rowptr int32, colidx int32, values) with
ascending columns in every row; for the symmetric / Hermitian stencils here the CSC arrays Eigen
would hold (reference: src/matrix/matrix.hpp:43-44) are ``(rowptr, colidx, conj(values))``, see
``csr_to_csc_hermitian``.  Every generator can emit a row block ``[row_begin, row_end)`` only, so each
rank of a sharded run builds just its own rows.
"""
from __future__ import annotations

import math

import numpy as np


def _stencil_rows(offsets, valid, diag_value, off_values, row_begin, row_end, dtype):
    """Assemble CSR rows from per-offset validity masks.

    offsets: ascending list of column offsets (0 included); valid[k]: bool array over the row block
    (entry r + offsets[k] exists); off_values[k]: scalar or array over the row block.
    """
    nrows = row_end - row_begin
    rows = np.arange(row_begin, row_end, dtype=np.int64)
    k = len(offsets)
    mask = np.empty((nrows, k), dtype=bool)
    cols = np.empty((nrows, k), dtype=np.int32)
    vals = np.empty((nrows, k), dtype=dtype)
    for j, off in enumerate(offsets):
        mask[:, j] = valid[j]
        cols[:, j] = (rows + off).astype(np.int32)
        vals[:, j] = diag_value if off == 0 else off_values[j]
    rowptr = np.zeros(nrows + 1, dtype=np.int64)
    np.cumsum(mask.sum(axis=1), out=rowptr[1:])
    if rowptr[-1] > np.iinfo(np.int32).max:
        raise ValueError("row block has more than 2^31-1 entries")
    return rowptr.astype(np.int32), np.ascontiguousarray(cols[mask]), np.ascontiguousarray(vals[mask])


def laplacian_2d(g: int, row_begin: int = 0, row_end: int | None = None):
    """5-point Dirichlet Laplacian on a g x g grid: diag 4, -1 to the 4 neighbours.
    lambda_max = 4 + 4 cos(pi / (g + 1))."""
    n = g * g
    row_end = n if row_end is None else row_end
    r = np.arange(row_begin, row_end, dtype=np.int64)
    i, j = r // g, r % g
    offsets = [-g, -1, 0, 1, g]
    valid = [i > 0, j > 0, np.ones_like(i, bool), j < g - 1, i < g - 1]
    return _stencil_rows(offsets, valid, 4.0, [-1.0] * 5, row_begin, row_end, np.float64)


def laplacian_2d_lambda_max(g: int) -> float:
    return 4.0 + 4.0 * math.cos(math.pi / (g + 1))


def laplacian_3d(g: int, row_begin: int = 0, row_end: int | None = None):
    """7-point Dirichlet Laplacian on a g^3 grid: diag 6, -1 to the 6 neighbours.
    lambda_max = 6 + 6 cos(pi / (g + 1))."""
    n = g * g * g
    row_end = n if row_end is None else row_end
    r = np.arange(row_begin, row_end, dtype=np.int64)
    k, rem = r // (g * g), r % (g * g)
    i, j = rem // g, rem % g
    offsets = [-g * g, -g, -1, 0, 1, g, g * g]
    valid = [k > 0, i > 0, j > 0, np.ones_like(i, bool), j < g - 1, i < g - 1, k < g - 1]
    return _stencil_rows(offsets, valid, 6.0, [-1.0] * 7, row_begin, row_end, np.float64)


def laplacian_3d_lambda_max(g: int) -> float:
    return 6.0 + 6.0 * math.cos(math.pi / (g + 1))


def hermitian_laplacian_2d(g: int, seed: int = 1234, gauge: bool = False, row_begin: int = 0,
                           row_end: int | None = None):
    """Complex Hermitian 2D Laplacian: diag 4 (real), A(r,c) = -exp(i theta_rc) for r < c on the
    5-point pattern and A(c,r) = conj(A(r,c)).

    gauge=False: theta_rc free uniform phases (spectrum inside [0, 8] by Gershgorin).
    gauge=True : theta_rc = phi_r - phi_c, a diagonal unitary similarity of the real Laplacian, so the
                 spectrum is exactly that of ``laplacian_2d`` (lambda_max = 4 + 4 cos(pi/(g+1))).
    Phases are functions of the GLOBAL index (counter-based hash of seed and index), so any row block
    generated alone matches the full matrix.
    """
    n = g * g
    row_end = n if row_end is None else row_end
    r = np.arange(row_begin, row_end, dtype=np.int64)
    i, j = r // g, r % g

    def uniform_phase(idx, stream):
        # splitmix64-style counter hash -> [0, 2 pi)
        z = (idx.astype(np.uint64) * np.uint64(2) + np.uint64(stream)) * np.uint64(0x9E3779B97F4A7C15) \
            + np.uint64(seed) * np.uint64(0xD1B54A32D192ED03)
        z ^= z >> np.uint64(30)
        z *= np.uint64(0xBF58476D1CE4E5B9)
        z ^= z >> np.uint64(27)
        z *= np.uint64(0x94D049BB133111EB)
        z ^= z >> np.uint64(31)
        return (z >> np.uint64(11)).astype(np.float64) * (2.0 * math.pi / float(1 << 53))

    with np.errstate(over="ignore"):
        if gauge:
            phi = lambda idx: uniform_phase(idx, 0)
            th_right = lambda idx: phi(idx) - phi(idx + 1)   # pair (idx, idx+1)
            th_down = lambda idx: phi(idx) - phi(idx + g)    # pair (idx, idx+g)
        else:
            th_right = lambda idx: uniform_phase(idx, 0)
            th_down = lambda idx: uniform_phase(idx, 1)
        up = -np.exp(-1j * th_down(np.maximum(r - g, 0)))     # A(r, r-g) = conj(A(r-g, r))
        left = -np.exp(-1j * th_right(np.maximum(r - 1, 0)))  # A(r, r-1) = conj(A(r-1, r))
        right = -np.exp(1j * th_right(r))
        down = -np.exp(1j * th_down(r))
    offsets = [-g, -1, 0, 1, g]
    valid = [i > 0, j > 0, np.ones_like(i, bool), j < g - 1, i < g - 1]
    return _stencil_rows(offsets, valid, 4.0 + 0.0j, [up, left, None, right, down], row_begin, row_end,
                         np.complex128)


def csr_to_csc_hermitian(rowptr, colidx, values):
    """CSC arrays of a (structurally symmetric) Hermitian matrix given its CSR arrays: A^T = conj(A),
    so column j of A holds conj of row j.  For real symmetric input this is the identity."""
    return rowptr, colidx, np.conj(values) if np.iscomplexobj(values) else values


def dense_random_symmetric(n: int, seed: int = 7):
    """A = (R + R^T)/2 with R uniform in [-1, 1); returned column-major (Fortran order)."""
    rng = np.random.default_rng(seed)
    a = rng.random((n, n)) * 2.0 - 1.0
    a = 0.5 * (a + a.T)
    return np.asfortranarray(a)


def dense_random_hermitian(n: int, seed: int = 7):
    rng = np.random.default_rng(seed)
    a = (rng.random((n, n)) * 2.0 - 1.0) + 1j * (rng.random((n, n)) * 2.0 - 1.0)
    a = 0.5 * (a + a.conj().T)
    return np.asfortranarray(a)


def random_sparse_csr(n: int, mean_row: float, seed: int = 0, complex_values: bool = False, skew: float = 0.0,
                      diag_boost: float = 0.0):
    """Irregular test matrix: row lengths ~ Poisson(mean_row) (plus a heavy tail when skew > 0), random
    distinct columns, values uniform in [-1, 1); optional dominant diagonal entry on row 0 so the power
    method converges quickly.  Columns ascending in every row."""
    rng = np.random.default_rng(seed)
    lens = rng.poisson(mean_row, n).astype(np.int64)
    if skew > 0:
        heavy = rng.random(n) < 0.01
        lens[heavy] = (lens[heavy] + 1) * int(skew)
    lens = np.minimum(lens, n)
    rowptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens, out=rowptr[1:])
    colidx = np.empty(rowptr[-1], dtype=np.int32)
    for r in range(n):
        k = lens[r]
        if k:
            colidx[rowptr[r]:rowptr[r + 1]] = np.sort(rng.choice(n, size=k, replace=False)).astype(np.int32)
    nnz = int(rowptr[-1])
    vals = rng.random(nnz) * 2.0 - 1.0
    if complex_values:
        vals = vals + 1j * (rng.random(nnz) * 2.0 - 1.0)
    if diag_boost and lens[0] > 0:
        vals[0] = diag_boost
    return rowptr.astype(np.int32), colidx, vals


def banded_random_csr(n: int, mean_row: float, band: int, seed: int = 0, regular: bool = False):
    """Large synthetic test matrix for tuning the kernel dispatch (vectorised, no Python row loop).
    regular=True : every row has k = int(mean_row) entries at the same k sorted offsets inside the band
                   (a k-point stencil, clipped at the borders);
    regular=False: row lengths ~ Poisson(mean_row), columns uniform in [r-band, r+band] (duplicates are
                   allowed: a CSR row may hold the same column twice, the products simply add)."""
    rng = np.random.default_rng(seed)
    if regular:
        k = int(mean_row)
        offs = np.sort(rng.choice(np.arange(-band, band + 1), size=k, replace=False))
        rows = np.arange(n, dtype=np.int64)
        valid = [(rows + o >= 0) & (rows + o < n) for o in offs]
        vals = [rng.random(n) * 2 - 1 for _ in offs]
        return _stencil_rows(list(offs), valid, 1.0, vals, 0, n, np.float64)
    lens = rng.poisson(mean_row, n).astype(np.int64)
    rowptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens, out=rowptr[1:])
    total = int(rowptr[-1])
    row_of = np.repeat(np.arange(n, dtype=np.int64), lens)
    cols = np.clip(row_of + rng.integers(-band, band + 1, total), 0, n - 1)
    order = np.argsort(row_of * n + cols, kind="stable")
    cols = cols[order].astype(np.int32)
    vals = rng.random(total) * 2 - 1
    return rowptr.astype(np.int32), cols, vals
