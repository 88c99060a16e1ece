"""Single-process multi-GPU group (EIGSOL_B200_NGPU=P, SURVEY.md 8e "Process model"): the reference-facing call
powerMethod(Matrix, SolverOptions) and the comm == NULL handle API shard themselves over P GPUs inside ONE process.
Same parity bar as the one-GPU path: eigenvalue within 1e-10 relative, eigenvector within 1e-8 up to phase, against
the oracle on the whole matrix.  Needs >= 2 devices (skipped on the single-GPU driver run); the host-side N > 1
logic is covered on CPU by tests/test_multi_rank_gloo.py and tests/test_host_plan.py."""
import numpy as np
import pytest

import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen
from _oracle import Oracle, phase_aligned_error

pytestmark = pytest.mark.gpu

LAMBDA_RTOL = 1e-10
VECTOR_ATOL = 1e-8


@pytest.fixture(scope="module")
def orc():
    return Oracle()


@pytest.fixture(params=[2, 4, 8])
def ngpu(request, monkeypatch):
    if es.device_count() < request.param:
        pytest.skip(f"needs {request.param} GPUs")
    monkeypatch.setenv("EIGSOL_B200_NGPU", str(request.param))
    monkeypatch.setenv("EIGSOL_B200_MIN_ROWS_PER_GPU", "1")
    monkeypatch.setenv("EIGSOL_B200_PEER_TIMEOUT_MS", "2000")
    return request.param


def check(r, ref, fixed=True):
    st, lam, x, iters, conv = ref
    assert st == 0
    assert r.iterations == iters if fixed else abs(r.iterations - iters) <= 1
    assert r.converged == conv
    assert abs(r.eigenvalue - lam) <= LAMBDA_RTOL * abs(lam), (r.eigenvalue, lam)
    assert phase_aligned_error(r.eigenvector, x) <= VECTOR_ATOL


def general_csc(n, rowptr, colidx, vals):
    import scipy.sparse as sp
    a = sp.csr_matrix((vals, colidx, rowptr), shape=(n, n)).tocsc()
    a.sort_indices()
    return a.indptr.astype(np.int32), a.indices.astype(np.int32), np.ascontiguousarray(a.data)


def test_power_method_csc_laplacian_3d(orc, ngpu):
    """The reference-facing one-shot call on Eigen's CSC arrays: column-slab upload, cross-GPU pull, fused kernel."""
    g = 24
    n = g ** 3
    rp, ci, v = gen.laplacian_3d(g)
    outer, inner, cv = gen.csr_to_csc_hermitian(rp, ci, v)
    x0 = orc.random_unit(n, np.float64, seed=1)
    M = es.Matrix(((n, n), outer, inner, cv))
    check(es.powerMethod(M, es.SolverOptions(60, -1.0), x0=x0), orc.power_csc((n, n), outer, inner, cv, x0, 60, -1.0))
    dm = es.DeviceMatrix.from_matrix(M)
    info = dm.info()
    assert info["nranks"] == ngpu and info["exchange"] == 2 and info["kernels_per_trip"] == 1, info
    y = dm.matvec(x0)
    yref = orc.spmv_csc((n, n), outer, inner, cv, x0)
    assert np.max(np.abs(y - yref)) <= 1e-14 * np.max(np.abs(yref))
    # a second solve on the resident group, then every kernel variant
    check(dm.power(x0, es.SolverOptions(25, -1.0)), orc.power_csc((n, n), outer, inner, cv, x0, 25, -1.0))
    for kernel, lanes in [(es.KERNEL_SELL, 2), (es.KERNEL_CSR_STREAM, 0), (es.KERNEL_CSR_VECTOR, 4)]:
        dm.set_kernel(kernel, lanes)
        check(dm.power(x0, es.SolverOptions(20, -1.0)), orc.power_csc((n, n), outer, inner, cv, x0, 20, -1.0))
    dm.close()


@pytest.mark.parametrize("cplx", [False, True])
def test_power_method_csc_irregular_nonsymmetric(orc, ngpu, cplx):
    """Unsymmetric pattern, skewed row lengths: the column-slab boundaries are reused as the row partition, rows pull
    entries from EVERY slab, the halo is scattered (list sends).  Real tolerance: converged flag and trip count."""
    n = 3000
    rp, ci, v = gen.random_sparse_csr(n, 6.0, seed=4, complex_values=cplx, skew=40.0, diag_boost=60.0)
    outer, inner, cv = general_csc(n, rp, ci, v)
    dt = np.complex128 if cplx else np.float64
    x0 = orc.random_unit(n, dt, seed=2)
    M = es.Matrix(((n, n), outer, inner, cv))
    check(es.powerMethod(M, es.SolverOptions(40, -1.0), x0=x0), orc.power_csc((n, n), outer, inner, cv, x0, 40, -1.0))
    check(es.powerMethod(M, es.SolverOptions(1000, 1e-10), x0=x0), orc.power_csc((n, n), outer, inner, cv, x0, 1000, 1e-10),
          fixed=False)


def test_power_method_dense(orc, ngpu):
    n = 700
    a = gen.dense_random_hermitian(n, seed=3)
    x0 = orc.random_unit(n, np.complex128, seed=2)
    check(es.powerMethod(es.Matrix(a), es.SolverOptions(40, -1.0), x0=x0), orc.power_dense(a, x0, 40, -1.0))
    a = gen.dense_random_symmetric(513, seed=5)
    x0 = orc.random_unit(513, np.float64, seed=6)
    check(es.powerMethod(es.Matrix(a), es.SolverOptions(30, -1.0), x0=x0), orc.power_dense(a, x0, 30, -1.0))


def test_csr_rows_entry_and_stepwise_api(orc, ngpu):
    """comm == NULL handle from whole-matrix CSR rows: nnz-balanced blocks uploaded by the ranks' own threads."""
    g = 40
    n = g * g
    rp, ci, v = gen.hermitian_laplacian_2d(g, 7, False)
    outer, inner, cv = gen.csr_to_csc_hermitian(rp, ci, v)
    x0 = orc.random_unit(n, np.complex128, seed=3)
    dm = es.DeviceMatrix.from_csr_rows(n, 0, rp, ci, v)
    assert dm.info()["nranks"] == ngpu
    dm.begin(x0, 30, -1.0)
    ms = dm.step(10) + dm.step(25)  # more launches than trips: the extra ones are no-ops
    assert ms > 0.0
    check(dm.end(), orc.power_csc((n, n), outer, inner, cv, x0, 30, -1.0))
    dot, nrm2 = dm.rayleigh(x0)
    yref = orc.spmv_csc((n, n), outer, inner, cv, x0)
    assert abs(dot - np.vdot(x0, yref)) <= 1e-12 * abs(np.vdot(x0, yref))
    assert abs(nrm2 - np.vdot(yref, yref).real) <= 1e-12 * np.vdot(yref, yref).real
    dm.close()


def test_tiny_matrices_and_edge_cases(orc, ngpu):
    """n smaller than the GPU count, zero iterations, the ||y|| == 0 exit (power_method.hpp:73-76) and the
    reference's error order -- all through a group."""
    a = np.array([[2.0, 0.0], [0.0, 1.0]])
    x0 = orc.random_unit(2, np.float64, seed=42)
    check(es.powerMethod(es.Matrix(a), es.SolverOptions(1000, 1e-10), x0=x0), orc.power_dense(a, x0, 1000, 1e-10), fixed=False)
    r = es.powerMethod(es.Matrix(a), es.SolverOptions(0, 1e-10), x0=x0)
    assert r.iterations == 0 and not r.converged and r.eigenvalue == 0.0 and np.allclose(r.eigenvector, x0)
    nil = np.array([[0.0, 1.0, 0.0], [0.0, 0.0, 1.0], [0.0, 0.0, 0.0]])
    x0 = orc.random_unit(3, np.float64, seed=5)
    check(es.powerMethod(es.Matrix(nil), es.SolverOptions(10, 1e-10), x0=x0), orc.power_dense(nil, x0, 10, 1e-10))
    with pytest.raises(RuntimeError, match="powerMethod: matrix must be square"):
        es.powerMethod(es.Matrix(np.zeros((2, 3))))


def test_group_is_the_default_row_partition_above_min_rows(orc, monkeypatch):
    """Without EIGSOL_B200_MIN_ROWS_PER_GPU a small matrix stays on one GPU even with EIGSOL_B200_NGPU set."""
    if es.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    monkeypatch.setenv("EIGSOL_B200_NGPU", "2")
    monkeypatch.delenv("EIGSOL_B200_MIN_ROWS_PER_GPU", raising=False)
    rp, ci, v = gen.laplacian_2d(16)
    dm = es.DeviceMatrix.from_matrix(es.Matrix(((256, 256), rp, ci, v)))
    assert dm.info()["nranks"] == 1
    dm.close()
    rp, ci, v = gen.laplacian_2d(64)
    dm = es.DeviceMatrix.from_matrix(es.Matrix(((4096, 4096), rp, ci, v)))
    assert dm.info()["nranks"] == 2
    dm.close()


@pytest.mark.parametrize("k", [1, 2, 4, 8])
@pytest.mark.parametrize("cplx", [False, True])
def test_block_power_iteration_on_a_group(orc, ngpu, k, cplx):
    """Block power iteration (k vectors per pass) row-sharded over the group: halo rows pushed over NVLink after every
    SpMM, Gram sums all-reduced in the control launch.  Fixed trips against the CPU restatement of the control step
    (tests/_blocksim.py) on the whole matrix, and -- k = 1 -- against the reference oracle."""
    import scipy.sparse as sp
    import _blocksim as bs
    g = 14 if cplx else 22
    n = g ** 3 if not cplx else g * g
    if cplx:
        rp, ci, v = gen.hermitian_laplacian_2d(g, 11, False)
    else:
        rp, ci, v = gen.laplacian_3d(g)
    a = sp.csr_matrix((v, ci, rp), shape=(n, n))
    dt = np.complex128 if cplx else np.float64
    x0 = es.random_block(n, k, dt, seed=3)
    dm = es.DeviceMatrix.from_matrix(es.Matrix(a.tocsc()))
    assert dm.info()["nranks"] == ngpu
    r = dm.block_power(x0, k, es.SolverOptions(30, -1.0))
    dm.close()
    th, X, it, conv, _ = bs.block_power_cpu(lambda q: a @ q, x0, k, 30, -1.0)
    assert r.iterations == it == 30 and not r.converged and not r.breakdown
    assert np.allclose(r.eigenvalues, th, rtol=1e-10), (r.eigenvalues, th)
    q1, _ = np.linalg.qr(r.eigenvectors)
    q2, _ = np.linalg.qr(X)
    assert np.linalg.norm(q1 @ (q1.conj().T @ q2) - q2) < 1e-8
    if k == 1:
        outer, inner, cv = gen.csr_to_csc_hermitian(rp, ci, v)
        st, lam, x, iters, _ = orc.power_csc((n, n), outer, inner, cv, np.ascontiguousarray(x0[:, 0]), 30, -1.0)
        assert st == 0 and iters == 30
        assert abs(r.eigenvalues[0] - lam) <= LAMBDA_RTOL * abs(lam)
        assert phase_aligned_error(r.eigenvectors[:, 0], x) <= VECTOR_ATOL


def test_group_rejects_malformed_input_and_recovers(orc, ngpu):
    """A failure inside one rank's thread (row index out of range found on the device) comes back as the same error a
    single GPU gives -- the other ranks leave their barriers instead of hanging -- and the group's workers are usable
    for the next matrix."""
    g = 12
    n = g ** 3
    rp, ci, v = gen.laplacian_3d(g)
    outer, inner, cv = gen.csr_to_csc_hermitian(rp, ci, v)
    bad = inner.copy()
    bad[len(bad) // 2] = n + 7
    with pytest.raises(ValueError, match="inner index out of range"):
        es.DeviceMatrix.from_matrix(es.Matrix(((n, n), outer, bad, cv)))
    bad_outer = outer.copy()
    bad_outer[n // 3], bad_outer[n // 3 + 1] = bad_outer[n // 3 + 1], bad_outer[n // 3]   # not monotone inside one slab
    with pytest.raises(ValueError, match="outer index not monotone"):
        es.DeviceMatrix.from_matrix(es.Matrix(((n, n), bad_outer, inner, cv)))
    x0 = orc.random_unit(n, np.float64, seed=9)
    check(es.powerMethod(es.Matrix(((n, n), outer, inner, cv)), es.SolverOptions(25, -1.0), x0=x0),
          orc.power_csc((n, n), outer, inner, cv, x0, 25, -1.0))
