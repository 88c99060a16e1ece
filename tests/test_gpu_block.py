"""GPU parity of the BLOCK power iteration (eigsol_b200_block_power_run), through the C ABI.
Checked against (a) the single-vector path / the oracle's restatement of the reference for block = 1 -- the block
path must reproduce powerMethod trip for trip -- (b) tests/_blocksim.py: the same k x k control code compiled
for the host + numpy products, on the same start block, and (c) numpy's eigenvalues of the test matrices.
Tolerances: Ritz values 1e-10 relative, Ritz vectors 1e-8 up to phase (the north star's), looser only where
the matrix itself is ill-conditioned (stated in the test)."""
import numpy as np
import pytest
import scipy.sparse as sp

import _blocksim as bs
import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen
from _oracle import Oracle, phase_aligned_error

pytestmark = pytest.mark.gpu


def lap2d(g):
    rp, ci, v = gen.laplacian_2d(g)
    return sp.csr_matrix((v, ci, rp), shape=(g * g, g * g))


def hermitian(n, seed, cplx=True, density=0.05):
    rng = np.random.default_rng(seed)
    a = sp.random(n, n, density=density, random_state=rng, format="csr")
    if cplx:
        a = a + 1j * sp.random(n, n, density=density, random_state=rng, format="csr")
    a = a + a.conj().T + sp.diags(np.linspace(1.0, 9.0, n))
    return a.tocsr()


def cpu(a, x0, nev, max_it, tol, sigma=1.0):
    return bs.block_power_cpu(lambda q: a @ q, x0, nev, max_it, tol, sigma)


def check_against_cpu(r, ref, a, nev, trips_exact=True, vec_tol=1e-8):
    th, x, it, conv, stop = ref
    if trips_exact:
        assert r.iterations == it
    else:
        assert abs(r.iterations - it) <= 1
    assert r.converged == conv and r.breakdown == (stop == 2)
    k = len(th)
    m = k if trips_exact else nev   # a one-trip difference still moves the unconverged tail of the block
    assert np.allclose(r.eigenvalues[:m], th[:m], rtol=1e-10, atol=1e-10 * abs(th[0]))
    for j in range(k):
        assert abs(np.linalg.norm(r.eigenvectors[:, j]) - 1.0) < 1e-10
    for j in range(nev):
        assert phase_aligned_error(r.eigenvectors[:, j], x[:, j]) <= vec_tol


@pytest.mark.parametrize("cplx", [False, True])
@pytest.mark.parametrize("dense", [False, True])
def test_block_of_one_reproduces_power_method(cplx, dense):
    n = 300
    a = hermitian(n, 3, cplx)
    M = es.Matrix(a.toarray() if dense else a.tocsc())
    dt = np.complex128 if cplx else np.float64
    x0 = es.random_unit(n, dt, seed=5)
    orc = Oracle()
    dm = es.DeviceMatrix.from_matrix(M)
    for max_it, tol in [(1, 1e-10), (25, -1.0), (2000, 1e-9)]:
        single = dm.power(x0, es.SolverOptions(max_it, tol))
        blk = dm.block_power(x0.reshape(n, 1), 1, es.SolverOptions(max_it, tol))
        _, lam, x, it, conv = orc.power_dense(a.toarray().astype(dt), x0, max_it, tol)
        exact = tol < 0 or max_it == 1   # near the threshold the two paths' rounding may move the stop by one trip
        assert abs(blk.iterations - single.iterations) <= (0 if exact else 1) and blk.converged == single.converged
        assert abs(blk.iterations - it) <= (0 if tol < 0 or max_it == 1 else 1) and blk.converged == conv
        assert abs(blk.eigenvalues[0] - lam) <= 1e-10 * abs(lam)
        assert abs(blk.eigenvalues[0] - single.eigenvalue) <= (1e-12 if exact else 1e-10) * abs(lam)
        assert phase_aligned_error(blk.eigenvectors[:, 0], x) <= 1e-8
    dm.close()


@pytest.mark.parametrize("k", [2, 4, 8])
@pytest.mark.parametrize("kind", ["real-sparse", "complex-sparse", "real-dense", "complex-dense"])
def test_fixed_trips_match_the_cpu_restatement(k, kind):
    cplx, dense = kind.startswith("complex"), kind.endswith("dense")
    n = 301 if dense else 700   # odd n: the dense kernel's two-rows-per-thread tail
    a = hermitian(n, 10 + k, cplx)
    dt = np.complex128 if cplx else np.float64
    x0 = es.random_block(n, k, dt, seed=2)
    dm = es.DeviceMatrix.from_matrix(es.Matrix(a.toarray() if dense else a.tocsc()))
    r = dm.block_power(x0, k, es.SolverOptions(40, -1.0))
    ref = cpu(a, x0, k, 40, -1.0)
    assert r.iterations == 40 and not r.converged and not r.breakdown
    assert np.allclose(r.eigenvalues, ref[0], rtol=1e-10)
    # the block spans the same subspace: projector difference
    q1, _ = np.linalg.qr(r.eigenvectors)
    q2, _ = np.linalg.qr(ref[1])
    assert np.linalg.norm(q1 @ (q1.conj().T @ q2) - q2) < 1e-8
    assert np.allclose(np.linalg.norm(r.eigenvectors, axis=0), 1.0, atol=1e-10)
    dm.close()


@pytest.mark.parametrize("k,nev", [(2, 1), (4, 2), (8, 4)])
def test_converged_run_finds_the_dominant_eigenpairs(k, nev):
    g = 20
    a = (lap2d(g) + sp.diags(np.linspace(0.0, 6.0, g * g) ** 2)).tocsr()   # well separated top of the spectrum
    n = g * g
    x0 = es.random_block(n, k, np.float64, seed=7)
    dm = es.DeviceMatrix.from_matrix(es.Matrix(a.tocsc()))
    r = dm.block_power(x0, nev, es.SolverOptions(5000, 1e-12))
    ref = cpu(a, x0, nev, 5000, 1e-12)
    # the test is on the eigenvalue (error ~ vector error squared): a one-trip difference moves vectors by ~1e-6
    check_against_cpu(r, ref, a, nev, trips_exact=False, vec_tol=1e-4)
    top = np.sort(np.linalg.eigvalsh(a.toarray()))[::-1]
    assert r.converged and np.allclose(r.eigenvalues[:nev].real, top[:nev], rtol=1e-9)
    for j in range(nev):
        v = r.eigenvectors[:, j]
        assert np.linalg.norm(a @ v - r.eigenvalues[j].real * v) < 1e-4
    single = dm.power(x0[:, 0].copy(), es.SolverOptions(5000, 1e-12))
    assert r.iterations < single.iterations    # convergence ratio |l_{k+1}/l_1| instead of |l_2/l_1|
    dm.close()


def test_nonsymmetric_complex_dense_known_spectrum():
    rng = np.random.default_rng(21)
    n, k = 120, 4
    d = np.concatenate([[10 + 3j, -9 + 1j, 8j, 7.0], 3.0 * (rng.standard_normal(n - 4) + 1j * rng.standard_normal(n - 4)) / 3])
    d[4:] *= 4.0 / np.abs(d[4:]).max()
    p = rng.standard_normal((n, n)) + 1j * rng.standard_normal((n, n))
    a = p @ np.diag(d) @ np.linalg.inv(p)
    x0 = es.random_block(n, k, np.complex128, seed=3)
    r = es.blockPowerMethod(es.Matrix(a), k, 4, es.SolverOptions(4000, 1e-12), x0)
    assert r.converged and np.allclose(r.eigenvalues, d[:4], rtol=1e-8)
    for j in range(4):
        v = r.eigenvectors[:, j]
        assert np.linalg.norm(a @ v - r.eigenvalues[j] * v) < 1e-6 * np.linalg.cond(p)


def test_real_matrix_with_a_dominant_complex_pair():
    """dtype f64, Ritz values 3 +- 4i: columns 0, 1 hold Re and Im of the pair's vector (dgeev convention)."""
    rng = np.random.default_rng(4)
    n = 60
    blocks = np.zeros((n, n))
    blocks[:2, :2] = [[3.0, -4.0], [4.0, 3.0]]
    blocks[2:, 2:] = np.diag(np.linspace(-2.0, 2.5, n - 2))
    p = rng.standard_normal((n, n))
    a = p @ blocks @ np.linalg.inv(p)
    x0 = es.random_block(n, 2, np.float64, seed=9)
    r = es.blockPowerMethod(es.Matrix(a), 2, 2, es.SolverOptions(3000, 1e-12), x0)
    assert r.converged
    assert np.allclose(r.eigenvalues, [3 + 4j, 3 - 4j], rtol=1e-9)
    v = r.eigenvectors[:, 0] + 1j * r.eigenvectors[:, 1]
    assert np.linalg.norm(a @ v - (3 + 4j) * v) < 1e-6 * np.linalg.cond(p)
    assert abs(np.linalg.norm(v) - 1.0) < 1e-9


def test_max_trips_zero_breakdown_and_argument_errors():
    n = 50
    a = sp.diags(np.linspace(1.0, 3.0, n)).tocsc()
    dm = es.DeviceMatrix.from_matrix(es.Matrix(a))
    x0 = es.random_block(n, 2, np.float64, seed=1)
    r = dm.block_power(x0, 2, es.SolverOptions(0, 1e-10))
    assert r.iterations == 0 and not r.converged and not r.breakdown and np.allclose(r.eigenvalues, 0)
    assert np.allclose(r.eigenvectors.T @ r.eigenvectors, np.eye(2), atol=1e-12)   # orth(x0)
    same = np.ones((n, 2))
    r = dm.block_power(same, 2, es.SolverOptions(10, 1e-10))
    assert r.breakdown and not r.converged and r.iterations == 0
    with pytest.raises(ValueError, match="block width must be 1, 2, 4 or 8"):
        dm.block_power(np.ones((n, 3)), 3)
    with pytest.raises(ValueError, match="nev must satisfy"):
        dm.block_power(x0, 3)
    dm.close()
    z = es.DeviceMatrix.from_matrix(es.Matrix(sp.csc_matrix((np.array([0.0]), (np.array([0]), np.array([0]))), shape=(n, n))))
    r = z.block_power(x0, 2, es.SolverOptions(10, 1e-10))
    assert r.breakdown and not r.converged   # A Q = 0: the block counterpart of normY == 0 (power_method.hpp:73-76)
    z.close()


@pytest.mark.parametrize("scale", [1e150, 1e-150])
def test_extreme_scales_use_the_prescale(scale):
    a = hermitian(200, 6, False)
    x0 = es.random_block(200, 4, np.float64, seed=4)
    dm = es.DeviceMatrix.from_matrix(es.Matrix(a.tocsc()))
    base = dm.block_power(x0, 2, es.SolverOptions(60, -1.0))
    dm.close()
    dm = es.DeviceMatrix.from_matrix(es.Matrix((a * scale).tocsc()))
    r = dm.block_power(x0, 2, es.SolverOptions(60, -1.0))
    dm.close()
    assert not r.breakdown and np.all(np.isfinite(r.eigenvalues.view(np.float64)))
    assert np.allclose(r.eigenvalues / scale, base.eigenvalues, rtol=1e-10)


def test_medium_3d_laplacian_block_of_four():
    """32^3 grid, fixed 30 trips: Ritz values against the CPU restatement (scipy products), repeated solve identical."""
    g = 32
    n = g ** 3
    rp, ci, v = gen.laplacian_3d(g, 0, n)
    a = sp.csr_matrix((v, ci, rp), shape=(n, n))
    x0 = es.random_block(n, 4, np.float64, seed=1)
    dm = es.DeviceMatrix.from_csr_rows(n, 0, rp, ci, v)
    r = dm.block_power(x0, 4, es.SolverOptions(30, -1.0))
    ref = cpu(a, x0, 4, 30, -1.0)
    assert r.iterations == 30 and np.allclose(r.eigenvalues, ref[0], rtol=1e-10)
    r2 = dm.block_power(x0, 4, es.SolverOptions(30, -1.0))
    assert np.array_equal(r.eigenvalues, r2.eigenvalues) and np.array_equal(r.eigenvectors, r2.eigenvectors)  # deterministic
    single = dm.power(x0[:, 0].copy(), es.SolverOptions(30, -1.0))   # the handle still serves the single-vector path
    assert single.iterations == 30
    dm.close()
