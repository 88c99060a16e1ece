"""The k x k algebra and the control step of the block power iteration (csrc/blockmath.hpp),
compiled for the host and checked against numpy; and the whole block iteration (control step + numpy products)
against the reference's powerMethod for k = 1 and against numpy's eigenvalues for k > 1."""
import numpy as np
import pytest

import _blocksim as bs
from _oracle import Oracle, phase_aligned_error


def rnd(rng, k, cplx=True):
    a = rng.standard_normal((k, k))
    return a + 1j * rng.standard_normal((k, k)) if cplx else a.astype(np.complex128)


@pytest.mark.parametrize("k", range(1, 9))
def test_cholesky_and_triangular_inverse(k):
    rng = np.random.default_rng(k)
    u = rng.standard_normal((50, k)) + 1j * rng.standard_normal((50, k))
    w = u.conj().T @ u
    ok, r = bs.chol(np.triu(w))  # only the upper triangle is read
    assert ok and np.allclose(np.tril(r, -1), 0) and np.all(r.diagonal().real > 0)
    assert np.allclose(r.conj().T @ r, w, rtol=1e-13, atol=1e-12)
    t = bs.triinv(r)
    assert np.allclose(t @ r, np.eye(k), atol=1e-12)
    q = u @ t
    assert np.allclose(q.conj().T @ q, np.eye(k), atol=1e-12)


def test_cholesky_reports_rank_loss():
    u = np.ones((10, 2), dtype=np.complex128)
    assert not bs.chol(u.conj().T @ u)[0]
    assert not bs.chol(np.zeros((3, 3)))[0]
    assert not bs.chol(np.full((2, 2), np.nan))[0]


def sort_like_device(vals):
    return sorted(vals, key=lambda z: (-round(abs(z), 9), -round(z.imag, 9), -round(z.real, 9)))


@pytest.mark.parametrize("k", range(1, 9))
@pytest.mark.parametrize("kind", ["complex", "real", "hermitian", "triangular", "defective-ish"])
def test_eig_general_against_numpy(k, kind):
    rng = np.random.default_rng(100 * k + len(kind))
    for _ in range(20):
        h = rnd(rng, k, kind != "real")
        if kind == "hermitian":
            h = h + h.conj().T
        elif kind == "triangular":
            h = np.triu(h)
        elif kind == "defective-ish":
            h = np.triu(h)
            np.fill_diagonal(h, 1.0 + 1e-9 * np.arange(k))
        ok, th, s = bs.eig(h)
        assert ok
        ref = np.linalg.eigvals(h)
        # same spectrum (matched greedily)
        left = list(ref)
        for t in th:
            j = int(np.argmin([abs(t - r) for r in left]))
            assert abs(t - left[j]) <= 1e-8 * max(1.0, abs(t)) if kind == "defective-ish" else abs(t - left[j]) <= 1e-11 * max(1.0, np.abs(ref).max())
            left.pop(j)
        # ordered by modulus
        mods = np.abs(th)
        assert np.all(mods[:-1] >= mods[1:] * (1 - 1e-9))
        if kind != "defective-ish":
            for j in range(k):
                v = s[:, j]
                assert abs(np.linalg.norm(v) - 1) < 1e-12
                assert np.linalg.norm(h @ v - th[j] * v) <= 1e-9 * max(1.0, np.abs(ref).max()) * max(1.0, np.linalg.cond(np.linalg.eig(h)[1]))
                big = v[np.argmax(np.abs(v))]
                assert abs(big.imag) < 1e-12 and big.real > 0


def test_schur_is_unitary_and_triangular():
    rng = np.random.default_rng(5)
    for k in (2, 3, 5, 8):
        h = rnd(rng, k)
        ok, t, z = bs.schur(h)
        assert ok and np.allclose(np.tril(t, -1), 0)
        assert np.allclose(z.conj().T @ z, np.eye(k), atol=1e-13)
        assert np.allclose(z @ t @ z.conj().T, h, atol=1e-12)


def test_real_matrix_conjugate_pairs_keep_a_stable_order():
    h = np.array([[0.0, -2.0, 0], [2.0, 0.0, 0], [0, 0, 1.0]])
    ok, th, _ = bs.eig(h)
    assert ok and np.allclose(th, [2j, -2j, 1.0])
    ok, th, _ = bs.eig(h + 1e-14 * np.arange(9).reshape(3, 3))
    assert ok and th[0].imag > 0 and th[1].imag < 0


# ---- the whole iteration on the CPU ------------------------------------------------------------------
def laplacian_1d2d(g):
    n = g * g
    a = np.zeros((n, n))
    for i in range(g):
        for j in range(g):
            r = i * g + j
            a[r, r] = 4
            if j: a[r, r - 1] = -1
            if j < g - 1: a[r, r + 1] = -1
            if i: a[r, r - g] = -1
            if i < g - 1: a[r, r + g] = -1
    return a


@pytest.mark.parametrize("cplx", [False, True])
def test_block_of_one_is_the_reference_power_method(cplx):
    """k = 1: same trips, flag, eigenvalue and vector as src/power_method/power_method.hpp:47-99 (oracle)."""
    rng = np.random.default_rng(3)
    n = 24
    a = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    a = a + a.conj().T + np.diag(np.arange(n) * 1.5)
    x0 = rng.standard_normal(n) + (1j * rng.standard_normal(n) if cplx else 0)
    x0 /= np.linalg.norm(x0)
    o = Oracle()
    for max_it, tol in [(1, 1e-10), (7, -1.0), (1000, 1e-10), (1000, 1e-6)]:
        _, lam, x, it, conv = o.power_dense(a, x0, max_it, tol)
        th, xb, itb, convb, _ = bs.block_power_cpu(lambda q: a @ q, x0.reshape(n, 1), 1, max_it, tol)
        assert (itb, convb) == (it, conv)
        assert abs(th[0] - lam) <= 1e-12 * abs(lam)
        assert phase_aligned_error(xb[:, 0], x) < 1e-10


@pytest.mark.parametrize("k,nev", [(2, 1), (4, 2), (4, 4), (8, 3)])
def test_block_iteration_finds_the_dominant_eigenvalues(k, nev):
    a = laplacian_1d2d(7) + np.diag(np.linspace(0, 3, 49) ** 2)
    rng = np.random.default_rng(k)
    x0 = rng.standard_normal((49, k))
    th, x, it, conv, stop = bs.block_power_cpu(lambda q: a @ q, x0, nev, 5000, 1e-13)
    assert conv and stop == 1
    ref = np.sort(np.linalg.eigvalsh(a))[::-1]
    assert np.allclose(th[:nev].real, ref[:nev], rtol=1e-9) and np.allclose(th.imag, 0, atol=1e-9)
    for j in range(nev):
        assert abs(np.linalg.norm(x[:, j]) - 1) < 1e-10
        assert np.linalg.norm(a @ x[:, j] - th[j] * x[:, j]) < 1e-4
    # the block converges in far fewer trips than the single vector (ratio |l_{k+1}/l_1| instead of |l_2/l_1|)
    _, _, it1, conv1, _ = bs.block_power_cpu(lambda q: a @ q, x0[:, :1], 1, 5000, 1e-13)
    assert conv1 and (nev > 1 or it < it1)


def test_block_iteration_nonsymmetric_complex_and_prescale():
    rng = np.random.default_rng(9)
    n, k = 30, 3
    d = np.array([9 + 2j, -8.5 + 1j, 7j] + list(rng.standard_normal(n - 3) + 1j * rng.standard_normal(n - 3)))
    p = rng.standard_normal((n, n)) + 1j * rng.standard_normal((n, n))
    a = p @ np.diag(d) @ np.linalg.inv(p)
    x0 = rng.standard_normal((n, k)) + 1j * rng.standard_normal((n, k))
    th, x, it, conv, _ = bs.block_power_cpu(lambda q: a @ q, x0, 3, 3000, 1e-13)
    assert conv and np.allclose(th, d[:3], rtol=1e-8)
    for j in range(3):
        assert np.linalg.norm(a @ x[:, j] - th[j] * x[:, j]) < 1e-5
    big = a * 2.0 ** 500
    th2, _, it2, conv2, _ = bs.block_power_cpu(lambda q: big @ q, x0, 3, 3000, 1e-13, sigma=2.0 ** -500)
    assert conv2 and it2 == it and np.allclose(th2 / 2.0 ** 500, th, rtol=1e-12)


def test_block_iteration_max_trips_zero_and_breakdown():
    a = np.diag([3.0, 2.0, 1.0, 0.5])
    x0 = np.random.default_rng(1).standard_normal((4, 2))
    th, x, it, conv, stop = bs.block_power_cpu(lambda q: a @ q, x0, 2, 0, 1e-10)
    assert it == 0 and not conv and stop == 3 and np.allclose(th, 0)
    assert np.allclose(x.conj().T @ x, np.eye(2), atol=1e-12)  # orth(X0), as the reference returns x0 normalised
    th, x, it, conv, stop = bs.block_power_cpu(lambda q: a @ q, x0, 2, 5, -1.0)
    assert it == 5 and not conv and stop == 3
    z = np.zeros((4, 4))
    th, x, it, conv, stop = bs.block_power_cpu(lambda q: z @ q, x0, 2, 5, 1e-10)
    assert stop == 2 and not conv  # A Q = 0: the block counterpart of normY == 0
    same = np.ones((4, 2))
    assert bs.block_power_cpu(lambda q: a @ q, same, 2, 5, 1e-10)[4] == 2  # rank-deficient start block


# ---- edge cases of the k x k eigen-solver (it runs unattended on the device between two launches) -------------
def test_eig_edge_cases_zero_identity_jordan_nan():
    ok, th, s = bs.eig(np.zeros((4, 4)))
    assert ok and np.all(th == 0) and np.allclose(np.abs(s).max(axis=0), 1.0)
    ok, th, s = bs.eig(np.eye(5) * (2 - 3j))
    assert ok and np.allclose(th, 2 - 3j) and np.allclose(s.conj().T @ s, np.eye(5), atol=1e-14)
    j = np.diag(np.full(6, 1.5)) + np.diag(np.ones(5), 1)          # one Jordan block: defective
    ok, th, s = bs.eig(j)
    assert ok and np.allclose(th, 1.5, atol=1e-2)                  # eigenvalues of a 6 x 6 Jordan block move by eps^(1/6)
    for bad in (np.nan, np.inf):
        h = np.eye(3)
        h[1, 2] = bad
        assert not bs.eig(h)[0]


@pytest.mark.parametrize("scale", [1e300, 1e-300, 2.0 ** 1000, 2.0 ** -1000])
def test_eig_is_scale_invariant(scale):
    rng = np.random.default_rng(8)
    h = rnd(rng, 6)
    ok0, th0, s0 = bs.eig(h)
    ok1, th1, s1 = bs.eig(h * scale)
    assert ok0 and ok1
    if scale in (2.0 ** 1000, 2.0 ** -1000):   # a power of two inside the clamp: bit-identical values
        assert np.array_equal(th1 / scale, th0) and np.array_equal(s1, s0)
    else:
        assert np.allclose(th1 / scale, th0, rtol=1e-12) and np.allclose(s1, s0, atol=1e-9)


@pytest.mark.parametrize("k", [2, 4, 8])
def test_eig_clustered_and_graded_spectra(k):
    rng = np.random.default_rng(40 + k)
    for _ in range(30):
        q, _ = np.linalg.qr(rnd(rng, k))
        d = np.concatenate([[1.0, 1.0 + 1e-13], 10.0 ** -rng.integers(0, 12, k - 2)])[:k] * np.exp(1j * rng.uniform(0, 6.28, k))
        h = q @ np.diag(d) @ q.conj().T                           # normal matrix: eigenvalues are well conditioned
        ok, th, s = bs.eig(h)
        assert ok
        ref = sorted(d, key=abs, reverse=True)
        assert np.allclose(np.sort(np.abs(th))[::-1], np.abs(ref), rtol=1e-9, atol=1e-13)
        for j in range(k):
            assert np.linalg.norm(h @ s[:, j] - th[j] * s[:, j]) < 1e-12


def test_values_only_path_matches_full_decomposition():
    """The loop uses eig_values_coop (no vectors, window-confined rotations); the result step uses eig_general.
    block_power_cpu exercises the former, bs.eig the latter: the Ritz values must agree bit for bit."""
    rng = np.random.default_rng(77)
    a = rng.standard_normal((40, 40)) + 1j * rng.standard_normal((40, 40))
    x0 = rng.standard_normal((40, 4)) + 1j * rng.standard_normal((40, 4))
    th, x, it, conv, stop = bs.block_power_cpu(lambda q: a @ q, x0, 4, 25, -1.0)
    q, _ = np.linalg.qr(x)                                          # the Ritz vectors span the final Q
    # recompute H of the final launch from the returned vectors' span and compare spectra (order-insensitive, loose):
    h = q.conj().T @ a @ q
    assert np.allclose(np.sort_complex(np.linalg.eigvals(h)), np.sort_complex(th), rtol=1e-8)
    for j in range(4):
        assert abs(np.linalg.norm(x[:, j]) - 1) < 1e-12


# ---- host-side behaviour of the Python entry point (no GPU needed) ---------------------------------------------
def test_block_power_method_checks_come_before_any_device_work():
    import pcsc_eigenvalue_solver_project_b200 as es
    with pytest.raises(RuntimeError, match="powerMethod: matrix must be square"):      # power_method.hpp:52-55
        es.blockPowerMethod(es.Matrix(np.zeros((2, 3))), 2)
    with pytest.raises(RuntimeError, match="powerMethod: matrix has zero size"):       # power_method.hpp:56-58
        es.blockPowerMethod(es.Matrix(np.zeros((0, 0))), 2)


def test_random_block_columns_are_the_reference_start_vectors():
    import pcsc_eigenvalue_solver_project_b200 as es
    xb = es.random_block(50, 4, np.float64, seed=3)
    assert xb.shape == (50, 4) and xb.flags.f_contiguous
    for j in range(4):
        assert np.array_equal(xb[:, j], es.random_unit(50, np.float64, seed=3 + j))   # power_method.hpp:62 after srand(seed + j)
        assert abs(np.linalg.norm(xb[:, j]) - 1) < 1e-14
    zb = es.random_block(20, 2, np.complex128, seed=1)
    assert zb.dtype == np.complex128 and np.linalg.matrix_rank(zb) == 2
