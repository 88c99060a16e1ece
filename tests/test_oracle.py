"""Pins the oracle (oracle/eigsol_oracle.c): against the golden vectors produced by the compiled,
unmodified reference (tests/golden/), against that compiled reference itself when oracle/_ref is
present, and against the reference's own known-answer tests (test/power_method_test.cpp:38-134).
CPU only."""
import numpy as np
import pytest

import _golden
from _oracle import Oracle, Reference, have_reference, phase_aligned_error

CASES, CLOSE_TABLE = _golden.load()


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def run_oracle(orc, c, x0=None, max_it=None, tol=None):
    x0 = c.x0 if x0 is None else x0
    max_it = c.max_it if max_it is None else max_it
    tol = c.tol if tol is None else tol
    if c.sparse:
        return orc.power_csc((c.n, c.n), c.outer, c.inner, c.vals, x0, max_it, tol)
    return orc.power_dense(c.a, x0, max_it, tol)


@pytest.mark.parametrize("c", CASES, ids=[c.name for c in CASES])
def test_x0_bit_exact_with_reference(orc, c):
    # Vector<Scalar>::Random(n) + normalize() after srand(seed): integer bit assembly, so bit-exact
    # up to the normalisation (Eigen's vectorised squaredNorm vs a sequential sum: <= 2 ulp)
    x0 = orc.random_unit(c.n, c.dtype, seed=c.seed)
    assert np.max(np.abs(x0 - c.x0)) <= 4 * np.finfo(float).eps


@pytest.mark.parametrize("c", CASES, ids=[c.name for c in CASES])
def test_matvec_matches_reference(orc, c):
    y = orc.spmv_csc((c.n, c.n), c.outer, c.inner, c.vals, c.x0) if c.sparse else orc.gemv(c.a, c.x0)
    scale = max(np.max(np.abs(c.y0)), 1e-300)
    assert np.max(np.abs(y - c.y0)) <= 8 * np.finfo(float).eps * scale * max(c.n, 1) ** 0.5


@pytest.mark.parametrize("c", CASES, ids=[c.name for c in CASES])
def test_power_matches_reference_golden(orc, c):
    st, lam, x, iters, conv = run_oracle(orc, c)
    assert st == 0
    assert conv == c.conv
    if c.tol < 0 or not c.conv:
        assert iters == c.iters
    else:
        assert abs(iters - c.iters) <= 1
    assert abs(lam - c.lam) <= 1e-10 * max(abs(c.lam), 1e-300) or (c.lam == 0 and lam == 0)
    assert phase_aligned_error(x, c.x) <= 1e-8


def test_close_relative_truth_table(orc):
    for are, aim, bre, bim, tol, want_c, want_r in CLOSE_TABLE:
        assert orc.is_close_relative(complex(are, aim), complex(bre, bim), tol) == bool(want_c)
        assert orc.is_close_relative(are, bre, tol) == bool(want_r)


def test_error_order(orc):
    # power_method.hpp:52-58: square check first, then zero size
    st = orc.power_dense(np.zeros((2, 3), order="F"), np.zeros(2), 5, 1e-10)[0]
    assert st == 1
    st = orc.power_dense(np.zeros((0, 0), order="F"), np.zeros(1), 5, 1e-10)[0]
    assert st == 2


def test_reference_known_answers(orc):
    # test/power_method_test.cpp: diag(2,1) -> 2 @1e-5; [[3,1],[0,2]] -> 3 @1e-6; maxIterations=1 -> 1 trip
    a = np.asfortranarray(np.array([[2.0, 0.0], [0.0, 1.0]]))
    x0 = orc.random_unit(2, np.float64, seed=5)
    st, lam, x, iters, conv = orc.power_dense(a, x0, 1000, 1e-10)
    assert conv and iters > 0 and orc.is_close_relative(2.0, lam, 1e-5)
    assert np.allclose(a @ x, lam * x, atol=1e-5)
    outer = np.array([0, 1, 3], dtype=np.int32)
    inner = np.array([0, 0, 1], dtype=np.int32)
    vals = np.array([3.0, 1.0, 2.0])
    st, lam, x, iters, conv = orc.power_csc((2, 2), outer, inner, vals, x0, 1000, 1e-8)
    assert conv and orc.is_close_relative(3.0, lam, 1e-6)
    st, lam, x, iters, conv = orc.power_dense(a, x0, 1, 1e-10)
    assert iters == 1 and not conv


@pytest.mark.skipif(not have_reference(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_against_live_reference(orc):
    ref = Reference()
    rng = np.random.default_rng(3)
    for n, cplx, seed in [(7, False, 21), (9, True, 22), (40, False, 23), (33, True, 24)]:
        a = rng.random((n, n)) * 2 - 1
        if cplx:
            a = a + 1j * (rng.random((n, n)) * 2 - 1)
        a = np.asfortranarray(a + np.diag(np.full(n, 2.0 * n ** 0.5)))
        dt = a.dtype
        h = ref.create_dense(a)
        x0 = ref.random_unit(n, dt, seed)
        assert np.max(np.abs(orc.random_unit(n, dt, seed=seed) - x0)) <= 4 * np.finfo(float).eps
        for max_it, tol in [(30, -1.0), (1000, 1e-10)]:
            st, lam, x, iters, conv, _ = ref.power(h, dt, n, seed, max_it, tol)
            assert st == 0
            s2, lam2, x2, it2, cv2 = orc.power_dense(a, x0, max_it, tol)
            assert cv2 == conv and abs(it2 - iters) <= (0 if tol < 0 else 1)
            assert abs(lam2 - lam) <= 1e-10 * abs(lam)
            assert phase_aligned_error(x2, x) <= 1e-8
        # mismatch / shape errors of the reference itself, and that they consume no rand()
        st = ref.power(h, np.float64 if cplx else np.complex128, n, seed, 5, 1e-10)[0]
        assert st == 1 and "scalar type mismatch" in ref.last_error()
        ref.destroy(h)
    h = ref.create_dense(np.zeros((2, 3), order="F"))
    assert ref.power(h, np.float64, 2, 1, 5, 1e-10)[0] == 1 and "must be square" in ref.last_error()
    ref.destroy(h)
