"""The device's control state machine, driven on the HOST (no GPU): tests/statemachine/sim.cu compiles the
`__host__ __device__` decision functions of csrc/kernels.cuh (control_wants_fold / control_apply / fold_launch /
next_sigma) and runs them through the engine's launch pipeline with a plain C++ matvec.  Checked against the
oracle's restatement of src/power_method/power_method.hpp:47-99 on the same inputs: the one-launch lag of the
convergence test, the three-buffer ring, the power-of-two rescaling, maxIterations, the ||y|| == 0 exit and the
host's poll cadence must give the reference's iteration count, flag, eigenvalue and vector."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest

from _oracle import Oracle, phase_aligned_error

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "statemachine")


@pytest.fixture(scope="module")
def sim():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    so = os.path.join(HERE, "libsim.so")
    src = os.path.join(HERE, "sim.cu")
    hdr = os.path.join(os.path.dirname(HERE), "..", "pcsc_eigenvalue_solver_project_b200", "csrc", "kernels.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.run([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-shared",
                        "-Xcompiler", "-fPIC", "-o", so, src], check=True, capture_output=True)
    lib = C.CDLL(so)
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
    lib.sim_power_dense.argtypes = [C.c_int, dp, dp, C.c_int, C.c_double, C.c_int, dp, dp, ip, ip, ip]
    lib.sim_power_dense.restype = C.c_int

    def run(a, x0, max_it, tol, poll=64):
        n = a.shape[0]
        ac = np.asfortranarray(a.astype(np.complex128))
        xc = np.ascontiguousarray(x0.astype(np.complex128))
        x = np.empty(n, dtype=np.complex128)
        lam = (C.c_double * 2)()
        it, cv, nl = C.c_int(), C.c_int(), C.c_int()
        rc = lib.sim_power_dense(n, ac.ctypes.data_as(dp), xc.ctypes.data_as(dp), max_it, tol, poll,
                                 x.ctypes.data_as(dp), lam, C.byref(it), C.byref(cv), C.byref(nl))
        assert rc == 0
        return complex(lam[0], lam[1]), x, it.value, bool(cv.value), nl.value

    return run


@pytest.fixture(scope="module")
def orc():
    return Oracle()


def _agree(res, ref, exact_trips=True):
    lam, x, it, cv, _ = res
    st, rlam, rx, rit, rcv = ref
    assert st == 0
    assert cv == rcv
    assert it == rit if exact_trips else abs(it - rit) <= 1
    assert abs(lam - rlam) <= 1e-10 * max(abs(rlam), 1e-300) or abs(lam - rlam) <= 1e-14
    assert phase_aligned_error(x, rx.astype(np.complex128)) <= 1e-8


@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("cplx", [False, True])
def test_fixed_trip_counts_match_the_reference(sim, orc, seed, cplx):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(2, 12))
    a = rng.standard_normal((n, n)) + (1j * rng.standard_normal((n, n)) if cplx else 0)
    a = a + a.conj().T
    a = a * 10.0 ** rng.integers(-3, 4)          # exercises the power-of-two rescaling
    x0 = orc.random_unit(n, np.complex128 if cplx else np.float64, seed=seed + 1)
    for max_it in (1, 2, 3, 7, 64, 65, 130):     # around the graph length and the poll cadence
        for poll in (1, 5, 64):
            ref = orc.power_dense(np.asfortranarray(a), x0, max_it, -1.0)
            res = sim(a, x0, max_it, -1.0, poll)
            _agree(res, ref)
            assert res[4] >= max_it + 1          # launch 0 is A*x0; the sums of launch max_it close the last trip


@pytest.mark.parametrize("seed", range(8))
def test_convergence_fires_on_the_reference_trip(sim, orc, seed):
    rng = np.random.default_rng(100 + seed)
    n = int(rng.integers(2, 9))
    q, _ = np.linalg.qr(rng.standard_normal((n, n)))
    ev = np.sort(rng.uniform(0.1, 1.0, n))
    ev[-1] = 2.0 + rng.uniform(0, 2)             # clear gap: converges in a few dozen trips
    a = (q * ev) @ q.T
    x0 = orc.random_unit(n, np.float64, seed=seed + 3)
    for tol in (1e-6, 1e-10, 1e-13):
        for poll in (1, 3, 64):
            ref = orc.power_dense(np.asfortranarray(a), x0, 1000, tol)
            res = sim(a, x0, 1000, tol, poll)
            _agree(res, ref, exact_trips=False)
            assert res[3] and res[2] >= 2        # the test cannot fire on the first trip (power_method.hpp:83)


def test_reference_edge_cases(sim, orc):
    # maxIterations <= 0: the loop never runs (power_method.hpp:61-68,98)
    a = np.diag([2.0, 1.0])
    x0 = orc.random_unit(2, np.float64, seed=1)
    for max_it in (0, -3):
        lam, x, it, cv, nl = sim(a, x0, max_it, 1e-10)
        assert (lam, it, cv, nl) == (0.0, 0, False, 0) and np.allclose(x, x0)
    # maxIterations = 1 -> iterations == 1, not converged (test/power_method_test.cpp:103-119)
    lam, x, it, cv, _ = sim(a, x0, 1, 1e-10)
    assert it == 1 and not cv
    # ||y|| == 0 (power_method.hpp:73-76): nilpotent matrix, x0 = e2 -> trip 0 gives e1, trip 1 gives 0
    n2 = np.zeros((4, 4))
    n2[0, 1] = 1.0
    e2 = np.array([0.0, 1.0, 0.0, 0.0])
    ref = orc.power_dense(np.asfortranarray(n2), e2, 10, 1e-10)
    res = sim(n2, e2, 10, 1e-10)
    assert (res[2], res[3]) == (ref[3], ref[4]) == (2, False)
    assert np.allclose(res[1], ref[2]) and abs(res[0] - ref[1]) <= 1e-15
    # zero matrix: ||y|| == 0 on the very first trip, x stays x0, lambda stays 0, one iteration
    z = np.zeros((3, 3))
    x0 = orc.random_unit(3, np.float64, seed=2)
    ref = orc.power_dense(np.asfortranarray(z), x0, 5, 1e-10)
    res = sim(z, x0, 5, 1e-10)
    assert (res[2], res[3]) == (ref[3], ref[4]) == (1, False) and np.allclose(res[1], x0) and res[0] == 0


def test_rescaling_keeps_huge_and_tiny_spectra_finite(sim, orc):
    for scale in (1e150, 1e-150):
        a = np.array([[3.0, 1.0], [1.0, 2.0]]) * scale
        x0 = orc.random_unit(2, np.float64, seed=5)
        ref = orc.power_dense(np.asfortranarray(a), x0, 400, -1.0)
        res = sim(a, x0, 400, -1.0)
        assert np.isfinite(res[0].real) and np.all(np.isfinite(res[1]))
        _agree(res, ref)
