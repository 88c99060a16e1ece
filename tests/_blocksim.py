"""CPU harness of the block power iteration (test infrastructure only): the control step of
csrc/blockmath.hpp compiled by g++ (tests/blockmath/sim.cpp) + a numpy SpMM.  `block_power_cpu` is the
restatement the GPU path (csrc/block.cuh) is compared with: the same k x k decisions, the n-sized products
done by numpy."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "blockmath")
HDR = os.path.join(os.path.dirname(HERE), "..", "pcsc_eigenvalue_solver_project_b200", "csrc", "blockmath.hpp")
_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    so, src = os.path.join(HERE, "libblocksim.so"), os.path.join(HERE, "sim.cpp")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(HDR)):
        subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-o", so, src], check=True, capture_output=True)
    L = C.CDLL(so)
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
    L.bm_chol.argtypes = [dp, C.c_int, dp]
    L.bm_triinv.argtypes = [dp, C.c_int, dp]
    L.bm_eig.argtypes = [dp, C.c_int, dp, dp]
    L.bm_schur.argtypes = [dp, C.c_int, dp, dp]
    L.bm_begin.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, dp]
    L.bm_step.argtypes = [dp, dp]
    L.bm_state.argtypes = [ip, dp, dp, dp]
    _lib = L
    return L


def _in(a):
    a = np.ascontiguousarray(a, dtype=np.complex128)
    return a, a.view(np.float64).ctypes.data_as(C.POINTER(C.c_double))


def _out(k, cols=None):
    a = np.zeros((k, k) if cols is None else (cols,), dtype=np.complex128)
    return a, a.view(np.float64).ctypes.data_as(C.POINTER(C.c_double))


def chol(w):
    k = w.shape[0]
    wi, wp = _in(w)
    r, rp = _out(k)
    ok = lib().bm_chol(wp, k, rp)
    return bool(ok), r


def triinv(r):
    k = r.shape[0]
    ri, rp = _in(r)
    t, tp = _out(k)
    lib().bm_triinv(rp, k, tp)
    return t


def eig(h):
    k = h.shape[0]
    hi, hp = _in(h)
    th, thp = _out(k, k)
    s, sp = _out(k)
    ok = lib().bm_eig(hp, k, thp, sp)
    return bool(ok), th, s


def schur(h):
    k = h.shape[0]
    hi, hp = _in(h)
    t, tp = _out(k)
    z, zp = _out(k)
    ok = lib().bm_schur(hp, k, tp, zp)
    return bool(ok), t, z


def state(k):
    ints = (C.c_int * 5)()
    th, thp = _out(k, k)
    t, tp = _out(k)
    m, mp = _out(k)
    lib().bm_state(ints, thp, tp, mp)
    return dict(launch_no=ints[0], iterations=ints[1], converged=bool(ints[2]), stop=ints[3], final_launch=ints[4],
                theta=th, T=t, M=m)


def block_power_cpu(matvec, x0, nev, max_it, tol, sigma=1.0):
    """matvec(Q) -> A @ Q for an n x k block.  Returns (theta[k], X[n x k], iterations, converged, stop)."""
    x0 = np.asarray(x0, dtype=np.complex128)
    n, k = x0.shape
    L = lib()
    u = [x0.copy(), None]
    w, wp = _in(u[0].conj().T @ u[0])
    L.bm_begin(k, nev, max_it, tol, sigma, wp)
    st = state(k)
    while st["stop"] == 0:
        m = st["launch_no"]
        q = u[m & 1] @ st["T"]
        unew = sigma * matvec(q)
        u[(m + 1) & 1] = unew
        c, cp = _in(q.conj().T @ unew)
        w, wp = _in(unew.conj().T @ unew)
        L.bm_step(cp, wp)
        st = state(k)
    f = max(st["final_launch"], 0)
    x = u[f & 1] @ st["M"]
    return st["theta"], x, st["iterations"], st["converged"], st["stop"]
