"""Generates tests/golden/power_method_golden.npz from the UNMODIFIED reference compiled here
(oracle/_ref/libeigsol_ref.so, built by oracle/Makefile from /root/reference).  Run in the authoring
container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

Each case stores the inputs (raw CSC / dense arrays, seed, options) and the outputs of
EigSol::powerMethod (eigenvalue, eigenvector, iterations, converged) plus the start vector the
reference drew for that seed and one product A*x0 -- everything needed to pin oracle/eigsol_oracle.c
and the CUDA path without the reference being present.
"""
import os
import sys

import numpy as np
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _oracle import Reference  # noqa: E402
from pcsc_eigenvalue_solver_project_b200 import generators as gen  # noqa: E402


def cases():
    out = []
    # the reference's own tests (test/power_method_test.cpp:38-83,103-119)
    out.append(("ref_dense_diag", "dense", np.array([[2.0, 0.0], [0.0, 1.0]]), 11, 1000, 1e-10))
    out.append(("ref_sparse_tri", "sparse", sp.csc_matrix(np.array([[3.0, 1.0], [0.0, 2.0]])), 12, 1000, 1e-8))
    out.append(("ref_maxit_one", "dense", np.array([[2.0, 0.0], [0.0, 1.0]]), 13, 1, 1e-10))
    # data/A.txt and data/B.txt as main.cpp:47,56,63 runs them
    a = np.array([[1 + 3j, 3 + 5j, 1 + 4j], [0, 2 + 4j, 3 + 2j], [0, 0, 5 - 1j]])
    out.append(("data_A", "dense", a, 1, 1000, 1e-10))
    rows = [0, 1, 2, 3, 4, 0, 1, 2]
    cols = [0, 1, 2, 3, 4, 1, 3, 3]
    v = [2 + 3j, 3 + 2j, 4 + 5j, 5 + 1j, 6 - 2j, -3 - 5j, 12 + 2j, -23 - 15j]
    out.append(("data_B", "sparse", sp.csc_matrix((v, (rows, cols)), shape=(5, 5)), 2, 1000, 1e-10))
    # synthetic families of BASELINE.json at sizes the CPU finishes instantly, fixed trip count
    g = 12
    rp, ci, vv = gen.laplacian_2d(g)
    out.append(("lap2d_12_fixed", "sparse", sp.csr_matrix((vv, ci, rp), shape=(g * g, g * g)).tocsc(), 3, 80, -1.0))
    g = 6
    rp, ci, vv = gen.laplacian_3d(g)
    out.append(("lap3d_6_fixed", "sparse", sp.csr_matrix((vv, ci, rp), shape=(g ** 3, g ** 3)).tocsc(), 4, 60, -1.0))
    g = 8
    rp, ci, vv = gen.hermitian_laplacian_2d(g, seed=5)
    out.append(("herm2d_8_fixed", "sparse", sp.csr_matrix((vv, ci, rp), shape=(g * g, g * g)).tocsc(), 5, 50, -1.0))
    out.append(("dense_sym_24_fixed", "dense", gen.dense_random_symmetric(24, seed=7), 6, 40, -1.0))
    out.append(("dense_herm_17_fixed", "dense", gen.dense_random_hermitian(17, seed=8), 7, 40, -1.0))
    # converging runs with the real tolerance
    rp, ci, vv = gen.laplacian_2d(10)
    vv = vv.copy()
    vv[0] += 25.0
    out.append(("lap2d_10_bump_conv", "sparse", sp.csr_matrix((vv, ci, rp), shape=(100, 100)).tocsc(), 8, 1000, 1e-10))
    # zero-norm exit (power_method.hpp:73-76) and zero iterations
    out.append(("nilpotent", "dense", np.array([[0.0, 1.0], [0.0, 0.0]]), 9, 10, 1e-10))
    out.append(("zero_iters", "dense", gen.dense_random_symmetric(5, seed=1), 10, 0, 1e-10))
    return out


def main():
    ref = Reference()
    store = {}
    names = []
    for name, kind, m, seed, max_it, tol in cases():
        if kind == "sparse":
            m = m.tocsc()
            m.sort_indices()
            n = m.shape[0]
            outer, inner = m.indptr.astype(np.int32), m.indices.astype(np.int32)
            vals = np.ascontiguousarray(m.data)
            h = ref.create_csc(m.shape, outer, inner, vals)
            store[f"{name}/outer"], store[f"{name}/inner"], store[f"{name}/vals"] = outer, inner, vals
            dt = vals.dtype
        else:
            a = np.asfortranarray(m)
            n = a.shape[0]
            h = ref.create_dense(a)
            store[f"{name}/a"] = a
            dt = a.dtype
        st, lam, x, iters, conv, _ = ref.power(h, dt, n, seed, max_it, tol)
        assert st == 0, ref.last_error()
        x0 = ref.random_unit(n, dt, seed)
        y0 = ref.matvec(h, x0, n)
        ref.destroy(h)
        store[f"{name}/meta"] = np.array([seed, max_it, iters, int(conv), 1 if kind == "sparse" else 0, n], dtype=np.int64)
        store[f"{name}/tol"] = np.array([tol])
        store[f"{name}/lambda"] = np.array([complex(lam).real, complex(lam).imag])
        store[f"{name}/x"] = x
        store[f"{name}/x0"] = x0
        store[f"{name}/y0"] = y0
        names.append(name)
        print(f"{name:24s} n={n:4d} iters={iters:4d} conv={conv} lambda={lam}")
    store["names"] = np.array(names)
    # is_close_relative truth table (src/core/tolerance.hpp:28-33)
    rng = np.random.default_rng(0)
    pairs = []
    for _ in range(64):
        a = complex(rng.normal(), rng.normal()) * 10.0 ** int(rng.integers(-3, 4))
        b = a * (1 + 10.0 ** int(rng.integers(-13, -1)) * complex(rng.normal(), rng.normal()))
        tol = 10.0 ** int(rng.integers(-12, -2))
        pairs.append((a.real, a.imag, b.real, b.imag, tol, ref.is_close_relative(np.complex128, a, b, tol),
                      ref.is_close_relative(np.float64, a.real, b.real, tol)))
    store["close_table"] = np.array(pairs, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "power_method_golden.npz"), **store)


if __name__ == "__main__":
    main()
