"""Worker of tests/test_multi_gpu.py: one rank per GPU (launched by torch.distributed.run).  Every rank
runs the row-sharded power iteration; rank 0 checks the result against the oracle run on the whole
matrix.  Exit code 0 = parity within the north-star tolerances on every case."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.distributed as dist

import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen
from _oracle import Oracle, phase_aligned_error


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local))
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(es.Comm.unique_id()), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    comm = es.Comm(rank, world, idt.cpu().numpy().tobytes(), local)
    orc = Oracle()
    failures = []

    def check(tag, r, ref, fixed=True):
        st, lam, x, iters, conv = ref
        ok = st == 0 and r.converged == conv and (r.iterations == iters if fixed else abs(r.iterations - iters) <= 1)
        ok = ok and abs(r.eigenvalue - lam) <= 1e-10 * abs(lam) and phase_aligned_error(r.eigenvector, x) <= 1e-8
        if rank == 0:
            print(f"[{tag}] world={world} iters={r.iterations}/{iters} lam={r.eigenvalue!r} ref={lam!r} "
                  f"xerr={phase_aligned_error(r.eigenvector, x):.2e} -> {'ok' if ok else 'FAIL'}", flush=True)
        if not ok:
            failures.append(tag)

    # 1. 3D Laplacian, slab partition by explicit row blocks (halo = neighbour planes, contiguous sends)
    g = 24
    n = g ** 3
    rp, ci, v = gen.laplacian_3d(g)
    lo, hi = n * rank // world, n * (rank + 1) // world
    rpl, cil, vl = gen.laplacian_3d(g, lo, hi)
    x0 = orc.random_unit(n, np.float64, seed=1)
    dm = es.DeviceMatrix.from_csr_rows(n, lo, rpl, cil, vl, comm)
    info = dm.info()
    assert info["nranks"] == world and info["local_rows"] == hi - lo
    expect = os.environ.get("EIGSOL_B200_EXPECT_EXCHANGE")
    if expect:  # 2 = peer memory (fused kernel), 1 = NCCL send/recv + all-reduce
        assert info["exchange"] == {"peer": 2, "nccl": 1}[expect], info
        assert info["kernels_per_trip"] == (1 if expect == "peer" else 2)
    y = dm.matvec(x0)
    yref = orc.spmv_csc((n, n), rp, ci, v, x0)
    if np.max(np.abs(y - yref)) > 1e-14 * np.max(np.abs(yref)):
        failures.append("lap3d matvec")
    check("lap3d fixed", dm.power(x0, es.SolverOptions(50, -1.0)), orc.power_csc((n, n), rp, ci, v, x0, 50, -1.0))
    for kernel, lanes in [(es.KERNEL_SELL, 2), (es.KERNEL_CSR_STREAM, 0), (es.KERNEL_CSR_VECTOR, 4), (es.KERNEL_CSR_WINDOW, 0)]:
        dm.set_kernel(kernel, lanes)
        check(f"lap3d k{kernel}", dm.power(x0, es.SolverOptions(20, -1.0)), orc.power_csc((n, n), rp, ci, v, x0, 20, -1.0))
    dm.close()

    # 2. irregular complex matrix through the global-CSC entry (nnz-balanced partition inside the library,
    #    scattered halo -> pack kernel), with the real tolerance
    import scipy.sparse as sp
    n = 2500
    rp, ci, v = gen.random_sparse_csr(n, 6.0, seed=4, complex_values=True, skew=40.0, diag_boost=60.0)
    a = sp.csr_matrix((v, ci, rp), shape=(n, n)).tocsc()
    a.sort_indices()
    outer, inner, cv = a.indptr.astype(np.int32), a.indices.astype(np.int32), np.ascontiguousarray(a.data)
    x0 = orc.random_unit(n, np.complex128, seed=2)
    dm = es.DeviceMatrix.from_matrix(es.Matrix(((n, n), outer, inner, cv)), comm)
    check("irregular c128 fixed", dm.power(x0, es.SolverOptions(30, -1.0)), orc.power_csc((n, n), outer, inner, cv, x0, 30, -1.0))
    check("irregular c128 tol", dm.power(x0, es.SolverOptions(1000, 1e-10)),
          orc.power_csc((n, n), outer, inner, cv, x0, 1000, 1e-10), fixed=False)
    dm.close()

    # 3. dense, row blocks + full exchange of x
    for n, cplx in [(300, False), (257, True)]:
        a = gen.dense_random_hermitian(n, 5) if cplx else gen.dense_random_symmetric(n, 5)
        x0 = orc.random_unit(n, a.dtype, seed=3)
        dm = es.DeviceMatrix.from_matrix(es.Matrix(a), comm)
        check(f"dense n={n}", dm.power(x0, es.SolverOptions(25, -1.0)), orc.power_dense(a, x0, 25, -1.0))
        dm.close()

    # 4. zero-norm exit and maxIterations = 0 behave identically on every rank
    a = np.zeros((4, 4))
    a[0, 1] = 1.0
    x0 = np.array([0.0, 1.0, 0.0, 0.0])
    dm = es.DeviceMatrix.from_matrix(es.Matrix(a), comm)
    r = dm.power(x0, es.SolverOptions(10, 1e-10))
    if (r.iterations, r.converged) != (2, False):
        failures.append(f"zero norm {r.iterations} {r.converged}")
    r = dm.power(x0, es.SolverOptions(0, 1e-10))
    if r.iterations != 0 or not np.array_equal(r.eigenvector, x0):
        failures.append("zero iterations")
    dm.close()

    # 5. Rayleigh step on the sharded matrix, and repeated solves on one handle (flag epochs keep counting)
    g = 20
    n = g ** 3
    rp, ci, v = gen.laplacian_3d(g)
    lo, hi = n * rank // world, n * (rank + 1) // world
    rpl, cil, vl = gen.laplacian_3d(g, lo, hi)
    x0 = orc.random_unit(n, np.float64, seed=7)
    dm = es.DeviceMatrix.from_csr_rows(n, lo, rpl, cil, vl, comm)
    y = orc.spmv_csc((n, n), rp, ci, v, x0)
    dot, nrm2 = dm.rayleigh(x0)
    if abs(dot - np.vdot(x0, y)) > 1e-12 or abs(nrm2 - np.vdot(y, y)) > 1e-12 * np.vdot(y, y):
        failures.append("rayleigh")
    ref = orc.power_csc((n, n), rp, ci, v, x0, 17, -1.0)
    for rep in range(3):
        check(f"repeat {rep}", dm.power(x0, es.SolverOptions(17, -1.0)), ref)

    # 5b. block power iteration on the sharded matrix (peer mode: CUDA IPC mappings of the block buffers, halo push after
    #     every SpMM, Gram sums all-reduced in the control launch) against the CPU restatement on the whole matrix
    if os.environ.get("EIGSOL_B200_EXPECT_EXCHANGE") == "peer":
        import _blocksim as bs
        a3 = sp.csr_matrix((v, ci, rp), shape=(n, n))
        for k in (1, 4):
            xb = es.random_block(n, k, np.float64, seed=5)
            rb = dm.block_power(xb, k, es.SolverOptions(25, -1.0))
            th, X, itb, _, _ = bs.block_power_cpu(lambda q: a3 @ q, xb, k, 25, -1.0)
            q1, _ = np.linalg.qr(rb.eigenvectors)
            q2, _ = np.linalg.qr(X)
            ok = rb.iterations == itb == 25 and not rb.breakdown and np.allclose(rb.eigenvalues, th, rtol=1e-10) \
                and np.linalg.norm(q1 @ (q1.conj().T @ q2) - q2) < 1e-8
            if rank == 0:
                print(f"[block k={k}] world={world} ritz0={rb.eigenvalues[0]!r} ref={th[0]!r} -> {'ok' if ok else 'FAIL'}", flush=True)
            if not ok:
                failures.append(f"block k={k}")

    # 6. a peer that stops taking part must surface as an error on the others, never as a hang (peer mode only)
    if os.environ.get("EIGSOL_B200_EXPECT_EXCHANGE") == "peer":
        dm.begin(x0, 1000, -1.0)  # collective
        if rank == 0:
            try:
                dm.step(40)
                dm.end(want_vector=False)
                failures.append("missing peer not detected")
            except es.EigsolB200Error as e:
                if "did not signal" not in str(e):
                    failures.append(f"unexpected error text: {e}")
                print(f"[missing peer] rank 0 got the expected error: {e}", flush=True)
        dist.barrier()
    dm.close()

    flag = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(flag)
    comm.close()
    dist.destroy_process_group()
    if failures:
        print(f"rank {rank} failures: {failures}", flush=True)
    sys.exit(1 if int(flag.item()) else 0)


if __name__ == "__main__":
    main()
