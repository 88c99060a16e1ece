"""Worker of tests/test_multi_rank_gloo.py (CPU, gloo, world_size 2+): exercises the HOST side of the
multi-GPU path in a real multi-process run -- the row partition and halo plan from the C ABI
(eigsol_b200_partition_rows / eigsol_b200_halo_plan), the exchange pattern derived from them (who sends
which entries to whom), the extended local numbering [owned | halo], and the all-reduced scalar state
machine -- with numpy standing in for the device kernels.  Rank 0 compares with the oracle."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.distributed as dist

from pcsc_eigenvalue_solver_project_b200 import _capi
from pcsc_eigenvalue_solver_project_b200 import generators as gen
from _oracle import Oracle, phase_aligned_error


def plan(rp_loc, ci_loc, rb, rank):
    nranks = rb.size - 1
    cnt = C.c_int64(0)
    L = _capi.lib()
    _capi.check(L.eigsol_b200_halo_plan(rp_loc.size - 1, _capi.ptr(rp_loc), _capi.ptr(ci_loc), nranks, _capi.ptr(rb), rank,
                                        C.byref(cnt), None, None))
    cols = np.zeros(max(cnt.value, 1), dtype=np.int32)
    peers = np.zeros(nranks, dtype=np.int64)
    _capi.check(L.eigsol_b200_halo_plan(rp_loc.size - 1, _capi.ptr(rp_loc), _capi.ptr(ci_loc), nranks, _capi.ptr(rb), rank,
                                        C.byref(cnt), _capi.ptr(cols), _capi.ptr(peers)))
    return cols[:cnt.value], peers


def run_case(name, n, rp, ci, v, x0, trips, rank, world, orc):
    rb = np.zeros(world + 1, dtype=np.int64)
    _capi.check(_capi.lib().eigsol_b200_partition_rows(n, _capi.ptr(rp), world, _capi.ptr(rb)))
    lo, hi = int(rb[rank]), int(rb[rank + 1])
    rpl = (rp[lo:hi + 1] - rp[lo]).astype(np.int32)
    cil = np.ascontiguousarray(ci[rp[lo]:rp[hi]])
    vl = v[rp[lo]:rp[hi]]
    halo, peers = plan(rpl, cil, rb, rank)
    # exchange plan: every rank learns which of its entries each peer reads (engine.cu plan_exchange)
    needs = [None] * world
    dist.all_gather_object(needs, halo)
    send_idx = [needs[q][(needs[q] >= lo) & (needs[q] < hi)] - lo if q != rank else np.zeros(0, np.int64) for q in range(world)]
    peer_off = np.concatenate([[0], np.cumsum(peers)])
    # extended local numbering [owned | halo]
    nloc = hi - lo
    local = (cil >= lo) & (cil < hi)
    cext = np.where(local, cil - lo, nloc + np.searchsorted(halo, cil)).astype(np.int64)
    rows = np.repeat(np.arange(nloc), np.diff(rpl))

    def exchange(u):
        ext = np.zeros(nloc + halo.size, dtype=u.dtype)
        ext[:nloc] = u
        reqs = []
        bufs = {}
        for q in range(world):
            if q == rank:
                continue
            if send_idx[q].size:
                reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(u[send_idx[q]]).view(np.float64).copy()), q))
            if peers[q]:
                bufs[q] = torch.zeros(int(peers[q]) * (2 if np.iscomplexobj(u) else 1), dtype=torch.float64)
                reqs.append(dist.irecv(bufs[q], q))
        for r in reqs:
            r.wait()
        for q, b in bufs.items():
            ext[nloc + peer_off[q]: nloc + peer_off[q + 1]] = b.numpy().view(u.dtype)
        return ext

    # the device loop, restated with numpy per rank + all-reduced scalars (kernels.cuh finalize_launch)
    u = x0[lo:hi].copy()
    qin, lam, used = 1.0, 0.0, 0
    prev_u, prev_q = u, 1.0
    for l in range(trips + 1):
        ext = exchange(u)
        s = 1.0 / np.sqrt(qin)
        z = np.zeros(nloc, dtype=u.dtype)
        np.add.at(z, rows, vl * ext[cext])
        z *= s
        part = np.array([np.vdot(u * s, z).real, np.vdot(u * s, z).imag, np.vdot(z, z).real])
        t = torch.from_numpy(part)
        dist.all_reduce(t)
        if l >= 1:
            lam = complex(part[0], part[1]) if np.iscomplexobj(u) else part[0]
            used = l
            prev_u, prev_q = u, qin
        u, qin = z, float(part[2])
    xloc = prev_u / np.sqrt(prev_q)
    parts = [None] * world
    dist.all_gather_object(parts, xloc)
    x = np.concatenate(parts)
    if rank == 0:
        import scipy.sparse as sp
        a = sp.csr_matrix((v, ci, rp), shape=(n, n)).tocsc()
        a.sort_indices()
        st, lref, xref, iters, conv = orc.power_csc((n, n), a.indptr.astype(np.int32), a.indices.astype(np.int32),
                                                    np.ascontiguousarray(a.data), x0, trips, -1.0)
        ok = used == iters and abs(lam - lref) <= 1e-10 * abs(lref) and phase_aligned_error(x, xref) <= 1e-8
        print(f"[{name}] world={world} halo={[int(p) for p in peers]} lam={lam!r} ref={lref!r} -> {'ok' if ok else 'FAIL'}", flush=True)
        return ok
    return True


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group(backend="gloo")
    orc = Oracle()
    ok = True
    g = 10
    rp, ci, v = gen.laplacian_3d(g)
    ok &= run_case("lap3d", g ** 3, rp, ci, v, orc.random_unit(g ** 3, np.float64, seed=1), 30, rank, world, orc)
    rp, ci, v = gen.random_sparse_csr(400, 5.0, seed=2, complex_values=True, skew=20.0, diag_boost=30.0)
    ok &= run_case("irregular c128", 400, rp, ci, v, orc.random_unit(400, np.complex128, seed=2), 20, rank, world, orc)
    flag = torch.tensor([0 if ok else 1])
    dist.all_reduce(flag)
    dist.destroy_process_group()
    sys.exit(int(flag.item() != 0))


if __name__ == "__main__":
    main()
