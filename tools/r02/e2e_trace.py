"""Phase trace (EIGSOL_B200_TRACE=1) of the one-shot reference-facing call on C5 at the driver's 20 trips: where the
e2e time goes.  Run twice so the second call sees warm pinned pools / contexts."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen

if os.environ.get("TRACE_TORCH"):  # diagnosis: does a torch CUDA context / an NCCL process group in the process change the upload?
    import torch
    torch.cuda.set_device(0)
    torch.zeros(1, device="cuda")
    if os.environ["TRACE_TORCH"] == "nccl":
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29655")
        dist.init_process_group(backend="nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
        t = torch.zeros(4, device="cuda")
        dist.all_reduce(t)
        torch.cuda.synchronize()
        dist.destroy_process_group()
g = int(os.environ.get("GRID", 256))
n = g ** 3
rp, ci, v = gen.laplacian_3d(g)
outer, inner, vals = gen.csr_to_csc_hermitian(rp, ci, v)
x0 = es.random_unit(n, np.float64, seed=1)
for rep in range(3):
    M = es.Matrix(((n, n), outer, inner, vals))  # a fresh Matrix per call: every call uploads (a Matrix keeps its device copy)
    t0 = time.perf_counter()
    r = es.powerMethod(M, es.SolverOptions(20, -1.0), x0=x0)
    dt = time.perf_counter() - t0
    print(f"== call {rep}: {dt * 1e3:.1f} ms, {20 / dt:.1f} it/s, lambda {r.eigenvalue!r}", file=sys.stderr, flush=True)
