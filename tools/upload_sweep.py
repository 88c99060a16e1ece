"""End-to-end one-shot call (host buffers) on C5 for several uploader thread counts (development tool)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen

g = int(sys.argv[1]) if len(sys.argv) > 1 else 256
n = g ** 3
rp, ci, v = gen.laplacian_3d(g, 0, n)
M = es.Matrix(((n, n), rp, ci, v))
x0 = es.random_unit(n, np.float64, seed=1)
es.powerMethod(M, es.SolverOptions(5, -1.0), x0=x0)
for thr in os.environ.get("SWEEP", "1,2,3,4,6,8").split(","):
    os.environ["EIGSOL_B200_UPLOAD_THREADS"] = thr
    best = 1e30
    for _ in range(3):
        t0 = time.perf_counter()
        r = es.powerMethod(M, es.SolverOptions(200, -1.0), x0=x0)
        best = min(best, time.perf_counter() - t0)
    print(f"threads={thr} seconds={best:.4f} e2e_it_per_s={200 / best:.1f} lambda={r.eigenvalue!r}", flush=True)
