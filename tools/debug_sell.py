import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen
for logn in (16, 18, 20):
    for k in (48, 64):
        n = 1 << logn
        rp, ci, v = gen.banded_random_csr(n, k, 4096, seed=1, regular=True)
        dm = es.DeviceMatrix.from_csr_rows(n, 0, rp, ci, v)
        x = np.random.default_rng(0).random(n)
        x /= np.linalg.norm(x)
        dm.set_kernel(es.KERNEL_CSR_SCALAR, 0)
        y0 = dm.matvec(x)
        dm.set_kernel(es.KERNEL_SELL, 2)
        y1 = dm.matvec(x)
        dm.begin(x)
        ms = dm.step(5)
        r = dm.end(want_vector=False)
        print(logn, k, dm.info()["kernel_name"], "maxdiff", np.max(np.abs(y0 - y1)), "norm", np.linalg.norm(y0), np.linalg.norm(y1),
              "iters", r.iterations, "lam", r.eigenvalue, "ms", ms, flush=True)
        dm.close()
