"""Block power iteration: device time per trip for block widths 1, 2, 4, 8 (development tool, one GPU).
    python tools/block_bench.py [c5|c2|c3|c4] [trips]
Prints one JSON line per width: us per trip, us per vector-trip, algorithmic GB/s
(B_alg = matrix stream + one read and one write of the n x k block)."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "c5"
    trips = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    if wl == "c3":
        n = 8192
        rng = np.random.default_rng(7)
        r = rng.uniform(-1, 1, (n, n))
        a = np.asfortranarray(0.5 * (r + r.T))
        dm = es.DeviceMatrix.from_matrix(es.Matrix(a))
        dt, mat_bytes = np.float64, 8 * n * n
    else:
        if wl == "c5":
            g = 256
            n = g ** 3
            rp, ci, v = gen.laplacian_3d(g, 0, n)
        elif wl == "c2":
            g = 1024
            n = g * g
            rp, ci, v = gen.laplacian_2d(g)
        else:
            g = 2048
            n = g * g
            rp, ci, v = gen.hermitian_laplacian_2d(g)
        dm = es.DeviceMatrix.from_csr_rows(n, 0, rp, ci, v)
        dt = v.dtype
        mat_bytes = len(ci) * (v.itemsize + 4) + 4 * (n + 1)
    ss = np.dtype(dt).itemsize
    widths = [int(w) for w in os.environ.get("BLOCK_WIDTHS", "1,2,4,8").split(",")]
    x0 = es.random_block(n, max(widths), dt, seed=1)
    single_us = None
    # BLOCK_CONFIGS="row;lanes:2;tile::8" (kernel[:unroll[:chunk]]) sweeps the SpMM variant (library knobs, read per solve)
    configs = [c for c in os.environ.get("BLOCK_CONFIGS", "default").split(";") if c]
    for cfg in configs:
      for key in ("EIGSOL_B200_BLOCK_KERNEL", "EIGSOL_B200_BLOCK_UNROLL", "EIGSOL_B200_BLOCK_CHUNK"):
          os.environ.pop(key, None)
      if cfg != "default":   # kernel[:unroll[:chunk]], empty fields keep the default
          parts = (cfg.split(":") + ["", ""])[:3]
          for key, val in zip(("EIGSOL_B200_BLOCK_KERNEL", "EIGSOL_B200_BLOCK_UNROLL", "EIGSOL_B200_BLOCK_CHUNK"), parts):
              if val:
                  os.environ[key] = val
      for k in widths:
        xk = np.asfortranarray(x0[:, :k])
        dm.block_power(xk, k, es.SolverOptions(5, -1.0), want_vectors=False)   # warm-up
        best = 1e30
        for _ in range(3):
            r = dm.block_power(xk, k, es.SolverOptions(trips, -1.0), want_vectors=False)
            best = min(best, r.device_ms / (trips + 1))   # trips + 1 launches (launch 0 is the set-up product)
        alg = mat_bytes + 2 * ss * k * n
        print(json.dumps({"workload": wl, "cfg": cfg, "block": k, "us_per_trip": best * 1e3, "us_per_vector_trip": best * 1e3 / k,
                          "alg_bytes": alg, "alg_GBps": alg / (best * 1e-3) / 1e9,
                          "ritz": [repr(complex(z)) for z in r.eigenvalues[:2]]}), flush=True)
    if os.environ.get("BLOCK_SINGLE", "1") == "1":
        x = np.ascontiguousarray(x0[:, 0])
        dm.begin(x, 2 ** 31 - 2, -1.0)
        dm.step(20)
        ms = min(dm.step(trips) for _ in range(3))
        dm.end(want_vector=False)
        print(json.dumps({"workload": wl, "single_vector_us_per_trip": ms / trips * 1e3}), flush=True)
    dm.close()


if __name__ == "__main__":
    main()
