"""Times every kernel variant on the BASELINE.json configs (one GPU).  Development tool: its numbers
pick the dispatch rule in plan.hpp; the numbers of record come from bench.py.

    python tools/variant_bench.py [c2 c5 c4 c3] > gpurun_out/variants.jsonl
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen

VARIANTS = [(es.KERNEL_CSR_WINDOW, 0), (es.KERNEL_SELL, 2), (es.KERNEL_SELL, 1), (es.KERNEL_CSR_STREAM, 0), (es.KERNEL_CSR_SCALAR, 0),
            (es.KERNEL_CSR_VECTOR, 2), (es.KERNEL_CSR_VECTOR, 4), (es.KERNEL_CSR_VECTOR, 8),
            (es.KERNEL_CSR_VECTOR, 16), (es.KERNEL_CSR_VECTOR, 32)]


def build(cfg):
    t = time.time()
    if cfg == "c2":
        g = 1024
        rp, ci, v = gen.laplacian_2d(g)
        n = g * g
        dm = es.DeviceMatrix.from_csr_rows(n, 0, rp, ci, v)
        dt = np.float64
    elif cfg == "c5":
        g = int(os.environ.get("C5_GRID", "256"))
        rp, ci, v = gen.laplacian_3d(g)
        n = g ** 3
        dm = es.DeviceMatrix.from_csr_rows(n, 0, rp, ci, v)
        dt = np.float64
    elif cfg == "c4":
        g = 2048
        rp, ci, v = gen.hermitian_laplacian_2d(g)
        n = g * g
        dm = es.DeviceMatrix.from_csr_rows(n, 0, rp, ci, v)
        dt = np.complex128
    elif cfg == "c3":
        n = 8192
        a = gen.dense_random_symmetric(n)
        dm = es.DeviceMatrix.from_matrix(es.Matrix(a))
        dt = np.float64
    elif cfg[0] in "si" and cfg[1:].rstrip("c").isdigit():
        # s<k>: regular k-point stencil rows; i<k>: irregular rows, Poisson(k) entries (dispatch tuning);
        # a trailing "c" makes the values complex (c128)
        n = 1 << 20
        cplx = cfg.endswith("c")
        rp, ci, v = gen.banded_random_csr(n, int(cfg[1:].rstrip("c")), 4096, seed=1, regular=cfg[0] == "s")
        if cplx:
            v = v + 1j * np.roll(v, 1)
        dm = es.DeviceMatrix.from_csr_rows(n, 0, rp, ci, v)
        dt = np.complex128 if cplx else np.float64
    else:
        raise SystemExit(f"unknown config {cfg}")
    rng = np.random.default_rng(1)
    x0 = rng.random(n) * 2 - 1
    if dt == np.complex128:
        x0 = x0 + 1j * (rng.random(n) * 2 - 1)
    x0 = (x0 / np.linalg.norm(x0)).astype(dt)
    print(f"# {cfg}: built in {time.time() - t:.1f}s", file=sys.stderr, flush=True)
    return dm, x0, n


def main():
    cfgs = sys.argv[1:] or ["c2", "c5", "c4", "c3"]
    trips = int(os.environ.get("TRIPS", "100"))
    for cfg in cfgs:
        dm, x0, n = build(cfg)
        variants = [(es.KERNEL_GEMV, 0)] if cfg == "c3" else VARIANTS
        for kernel, lanes in variants:
            if cfg != "c3":
                t = time.time()
                dm.set_kernel(kernel, lanes)
                tset = time.time() - t
            else:
                tset = 0.0
            info = dm.info()
            dm.begin(x0)
            dm.step(10)
            best = min(dm.step(trips) for _ in range(3)) / trips
            r = dm.end(want_vector=False)
            out = {"cfg": cfg, "row_len_mean": info["row_len_mean"], "row_len_max": info["row_len_max"],
                   "kernel": info["kernel_name"], "lanes": lanes, "ms_per_trip": best,
                   "it_per_s": 1e3 / best, "alg_GBps": info["alg_bytes"] / best / 1e6,
                   "stored_GBps": info["stored_bytes"] / best / 1e6, "grid": info["grid"], "lambda": repr(r.eigenvalue),
                   "set_kernel_s": round(tset, 2)}
            print(json.dumps(out), flush=True)
        dm.close()


if __name__ == "__main__":
    main()
