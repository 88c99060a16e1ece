"""Times the fast text reader against the reference's own readMatrixFromFile (oracle/_ref) on a generated file
of the C2 matrix (2D Laplacian 1024^2, 5.2M entries in "r c value" triples), and the binary container.
CPU only.    python tools/reader_bench.py [grid]"""
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import pcsc_eigenvalue_solver_project_b200 as es
from pcsc_eigenvalue_solver_project_b200 import generators as gen
from _oracle import Reference, have_reference

g = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
rp, ci, v = gen.laplacian_2d(g)
n = g * g
rows_of = np.repeat(np.arange(n), np.diff(rp))
d = tempfile.mkdtemp()
p = os.path.join(d, "c2.txt")
with open(p, "w") as f:
    f.write(f"sparse {n} {n} {len(ci)}\n")
    for lo in range(0, len(ci), 1 << 20):
        hi = min(lo + (1 << 20), len(ci))
        f.write("".join(f"{r} {c} {x!r}\n" for r, c, x in zip(rows_of[lo:hi].tolist(), ci[lo:hi].tolist(), v[lo:hi].tolist())))
size = os.path.getsize(p)
out = {"grid": g, "entries": int(len(ci)), "file_MB": size / 1e6, "host_threads": os.cpu_count()}
for threads in (1, os.cpu_count()):
    os.environ["EIGSOL_B200_IO_THREADS"] = str(threads)
    t = time.time()
    m = es.readMatrixFromFile(p)
    out[f"fast_reader_{threads}thr_s"] = time.time() - t
if have_reference() and len(ci) <= 500_000:  # the reference's insert() per entry is quadratic-ish: 20 s at 327k entries
    ref = Reference()
    if hasattr(ref.L, "ref_read_matrix_file"):
        isd, rows, cols, outer, inner, vals, sec = ref.read_matrix_file(p, np.float64)
        out["reference_reader_s"] = sec
        out["identical_arrays"] = bool(np.array_equal(outer, m._outer) and np.array_equal(inner, m._inner) and np.array_equal(vals, m._values))
b = os.path.join(d, "c2.bin")
t = time.time()
es.writeMatrixBinary(m, b)
out["binary_write_s"] = time.time() - t
t = time.time()
m2 = es.readMatrixBinary(b)
out["binary_map_and_copy_s"] = time.time() - t
out["binary_MB"] = os.path.getsize(b) / 1e6
print(json.dumps(out))
