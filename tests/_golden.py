"""Loader for tests/golden/power_method_golden.npz (made by tests/golden/make_golden.py from the
compiled, unmodified reference)."""
import os

import numpy as np

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "power_method_golden.npz")


class Case:
    def __init__(self, z, name):
        self.name = name
        seed, max_it, iters, conv, sparse, n = (int(v) for v in z[f"{name}/meta"])
        self.seed, self.max_it, self.iters, self.conv, self.sparse, self.n = seed, max_it, iters, bool(conv), bool(sparse), n
        self.tol = float(z[f"{name}/tol"][0])
        lam = z[f"{name}/lambda"]
        self.x, self.x0, self.y0 = z[f"{name}/x"], z[f"{name}/x0"], z[f"{name}/y0"]
        self.dtype = self.x.dtype
        self.lam = complex(lam[0], lam[1]) if self.dtype == np.complex128 else float(lam[0])
        if self.sparse:
            self.outer, self.inner, self.vals = z[f"{name}/outer"], z[f"{name}/inner"], z[f"{name}/vals"]
        else:
            self.a = np.asfortranarray(z[f"{name}/a"])


def load():
    z = np.load(PATH)
    return [Case(z, str(nm)) for nm in z["names"]], z["close_table"]
