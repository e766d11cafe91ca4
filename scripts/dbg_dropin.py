"""Debug: scalarPotential at Eigen's default tolerance -- true residual and null-space drift, old vs new library."""
import ctypes as C
import os
import sys

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_golden  # noqa: E402
from femeko_jl_b200 import lib as L  # noqa: E402

if len(sys.argv) > 1:
    L.LIB_PATH = sys.argv[1]
lib = L.load()
import femeko_jl_b200 as fk  # noqa: E402
for name in ["kuhn6", "delaunay300"]:
    m, g = load_golden(name)
    u = np.zeros(m.nv)
    p = np.asfortranarray(m.p); t = np.asfortranarray(m.t); sT = np.asfortranarray(m.surfaceT); nrm = np.asfortranarray(m.normal)
    lib.scalarPotential(L.dptr(u), L.dptr(p), L.iptr(t), L.iptr(sT), L.dptr(nrm), m.nv, m.nt, m.ns, L.dptr(m.VE), L.dptr(g["mu"]), L.dptr(g["F"]), 0)
    A = sp.csc_matrix((g["csc_values_mu"], g["csc_inner"], g["csc_outer"]), shape=(m.nv, m.nv))
    b = -g["boundary_integral"]
    ref = g["cg_u"]
    rl = lambda a, c: np.linalg.norm(a - c) / np.linalg.norm(c)
    print(name, "dropin: mean", u.mean(), "true relres", np.linalg.norm(b - A @ u) / np.linalg.norm(b), "vs golden", rl(u - u.mean(), ref - ref.mean()))
    h = fk.Femeko(m)
    h.assemble_stiffness(g["mu"], want_vals=False)
    for tol, mi in [(1e-10, 0), (1e-14, 0), (0.0, 0), (0.0, 100), (0.0, 200), (0.0, 400)]:
        x, it, rr = h.pcg(b, tol=tol, max_iter=mi)
        print("   tol", tol, "max_iter", mi, "iters", it, "rr", rr, "mean", x.mean(), "true relres", np.linalg.norm(b - A @ x) / np.linalg.norm(b),
              "vs golden", rl(x - x.mean(), ref - ref.mean()))
    h.close()
