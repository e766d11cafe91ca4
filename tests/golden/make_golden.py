"""Generate the golden fixtures under tests/golden/ by RUNNING THE COMPILED REFERENCE
(oracle/_ref/libfemeko_ref.so = unmodified /root/reference/src/FEMc.cpp + vendored Eigen, see
oracle/Makefile).  The reference ships no golden vectors of its own, so these outputs of
the reference itself are what pins the oracle and the CUDA path.  Run in the dev container only:

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import femeko_jl_b200 as fk  # noqa: E402
from oracle import oracle as orc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def make(name, mesh):
    mu = np.full(mesh.nt, fk.mu0)
    mu[mesh.InsideElements] = 3 * fk.mu0          # magneticField.jl:91-92
    F = np.array([1.0, 0.0, 0.0])                 # Hext of run.jl:23
    outer, inner, values = orc.ref_stiffness(mesh, mu)
    outer1, inner1, values1 = orc.ref_stiffness(mesh, None)
    assert np.array_equal(outer, outer1) and np.array_equal(inner, inner1)
    Ak = orc.ref_local_stiffness(mesh)
    lag = orc.ref_lagrange(mesh)
    rhs = orc.ref_boundary_integral(mesh, F, 0)
    u, iters, err, info, _ = orc.ref_cg(outer, inner, values, -rhs, 1e-10, 0)
    assert info == 0
    rng = np.random.default_rng(7)
    samples = rng.integers(0, mesh.nt, size=16)
    abcd = np.array([[orc.ref_abcd(mesh.p, mesh.t[:, k], mesh.t[i, k]) for i in range(4)] for k in samples])
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"),
        p=mesh.p, t=mesh.t, surfaceT=mesh.surfaceT, normal=mesh.normal, VE=mesh.VE, AE=mesh.AE,
        InsideElements=mesh.InsideElements, mu=mu, F=F,
        csc_outer=outer, csc_inner=inner, csc_values_mu=values, csc_values_nomu=values1,
        Ak=Ak, lagrange=lag, boundary_integral=rhs, cg_u=u, cg_iters=iters, cg_err=err,
        abcd_elems=samples, abcd=abcd)
    print(name, "nv", mesh.nv, "nt", mesh.nt, "nnz", len(inner), "cg iters", iters)


if __name__ == "__main__":
    assert orc.have_ref(), "build oracle/_ref first (make -C oracle ref)"
    make("kuhn3", fk.kuhn_cube(3, inside_edge=0.4))
    make("kuhn6", fk.kuhn_cube(6, inside_edge=0.34))
    make("delaunay300", fk.delaunay_ball(300, seed=1234))
