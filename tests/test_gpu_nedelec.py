"""GPU: Nedelec (lowest-order edge) elements -- nedelecLocalStiffness / nedelecTangentialStiffness (src/FEM3D.jl:153-271) and
the vector-potential Newton loop of Magnetostatics/nedelec3D.jl:104-250 -- through the C ABI (fk_create_edge, fk_nedelec_*,
fk_assemble_nedelec_*) against the oracle's restatement.  Tolerances: pattern bit-exact, matrix / vector values 1e-12 of the
largest magnitude, fields of a converged solve 1e-8 (linear) and 1e-6 (Newton loop, two CG solvers) relative L2."""
import os

import numpy as np
import pytest

import femeko_jl_b200 as fk
from conftest import GOLDEN
from femeko_jl_b200 import mesh as M
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def meshes(name):
    if name == "kuhn5":
        m = fk.kuhn_cube_p2(5, inside_edge=0.4)
    else:
        m = fk.to_p2(fk.delaunay_ball(1500, seed=4))
    return M.edgeNumbering(m)


def nu_splines():
    g = np.load(os.path.join(GOLDEN, "gd_270K.npz"))
    B, nu = g["B"], g["nu"]
    return fk.Spline1D(B, nu, bc="nearest"), fk.Spline1D(B, fk.gradient(B, nu), bc="nearest")


@pytest.mark.parametrize("mesh", ["kuhn5", "delaunay"])
def test_nedelec_matrices_load_and_field_against_oracle(mesh):
    m = meshes(mesh)
    rng = np.random.default_rng(7)
    h = fk.Femeko(m, nedelec=True)
    assert h.nv == m.ne and h.nt == m.nt
    rp, ci = h.pattern()
    rpo, cio, v1 = orc.nedelec_stiffness(m)
    assert np.array_equal(rp, rpo) and np.array_equal(ci, cio)
    # local matrices (36 x nt)
    Ak = h.nedelec_local_stiffness()
    Ako = orc.nedelec_local_stiffness(m)
    assert np.max(np.abs(Ak - Ako)) < 1e-12 * np.max(np.abs(Ako))
    # global curl-curl matrix, unscaled and scaled by nu; bitwise run to run
    nu = rng.uniform(0.2, 1.0, m.nt)
    v = h.assemble_nedelec_stiffness(None, slot=0)
    assert np.max(np.abs(v - v1)) < 1e-12 * np.max(np.abs(v1))
    vn = h.assemble_nedelec_stiffness(nu, slot=1)
    _, _, vno = orc.nedelec_stiffness(m, nu)
    assert np.max(np.abs(vn - vno)) < 1e-12 * np.max(np.abs(vno))
    assert np.array_equal(vn, h.assemble_nedelec_stiffness(nu, slot=3))
    items, contrib, slots, _ = h.geo_stats()
    assert contrib == 21 * m.nt and items == (h.nnz + m.ne) // 2
    # tangential matrix on a random state
    dnu = rng.uniform(0.1, 1.0, m.nt)
    Bf = rng.normal(size=(3, m.nt))
    Bn = np.linalg.norm(Bf, axis=0)
    vt = h.assemble_nedelec_tangential(dnu, Bf, Bn, slot=2)
    _, _, vto = orc.nedelec_tangential(m, dnu, Bf, Bn)
    assert np.max(np.abs(vt - vto)) < 1e-12 * np.max(np.abs(vto))
    # the stiffness slot is untouched by the tangential assembly, A + At is a value operation
    h.matrix_axpby(3, 1.0, 1, 1.0, 2)
    assert np.max(np.abs(h.matrix_get(3) - (vno + vto))) < 1e-12 * np.max(np.abs(vno + vto))
    # load vector and field recovery
    Bext = np.array([0.2, -0.5, 1.0])
    rhs = h.nedelec_load_vector(Bext)
    rhso = orc.nedelec_load(m, Bext)
    assert np.max(np.abs(rhs - rhso)) < 1e-12 * np.max(np.abs(rhso))
    assert np.array_equal(rhs, h.nedelec_load_vector(Bext))
    u = rng.normal(size=m.ne)
    Bfg, Bg = h.nedelec_recover_field(u)
    Bfo, Bo = orc.nedelec_recover_B(m, u)
    assert np.max(np.abs(Bfg - Bfo)) < 1e-12 * np.max(np.abs(Bfo)) and np.max(np.abs(Bg - Bo)) < 1e-12 * np.max(Bo)
    # SpMV on the edge matrix == scipy on the oracle matrix
    import scipy.sparse as sp

    y = h.spmv(u, slot=1)
    yo = sp.csr_matrix((vno, cio, rpo), shape=(m.ne, m.ne)) @ u
    assert np.max(np.abs(y - yo)) < 1e-12 * np.max(np.abs(yo))
    h.close()


def test_nedelec_drop_in_functions():
    """nedelecLocalStiffness(mesh), nedelecTangentialStiffness(mesh, dnu, Bfield, B) with the reference's signatures."""
    m = meshes("kuhn5")
    rng = np.random.default_rng(1)
    Ak = fk.nedelecLocalStiffness(m)
    assert Ak.shape == (36, m.nt)
    dnu, Bf = rng.uniform(0.1, 1.0, m.nt), rng.normal(size=(3, m.nt))
    At = fk.nedelecTangentialStiffness(m, dnu, Bf, np.linalg.norm(Bf, axis=0))
    _, _, vto = orc.nedelec_tangential(m, dnu, Bf, np.linalg.norm(Bf, axis=0))
    assert At.shape == (m.ne, m.ne) and np.max(np.abs(At.vals - vto)) < 1e-12 * np.max(np.abs(vto))
    A = fk.nedelecStiffnessMatrix(m)
    assert abs(A.toscipy() - A.toscipy().T).max() < 1e-13 * np.max(np.abs(A.vals))
    m._fk_edge.close()


def test_nedelec_uniform_medium_returns_the_applied_field():
    """nu = 1 everywhere: the solution of A u = RHS(Bext) is the interpolant of Bext x r / 2 (up to a gradient), so the
    recovered field is Bext on every element -- what the first pass of nedelec3D.jl computes."""
    m = meshes("delaunay")
    h = fk.Femeko(m, nedelec=True)
    Bext = np.array([0.3, -0.2, 1.0])
    h.assemble_nedelec_stiffness(None, slot=0, want_vals=False)
    rhs = h.nedelec_load_vector(Bext)
    u, it, rr = h.pcg(rhs, slot=0, tol=1e-12, max_iter=20000)
    assert rr < 1e-11 and it < 20000
    Bf, B = h.nedelec_recover_field(u)
    assert np.max(np.abs(Bf - Bext[:, None])) < 1e-8 and np.max(np.abs(B - np.linalg.norm(Bext))) < 1e-8
    _, nr = h.residual(u, rhs, slot=0)
    assert nr < 1e-10 * np.linalg.norm(rhs)
    h.close()


def test_nedelec_newton_loop_against_oracle():
    """The whole loop of nedelec3D.jl:148-250 (GPU: matrices, Jacobi-CG, lineSearch, B recovery, nu / dnu splines) against the
    oracle's restatement with scipy's CG: same number of passes, same B, nu, convergence metric."""
    m = M.edgeNumbering(fk.kuhn_cube_p2(4, inside_edge=0.5))
    spl, spld = nu_splines()
    Bext = [0.0, 0.0, 1.0]
    ref = orc.nedelec_newton_loop(m, Bext, (spl.t, spl.c), (spld.t, spld.c), tol=1e-6, max_att=30, cg_tol=1e-12)
    res = fk.nedelec3D(m, Bext, spl, spld, tol=1e-6, maxAtt=30, cg_tol=1e-12)
    assert res.att == ref["att"] and res.att < 30
    assert np.linalg.norm(res.Bfield - ref["Bfield"]) < 1e-6 * np.linalg.norm(ref["Bfield"])
    ins = m.InsideElements
    assert np.max(np.abs(res.nu[ins] - ref["nu"][ins])) < 1e-6 * np.max(ref["nu"][ins])
    assert np.all(res.nu[np.setdiff1d(np.arange(m.nt), ins)] == 1.0)
    for (a, d, *_), (ao, do, *_) in zip(res.history, ref["history"]):
        assert a == ao and abs(d - do) < 1e-5 * max(1.0, do)
    # the first pass is the uniform medium: B = Bext
    assert abs(res.history[0][1] - 1.0) < 1e-8
