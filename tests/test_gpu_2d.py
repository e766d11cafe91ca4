"""GPU: 2-D P1 triangles (src/FEM2D.jl:48-76 stiffnessMatrix2D, :139-161 massMatrix2D) through the tet engine
(elements (n0, n1, n2, n2)): pattern and values against the oracle restatement, determinism, and a heat-type solve
(heat2D: T <- (M + dt A)^-1 (M T)) against a sparse direct solve."""
import numpy as np
import pytest

import femeko_jl_b200 as fk
from conftest import rel_l2
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mesh", ["grid", "disc"])
def test_2d_stiffness_and_mass_against_oracle(mesh):
    m = fk.tri_grid(24) if mesh == "grid" else fk.delaunay_disc(1500, seed=11)
    mu = np.random.default_rng(1).uniform(0.5, 3.0, m.nt)
    h = fk.Femeko(m)
    rp, ci = h.pattern()
    orp, oci, vref = orc.stiffness2d(m, mu)
    assert np.array_equal(rp, orp) and np.array_equal(ci, oci)
    v = h.assemble_stiffness2d(mu, slot=0)
    scale = np.abs(vref).max()
    assert np.max(np.abs(v - vref)) < 1e-12 * scale
    assert np.array_equal(h.assemble_stiffness2d(mu, slot=2), v)          # bitwise run to run
    _, _, mref = orc.mass2d(m)
    vm = h.assemble_mass2d(slot=1)
    assert np.max(np.abs(vm - mref)) < 1e-13 * np.abs(mref).max() and np.isclose(vm.sum(), m.VE.sum(), rtol=1e-13)
    # computed areas (no VE handed over) give the same matrices to rounding
    h2 = fk.Femeko(m, use_mesh_VE=False)
    assert np.max(np.abs(h2.assemble_stiffness2d(mu, slot=0) - vref)) < 1e-12 * scale
    h2.close()
    h.close()


def test_2d_heat_steps_against_direct_solve():
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla

    m = fk.delaunay_disc(3000, seed=2)
    k = np.ones(m.nt)
    k[m.InsideElements] = 2.0
    A = fk.stiffnessMatrix2D(m, k)
    M = fk.massMatrix2D(m)
    h = A.fk
    rp, ci = h.pattern()
    As = sp.csr_matrix((A.vals if hasattr(A, "vals") else h.matrix_get(0), ci, rp), shape=(m.nv, m.nv))
    Ms = sp.csr_matrix((h.matrix_get(1), ci, rp), shape=(m.nv, m.nv))
    dt = 1e-3
    h.matrix_axpby(3, 1.0, 1, dt, 0)
    T = np.zeros(m.nv)
    T[m.InsideNodes] = 10.0
    Tref = T.copy()
    lu = spla.splu((Ms + dt * As).tocsc())
    for _ in range(5):
        T, it, rr = h.heat_step(T, slot_mass=1, slot_lhs=3, tol=1e-12)
        Tref = lu.solve(Ms @ Tref)
    assert rel_l2(T, Tref) < 1e-9
    assert abs((Ms @ T).sum() - (Ms @ Tref).sum()) < 1e-9 * abs((Ms @ Tref).sum())
