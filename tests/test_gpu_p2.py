"""GPU: P2 (10-node) tetrahedra -- quadraticStiffnessMatrix (src/FEM3D.jl:324-395) and quadraticMassMatrix
(src/FEM3D.jl:494-561) through the contribution plan, against the oracle restatement (fitted basis functions, literal
loops), plus a Poisson-type solve with Dirichlet data on the P2 matrix against a sparse direct solve."""
import numpy as np
import pytest

import femeko_jl_b200 as fk
from conftest import rel_l2
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mesh", ["kuhn3", "delaunay"])
def test_p2_stiffness_and_mass_against_oracle(mesh):
    m = fk.kuhn_cube_p2(3, inside_edge=0.5) if mesh == "kuhn3" else fk.to_p2(fk.delaunay_ball(120, seed=3))
    rng = np.random.default_rng(4)
    mu = rng.uniform(0.5, 3.0, m.nt)
    h = fk.Femeko(m)
    rp, ci = h.pattern()
    orp, oci = orc.pattern_npe(m.t, m.nv)
    assert np.array_equal(rp, orp) and np.array_equal(ci, oci)
    S = orc.quadratic_basis_all(m)
    _, _, vref = orc.quadratic_stiffness(m, S, mu)
    v = h.assemble_stiffness_p2(mu, slot=0)
    print(f"P2 stiffness {mesh}: max |dA| / max |A| = {np.max(np.abs(v - vref)) / np.abs(vref).max():.2e}")
    assert np.max(np.abs(v - vref)) < 1e-10 * np.abs(vref).max()       # the reference's 10 x 10 fits carry ~1e-12 themselves
    assert np.array_equal(h.assemble_stiffness_p2(mu, slot=2), v)       # bitwise run to run
    F = rng.uniform(1.0, 2.0, m.nt)
    _, _, mref = orc.quadratic_mass(m, S, F)
    vm = h.assemble_mass_p2(F, slot=1)
    print(f"P2 mass {mesh}: max |dM| / max |M| = {np.max(np.abs(vm - mref)) / np.abs(mref).max():.2e}")
    assert np.max(np.abs(vm - mref)) < 1e-10 * np.abs(mref).max()
    items, contrib, slots, _ = h.geo_stats()
    assert contrib == 55 * m.nt and items == (h.nnz + m.nv) // 2
    h.close()


def test_p2_poisson_solve_against_direct():
    """-div grad u = f on the unit cube with u = g on the boundary, P2 elements: PCG on the GPU matrix (Dirichlet
    elimination on the pattern) against scipy on the same matrix; quadratic exact solutions are reproduced."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla

    m = fk.kuhn_cube_p2(6)
    h = fk.Femeko(m)
    rp, ci = h.pattern()
    A = sp.csr_matrix((h.assemble_stiffness_p2(None, slot=0), ci, rp), shape=(m.nv, m.nv))
    x, y, z = m.p
    uex = x * x - y * y + 0.5 * x * z + 0.3 * y                         # harmonic quadratic: in the P2 space, f = 0
    fixed = np.flatnonzero((np.abs(np.abs(m.p) - 0.5) < 1e-12).any(axis=0)).astype(np.int32)
    free = np.setdiff1d(np.arange(m.nv), fixed)
    h.set_dirichlet(fixed, uex[fixed])
    b = h.apply_dirichlet(0, np.zeros(m.nv))
    u, it, rr = h.pcg(b, slot=0, tol=1e-12)
    ud = uex.copy()
    ud[free] = spla.spsolve(A[free][:, free].tocsc(), -A[free][:, fixed] @ uex[fixed])
    assert rel_l2(u, ud) < 1e-8 and np.array_equal(u[fixed], uex[fixed])
    # the reference's stiffness uses the mean over the 10 nodes as quadrature: not exact for P2 x P2, so the discrete
    # solution is only close to the quadratic (the direct solve of the same matrix is the parity check above)
    print(f"P2 Poisson: {it} PCG iterations, |u - u_exact| / |u_exact| = {rel_l2(u, uex):.2e}")
    h.close()


@pytest.mark.parametrize("mesh", ["kuhn4", "delaunay"])
@pytest.mark.parametrize("compat", [True, False])
def test_p2_convection_matrix_against_oracle(compat, mesh):
    """quadraticConvectionMatrix (src/FEM3D.jl:563-648) through fk_assemble_convection_p2 (non-symmetric values through the
    two-factor contribution sums) against the oracle's literal restatement -- including the reference's transposed assembly."""
    from oracle import oracle as orc

    m = fk.kuhn_cube_p2(4, inside_edge=0.5) if mesh == "kuhn4" else fk.to_p2(fk.delaunay_ball(800, seed=6))
    tol = 1e-12 if mesh == "kuhn4" else 1e-10      # the reference's 10 x 10 fits lose digits on the slivers of a Delaunay mesh
    rng = np.random.default_rng(3)
    u = rng.normal(size=(3, m.nv))
    F = rng.uniform(0.5, 2.0, m.nt)
    h = fk.Femeko(m)
    v = h.assemble_convection_p2(u, F, slot=2, compat_bug=compat)
    rp, ci, vo = orc.quadratic_convection(m, u, None, F, compat_bug=compat)
    hp = h.pattern()
    assert np.array_equal(hp[0], rp) and np.array_equal(hp[1], ci)
    assert np.max(np.abs(v - vo)) < tol * np.max(np.abs(vo))
    assert np.array_equal(v, h.assemble_convection_p2(u, F, slot=3, compat_bug=compat))     # bitwise run to run
    v1 = h.assemble_convection_p2(u, None, slot=2, compat_bug=compat)
    _, _, v1o = orc.quadratic_convection(m, u, None, None, compat_bug=compat)
    assert np.max(np.abs(v1 - v1o)) < tol * np.max(np.abs(v1o))
    if not compat:                                                          # rows sum to zero: C 1 = 0 through the device SpMV
        y = h.spmv(np.ones(m.nv), slot=2)
        assert np.max(np.abs(y)) < 1e-12 * np.max(np.abs(v1))
    # the stiffness / mass factors of the handle are untouched by the convection scratch
    K = h.assemble_stiffness_p2(None, slot=0)
    _, _, Ko = orc.quadratic_stiffness(m)
    assert np.max(np.abs(K - Ko)) < tol * np.max(np.abs(Ko))
    A = fk.quadraticConvectionMatrix(m, None, u, F, compat_bug=compat)
    assert np.array_equal(A.vals, v)
    h.close()
