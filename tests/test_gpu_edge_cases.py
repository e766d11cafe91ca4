"""GPU: edge cases of the hot path through the C ABI -- the smallest mesh, an unused node, an empty material list, a zero
right-hand side, iteration limits, and the error paths (every one a message through fk_last_error, never a crash)."""
import os

import numpy as np
import pytest

import femeko_jl_b200 as fk
from conftest import GOLDEN
from femeko_jl_b200 import lib as L
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
MU0 = 4e-7 * np.pi


def tiny_mesh(extra_nodes=0):
    """One tetrahedron (plus `extra_nodes` points no element uses)."""
    m = fk.MESH()
    pts = np.array([[0.0, 0.0, 0.0], [1.0, 0.1, 0.0], [0.2, 0.9, 0.1], [0.1, 0.2, 0.8]] + [[2.0 + i, 2.0, 2.0] for i in range(extra_nodes)]).T
    m.p = np.asfortranarray(pts)
    m.t = np.asfortranarray(np.array([[0], [1], [2], [3]], dtype=np.int32))
    m.nv, m.nt = m.p.shape[1], 1
    m.VE = fk.elementVolume(m.p, m.t)
    m.InsideElements = np.array([0], dtype=np.int32)
    m.InsideNodes = np.arange(4, dtype=np.int32)
    m.nInside, m.nInsideNodes = 1, 4
    m.surfaceT = np.zeros((4, 0), dtype=np.int32, order="F")
    m.ns = 0
    return m


@pytest.mark.parametrize("geo_mode", [0, 2])
def test_single_tetrahedron(geo_mode):
    m = tiny_mesh()
    h = fk.Femeko(m)
    h.tune(10, geo_mode)
    rp, ci = h.pattern()
    assert np.array_equal(rp, [0, 4, 8, 12, 16]) and np.array_equal(ci, np.tile(np.arange(4), 4))
    f = np.array([2.5])
    v = h.assemble_stiffness(f, slot=0)
    _, _, vo = orc.stiffness(m, f, order="eigen")
    assert np.max(np.abs(v - vo)) < 1e-14 * np.max(np.abs(vo))
    A = v.reshape(4, 4)
    assert np.max(np.abs(A.sum(axis=1))) < 1e-14 * np.max(np.abs(A)) and np.max(np.abs(A - A.T)) == 0.0
    M = h.assemble_mass(None, slot=1)
    assert abs(M.sum() - m.VE[0]) < 1e-15
    x = np.array([0.3, -1.0, 0.5, 0.2])
    y = h.spmv(x, slot=0)
    assert np.max(np.abs(y - A @ x)) < 1e-14 * np.max(np.abs(A))
    b = A @ x                                                   # consistent right-hand side of the singular system
    u, it, rr = h.pcg(b, slot=0, tol=1e-12)
    assert it <= 4 and np.linalg.norm(b - A @ u) < 1e-10 * np.linalg.norm(b)
    Hf, Hn, _ = h.recover_field(x)
    Hfo, Hno = orc.recover_field(m, x)
    assert np.max(np.abs(Hf - Hfo)) < 1e-13 * np.max(np.abs(Hfo)) and abs(Hn[0] - Hno[0]) < 1e-13 * Hno[0]
    h.close()


def test_unused_nodes_get_an_empty_row_that_the_solver_tolerates():
    m = tiny_mesh(extra_nodes=2)
    h = fk.Femeko(m)
    v = h.assemble_stiffness(None, slot=0)
    rp, ci = h.pattern()
    import scipy.sparse as sp

    A = sp.csr_matrix((v, ci, rp), shape=(m.nv, m.nv)).toarray()
    assert np.all(A[4:] == 0.0) and np.all(A[:, 4:] == 0.0)
    x = np.array([0.3, -1.0, 0.5, 0.2, 0.0, 0.0])
    b = A @ x
    u, it, rr = h.pcg(b, slot=0, tol=1e-12)                     # Eigen's Jacobi: 1 where the diagonal is 0 (BasicPreconditioners.h:62-75)
    assert np.all(np.isfinite(u)) and np.all(u[4:] == 0.0) and np.linalg.norm(b - A @ u) < 1e-10 * np.linalg.norm(b)
    h.close()


def test_zero_rhs_iteration_limits_and_empty_material_list():
    m = fk.kuhn_cube(6, inside_edge=0.5)
    h = fk.Femeko(m)
    mu = np.full(m.nt, MU0)
    v = h.assemble_stiffness(mu, slot=0)
    # b = 0: Eigen returns x = 0 after 0 iterations (ConjugateGradient.h:43-48)
    u, it, rr = h.pcg(np.zeros(m.nv), slot=0)
    assert it == 0 and np.all(u == 0.0)
    rng = np.random.default_rng(0)
    b = rng.normal(size=m.nv)
    b -= b.mean()
    # exactly max_iter iterations when the tolerance cannot be met, exactly that many with tol_disable
    _, it1, _ = h.pcg(b, slot=0, tol=1e-300, max_iter=1)
    _, it7, _ = h.pcg(b, slot=0, tol_disable=True, max_iter=7)
    _, it40, _ = h.pcg(b, slot=0, tol_disable=True, max_iter=40, check_every=16)
    assert (it1, it7, it40) == (1, 7, 40)
    # no element carries a B-H curve: the sweep is a plain stiffness assembly with the stored mu, div = 0
    g = np.load(os.path.join(GOLDEN, "gd_270K.npz"))
    h.set_spline(0, fk.Spline1D(g["HofM"], g["mu"], bc="nearest"))
    h.set_inside(np.zeros(0, dtype=np.int32))
    h.set_mu(mu)
    for sweep in range(3):                                      # tile kernel first, rescale path from the second sweep on
        div = h.picard_step(rng.normal(size=m.nv) * 1e4, slot=1)
        assert div == 0.0
        v1 = h.matrix_get(1)
        assert np.max(np.abs(v1 - v)) <= 1e-14 * np.max(np.abs(v)) and np.array_equal(h.get_mu(), mu)
    h.close()


def test_error_paths_report_and_do_not_crash():
    m = tiny_mesh()
    bad = tiny_mesh()
    bad.t = np.asfortranarray(np.array([[0], [1], [2], [7]], dtype=np.int32))
    with pytest.raises(L.FemekoError, match="out of range"):
        fk.Femeko(bad, use_mesh_VE=False)
    h = fk.Femeko(m)
    with pytest.raises(L.FemekoError, match="not been assembled|empty"):
        h.pcg(np.ones(4), slot=2)
    with pytest.raises(L.FemekoError, match="slot"):
        h.assemble_stiffness(None, slot=17)
    with pytest.raises(ValueError, match="length nt"):
        h.assemble_stiffness(np.ones(3), slot=0)
    with pytest.raises(L.FemekoError, match="P2"):
        h.assemble_stiffness_p2(None, slot=0)
    with pytest.raises(L.FemekoError, match="edge elements"):
        h.nedelec_local_stiffness()
    # a spline evaluated outside its knots with bc = "error" reports, the handle stays usable
    spl = fk.Spline1D(np.linspace(0.0, 1.0, 8), np.linspace(1.0, 2.0, 8) ** 2, bc="error")
    h.set_spline(0, spl)
    with pytest.raises(L.FemekoError, match="knot range"):
        h.spline_apply(0, np.array([0.5, 3.0]))
    assert np.allclose(h.spline_apply(0, np.array([0.5])), spl(np.array([0.5])), rtol=1e-13)
    v = h.assemble_stiffness(None, slot=0)
    assert np.all(np.isfinite(v))
    h.close()


def test_host_register_keeps_results_and_reports_misuse():
    """fk_host_register / fk_host_unregister: page-locking a caller's (pageable) arrays in place changes copy speed only."""
    import ctypes as C

    m = fk.kuhn_cube(8, inside_edge=0.5)
    h = fk.Femeko(m)
    h.assemble_stiffness(None, slot=0, want_vals=False)
    rng = np.random.default_rng(1)
    b = rng.normal(size=m.nv)
    b -= b.mean()
    x0, it0, _ = h.pcg(b, slot=0, tol=1e-10)
    lib = L.load()
    bb, xx = b.copy(), np.zeros(m.nv)
    for a in (bb, xx):
        L.check(lib.fk_host_register(C.c_void_p(a.ctypes.data), a.nbytes))
    o = h.pcg_opts(tol=1e-10)
    it, rr = C.c_int(), C.c_double()
    L.check(lib.fk_pcg(h.h, 0, L.dptr(bb), L.dptr(xx), C.byref(o), C.byref(it), C.byref(rr)))
    assert it.value == it0 and np.array_equal(xx, x0)
    for a in (bb, xx):
        L.check(lib.fk_host_unregister(C.c_void_p(a.ctypes.data)))
    with pytest.raises(L.FemekoError, match="CUDA error"):
        L.check(lib.fk_host_unregister(C.c_void_p(bb.ctypes.data)))      # not registered any more
    with pytest.raises(L.FemekoError, match="NULL"):
        L.check(lib.fk_host_register(None, 8))
    x1, it1, _ = h.pcg(b, slot=0, tol=1e-10)                             # the handle is still fine after the reported errors
    assert it1 == it0 and np.array_equal(x1, x0)
    h.close()
