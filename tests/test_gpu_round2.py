"""GPU parity tests added in round 2 (all through the C ABI):
  * the single-reduction PCG recurrence against Eigen's recurrence and the oracle;
  * Dirichlet elimination against scipy elimination on the oracle-assembled matrix (Fluids/fluids3D.jl:142-174);
  * null-space projection against the reference's bordered solve (magneticField.jl:99-107);
  * lineSearch (src/FEM3D.jl:123-152), the Picard loop (manyBodyIntersecting.jl:141-210) and the Newton-Raphson loop
    (NewtonRaphson.jl:148-214) end to end against the oracle's restated loops, real Gd table;
  * the device spline against the scipy-FITPACK samples of tests/golden/gd_270K.npz;
  * the exact step bench.py times (Kuhn N=119, Gd table, 100 fixed PCG iterations) against the oracle, the compiled
    reference live at N=48, and an unordered ~0.2 M-tet Delaunay mesh (configs[0] size).
"""
import os

import numpy as np
import pytest

import femeko_jl_b200 as fk
from conftest import GOLDEN, abs_sum_scale, rel_err_matrix, rel_l2
from femeko_jl_b200 import lib as L
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def gd():
    return dict(np.load(os.path.join(GOLDEN, "gd_270K.npz")))


def gd_splines():
    g = gd()
    return fk.Spline1D(g["HofM"], g["mu"], bc="nearest"), fk.Spline1D(g["HofM"], g["dmu"], bc="nearest")


def plain_max_err(v, ref, rp, ci):
    """max |delta a_ij| / |a_ii| (the SURVEY 8c metric, printed beside the Cauchy-Schwarz one for comparison)."""
    rows = np.repeat(np.arange(len(rp) - 1, dtype=np.int64), np.diff(rp))
    diag = np.zeros(len(rp) - 1)
    on = ci == rows
    diag[rows[on]] = np.abs(ref[on])
    return float(np.max(np.abs(v - ref) / diag[rows]))


# --------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("mesh", ["kuhn16", "delaunay"])
def test_single_reduction_recurrence(mesh):
    m = fk.kuhn_cube(16) if mesh == "kuhn16" else fk.delaunay_ball(3000, seed=21)
    mu = np.full(m.nt, fk.mu0)
    mu[m.InsideElements] = 3 * fk.mu0
    h = fk.Femeko(m)
    rp, ci = h.pattern()
    v = h.assemble_stiffness(mu)
    b = -h.boundary_integral(m.surfaceT, m.normal, m.AE, [1.0, 0.2, -0.3], 0)
    xe, ite, rre = h.pcg(b, tol=1e-10)                              # Eigen's recurrence (ConjugateGradient.h:59-82)
    xs, its, rrs = h.pcg(b, tol=1e-10, recurrence=1)
    assert abs(its - ite) <= 2 and rrs < 1e-10
    assert rel_l2(xs - xs.mean(), xe - xe.mean()) < 1e-8
    assert np.array_equal(xs, h.pcg(b, tol=1e-10, recurrence=1)[0])                 # bitwise run to run
    x2, it2, _ = h.pcg(b, tol=1e-10, recurrence=1, check_every=2)
    assert it2 == its and np.array_equal(x2, xs)
    # fixed count (bench mode) against the oracle's Eigen loop: same Krylov iterate
    x3, it3, _ = h.pcg(b, max_iter=25, tol_disable=True, recurrence=1)
    xo3, _, _, _ = orc.pcg(rp, ci, v, b, 0.0, 25)
    assert it3 == 25 and rel_l2(x3, xo3) < 1e-9
    # odd fixed count, max_iter hit exactly at a chunk end, warm start, zero right-hand side
    x4, it4, rr4 = h.pcg(b, tol=1e-14, max_iter=7, recurrence=1)
    xo4, _, _, _ = orc.pcg(rp, ci, v, b, 0.0, 7)
    assert it4 == 7 and rr4 > 1e-14 and rel_l2(x4, xo4) < 1e-10
    _, it5, _ = h.pcg(b, x0=xs, tol=1e-9, recurrence=1)
    assert it5 <= 1
    xz, itz, rz = h.pcg(np.zeros(m.nv), recurrence=1)
    assert itz == 0 and rz == 0.0 and not xz.any()
    # the true residual of the converged iterate
    import scipy.sparse as sp
    A = sp.csr_matrix((v, ci, rp), shape=(m.nv, m.nv))
    assert np.linalg.norm(b - A @ xs) < 5e-10 * np.linalg.norm(b)
    h.close()


def test_dirichlet_elimination_against_scipy():
    """LHS[free,free] \\ (-LHS[free,fixed] gD) with UP[fixed] = gD (Fluids/fluids3D.jl:142-174), plus a source term."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla

    m = fk.delaunay_ball(2500, seed=5)
    rng = np.random.default_rng(2)
    f = rng.uniform(0.5, 2.0, m.nt)
    rp, ci, vref = orc.stiffness(m, f, order="julia")
    A = sp.csr_matrix((vref, ci, rp), shape=(m.nv, m.nv))
    fixed = np.unique(m.surfaceT[:3]).astype(np.int32)              # the outer shell (and cavity walls)
    gD = np.sin(m.p[0, fixed]) + 0.3 * m.p[2, fixed]
    b = rng.normal(size=m.nv) * 1e-2
    free = np.setdiff1d(np.arange(m.nv), fixed)
    xr = np.zeros(m.nv)
    xr[fixed] = gD
    xr[free] = spla.spsolve(A[free][:, free].tocsc(), b[free] - A[free][:, fixed] @ gD)
    h = fk.Femeko(m)
    h.assemble_stiffness(f, slot=0, want_vals=False)
    h.set_dirichlet(fixed, gD)
    bl = h.apply_dirichlet(0, b)
    assert np.array_equal(bl[fixed], gD) and rel_l2(bl[free], b[free] - A[free][:, fixed] @ gD) < 1e-13
    vm = h.matrix_get(0)
    Am = sp.csr_matrix((vm, ci, rp), shape=(m.nv, m.nv))
    Ag = sp.csr_matrix((h.assemble_stiffness(f, slot=1), ci, rp), shape=(m.nv, m.nv))   # the handle's own unconstrained matrix
    assert abs(Am[fixed][:, free]).sum() == 0 and abs(Am[free][:, fixed]).sum() == 0
    assert np.array_equal(Am[fixed][:, fixed].diagonal(), np.ones(len(fixed)))
    assert np.array_equal(Am[free][:, free].data, Ag[free][:, free].data)          # the free block is untouched, bit for bit
    for rec in (0, 1):
        x, it, rr = h.pcg(bl, slot=0, tol=1e-12, recurrence=rec)
        assert np.array_equal(x[fixed], gD) and rel_l2(x, xr) < 1e-8
    # homogeneous values (gD = None) and clearing the set
    h.assemble_stiffness(f, slot=0, want_vals=False)
    h.set_dirichlet(fixed)
    bl0 = h.apply_dirichlet(0, b)
    x0, _, _ = h.pcg(bl0, slot=0, tol=1e-12)
    xr0 = np.zeros(m.nv)
    xr0[free] = spla.spsolve(A[free][:, free].tocsc(), b[free])
    assert rel_l2(x0, xr0) < 1e-8
    h.set_dirichlet([])
    h.assemble_stiffness(f, slot=0, want_vals=False)
    h.apply_dirichlet(0, b)
    assert np.array_equal(h.matrix_get(0), h.assemble_stiffness(f, slot=1))
    h.close()


@pytest.mark.parametrize("mesh", ["kuhn12", "delaunay"])
def test_nullspace_projection_against_bordered_solve(mesh):
    """u = ([A Lag; Lag' 0] \\ [-RHS; 0])[1:nv] (magneticField.jl:99-107) = PCG with project_nullspace, w = lagrange."""
    m = fk.kuhn_cube(12, inside_edge=0.4) if mesh == "kuhn12" else fk.delaunay_ball(2500, seed=6)
    mu = np.full(m.nt, fk.mu0)
    mu[m.InsideElements] = 5 * fk.mu0
    rp, ci, vref = orc.stiffness(m, mu, order="julia")
    Lag = orc.lagrange(m)
    RHS = orc.boundary_integral(m, [0.3, -0.2, 1.0], [0], "julia")
    ur = orc.bordered_solve(orc._csr(rp, ci, vref, m.nv), Lag, -RHS)
    h = fk.Femeko(m)
    h.assemble_stiffness(mu, want_vals=False)
    h.set_nullspace(h.lagrange())
    for rec in (0, 1):
        u, it, rr = h.pcg(-RHS, tol=1e-13, project_nullspace=True, recurrence=rec)
        assert abs(Lag @ u) < 1e-12 * np.abs(Lag).sum() * np.abs(u).max()
        assert rel_l2(u, ur) < 1e-8
    # an inconsistent right-hand side: the multiplier term lambda*Lag makes it consistent, as the bordered system does
    b2 = -RHS + 1e-3 * np.abs(RHS).max()
    ur2 = orc.bordered_solve(orc._csr(rp, ci, vref, m.nv), Lag, b2)
    u2, _, _ = h.pcg(b2, tol=1e-13, project_nullspace=True)
    assert rel_l2(u2, ur2) < 1e-8
    # Eigen's default tolerance (machine epsilon) is reachable on the singular system once r is kept in range(A)
    u3, it3, rr3 = h.pcg(-RHS, project_nullspace=True)
    assert rr3 <= np.finfo(float).eps and it3 < 2 * m.nv and rel_l2(u3, ur) < 1e-8
    h.close()


def test_line_search_against_restatement():
    """src/FEM3D.jl:123-152: same trial sequence, same accepted step, same residue."""
    m = fk.kuhn_cube(10, inside_edge=0.4)
    rng = np.random.default_rng(12)
    f = rng.uniform(1.0, 3.0, m.nt)
    h = fk.Femeko(m)
    rp, ci = h.pattern()
    v = h.assemble_stiffness(f)
    A = orc._csr(rp, ci, v, m.nv)
    RHS = rng.normal(size=m.nv)
    RHS -= RHS.mean()
    xs, _, _ = h.pcg(RHS, tol=1e-12)
    for scale, expect_backtrack in ((1.0, False), (2.5, True), (40.0, True)):
        u0 = rng.normal(size=m.nv) * 1e-3
        du = scale * (xs - u0)                 # scale 1 lands on the solution; > 2 overshoots and must backtrack
        ur = u0.copy()
        res_r, alf_r = orc.line_search(ur, du, A, RHS)
        ug = u0.copy()
        res_g, alf_g = h.line_search(ug, du, RHS, slot=0)
        assert alf_g == alf_r and (alf_r < 1.0) == expect_backtrack
        assert np.isclose(res_g, res_r, rtol=1e-9) and rel_l2(ug, ur) < 1e-13
        A_obj = fk.SparseCSR(h, 0, v)
        uj = u0.copy()
        assert np.isclose(fk.lineSearch(uj, du, A_obj, RHS), res_r, rtol=1e-9) and np.array_equal(uj, ug)
    # a step that cannot decrease the residue: u stays untouched (the reference's warning branch)
    ub = xs.copy()
    res_b, alf_b = h.line_search(ub, rng.normal(size=m.nv), RHS, slot=0)
    assert alf_b == 0.0 and np.array_equal(ub, xs)
    r, nr = h.residual(u0, RHS, slot=0)
    assert rel_l2(r, RHS - A @ u0) < 1e-13 and np.isclose(nr, np.linalg.norm(RHS - A @ u0), rtol=1e-12)
    h.close()


@pytest.mark.parametrize("mesh", ["kuhn16", "delaunay"])
def test_picard_loop_against_oracle(mesh):
    """manyBodyIntersecting.jl:141-210 end to end, real Gd table: converged potential / field and the pass count."""
    m = fk.kuhn_cube(16, inside_edge=0.4) if mesh == "kuhn16" else fk.delaunay_ball(3000, seed=31)
    spl, _ = gd_splines()
    ref = orc.picard_loop(m, [0.5, 0.0, 0.0], (spl.t, spl.c), tol=1e-4, solver="direct")
    for rec in (0, 1):
        out = fk.picard(m, [0.5, 0.0, 0.0], spl, picardDeviation=1e-4, recurrence=rec)
        assert out.att == ref["att"]
        assert rel_l2(out.u, ref["u"]) < 1e-8 and rel_l2(out.Hfield, ref["Hfield"]) < 1e-8 and rel_l2(out.mu, ref["mu"]) < 1e-8
        assert np.isclose(out.div, ref["div"], rtol=1e-6)


@pytest.mark.parametrize("mesh,compat", [("kuhn12", True), ("kuhn12", False), ("delaunay", False)])
def test_newton_loop_against_oracle(mesh, compat):
    """NewtonRaphson.jl:148-214 end to end (tangential matrix, dmu/dH spline, lineSearch; compat = the row-2 defect of
    :187-189 reproduced on both sides)."""
    m = fk.kuhn_cube(12, inside_edge=0.4) if mesh == "kuhn12" else fk.delaunay_ball(2500, seed=32)
    spl, dspl = gd_splines()
    ref = orc.newton_loop(m, [0.5, 0.0, 0.0], (spl.t, spl.c), (dspl.t, dspl.c), tol=1e-4, compat_bug=compat)
    out = fk.newton_raphson(m, [0.5, 0.0, 0.0], spl, dspl, maxDeviation=1e-4, compat_bug=compat)
    assert out.att == ref["att"]
    assert [hh[4] for hh in out.history] == [hh[3] for hh in ref["history"]]       # the accepted line-search steps
    assert rel_l2(out.u, ref["u"]) < 1e-8 and rel_l2(out.Hfield, ref["Hfield"]) < 1e-8
    assert np.isclose(out.div, ref["div"], rtol=1e-5)


def test_device_spline_against_fitpack_fixture():
    """The real Gd table (tests/golden/gd_270K.npz: knots, coefficients and samples from scipy's FITPACK, the Fortran
    Dierckx.jl wraps): host fit == fixture, device splev == fixture samples."""
    g = gd()
    spl = fk.Spline1D(g["HofM"], g["mu"], bc="nearest")
    assert np.allclose(spl.t, g["spl_t"], rtol=0, atol=0) and np.allclose(spl.c[:len(g["HofM"])], g["spl_c"][:len(g["HofM"])], rtol=1e-11)
    m = fk.kuhn_cube(3)
    h = fk.Femeko(m)
    h.set_spline(0, spl)
    y = h.spline_apply(0, g["H_samples"])
    assert rel_l2(y, g["mu_samples"]) < 1e-12 and np.max(np.abs(y - g["mu_samples"]) / np.abs(g["mu_samples"])) < 1e-11
    # the fixture's own knots/coefficients on the device, and the bc variants
    class Tck:
        t, c, k, bc = g["spl_t"], np.concatenate([g["spl_c"], np.zeros(len(g["spl_t"]) - len(g["spl_c"]))]), 3, "nearest"
    h.set_spline(1, Tck)
    assert rel_l2(h.spline_apply(1, g["H_samples"]), g["mu_samples"]) < 1e-13
    hi = g["HofM"][-1]
    assert h.spline_apply(0, [2 * hi])[0] == h.spline_apply(0, [hi])[0]              # nearest
    h.set_spline(2, fk.Spline1D(g["HofM"], g["mu"], bc="error"))
    with pytest.raises(L.FemekoError, match="outside the knot range"):
        h.spline_apply(2, [2 * hi])
    h.close()


# --------------------------------------------------------------------------------------------------
# parity where the numbers are taken
# --------------------------------------------------------------------------------------------------
def test_bench_step_against_oracle_at_full_size():
    """BASELINE configs[1] exactly as bench.py runs it (Kuhn N=119, Gd table @270 K, B_ext = 0.5 T, fused Picard sweep +
    100 fixed Jacobi-PCG iterations), every output against the oracle."""
    N = 119
    m = fk.kuhn_cube(N, inside_edge=0.2)
    spl, dspl = gd_splines()
    h = fk.Femeko(m, use_mesh_VE=False)
    rp, ci = h.pattern()
    orp, oci = orc.pattern(m.t, m.nv)
    assert np.array_equal(rp, orp) and np.array_equal(ci, oci)                      # bit-exact
    e2n = h.elem2nnz()
    assert np.array_equal(e2n, orc.elem2nnz(m.t, rp, ci))                           # bit-exact
    nsell, nsl, n16, cb = h.sell_stats()
    assert n16 == nsl
    # linear stiffness, all 25.6 M values, both reference summation orders
    mu_lin = np.full(m.nt, fk.mu0)
    mu_lin[m.InsideElements] = 3 * fk.mu0
    Ak1 = orc.local_stiffness(m.p, m.t, m.VE, None)
    scale = None
    v = h.assemble_stiffness(mu_lin)
    ref_e = orc.assemble_eigen(e2n, h.nnz, Ak1, mu_lin)
    scale = orc.assemble_eigen(e2n, h.nnz, _cs_scale(Ak1), np.abs(mu_lin))
    assert rel_err_matrix(v, ref_e, scale) < 1e-12
    print("stiffness N=119: max |da_ij|/|a_ii| =", plain_max_err(v, ref_e, rp, ci))
    ref_j = orc.assemble_julia(e2n, h.nnz, Ak1 * mu_lin)
    assert rel_err_matrix(v, ref_j, scale) < 1e-12
    del ref_j
    # the bench's potential: linear presolve, then the fused sweep with the real table
    h.set_spline(0, spl)
    h.set_spline(L.FK_MAX_SPLINES // 2, dspl)
    h.set_inside(m.InsideElements)
    mu0v = np.full(m.nt, fk.mu0)
    h.set_mu(mu0v)
    h.assemble_stiffness(mu0v, want_vals=False)
    b = -h.boundary_integral(m.surfaceT, m.normal, m.AE, [0.5, 0.0, 0.0], 0)
    assert rel_l2(b, -orc.boundary_integral(m, [0.5, 0.0, 0.0], [0], "julia")) < 1e-13
    u, it, rr = h.pcg(b, tol=1e-8, max_iter=20000)
    div = h.picard_step(u)
    Hf, Hn = orc.recover_field(m, u)
    ins = m.InsideElements
    mu_ref = mu0v.copy()
    mu_ref[ins] = orc.splev(spl.t, spl.c, 3, Hn[ins], "nearest")
    Hfd, Hnd = h.get_field()
    mud, dmud = h.get_mu(want_dmu=True)
    assert rel_l2(Hfd, Hf) < 1e-10 and rel_l2(Hnd, Hn) < 1e-10
    assert np.max(np.abs(mud - mu_ref) / mu_ref) < 1e-12
    assert rel_l2(dmud[ins], orc.splev(dspl.t, dspl.c, 3, Hn[ins], "nearest")) < 1e-10
    assert np.isclose(div, fk.mu0 * np.max(np.abs(Hn[ins])), rtol=1e-12)
    vp = h.matrix_get(0)
    ref_p = orc.assemble_eigen(e2n, h.nnz, Ak1, mu_ref)
    assert rel_err_matrix(vp, ref_p, orc.assemble_eigen(e2n, h.nnz, _cs_scale(Ak1), mu_ref)) < 1e-12
    del Ak1
    # 100 fixed PCG iterations on the swept matrix: x against the oracle's Eigen loop, both recurrences
    xo, _, _, _ = orc.pcg(rp, ci, ref_p, b, 0.0, 100)
    for rec in (0, 1):
        x, itx, _ = h.pcg(b, slot=0, max_iter=100, tol_disable=True, check_every=100, recurrence=rec)
        assert itx == 100 and rel_l2(x, xo) < 1e-8
    h.close()


def _cs_scale(Ak):
    d = np.abs(np.stack([Ak[5 * i] for i in range(4)]))
    S = np.empty_like(Ak)
    for i in range(4):
        for j in range(4):
            S[4 * j + i] = np.sqrt(d[i] * d[j])
    return S


@pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref (compiled reference) not present")
def test_live_against_compiled_reference_n48():
    """The compiled, unmodified reference (src/FEMc.cpp stiffnessMatrix + Eigen ConjugateGradient) run live at the size
    of the bench's CPU arm (Kuhn N=48): pattern bit-exact, 1.7 M values at 1e-12, CG potential at 1e-8."""
    m = fk.kuhn_cube(48, inside_edge=0.2)
    rng = np.random.default_rng(48)
    mu = fk.mu0 * rng.uniform(1.0, 40.0, m.nt)
    outer, inner, vals = orc.ref_stiffness(m, mu)
    h = fk.Femeko(m)                                         # the reference takes VE from the caller: so does the handle
    rp, ci = h.pattern()
    assert np.array_equal(rp, outer) and np.array_equal(ci, inner)
    v = h.assemble_stiffness(mu)
    Ak = orc.ref_local_stiffness(m)
    assert rel_err_matrix(v, vals, abs_sum_scale(m, Ak, mu)) < 1e-12
    print("stiffness N=48 vs reference: max |da_ij|/|a_ii| =", plain_max_err(v, vals, rp, ci))
    b = -orc.ref_boundary_integral(m, [0.5, 0.0, 0.0], 0)
    assert rel_l2(-h.boundary_integral(m.surfaceT, m.normal, None, [0.5, 0.0, 0.0], 0), b) < 1e-14
    xr, itr, err, info, _ = orc.ref_cg(outer, inner, vals, b, 1e-10, 0)
    x, it, rr = h.pcg(b, tol=1e-10)
    assert info == 0 and abs(it - itr) <= 1 and rel_l2(x - x.mean(), xr - xr.mean()) < 1e-8
    x1, it1, _ = h.pcg(b, tol=1e-10, recurrence=1)
    assert abs(it1 - itr) <= 2 and rel_l2(x1 - x1.mean(), xr - xr.mean()) < 1e-8
    # fixed 100 iterations, as the bench's CPU arm runs Eigen
    xr100 = orc.ref_cg(outer, inner, vals, b, 1e-300, 100)[0]
    x100, _, _ = h.pcg(b, max_iter=100, tol_disable=True)
    assert rel_l2(x100, xr100) < 1e-8
    h.close()


def test_unordered_mesh_at_config0_size():
    """configs[0] size (~0.2 M tets) on an unordered 'Gmsh-like' mesh: Delaunay ball of 36 k seeded points, nodes in
    random order (no locality in the numbering).  Parity vs the oracle, and how the packed column stream copes."""
    m = fk.delaunay_ball(36000, seed=1234)
    assert m.nt > 180000
    rng = np.random.default_rng(7)
    f = fk.mu0 * rng.uniform(1.0, 30.0, m.nt)
    h = fk.Femeko(m)
    rp, ci = h.pattern()
    orp, oci = orc.pattern(m.t, m.nv)
    assert np.array_equal(rp, orp) and np.array_equal(ci, oci)
    e2n = h.elem2nnz()
    assert np.array_equal(e2n, orc.elem2nnz(m.t, rp, ci))
    v = h.assemble_stiffness(f)
    Ak1 = orc.local_stiffness(m.p, m.t, m.VE, None)
    ref = orc.assemble_julia(e2n, h.nnz, orc.local_stiffness(m.p, m.t, m.VE, f))
    assert rel_err_matrix(v, ref, abs_sum_scale(m, Ak1, f)) < 1e-12
    nsell, nsl, n16, cb = h.sell_stats()
    print(f"delaunay 36k: nt={m.nt} nv={m.nv} nnz={h.nnz} slices={nsl} 16-bit slices={n16} "
          f"({100.0 * n16 / nsl:.1f} %), padding {nsell / h.nnz:.3f}x, {h.tile_stats()}")
    assert n16 == nsl                       # 36 k nodes < 65536: every slice fits 16-bit offsets even without locality
    b = -h.boundary_integral(m.surfaceT, m.normal, m.AE, [0.0, 0.0, 1.0], 0)
    xo, ito, _, info = orc.pcg(rp, ci, ref, b, 1e-10)
    for rec in (0, 1):
        x, it, rr = h.pcg(b, tol=1e-10, recurrence=rec)
        assert info == 0 and abs(it - ito) <= 2 and rel_l2(x - x.mean(), xo - xo.mean()) < 1e-8
    Hf, Hn, _ = h.recover_field(x)
    Hfo, Hno = orc.recover_field(m, x)
    assert rel_l2(Hf, Hfo) < 1e-10 and rel_l2(Hn, Hno) < 1e-10
    h.close()
    # a mesh big enough that unordered numbering forces 32-bit slices: same bits as the 16-bit layout of a sorted copy
    m2 = fk.delaunay_ball(90000, seed=99)
    h2 = fk.Femeko(m2)
    ns2, nsl2, n16b, _ = h2.sell_stats()
    print(f"delaunay 90k unordered: slices={nsl2} 16-bit={n16b} ({100.0 * n16b / nsl2:.1f} %)")
    f2 = fk.mu0 * np.random.default_rng(3).uniform(1, 5, m2.nt)
    v2 = h2.assemble_stiffness(f2)
    rp2, ci2 = h2.pattern()
    x2 = np.random.default_rng(4).normal(size=m2.nv)
    assert rel_l2(h2.spmv(x2), orc.spmv(rp2, ci2, v2, x2)) < 1e-13
    ref2 = orc.assemble_julia(h2.elem2nnz(), h2.nnz, orc.local_stiffness(m2.p, m2.t, m2.VE, f2))
    assert rel_err_matrix(v2, ref2, abs_sum_scale(m2, orc.local_stiffness(m2.p, m2.t, m2.VE, None), f2)) < 1e-12
    h2.close()
