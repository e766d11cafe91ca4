"""CPU: pin the oracle (oracle/femeko_oracle.c) against the reference.

* against the committed golden fixtures = outputs of the compiled, unmodified reference C++
  (tests/golden/make_golden.py) -- runs everywhere;
* against the compiled reference itself (oracle/_ref) on fresh meshes -- runs where it was built;
* against scipy's FITPACK wrapper for the Dierckx spline pieces (Dierckx.jl wraps the same Fortran);
* analytic known answers: patch test, zero row sums, mass = volume.
"""
import numpy as np
import pytest

import femeko_jl_b200 as fk
from conftest import abs_sum_scale, load_golden, rel_err_matrix, rel_l2
from oracle import oracle as orc

GOLD = ["kuhn3", "kuhn6", "delaunay300"]


@pytest.mark.parametrize("name", GOLD)
def test_oracle_matches_golden_reference_outputs(name):
    m, g = load_golden(name)
    rp, ci = orc.pattern(m.t, m.nv)
    assert np.array_equal(rp, g["csc_outer"]) and np.array_equal(ci, g["csc_inner"])  # pattern bit-exact
    # local matrices (FEMc.cpp:76-99 uses a QR solve, the oracle an LU solve: agreement to rounding)
    Ak = orc.local_stiffness(m.p, m.t, m.VE, None)
    assert np.max(np.abs(Ak - g["Ak"])) <= 1e-12 * np.max(np.abs(g["Ak"]))
    # abcd samples
    for row, k in zip(g["abcd"], g["abcd_elems"]):
        for i in range(4):
            r = orc.abcd(m.p, m.t[:, k], m.t[i, k])
            assert np.allclose(r, row[i], rtol=1e-11, atol=1e-11 * np.abs(row[i]).max())
    e2n = orc.elem2nnz(m.t, rp, ci)
    sc_mu, sc_1 = abs_sum_scale(m, g["Ak"], g["mu"]), abs_sum_scale(m, g["Ak"])
    v = orc.assemble_eigen(e2n, len(ci), Ak, g["mu"])
    assert rel_err_matrix(v, g["csc_values_mu"], sc_mu) < 1e-12
    v = orc.assemble_julia(e2n, len(ci), orc.local_stiffness(m.p, m.t, m.VE, g["mu"]))
    assert rel_err_matrix(v, g["csc_values_mu"], sc_mu) < 1e-12      # Julia order == C++ order to 1e-12
    v1 = orc.assemble_eigen(e2n, len(ci), Ak, None)
    assert rel_err_matrix(v1, g["csc_values_nomu"], sc_1) < 1e-12
    assert np.array_equal(orc.lagrange(m), g["lagrange"])            # same loop, same order: bit-exact
    assert np.array_equal(orc.boundary_integral(m, g["F"], 0, "cpp"), g["boundary_integral"])
    assert rel_l2(orc.boundary_integral(m, g["F"], [0], "julia"), g["boundary_integral"]) < 1e-14
    x, it, err, info = orc.pcg(rp, ci, g["csc_values_mu"], -g["boundary_integral"], 1e-10)
    assert info == 0 and abs(it - int(g["cg_iters"])) <= 1
    assert rel_l2(x, g["cg_u"]) < 1e-8


@pytest.mark.skipif(not orc.have_ref(), reason="compiled reference (oracle/_ref) not present")
@pytest.mark.parametrize("mesh", ["kuhn5", "delaunay"])
def test_oracle_matches_compiled_reference(mesh):
    m = fk.kuhn_cube(5) if mesh == "kuhn5" else fk.delaunay_ball(800, seed=3)
    mu = np.random.default_rng(1).uniform(1.0, 5.0, m.nt) * fk.mu0
    outer, inner, vals = orc.ref_stiffness(m, mu)
    rp, ci, v = orc.stiffness(m, mu, order="eigen")
    assert np.array_equal(outer, rp) and np.array_equal(inner, ci)
    sc = abs_sum_scale(m, orc.ref_local_stiffness(m), mu)
    assert rel_err_matrix(v, vals, sc) < 1e-12
    _, _, vj = orc.stiffness(m, mu, order="julia")
    assert rel_err_matrix(vj, vals, sc) < 1e-12
    b = -orc.ref_boundary_integral(m, [0.3, -1.0, 0.5], 0)
    x1, it1, e1, info, _ = orc.ref_cg(outer, inner, vals, b, 1e-9, 0)
    x2, it2, e2, _ = orc.pcg(rp, ci, vals, b, 1e-9)
    assert info == 0 and abs(it1 - it2) <= 1 and rel_l2(x2, x1) < 1e-8


def test_patch_test_and_row_sums():
    """Linear u = g.x: recovered H = -g exactly; A u vanishes at interior nodes; A 1 = 0; sum(M) = volume."""
    m = fk.delaunay_ball(600, seed=5)
    g = np.array([0.7, -1.3, 0.4])
    u = g @ m.p
    Hf, H = orc.recover_field(m, u)
    assert np.allclose(Hf, -g[:, None], atol=1e-10)
    rp, ci, v = orc.stiffness(m, None)
    interior = np.ones(m.nv, bool)
    interior[np.unique(m.surfaceT[:3])] = False
    Au = orc.spmv(rp, ci, v, u)
    assert np.max(np.abs(Au[interior])) < 1e-9 * np.max(np.abs(v)) * np.max(np.abs(u))
    assert np.max(np.abs(orc.spmv(rp, ci, v, np.ones(m.nv)))) < 1e-10 * np.max(np.abs(v))
    _, _, mv = orc.mass(m)
    assert abs(mv.sum() - m.VE.sum()) < 1e-10 * m.VE.sum()
    Hb, _ = orc.recover_field(m, u, compat_bug=True)            # NewtonRaphson.jl:187-189
    assert np.allclose(Hb[0], -g[0]) and np.allclose(Hb[1], -(g[1] + g[2])) and np.all(Hb[2] == 0)


def test_tangential_oracle_definition():
    m = fk.kuhn_cube(3)
    rng = np.random.default_rng(2)
    Hv = np.asfortranarray(rng.normal(size=(3, m.nt)))
    dmu = rng.random(m.nt)
    Ak = orc.local_tangential(m.p, m.t, m.VE, Hv, dmu)
    k = 17
    G = np.array([orc.abcd(m.p, m.t[:, k], m.t[i, k])[1:] for i in range(4)])
    gh = G @ Hv[:, k]
    ref = m.VE[k] * dmu[k] / np.linalg.norm(Hv[:, k]) * np.outer(gh, gh)
    assert np.allclose(Ak[:, k].reshape(4, 4), ref, rtol=1e-12)


def test_splev_matches_fitpack():
    from scipy import interpolate

    x = np.array([0.0, 20, 22.4, 30, 55, 80, 120, 200, 400, 900, 2000, 5000.0])
    y = 1.0 / (1.0 + x / 300.0) + 0.2
    tck = interpolate.splrep(x, y, k=3, s=0)
    xs = np.linspace(-50, 6000, 777)
    inside = (xs >= x[0]) & (xs <= x[-1])
    assert np.allclose(orc.splev(tck[0], tck[1], 3, xs[inside], "error"), interpolate.splev(xs[inside], tck, ext=2), rtol=1e-13)
    assert np.allclose(orc.splev(tck[0], tck[1], 3, xs, "nearest"), interpolate.splev(xs, tck, ext=3), rtol=1e-13)
    assert np.allclose(orc.splev(tck[0], tck[1], 3, xs, "extrapolate"), interpolate.splev(xs, tck, ext=0), rtol=1e-12)
    assert np.allclose(orc.splev(tck[0], tck[1], 3, xs, "zero"), interpolate.splev(xs, tck, ext=1), rtol=1e-13)
    with pytest.raises(ValueError):
        orc.splev(tck[0], tck[1], 3, xs, "error")


def test_material_permeability_rules():
    """magneticProperties.jl:157-188: mu = B/H with non-finite -> max; dmu central differences."""
    H = np.array([0.0, 10.0, 20.0, 40.0, 80.0])
    B = np.array([0.0, 0.5, 0.8, 1.0, 1.1])
    mu, dmu, nu = orc.material_permeability(H, B)
    assert mu[0] == np.max(B[1:] / H[1:]) and np.allclose(mu[1:], B[1:] / H[1:])
    assert np.isclose(dmu[2], (mu[3] - mu[1]) / (H[3] - H[1])) and np.isclose(dmu[0], (mu[1] - mu[0]) / 10.0)
    assert nu[0] == 0.0 and np.allclose(nu[1:], fk.mu0 * H[1:] / B[1:])
    assert np.allclose(orc.gradient(H, B)[1:-1], (B[2:] - B[:-2]) / (H[2:] - H[:-2]))


@pytest.mark.parametrize("name", ["kuhn6", "delaunay300"])
def test_oracle_demag_field_matches_golden_reference(name):
    """oracle restatement of LL::magnetostaticField vs the compiled reference class (golden)."""
    m, g = load_golden(name)
    d = dict(np.load(__import__("os").path.join(__import__("conftest").GOLDEN, "demag_" + name + ".npz")))
    ins = np.asarray(g["InsideElements"], dtype=np.int32)
    Hd, u, it, rhs, vol = orc.demag_field(m, ins, d["M"])
    assert np.array_equal(vol, d["volumes"])                      # same loop order: bit-exact
    fin = np.isfinite(d["Hd"])
    assert np.array_equal(fin, np.isfinite(Hd))                   # NaN exactly where the node volume is 0
    assert np.array_equal(np.flatnonzero(fin[0]), d["InsideNodes"])
    assert rel_l2(Hd[fin], d["Hd"][fin]) < 1e-10
    assert abs(rhs.sum()) < 1e-12 * np.abs(rhs).sum()              # sum_i grad(lambda_i) = 0: the load vector has zero mean


def test_nonlinear_loop_restatements():
    """The restated Picard loop (manyBodyIntersecting.jl:141-210) gives the same converged state whichever linear solve
    runs inside it -- the bordered sparse direct solve the Julia driver uses, the C port of Eigen's Jacobi-CG, or the
    compiled reference (FEMc.cpp stiffnessMatrix + Eigen ConjugateGradient) -- and the restated Newton loop
    (NewtonRaphson.jl:148-214) converges to the same fixed point in fewer passes."""
    import os
    from scipy.interpolate import splrep
    import femeko_jl_b200 as fk
    from conftest import GOLDEN, rel_l2

    g = np.load(os.path.join(GOLDEN, "gd_270K.npz"))
    tck = splrep(g["HofM"], g["mu"], k=3, s=0)
    tckd = splrep(g["HofM"], g["dmu"], k=3, s=0)
    m = fk.kuhn_cube(8, inside_edge=0.5)
    a = orc.picard_loop(m, [0.5, 0, 0], (tck[0], tck[1]), solver="direct")
    b = orc.picard_loop(m, [0.5, 0, 0], (tck[0], tck[1]), solver="pcg")
    assert a["att"] == b["att"] and rel_l2(a["Hfield"], b["Hfield"]) < 1e-9 and rel_l2(a["mu"], b["mu"]) < 1e-9
    if orc.have_ref():
        c = orc.picard_loop(m, [0.5, 0, 0], (tck[0], tck[1]), solver="ref", cg_tol=1e-12)
        assert a["att"] == c["att"] and rel_l2(a["Hfield"], c["Hfield"]) < 1e-9
    assert a["div"] < 1e-4 and a["history"][0][1] > 0.1           # starts far from the fixed point, ends converged
    n = orc.newton_loop(m, [0.5, 0, 0], (tck[0], tck[1]), (tckd[0], tckd[1]), compat_bug=False)
    assert n["att"] < a["att"] and rel_l2(n["Hfield"], a["Hfield"]) < 2e-3
    # line search: a full step that lands on the solution is accepted, an overshoot backtracks by powers of 0.9
    rp, ci, v = orc.stiffness(m, None)
    A = orc._csr(rp, ci, v, m.nv)
    xs = np.random.default_rng(0).normal(size=m.nv)
    RHS = A @ xs
    u = np.zeros(m.nv)
    res, alf = orc.line_search(u, xs, A, RHS)
    assert alf == 1.0 and res < 1e-10 * np.linalg.norm(RHS)
    u = np.zeros(m.nv)
    res, alf = orc.line_search(u, 2.5 * xs, A, RHS)
    assert 0 < alf < 1.0 and np.isclose(np.log(alf) / np.log(0.9), round(np.log(alf) / np.log(0.9)))


def test_2d_p1_oracle_properties():
    """src/FEM2D.jl:48-76,139-161 restated: zero row sums and symmetry of the stiffness matrix, the patch test (linear
    fields are in the kernel away from the boundary), sum(M) = area, and the local matrix against the closed form."""
    import scipy.sparse as sp

    for m in (fk.tri_grid(7), fk.delaunay_disc(400, seed=3)):
        mu = np.random.default_rng(0).uniform(0.5, 2.0, m.nt)
        rp, ci, v = orc.stiffness2d(m, mu)
        A = sp.csr_matrix((v, ci, rp), shape=(m.nv, m.nv))
        assert abs(A.sum(axis=1)).max() < 1e-12 * abs(v).max() and abs(A - A.T).max() < 1e-13 * abs(v).max()
        _, _, vm = orc.mass2d(m)
        assert np.isclose(vm.sum(), m.VE.sum(), rtol=1e-13) and vm.min() > 0
        # one element against the closed-form gradients
        k = 5
        nds = m.t[:, k]
        x, y = m.p[0, nds], m.p[1, nds]
        det = (x[1] - x[0]) * (y[2] - y[0]) - (x[2] - x[0]) * (y[1] - y[0])
        b = np.array([y[1] - y[2], y[2] - y[0], y[0] - y[1]]) / det
        c = np.array([x[2] - x[1], x[0] - x[2], x[1] - x[0]]) / det
        for i in range(3):
            _, bi, ci_ = orc.abc(m.p, nds, int(nds[i]))
            assert np.isclose(bi, b[i], rtol=1e-12, atol=1e-14) and np.isclose(ci_, c[i], rtol=1e-12, atol=1e-14)
    g = fk.tri_grid(6)
    rp, ci, v = orc.stiffness2d(g)
    A = sp.csr_matrix((v, ci, rp), shape=(g.nv, g.nv))
    r = A @ (2 * g.p[0] + 3 * g.p[1])
    inner = (np.abs(g.p[0]) < 0.49) & (np.abs(g.p[1]) < 0.49)
    assert abs(r[inner]).max() < 1e-13


def test_p2_oracle_properties():
    """src/FEM3D.jl:21-37,324-395,494-561 restated: the fitted basis functions are the Lagrange P2 functions on
    straight-sided elements (so the kernel's closed forms are legitimate), the stiffness matrix kills constants, the
    mass matrix sums to V/6 (the reference multiplies VE by weights that sum to 1/6) and is exact for quadratics."""
    import scipy.sparse as sp

    m = fk.kuhn_cube_p2(2)
    S = orc.quadratic_basis_all(m)
    # S reproduces phi = l(2l - 1) / 4 l_a l_b at an interior point of element 3
    k = 3
    v = m.p[:, m.t[:4, k]]
    lam = np.array([0.1, 0.2, 0.3, 0.4])
    x, y, z = v @ lam
    mono = np.array([1, x, y, z, x * x, x * y, x * z, y * y, y * z, z * z])
    phi_fit = S[:, :, k].T @ mono
    edges = ((0, 1), (1, 2), (0, 2), (0, 3), (2, 3), (1, 3))
    phi_cf = np.array([lam[i] * (2 * lam[i] - 1) for i in range(4)] + [4 * lam[a] * lam[b] for a, b in edges])
    assert np.max(np.abs(phi_fit - phi_cf)) < 1e-12
    rp, ci, v = orc.quadratic_stiffness(m, S, np.linspace(1, 2, m.nt))
    A = sp.csr_matrix((v, ci, rp), shape=(m.nv, m.nv))
    assert abs(A @ np.ones(m.nv)).max() < 1e-11 * abs(v).max() and abs(A - A.T).max() < 1e-12 * abs(v).max()
    _, _, vm = orc.quadratic_mass(m, S)
    M = sp.csr_matrix((vm, ci, rp), shape=(m.nv, m.nv))
    assert np.isclose(vm.sum(), m.VE.sum() / 6.0, rtol=1e-12)
    q = m.p[0] ** 2 + m.p[1] * m.p[2]                    # in the P2 space: 6 q'Mq = int q^2 over the unit cube
    exact = 1 / 80 + 1 / 144                              # int x^4 + int y^2 z^2 over [-1/2, 1/2]^3 (cross term vanishes)
    assert np.isclose(6.0 * (q @ (M @ q)), exact, rtol=1e-10)


def test_nedelec_restatement_properties():
    """Edge elements (src/FEM3D.jl:153-271, nedelec3D.jl): no C++ twin and no Julia here, so the restatement is pinned by what the
    lowest-order Nedelec space must satisfy with the reference's conventions (edge order and orientation of
    NodesFromLocalEdge, curl N = 2 L grad l_i x grad l_j): gradients lie in the null space of the curl-curl matrix, the
    interpolant of A = B0 x r / 2 recovers the constant B0 exactly, the matrix is symmetric, the tangential matrix follows
    its definition entry by entry, and the load vector equals A_1 u for that interpolant (integration by parts on constants)."""
    import femeko_jl_b200 as fk
    from femeko_jl_b200 import mesh as M
    import scipy.sparse as sp

    m = fk.to_p2(fk.delaunay_ball(250, seed=3))
    M.edgeNumbering(m)
    E = M.edgeDofs(m)
    t = m.t[:4].astype(np.int64)
    rp, ci, v = orc.nedelec_stiffness(m)
    A = sp.csr_matrix((v, ci, rp), shape=(m.ne, m.ne))
    assert abs(A - A.T).max() == 0.0
    rng = np.random.default_rng(0)
    phi = rng.normal(size=m.nv)
    B0 = np.array([0.3, -1.2, 0.7])
    ug, ub = np.zeros(m.ne), np.zeros(m.ne)
    for ie, (a, b) in enumerate(orc.NED_EDGES):
        lo, hi = np.minimum(t[a], t[b]), np.maximum(t[a], t[b])
        d = m.p[:, hi] - m.p[:, lo]
        Ln = np.linalg.norm(d, axis=0)
        ug[E[ie]] = (phi[hi] - phi[lo]) / Ln
        Am = 0.5 * np.cross(B0, (0.5 * (m.p[:, hi] + m.p[:, lo])).T).T
        ub[E[ie]] = np.einsum("dk,dk->k", Am, d) / Ln
    scale = np.abs(A).max()
    assert np.abs(A @ ug).max() < 1e-12 * scale * np.abs(ug).max()
    Bf, B = orc.nedelec_recover_B(m, ub)
    assert np.abs(Bf - B0[:, None]).max() < 1e-12 and np.abs(B - np.linalg.norm(B0)).max() < 1e-12
    assert np.abs(orc.nedelec_recover_B(m, ug)[0]).max() < 1e-11 * np.abs(ug).max()
    # sum_k VE B0 . curlN_e = (A u_B0)_e
    rhs = orc.nedelec_load(m, B0)
    assert np.abs(rhs - A @ ub).max() < 1e-11 * np.abs(rhs).max()
    # local matrices: symmetric, 36 x nt, equal to VE c_i . c_j
    Ak = orc.nedelec_local_stiffness(m)
    c = orc.nedelec_curls(m)
    assert np.allclose(Ak[6 * 2 + 5], m.VE * np.einsum("kd,kd->k", c[:, 2], c[:, 5]), rtol=1e-14, atol=0)
    assert np.array_equal(Ak[6 * 2 + 5], Ak[6 * 5 + 2])
    # tangential matrix: entry by entry against its definition on a random state
    dnu = rng.uniform(0.1, 1.0, m.nt)
    Bfr = rng.normal(size=(3, m.nt))
    Bn = np.linalg.norm(Bfr, axis=0)
    rpt, cit, vt = orc.nedelec_tangential(m, dnu, Bfr, Bn)
    At = sp.csr_matrix((vt, cit, rpt), shape=(m.ne, m.ne)).toarray()
    ref = np.zeros((m.ne, m.ne))
    d = np.einsum("ked,dk->ke", c, Bfr)
    for k in range(m.nt):
        ref[np.ix_(E[:, k], E[:, k])] += m.VE[k] * dnu[k] / Bn[k] * np.outer(d[k], d[k])
    assert np.abs(At - ref).max() < 1e-12 * np.abs(ref).max()


def test_quadratic_convection_restatement_properties():
    """quadraticConvectionMatrix (src/FEM3D.jl:563-648): the literal restatement assembles the TRANSPOSE of the local matrices
    (column-major Ck[:] read row-major), so its COLUMN sums vanish (sum_j grad phi_j = 0); the untransposed form has zero row
    sums and integrates u . grad f exactly for a linear f (times the reference's 1/6)."""
    import femeko_jl_b200 as fk
    import scipy.sparse as sp

    m = fk.to_p2(fk.delaunay_ball(220, seed=5))
    rng = np.random.default_rng(2)
    u = rng.normal(size=(3, m.nv))
    F = rng.uniform(0.5, 2.0, m.nt)
    S = orc.quadratic_basis_all(m)
    rp, ci, vT = orc.quadratic_convection(m, u, S, F, compat_bug=True)
    _, _, v = orc.quadratic_convection(m, u, S, F, compat_bug=False)
    CT = sp.csr_matrix((vT, ci, rp), shape=(m.nv, m.nv))
    Cm = sp.csr_matrix((v, ci, rp), shape=(m.nv, m.nv))
    assert abs(CT - Cm.T).max() < 1e-15 * abs(Cm).max()
    assert abs(Cm - Cm.T).max() > 1e-3 * abs(Cm).max()                    # genuinely non-symmetric
    one = np.ones(m.nv)
    assert np.abs(Cm @ one).max() < 1e-12 * abs(Cm).max() and np.abs(CT.T @ one).max() < 1e-12 * abs(Cm).max()
    a = np.array([0.7, -0.4, 1.1])
    f = a @ m.p
    ua = u[:, m.t[:4]].mean(axis=1)                                        # 3 x nt
    # (the reference's weights sum to 1/6 AND multiply VE, as in quadraticMassMatrix: the integral carries that 1/6)
    assert abs(one @ (Cm @ f) - np.sum(F * m.VE * (a @ ua)) / 6.0) < 1e-11 * np.sum(np.abs(F * m.VE * (a @ ua)))


def test_nedelec_newton_loop_restatement_converges():
    """The restated loop of nedelec3D.jl:148-250 on a small cube with the real Gd nu(B) table: the first pass is the uniform
    medium (B = Bext on every element), the passes converge quadratically once the tangential matrix is in, nu stays 1 outside."""
    import femeko_jl_b200 as fk
    from femeko_jl_b200 import mesh as M
    from conftest import GOLDEN
    import os

    g = np.load(os.path.join(GOLDEN, "gd_270K.npz"))
    B, nu = g["B"], g["nu"]
    # FITPACK fit through scipy (no product code on this path)
    from scipy.interpolate import splrep

    tck = splrep(B, nu, k=3, s=0)
    tckd = splrep(B, orc.gradient(B, nu), k=3, s=0)
    m = M.edgeNumbering(fk.kuhn_cube_p2(3, inside_edge=0.5))
    r = orc.nedelec_newton_loop(m, [0.0, 0.0, 1.0], (tck[0], tck[1]), (tckd[0], tckd[1]), tol=1e-6, max_att=30)
    assert r["att"] < 12 and r["div"] < 1e-6
    assert abs(r["history"][0][1] - 1.0) < 1e-6                        # |B - 0| after the first pass = |Bext| (CG at sqrt(eps))
    divs = [h[1] for h in r["history"]]
    assert divs[-1] < 1e-2 * divs[-2]                                   # Newton's fast tail once the tangential matrix is in
    out = np.setdiff1d(np.arange(m.nt), m.InsideElements)
    assert np.all(r["nu"][out] == 1.0) and np.all(r["nu"][m.InsideElements] < 1.0)
    # the converged state satisfies the nonlinear equation: A(nu(B(u))) u = RHS
    import scipy.sparse as sp

    rp, ci, v = orc.nedelec_stiffness(m, r["nu"])
    A = sp.csr_matrix((v, ci, rp), shape=(m.ne, m.ne))
    rhs = orc.nedelec_load(m, [0.0, 0.0, 1.0])
    assert np.linalg.norm(rhs - A @ r["u"]) < 1e-5 * np.linalg.norm(rhs)
    # (u itself is fixed up to a gradient only: the reference solves the singular system without a gauge; B = curl u is what counts)
    Bf, Bn = orc.nedelec_recover_B(m, r["u"])
    assert np.array_equal(Bf, r["Bfield"]) and np.all(Bn[m.InsideElements] > 1.0)     # the magnet concentrates the flux
