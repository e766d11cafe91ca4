"""GPU parity tests proper: everything goes through the C ABI of libfemeko_cuda.so and is compared
with (a) the golden outputs of the compiled reference, (b) the CPU oracle on the same seeded inputs,
(c) size-independent properties at the benchmark's full size.  Tolerances follow the north star:
pattern / element->nnz map bit-exact, matrix values 1e-12 (metric in conftest.rel_err_matrix),
potential and field 1e-8 relative L2 at equal CG tolerance."""
import numpy as np
import pytest

import femeko_jl_b200 as fk
from conftest import abs_sum_scale, load_golden, rel_err_matrix, rel_l2
from femeko_jl_b200 import lib as L
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
GOLD = ["kuhn3", "kuhn6", "delaunay300"]


@pytest.mark.parametrize("name", GOLD)
def test_stiffness_against_reference_golden(name):
    m, g = load_golden(name)
    h = fk.Femeko(m)
    rp, ci = h.pattern()
    assert np.array_equal(rp, g["csc_outer"]) and np.array_equal(ci, g["csc_inner"])      # bit-exact
    assert np.array_equal(h.elem2nnz(), orc.elem2nnz(m.t, rp, ci))                        # bit-exact
    v = h.assemble_stiffness(g["mu"])
    assert rel_err_matrix(v, g["csc_values_mu"], abs_sum_scale(m, g["Ak"], g["mu"])) < 1e-12
    v1 = h.assemble_stiffness(None)
    assert rel_err_matrix(v1, g["csc_values_nomu"], abs_sum_scale(m, g["Ak"])) < 1e-12
    Ak = h.local_stiffness(None)
    assert np.max(np.abs(Ak - g["Ak"])) <= 1e-12 * np.max(np.abs(g["Ak"]))
    assert np.array_equal(h.lagrange(), g["lagrange"])                                     # same order: bit-exact
    rhs = h.boundary_integral(m.surfaceT, m.normal, None, g["F"], 0)                       # C++ form (area recomputed)
    assert rel_l2(rhs, g["boundary_integral"]) < 1e-14
    rhs2 = h.boundary_integral(m.surfaceT, m.normal, m.AE, g["F"], [0])                    # Julia form (stored AE)
    assert rel_l2(rhs2, g["boundary_integral"]) < 1e-13
    # Jacobi-PCG at the tolerance the fixture was solved with
    x, it, rr = h.pcg(-g["boundary_integral"], slot=0, tol=1e-10)   # slot 0 holds the f=None matrix now
    h.assemble_stiffness(g["mu"], want_vals=False)
    x, it, rr = h.pcg(-g["boundary_integral"], slot=0, tol=1e-10)
    assert abs(it - int(g["cg_iters"])) <= 1 and rr < 1e-10
    assert rel_l2(x, g["cg_u"]) < 1e-8
    h.close()


@pytest.mark.parametrize("mesh,opts", [("kuhn12", {}), ("kuhn12", {"natural_order": 1}),
                                       ("kuhn12", {"tile_max_elems": 256, "tile_max_rows": 32, "sell_sigma": 64}),
                                       ("delaunay", {}), ("delaunay", {"tile_max_elems": 1024, "tile_max_rows": 4096})])
def test_assembly_family_against_oracle(mesh, opts):
    m = fk.kuhn_cube(12) if mesh == "kuhn12" else fk.delaunay_ball(3000, seed=11)
    rng = np.random.default_rng(5)
    f = rng.uniform(0.5, 4.0, m.nt)
    for use_ve in (True, False):
        h = fk.Femeko(m, use_mesh_VE=use_ve, **opts)
        rp, ci = h.pattern()
        orp, oci = orc.pattern(m.t, m.nv)
        assert np.array_equal(rp, orp) and np.array_equal(ci, oci)
        Ak1 = orc.local_stiffness(m.p, m.t, m.VE, None)
        ref = orc.assemble_julia(h.elem2nnz(), h.nnz, orc.local_stiffness(m.p, m.t, m.VE, f))
        v = h.assemble_stiffness(f)
        assert rel_err_matrix(v, ref, abs_sum_scale(m, Ak1, f)) < 1e-12
        assert np.array_equal(v, h.assemble_stiffness(f))                 # bitwise run-to-run determinism
        assert np.allclose(h.volumes(), m.VE, rtol=1e-14, atol=0)
        # mass (FEM3D.jl:466-492)
        _, _, mref = orc.mass(m)
        mv = h.assemble_mass()
        assert rel_err_matrix(mv, mref) < 1e-13
        # tangential (FEM3D.jl:79-119)
        Hv = np.asfortranarray(rng.normal(size=(3, m.nt)))
        dmu = rng.uniform(-1, 1, m.nt)
        _, _, tref = orc.tangential(m, Hv, dmu)
        tv = h.assemble_tangential(Hv, dmu)
        tscale = abs_sum_scale(m, np.abs(orc.local_tangential(m.p, m.t, m.VE, Hv, np.abs(dmu))))
        assert rel_err_matrix(tv, tref, tscale) < 1e-11
        # slot algebra: M + dt*A (heat3D.jl:85)
        h.assemble_stiffness(f, slot=0, want_vals=False)
        h.matrix_axpby(3, 1.0, 1, 1e-3, 0)
        assert np.allclose(h.matrix_get(3), mv + 1e-3 * v, rtol=1e-15, atol=0)
        # SpMV
        x = rng.normal(size=m.nv)
        y = h.spmv(x, slot=0)
        yref = orc.spmv(rp, ci, v, x)
        assert rel_l2(y, yref) < 1e-14
        h.close()


def test_pcg_against_oracle_and_eigen_semantics():
    m = fk.kuhn_cube(16)
    mu = np.full(m.nt, fk.mu0)
    mu[m.InsideElements] = 3 * fk.mu0
    h = fk.Femeko(m)
    rp, ci = h.pattern()
    v = h.assemble_stiffness(mu)
    b = -h.boundary_integral(m.surfaceT, m.normal, m.AE, [1.0, 0, 0], 0)
    xo, ito, erro, info = orc.pcg(rp, ci, v, b, 1e-10)
    x, it, rr = h.pcg(b, tol=1e-10)
    assert info == 0 and abs(it - ito) <= 1 and rr < 1e-10 and rel_l2(x, xo) < 1e-8
    assert np.array_equal(x, h.pcg(b, tol=1e-10)[0])                      # deterministic
    # small polling chunks stop at the same iteration
    x2, it2, _ = h.pcg(b, tol=1e-10, check_every=2)
    assert it2 == it and np.array_equal(x2, x)
    # fixed iteration count (bench mode): exactly max_iter iterations
    x3, it3, rr3 = h.pcg(b, max_iter=25, tol_disable=True)
    xo3, ito3, _, _ = orc.pcg(rp, ci, v, b, 0.0, 25)
    assert it3 == 25 and rel_l2(x3, xo3) < 1e-10
    # max_iter reached reports max_iter like Eigen (iters = i after the loop)
    _, it4, rr4 = h.pcg(b, tol=1e-14, max_iter=7)
    assert it4 == 7 and rr4 > 1e-14
    # warm start from the solution converges immediately; zero rhs returns zeros
    _, it5, _ = h.pcg(b, x0=x, tol=1e-9)
    assert it5 <= 1
    xz, itz, rz = h.pcg(np.zeros(m.nv))
    assert itz == 0 and rz == 0.0 and not xz.any()
    # unpreconditioned variant = IterativeSolvers.cg call sites (run.jl:141)
    A = fk.SparseCSR(h, 0, v)
    xu, itu, rru = fk.cg(A, b, Pl=None, log=True)
    assert rru < 1.5e-8 and rel_l2(xu - xu.mean(), x - x.mean()) < 1e-6
    h.close()


@pytest.mark.parametrize("name", ["kuhn6", "delaunay300"])
def test_scalar_potential_dropin(name):
    """ABI-identical legacy entry (magnetostatics.cpp:49-60) against the reference's CG solution;
    compared up to the constant null-space mode of the pure Neumann problem."""
    import ctypes as C

    m, g = load_golden(name)
    u = np.zeros(m.nv)
    p = np.asfortranarray(m.p)
    t = np.asfortranarray(m.t)
    sT = np.asfortranarray(m.surfaceT)
    nrm = np.asfortranarray(m.normal)
    L.load().scalarPotential(L.dptr(u), L.dptr(p), L.iptr(t), L.iptr(sT), L.dptr(nrm), m.nv, m.nt, m.ns, L.dptr(m.VE),
                             L.dptr(g["mu"]), L.dptr(g["F"]), 0)
    ref = g["cg_u"]
    # The golden potential is Eigen's CG stopped at a relative residual of 1e-10 (cg_err in the fixture).  The drop-in
    # runs to Eigen's default tolerance like magnetostatics.cpp:26-38: machine epsilon, which the singular Neumann
    # system never reaches within 2n iterations -- the reference throws there (:36-38), the drop-in warns and returns
    # the last iterate.  Past convergence the recursion wanders (true residual between 1e-16 and 1e-7 depending on
    # the rounding of the dot products, scripts/dbg_dropin.py shows it for every version of the kernels), so this
    # entry point is pinned at 1e-6 on both the solution and its true residual against the reference's own matrix;
    # the solver itself is pinned at equal tolerance in test_pcg_against_oracle_and_eigen_semantics and below.
    import scipy.sparse as sp
    A = sp.csc_matrix((g["csc_values_mu"], g["csc_inner"], g["csc_outer"]), shape=(m.nv, m.nv))
    b = -g["boundary_integral"]
    assert np.linalg.norm(b - A @ u) / np.linalg.norm(b) < 1e-6
    assert rel_l2(u - u.mean(), ref - ref.mean()) < 1e-6
    # round 2: the drop-in projects the constant mode out of the residual (fk_pcg_opts.project_nullspace = 2), so the
    # recursion no longer wanders: against a direct solve of the reference's own matrix (one node grounded, means
    # removed) the potential holds the north star's 1e-8, and the true residual of the consistent system is tiny
    import scipy.sparse.linalg as spla
    bc = b - b.mean()
    xd = np.zeros(m.nv)
    xd[1:] = spla.spsolve(A.tocsc()[1:, 1:], bc[1:])
    print(f"scalarPotential {name}: rel-L2 vs direct solve {rel_l2(u - u.mean(), xd - xd.mean()):.2e}, "
          f"true residual {np.linalg.norm(bc - A @ u) / np.linalg.norm(bc):.2e}")
    assert rel_l2(u - u.mean(), xd - xd.mean()) < 1e-8
    assert np.linalg.norm(bc - A @ u) / np.linalg.norm(bc) < 1e-9
    # same matrix, same right-hand side, the golden's tolerance: iteration count and potential of Eigen itself
    h = fk.Femeko(m)
    h.assemble_stiffness(g["mu"], want_vals=False)
    x, it, rr = h.pcg(b, tol=1e-10)
    assert abs(it - int(g["cg_iters"])) <= 1 and rel_l2(x - x.mean(), ref - ref.mean()) < 1e-8
    h.close()


def test_field_recovery_and_patch_test():
    m = fk.delaunay_ball(2500, seed=8)
    h = fk.Femeko(m)
    rng = np.random.default_rng(3)
    u = rng.normal(size=m.nv)
    mu = rng.uniform(1, 3, m.nt) * fk.mu0
    for compat in (False, True):
        Hf, Hn, B = h.recover_field(u, mu, compat_bug=compat)
        Hfo, Hno = orc.recover_field(m, u, compat_bug=compat)
        assert rel_l2(Hf, Hfo) < 1e-10 and rel_l2(Hn, Hno) < 1e-10 and rel_l2(B, mu * Hno) < 1e-10
    g = np.array([0.7, -1.3, 0.4])
    Hf, Hn, _ = h.recover_field(g @ m.p)
    assert np.allclose(Hf, -g[:, None], atol=1e-10) and np.allclose(Hn, np.linalg.norm(g), rtol=1e-10)
    h.close()


def _picard_reference_step(m, u, mu, spl_tck, inside, compat=False):
    Hf, Hn = orc.recover_field(m, u, compat_bug=compat)
    mu = mu.copy()
    mu[inside] = orc.splev(spl_tck[0], spl_tck[1], 3, Hn[inside], "nearest")
    return Hf, Hn, mu


@pytest.mark.parametrize("mesh", ["kuhn10", "delaunay"])
def test_fused_picard_step_against_oracle(mesh):
    m = fk.kuhn_cube(10, inside_edge=0.4) if mesh == "kuhn10" else fk.delaunay_ball(2500, seed=4)
    data = fk.synthetic_gd_like()
    spl = fk.Spline1D(data.HofM, data.mu, bc="nearest")
    dspl = fk.Spline1D(data.HofM, data.dmu, bc="nearest")
    h = fk.Femeko(m)
    h.set_spline(0, spl)
    h.set_spline(L.FK_MAX_SPLINES // 2, dspl)
    h.set_inside(m.InsideElements)
    rng = np.random.default_rng(9)
    u = 2e5 * (m.p[0] + 0.1 * rng.normal(size=m.nv))
    mu0v = np.full(m.nt, fk.mu0)
    h.set_mu(mu0v)
    div = h.picard_step(u)
    Hf, Hn, mu_ref = _picard_reference_step(m, u, mu0v, (spl.t, spl.c), m.InsideElements)
    mu_dev, dmu_dev = h.get_mu(want_dmu=True)
    assert rel_l2(mu_dev, mu_ref) < 1e-12
    assert rel_l2(dmu_dev[m.InsideElements], orc.splev(dspl.t, dspl.c, 3, Hn[m.InsideElements], "nearest")) < 1e-10
    Hfd, Hnd = h.get_field()
    assert rel_l2(Hfd, Hf) < 1e-10 and rel_l2(Hnd, Hn) < 1e-10
    assert np.isclose(div, fk.mu0 * np.max(np.abs(Hn[m.InsideElements] - 0.0)), rtol=1e-10)
    v = h.matrix_get(0)
    ref = orc.assemble_julia(h.elem2nnz(), h.nnz, orc.local_stiffness(m.p, m.t, m.VE, mu_ref))
    assert rel_err_matrix(v, ref, abs_sum_scale(m, orc.local_stiffness(m.p, m.t, m.VE, None), mu_ref)) < 1e-12
    # second sweep: div is measured against the stored |H|
    u2 = 1.05 * u
    div2 = h.picard_step(u2)
    _, Hn2 = orc.recover_field(m, u2)
    assert np.isclose(div2, fk.mu0 * np.max(np.abs(Hn2[m.InsideElements] - Hn[m.InsideElements])), rtol=1e-9)
    # one-call nonlinear iteration (fk_picard_pcg) = picard_step + pcg, bit for bit (b rides up on a second stream)
    b = rng.normal(size=m.nv)
    b -= b.mean()
    h.set_mu(mu0v)
    h.reset_field()
    d_a = h.picard_step(u)
    x_a, it_a, rr_a = h.pcg(b, tol=1e-9)
    h.set_mu(mu0v)
    h.reset_field()
    x_b, d_b, it_b, rr_b = h.picard_pcg(u, b, tol=1e-9)
    assert d_a == d_b and it_a == it_b and rr_a == rr_b and np.array_equal(x_a, x_b)
    x_c, _, it_c, _ = h.picard_pcg(u, b, x0=x_b, tol=1e-8)
    assert it_c <= 1 and rel_l2(x_c, x_b) < 1e-7
    # bc = "error" semantics (NewtonRaphson.jl:58): |H| beyond the table raises
    h.set_spline(0, fk.Spline1D(data.HofM, data.mu, bc="error"))
    with pytest.raises(L.FemekoError, match="outside the spline"):
        h.picard_step(1e4 * u)
    h.close()


def test_heat_step_against_direct_solve():
    """heat3D.jl:83-86: T+ = (M + dt A)^-1 (M T) with PCG; checked against a sparse direct solve."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla

    m = fk.kuhn_cube(8, inside_edge=0.5)
    h = fk.Femeko(m)
    k = np.ones(m.nt)
    k[m.InsideElements] = 2.0
    rp, ci = h.pattern()
    A = sp.csr_matrix((h.assemble_stiffness(k, slot=0), ci, rp), shape=(m.nv, m.nv))
    M = sp.csr_matrix((h.assemble_mass(slot=1), ci, rp), shape=(m.nv, m.nv))
    dt = 1e-3
    h.matrix_axpby(3, 1.0, 1, dt, 0)
    T = np.zeros(m.nv)
    T[m.InsideNodes] = 10.0
    Tref = T.copy()
    lu = spla.splu((M + dt * A).tocsc())
    for _ in range(5):
        rhs = h.spmv(T, slot=1)
        T, it, rr = h.pcg(rhs, slot=3, x0=T, tol=1e-12)
        Tref = lu.solve(M @ Tref)
    assert rel_l2(T, Tref) < 1e-9
    assert abs((M @ T).sum() - (M @ Tref).sum()) < 1e-9 * abs((M @ Tref).sum())   # heat is conserved (natural BC)
    # the one-call form of the step (fk_heat_step) takes the same steps
    T2 = np.zeros(m.nv)
    T2[m.InsideNodes] = 10.0
    for _ in range(5):
        T2, it2, _ = h.heat_step(T2, slot_mass=1, slot_lhs=3, tol=1e-12)
    assert np.array_equal(T2, T)
    h.close()


def test_full_size_properties():
    """BASELINE config 2 size (Kuhn N=119: 10.1 M tets, 1.73 M nodes): size-independent properties."""
    m = fk.kuhn_cube(119)
    h = fk.Femeko(m, use_mesh_VE=False)
    assert h.nnz == 25575838 and h.nt == 10110954
    mu = np.full(m.nt, fk.mu0)
    mu[m.InsideElements] = 3 * fk.mu0
    v = h.assemble_stiffness(mu)
    assert np.array_equal(v, h.assemble_stiffness(mu))                         # determinism at scale
    rp, ci = h.pattern()
    rows = np.repeat(np.arange(m.nv, dtype=np.int64), np.diff(rp))
    rs = np.bincount(rows, weights=v, minlength=m.nv)
    diag = v[ci == rows]
    assert np.max(np.abs(rs)) < 1e-12 * diag.max()                             # constants in the null space
    e2n = h.elem2nnz().reshape(m.nt, 4, 4)
    assert np.array_equal(v[e2n[:, 0, 1]], v[e2n[:, 1, 0]]) and np.array_equal(v[e2n[:, 2, 3]], v[e2n[:, 3, 2]])  # symmetric
    mv = h.assemble_mass()
    assert abs(mv.sum() - 1.0) < 1e-10                                          # sum(M) = volume of the unit cube
    g = np.array([0.3, -0.8, 0.5])
    u = g @ m.p
    Hf, Hn, _ = h.recover_field(u)
    assert np.allclose(Hf, -g[:, None], atol=1e-9)                             # patch test: exact gradient
    Au = h.spmv(u, slot=0)
    interior = np.ones(m.nv, bool)
    interior[np.unique(m.surfaceT[:3])] = False
    inside_nodes = np.zeros(m.nv, bool)
    inside_nodes[np.unique(m.t[:, np.concatenate([m.InsideElements, _neighbours(m)])])] = True
    far = interior & ~inside_nodes
    assert np.max(np.abs(Au[far])) < 1e-9 * diag.max() * np.abs(u).max()        # linear fields are discretely harmonic
    # packed column stream: every slice of the lattice numbering is 16-bit; 32-bit columns give the same bits
    nsell, nsl, n16, cb = h.sell_stats()
    assert n16 == nsl and 2 * nsell <= cb <= 2 * nsell + 2 * 96 * nsl
    b = np.random.default_rng(5).standard_normal(m.nv)
    b -= b.mean()
    xs, its, _ = h.pcg(b, max_iter=20, tol_disable=True)
    h.close()
    hw = fk.Femeko(m, use_mesh_VE=False, sell_wide_cols=1)
    assert hw.sell_stats()[2] == 0
    hw.assemble_stiffness(mu, want_vals=False)
    assert np.array_equal(hw.spmv(u, slot=0), Au)
    assert np.array_equal(hw.pcg(b, max_iter=20, tol_disable=True)[0], xs)
    hw.close()


def _neighbours(m):
    """elements sharing a node with the material region (where mu jumps, A u != 0 for linear u)"""
    mark = np.zeros(m.nv, bool)
    mark[m.InsideNodes] = True
    return np.flatnonzero(mark[m.t].any(axis=0)).astype(np.int32)


@pytest.mark.parametrize("name", ["kuhn6", "delaunay300"])
def test_demag_field_against_reference_golden(name):
    """fk_demag_setup / fk_demag_field vs the compiled reference class LL (golden) and the oracle."""
    import os
    from conftest import GOLDEN
    m, g = load_golden(name)
    d = dict(np.load(os.path.join(GOLDEN, "demag_" + name + ".npz")))
    ins = np.asarray(g["InsideElements"], dtype=np.int32)
    h = fk.Femeko(m)
    vol = h.demag_setup(ins, slot=0)
    assert np.array_equal(vol, d["volumes"])                                          # gathered in list order: bit-exact
    Hd, it, u = h.demag_field(d["M"], want_u=True)                                    # Eigen defaults, like CG.solve(RHS)
    fin = np.isfinite(d["Hd"])
    assert np.array_equal(fin, np.isfinite(Hd))
    assert rel_l2(Hd[fin], d["Hd"][fin]) < 1e-8
    Ho, uo, ito, rhs, _ = orc.demag_field(m, ins, d["M"])
    assert abs(it - ito) <= 1 and rel_l2(u - u.mean(), uo - uo.mean()) < 1e-8
    Hz, _ = h.demag_field(d["M"], nan_outside=False)
    assert np.all(Hz[~fin] == 0.0) and np.array_equal(Hz[fin], Hd[fin])
    # warm start from the converged potential: nothing left to do
    _, it2 = h.demag_field(d["M"], warm_start=True, tol=1e-10)
    assert it2 <= 1
    h.close()


def test_demag_field_oracle_and_properties_at_size():
    """Load vector + box method bit-for-bit against the oracle on a seeded mesh (same summation order), linearity in M,
    and the repeated-solve loop through device pointers."""
    import torch
    m = fk.kuhn_cube(24, inside_edge=0.5)
    ins = m.InsideElements.astype(np.int32)
    insn = np.unique(m.t[:, ins])
    rng = np.random.default_rng(9)
    M = np.zeros((3, m.nv), order="F")
    M[:, insn] = rng.normal(size=(3, len(insn)))
    h = fk.Femeko(m)
    vol = h.demag_setup(ins)
    assert np.array_equal(vol, orc.node_volumes(m, ins))
    Hd, it, u = h.demag_field(M, want_u=True, tol=1e-12)
    rhs = orc.demag_rhs(m, ins, M)
    rp, ci = h.pattern()
    A = h.matrix_get(0)
    import scipy.sparse as sp
    r = rhs - sp.csr_matrix((A, ci, rp), shape=(m.nv, m.nv)) @ u
    assert np.linalg.norm(r) <= 2e-12 * np.linalg.norm(rhs)                            # u solves A u = RHS(oracle)
    Hb = orc.demag_box(m, ins, u, vol)
    fin = np.isfinite(Hb)
    assert np.array_equal(fin, np.isfinite(Hd)) and rel_l2(Hd[fin], Hb[fin]) < 1e-13  # box method on the same u
    H2, _ = h.demag_field(2.5 * M, tol=1e-12)
    assert rel_l2(H2[fin], 2.5 * Hd[fin]) < 1e-9                                       # linear in M
    # device-pointer form, repeated solves with warm start (the LL time loop's call pattern)
    Md = torch.from_numpy(np.ascontiguousarray(M.T)).cuda()                            # 3 x nv column-major == nv x 3 row-major
    Hdd = torch.empty_like(Md)
    it0, _ = h.demag_field_dev(Md.data_ptr(), Hdd.data_ptr(), tol=1e-12)
    it1, _ = h.demag_field_dev(Md.data_ptr(), Hdd.data_ptr(), tol=1e-12, warm_start=True)
    torch.cuda.synchronize()
    assert it0 == it and it1 <= 1
    assert np.array_equal(np.isfinite(Hdd.cpu().numpy().T), fin) and rel_l2(Hdd.cpu().numpy().T[fin], Hd[fin]) < 1e-9
    h.close()
