"""CPU: the C-ABI library loads and exports every symbol include/femeko_cuda.h declares; the
symbolic phase (pattern, element->nnz map, assembly gather plan, SpMV row blocks) is validated
against the oracle by executing the plan in numpy; host spline fit/eval against scipy FITPACK.
No compute calls that need a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import femeko_jl_b200 as fk
from conftest import ROOT, load_golden, rel_err_matrix, abs_sum_scale
from femeko_jl_b200 import lib as L
from oracle import oracle as orc
from plan_emulator import SYM, export_symbolic, run_plan, run_rescale_plan, sell_spmv


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "femeko_cuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(fk_[a-z0-9_]+|scalarPotential)\s*\(", hdr))
    assert len(declared) > 40
    lib = C.CDLL(L.LIB_PATH)
    missing = [n for n in sorted(declared) if not hasattr(lib, n)]
    assert not missing, f"declared in the header but not exported: {missing}"
    # and the Python binding covers the same set
    assert declared == set(L._SIGS), (declared ^ set(L._SIGS))
    assert L.load().fk_version() >= 100


def test_create_without_gpu_fails_loudly():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    m = fk.kuhn_cube(2)
    with pytest.raises(L.FemekoError, match="no CUDA device"):
        fk.Femeko(m)


@pytest.mark.parametrize("name", ["kuhn3", "kuhn6", "delaunay300"])
@pytest.mark.parametrize("opts", [{}, {"natural_order": 1}, {"tile_max_elems": 128, "tile_max_rows": 16}])
def test_symbolic_pattern_map_and_plan_against_reference(name, opts):
    m, g = load_golden(name)
    plan = export_symbolic(m, **opts)
    # CSR pattern bit-exact with the reference's Eigen pattern (CSC == CSR by symmetry)
    assert np.array_equal(plan["rowptr"], g["csc_outer"]) and np.array_equal(plan["colind"], g["csc_inner"])
    rp, ci = plan["rowptr"], plan["colind"]
    e2n = plan["e2nnz"]
    assert np.array_equal(e2n, orc.elem2nnz(m.t, rp, ci))
    # the map really addresses (t[i,k], t[j,k])
    rows = np.repeat(np.arange(m.nv), np.diff(rp))
    for i in range(4):
        for j in range(4):
            idx = e2n[4 * i + j::16]
            assert np.array_equal(rows[idx], m.t[i]) and np.array_equal(ci[idx], m.t[j])
    assert np.array_equal(ci[plan["diag"]], np.arange(m.nv))
    # executing the gather plan on the REFERENCE's local matrices reproduces the reference's matrix
    vals, written = run_plan(plan, g["Ak"] * g["mu"][None, :])
    assert np.all(written == 1)
    assert rel_err_matrix(vals, g["csc_values_mu"], abs_sum_scale(m, g["Ak"], g["mu"])) < 1e-14
    # the rescale plan (one item per entry pair, contributions in ascending element order) on the same local matrices:
    # every stored entry written exactly once, 10 contributions per element, values == the reference matrix
    G10 = np.zeros((10, m.nt))
    for i in range(4):
        for j in range(i, 4):
            G10[SYM[i, j]] = g["Ak"][4 * j + i]
    rvals, rwritten = run_rescale_plan(plan, G10, g["mu"])
    assert np.all(rwritten == 1) and plan["r_items"] == (plan["nnz"] + m.nv) // 2
    assert int((plan["r_cref"] != 0xFFFFFFFF).sum()) == 10 * m.nt
    assert rel_err_matrix(rvals, g["csc_values_mu"], abs_sum_scale(m, g["Ak"], g["mu"])) < 1e-14
    # every element owned by exactly one tile; every (k,i,j) contribution listed exactly once
    assert int((plan["tile_elems"] >> 31).sum()) == m.nt
    assert plan["tile_ents"].size == 16 * m.nt
    # SELL-32-sigma device layout: every row in exactly one lane, csr2sell injective, padding inert, and
    # walking it the way spmv_sell_kernel does reproduces the CSR product of the reference's matrix
    sr = plan["srow"]
    assert np.array_equal(np.sort(sr[sr >= 0]), np.arange(m.nv))
    assert np.unique(plan["csr2sell"]).size == plan["nnz"] and plan["nnz_sell"] >= plan["nnz"]
    assert np.array_equal(plan["sell_col"][plan["csr2sell"]], ci)
    x = np.random.default_rng(3).standard_normal(m.nv)
    import scipy.sparse as sp
    A = sp.csr_matrix((g["csc_values_mu"], ci, rp), shape=(m.nv, m.nv))
    y = sell_spmv(plan, g["csc_values_mu"], x)
    assert np.max(np.abs(y - A @ x)) <= 1e-14 * np.max(np.abs(A @ x))


def test_sell_packed_columns_mixed_widths():
    """Packed column stream of the SpMV: slices whose columns span < 65536 store 16-bit offsets, the others 32-bit
    columns.  A 74 k-node Kuhn cube in lattice order is all-narrow; swapping two far-apart nodes makes a few slices
    wide; sell_wide_cols=1 forces every slice wide.  All three walk to the same product."""
    import scipy.sparse as sp
    m = fk.kuhn_cube(41)
    assert m.nv > 65536
    rng = np.random.default_rng(11)
    x = rng.standard_normal(m.nv)
    for case in ("lattice", "swapped", "forced"):
        t = m.t.copy()
        p = m.p.copy()
        if case == "swapped":
            a, b = 5, m.nv - 7
            t[m.t == a], t[m.t == b] = b, a
            p[:, [a, b]] = p[:, [b, a]]
        mm = fk.MESH(p=p, t=t, surfaceT=m.surfaceT, nv=m.nv, nt=m.nt) if case == "swapped" else m
        plan = export_symbolic(mm, natural_order=1, **({"sell_wide_cols": 1} if case == "forced" else {}))
        n16, nsl = plan["n_slices16"], plan["nslices"]
        if case == "lattice":
            assert n16 == nsl
        elif case == "swapped":
            assert 0 < nsl - n16 < 64
        else:
            assert n16 == 0 and plan["colz"].size == 2 * plan["nnz_sell"]
        vals = rng.standard_normal(plan["nnz"])
        A = sp.csr_matrix((vals, plan["colind"], plan["rowptr"]), shape=(m.nv, m.nv))
        y = sell_spmv(plan, vals, x)
        ref = A @ x
        assert np.max(np.abs(y - ref)) <= 1e-13 * np.max(np.abs(ref))


def test_symbolic_owned_rows_only():
    """Multi-GPU partitions: only rows < n_owned get a pattern; halo rows are empty."""
    m = fk.kuhn_cube(4)
    n_owned = 60
    plan = export_symbolic(m, n_owned=n_owned)
    full = export_symbolic(m)
    assert np.array_equal(plan["rowptr"][:n_owned + 1], full["rowptr"][:n_owned + 1])
    assert np.all(plan["rowptr"][n_owned:] == plan["rowptr"][n_owned])
    assert np.array_equal(plan["colind"], full["colind"][:plan["nnz"]])
    Ak = orc.local_stiffness(m.p, m.t, m.VE, None)
    v, w = run_plan(plan, Ak)
    vf, _ = run_plan(full, Ak)
    assert np.all(w == 1) and np.allclose(v, vf[:plan["nnz"]], rtol=0, atol=1e-15)
    rows_of = np.repeat(np.arange(m.nv), np.diff(full["rowptr"]))
    e = plan["e2nnz"].reshape(m.nt, 4, 4)
    halo = m.t.T >= n_owned
    assert np.all(e[halo] == -1) and np.all(e[~halo] >= 0)


def test_symbolic_rejects_bad_input():
    m = fk.kuhn_cube(2)
    bad = m.t.copy()
    bad[0, 0] = m.nv
    m2 = fk.MESH(p=m.p, t=bad, nv=m.nv, nt=m.nt)
    with pytest.raises(L.FemekoError, match="out of range"):
        export_symbolic(m2)
    with pytest.raises(L.FemekoError, match="more than one assembly tile"):
        export_symbolic(m, tile_max_elems=4)


def test_spline1d_matches_fitpack():
    from scipy import interpolate

    H = np.concatenate([[0.0], np.geomspace(20, 5e5, 60)])
    mu = fk.mu0 * (1 + 700.0 / (1 + (H / 4000.0) ** 1.2))
    spl = fk.Spline1D(H, mu, bc="nearest")
    tck = interpolate.splrep(H, mu, k=3, s=0)
    assert np.array_equal(spl.t, tck[0])                                      # FITPACK's knot choice
    assert np.allclose(spl.c, tck[1], rtol=1e-9, atol=1e-12 * np.abs(tck[1]).max())
    xs = np.concatenate([H, np.linspace(-10, 6e5, 2001)])
    ref = interpolate.splev(xs, tck, ext=3)
    assert np.allclose(spl(xs), ref, rtol=1e-10)
    assert np.allclose(spl(H), mu, rtol=1e-12)                                # interpolates the table
    with pytest.raises(L.FemekoError, match="outside the knot range"):
        fk.Spline1D(H, mu, bc="error")(np.array([6e5]))
    assert np.allclose(fk.Spline1D(H, mu, bc="error")(H[3:9]), mu[3:9], rtol=1e-12)
    with pytest.raises(L.FemekoError):
        fk.Spline1D(H[::-1].copy(), mu)


def test_material_permeability_matches_oracle():
    H = np.concatenate([[0.0], np.geomspace(5, 1e5, 40)])
    B = 1.6 * np.tanh(H / 800.0) + fk.mu0 * H
    d = fk.DATA()
    d.HofM, d.B = H.copy(), B.copy()
    fk.materialPermeability(d)
    mu, dmu, nu = orc.material_permeability(H, B)
    assert np.array_equal(d.mu, mu) and np.array_equal(d.dmu, dmu) and np.array_equal(d.nu, nu)
    assert np.array_equal(fk.gradient(H, B), orc.gradient(H, B))


def test_mesh_generators_geometry():
    m = fk.kuhn_cube(5)
    assert m.nt == 6 * 125 and m.nv == 216 and m.ns == 12 * 25
    assert np.isclose(m.VE.sum(), 1.0) and np.isclose(m.AE.sum(), 6.0) and np.all(m.VE > 0)
    # outward unit normals
    c = (m.p[:, m.surfaceT[0]] + m.p[:, m.surfaceT[1]] + m.p[:, m.surfaceT[2]]) / 3
    assert np.all((m.normal * c).sum(0) > 0) and np.allclose(np.linalg.norm(m.normal, axis=0), 1)
    d = fk.delaunay_ball(500, seed=2)
    assert np.all(d.VE > 0) and d.nInside > 0 and len(np.unique(d.t)) == d.nv
    F = np.array([0.3, -0.2, 0.9])          # closed boundary: the Neumann load sums to zero
    assert abs(orc.boundary_integral(d, F, [0]).sum()) < 1e-9 * d.AE.sum()


def test_2d_triangles_through_the_tet_plan():
    """2-D P1 (src/FEM2D.jl) rides the same symbolic phase as the elements (n0, n1, n2, n2): the pattern is the triangle
    mesh's node adjacency, and executing the gather plan on 4 x 4 local matrices whose fourth row / column is zero
    reproduces the oracle's stiffnessMatrix2D / massMatrix2D."""
    for m in (fk.tri_grid(9), fk.delaunay_disc(600, seed=5)):
        m4 = fk.MESH()
        m4.p, m4.nv, m4.nt = m.p, m.nv, m.nt
        m4.t = np.asfortranarray(np.vstack([m.t, m.t[2:3]]).astype(np.int32))
        plan = export_symbolic(m4, _rescale=False)
        rp, ci, vref = orc.stiffness2d(m, np.linspace(1.0, 2.0, m.nt))
        assert np.array_equal(plan["rowptr"], rp) and np.array_equal(plan["colind"], ci)
        Ak = np.zeros((16, m.nt))
        for k in range(m.nt):
            nds = m.t[:, k]
            bc = np.array([orc.abc(m.p, nds, int(nds[i]))[1:] for i in range(3)])
            loc = np.zeros((4, 4))
            loc[:3, :3] = m.VE[k] * (1.0 + k / (m.nt - 1.0)) * (bc @ bc.T)
            Ak[:, k] = loc.ravel(order="F")
        vals, written = run_plan(plan, Ak)
        assert np.all(written == 1)
        assert np.max(np.abs(vals - vref)) < 1e-13 * np.max(np.abs(vref))


def test_p2_symbolic_and_contribution_plan():
    """10-node (P2) elements: the generalised symbolic phase gives the clique pattern of every element, the 100-entry
    element -> nnz map, and a contribution plan (55 local pairs) whose execution on the oracle's local matrices
    reproduces quadraticStiffnessMatrix (src/FEM3D.jl:378-395)."""
    m = fk.kuhn_cube_p2(2)
    plan = export_symbolic(m, nodes_per_element=10)
    rp, ci = orc.pattern_npe(m.t, m.nv)
    assert np.array_equal(plan["rowptr"], rp) and np.array_equal(plan["colind"], ci) and plan["ntiles"] == 0
    rows = np.repeat(np.arange(m.nv), np.diff(rp))
    e2n = plan["e2nnz"].reshape(m.nt, 10, 10)
    for i in (0, 3, 7, 9):
        for j in (0, 4, 9):
            assert np.array_equal(rows[e2n[:, i, j]], m.t[i]) and np.array_equal(ci[e2n[:, i, j]], m.t[j])
    S = orc.quadratic_basis_all(m)
    mu = np.linspace(1.0, 2.0, m.nt)
    Ak = orc.quadratic_local_stiffness(m, S)            # unscaled (mu = 1): the plan's factors
    G55 = np.zeros((55, m.nt))
    n = 0
    for a in range(10):
        for b in range(a, 10):
            G55[n] = Ak[10 * b + a]
            n += 1
    vals, written = run_rescale_plan(plan, G55, mu, shift=26)
    assert np.all(written == 1) and int((plan["r_cref"] != 0xFFFFFFFF).sum()) == 55 * m.nt
    _, _, vref = orc.quadratic_stiffness(m, S, mu)
    assert np.max(np.abs(vals - vref)) < 1e-12 * np.max(np.abs(vref))


def test_nedelec_symbolic_and_contribution_plan():
    """Edge elements: the symbolic phase on the 6 edge labels of every element (edge2localMap[t[5:10, :]], src/mesh.jl:170-185)
    gives the pattern of the reference's sum of sparse(edge1, edge2, ...) and a contribution plan (21 local pairs) whose
    execution on the oracle's nedelecLocalStiffness reproduces the global curl-curl matrix (nedelec3D.jl:163-178)."""
    from femeko_jl_b200 import mesh as M

    m = fk.to_p2(fk.delaunay_ball(200, seed=2))
    M.edgeNumbering(m)
    ed, mp = orc.edge_numbering(m.t)                   # the oracle's literal first-appearance walk
    assert np.array_equal(ed, m.edges) and np.array_equal(mp, m.edge2localMap)
    assert not np.array_equal(m.edges, np.sort(m.edges)), "Julia's unique() keeps first appearance, not sorted order"

    class Shim:
        pass
    e = Shim()
    e.t, e.nv, e.nt, e.p = M.edgeDofs(m), m.ne, m.nt, np.zeros((3, m.ne), order="F")
    plan = export_symbolic(e, nodes_per_element=6)
    rp, ci = orc.pattern_npe(e.t, e.nv)
    assert np.array_equal(plan["rowptr"], rp) and np.array_equal(plan["colind"], ci) and plan["ntiles"] == 0
    nu = np.linspace(1.0, 3.0, m.nt)
    Ak = orc.nedelec_local_stiffness(m)
    G21 = np.zeros((21, m.nt))
    n = 0
    for a in range(6):
        for b in range(a, 6):
            G21[n] = Ak[6 * a + b]
            n += 1
    vals, written = run_rescale_plan(plan, G21, nu, shift=26)
    assert np.all(written == 1) and int((plan["r_cref"] != 0xFFFFFFFF).sum()) == 21 * m.nt
    _, _, vref = orc.nedelec_stiffness(m, nu)
    assert np.max(np.abs(vals - vref)) < 1e-12 * np.max(np.abs(vref))
