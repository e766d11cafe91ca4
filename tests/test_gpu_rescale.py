"""GPU: the rescale assembly (csrc/rescale.cuh -- geometric factors computed once, every later stiffness assembly /
Picard sweep streams sum_k VE[k] f[k] g the way NewtonRaphson.jl:114-127,153-161 keeps Ak and rescales it by mu)
against the tile kernel (bit for bit: same products, same summation order) and against the oracle."""
import os

import numpy as np
import pytest

import femeko_jl_b200 as fk
from conftest import GOLDEN, abs_sum_scale, rel_err_matrix
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
MU0 = 4e-7 * np.pi


def gd_spline():
    g = np.load(os.path.join(GOLDEN, "gd_270K.npz"))
    return fk.Spline1D(g["HofM"], g["mu"], bc="nearest")


def same_matrix(h, v_geo, v_tile, rowsum):
    """rowsum = False: bit for bit.  rowsum = True (default): every off-diagonal entry bit for bit, the diagonal -- taken from the
    row sums instead of its own contributions -- equal to rounding (a few ulp of the row's absolute sum)."""
    if not rowsum:
        return np.array_equal(v_geo, v_tile)
    rp, ci = h.pattern()
    rows = np.repeat(np.arange(len(rp) - 1), np.diff(rp))
    off = ci != rows
    absrow = np.zeros(len(rp) - 1)
    np.add.at(absrow, rows, np.abs(v_tile))
    return bool(np.array_equal(v_geo[off], v_tile[off]) and np.all(np.abs(v_geo[~off] - v_tile[~off]) <= 1e-14 * absrow[rows[~off]]))


def meshes(name):
    if name == "kuhn12":
        return fk.kuhn_cube(12, inside_edge=0.5)
    if name == "kuhn40":
        return fk.kuhn_cube(40, inside_edge=0.3)
    return fk.delaunay_ball(4000, seed=11)


@pytest.mark.parametrize("mesh", ["kuhn12", "delaunay", "kuhn40"])
@pytest.mark.parametrize("use_ve", [False, True])
@pytest.mark.parametrize("rowsum", [False, True])
def test_stiffness_rescale_bitwise_and_oracle(mesh, use_ve, rowsum):
    """fk_assemble_stiffness through the rescale path == through the tile kernel, bit for bit; both == the oracle."""
    m = meshes(mesh)
    rng = np.random.default_rng(5)
    f = rng.uniform(0.5, 3.0, m.nt)
    h = fk.Femeko(m, use_mesh_VE=use_ve)
    h.tune(10, 0)
    v_tile = h.assemble_stiffness(f, slot=0)
    v_tile1 = h.assemble_stiffness(None, slot=0)
    h.tune(10, 2)
    h.tune(18, int(rowsum))
    v_geo = h.assemble_stiffness(f, slot=1)
    v_geo1 = h.assemble_stiffness(None, slot=1)
    items, contrib, slots, nbytes = h.geo_stats()
    rp, ci = h.pattern()
    if rowsum:
        assert contrib == 6 * m.nt and items == (h.nnz - m.nv) // 2 and slots >= contrib
    else:
        assert contrib == 10 * m.nt and items == (h.nnz + m.nv) // 2 and slots >= contrib
    assert slots < 1.12 * contrib, "padding of the contribution slices above 12 %"
    assert same_matrix(h, v_geo, v_tile, rowsum) and same_matrix(h, v_geo1, v_tile1, rowsum)
    _, _, ref = orc.stiffness(m, f, order="eigen")
    scale = abs_sum_scale(m, orc.local_stiffness(m.p, m.t, m.VE), f)
    assert rel_err_matrix(v_geo, ref, scale) < 1e-12
    h.close()


@pytest.mark.parametrize("mesh", ["kuhn12", "delaunay"])
@pytest.mark.parametrize("compat", [False, True])
@pytest.mark.parametrize("rowsum", [False, True])
def test_picard_sweep_rescale_bitwise(mesh, compat, rowsum):
    """A Picard sweep (H recovery of every element + mu(|H|) on the material + stiffness) through the rescale path
    == the fused tile kernel: matrix, mu, |H|, H and the convergence metric, bit for bit; and == the oracle."""
    m = meshes(mesh)
    spl = gd_spline()
    rng = np.random.default_rng(9)
    u = rng.normal(size=m.nv) * 2e4           # |H| ~ 1e5 A/m on a unit-size mesh: the nonlinear part of the table
    out = []
    for mode in (0, 2):
        h = fk.Femeko(m)
        h.tune(10, mode)
        h.tune(18, int(rowsum))
        h.set_spline(0, spl)
        h.set_inside(m.InsideElements)
        h.set_mu(np.full(m.nt, MU0))
        d1 = h.picard_step(u, slot=0, compat_bug=compat)
        d2 = h.picard_step(0.5 * u, slot=0, compat_bug=compat)    # second sweep: |H|old is the first one's
        Hf, Hn = h.get_field()
        out.append((h.matrix_get(0), h.get_mu(), Hf.copy(), Hn.copy(), d1, d2))
        ok = same_matrix(h, out[-1][0], out[0][0], rowsum)
        h.close()
    (v0, mu0, Hf0, Hn0, d10, d20), (v1, mu1, Hf1, Hn1, d11, d21) = out
    assert np.array_equal(mu0, mu1) and np.array_equal(Hn0, Hn1) and np.array_equal(Hf0, Hf1)
    assert d10 == d11 and d20 == d21
    assert ok
    _, Hno = orc.recover_field(m, 0.5 * u, compat_bug=compat)
    mu_ref = np.full(m.nt, MU0)
    mu_ref[m.InsideElements] = orc.splev(spl.t, spl.c, 3, Hno[m.InsideElements], "nearest")
    assert np.max(np.abs(mu1 - mu_ref) / mu_ref) < 1e-11
    _, _, ref = orc.stiffness(m, mu_ref, order="eigen")
    scale = abs_sum_scale(m, orc.local_stiffness(m.p, m.t, m.VE), mu_ref)
    assert rel_err_matrix(v1, ref, scale) < 1e-11


def test_default_mode_switches_on_second_sweep():
    """Default policy (fk_tune 10 = 1): the first sweep of a handle runs the tile kernel, every later one the rescale
    path -- a Picard loop gets the streamed sweep without asking, with unchanged bits."""
    m = fk.kuhn_cube(10, inside_edge=0.5)
    spl = gd_spline()
    u = np.random.default_rng(1).normal(size=m.nv) * 2e4
    h = fk.Femeko(m)
    h.set_spline(0, spl)
    h.set_inside(m.InsideElements)
    h.set_mu(np.full(m.nt, MU0))
    h.picard_step(u, slot=0)
    v_first = h.matrix_get(0)
    mu_first = h.get_mu()
    h.set_mu(np.full(m.nt, MU0))
    h.reset_field()
    l0 = h.launch_count()
    h.picard_step(u, slot=0)
    assert same_matrix(h, h.matrix_get(0), v_first, True) and np.array_equal(h.get_mu(), mu_first)
    # a later plain stiffness assembly re-uses the factors as well
    f = np.random.default_rng(2).uniform(1, 2, m.nt)
    v = h.assemble_stiffness(f, slot=1)
    h.tune(10, 0)
    assert same_matrix(h, v, h.assemble_stiffness(f, slot=2), True)
    assert h.launch_count() > l0
    h.close()
