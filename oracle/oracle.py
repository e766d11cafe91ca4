"""TEST INFRASTRUCTURE ONLY -- ctypes front end of the CPU oracle.

Two libraries:
  * ``oracle/_build/liboracle.so``   plain-C restatement (femeko_oracle.c), always buildable (gcc);
  * ``oracle/_ref/libfemeko_ref.so`` the UNMODIFIED reference C++ (src/FEMc.cpp + Eigen) behind
    ref_shim.cpp, and ``oracle/_ref/magnetostatics.so`` (the reference's own ccall library,
    Femeko_jl_C++/Magnetostatics/magnetostatics.cpp:49-95).  Built in the dev container where
    /root/reference exists; they travel to the GPU box as prebuilt files.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  Parity status: pinned against outputs of the compiled reference (no golden vectors
exist upstream); Julia-only orderings and Dierckx are pinned via scipy's FITPACK.
The numpy restatements further down (2-D P1, P2 stiffness / mass / convection, Nedelec elements and its Newton loop) follow code
that exists in Julia only (no C++ twin, no Julia here): PARITY UNPINNED against Julia itself for those; tests/test_oracle.py pins
them by the properties the elements must satisfy.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def build(ref: bool = True) -> None:
    """Compile the oracle (and, when /root/reference is present, the reference shim)."""
    subprocess.run(["make", "-s", "-C", _HERE, "_build/liboracle.so"] + (["ref"] if ref else []),
                   check=True, env={**os.environ, "MAKEFLAGS": ""})


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _i(a):
    return None if a is None else a.ctypes.data_as(_ip)


def _f64(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float64).ravel(order="K"))


def _i32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.int32).ravel(order="K"))


def _flat(a, dtype):
    """Column-major flattening of a (possibly F-ordered) matrix: the bytes Julia would pass."""
    if a is None:
        return None
    return np.ascontiguousarray(np.asarray(a, dtype=dtype).ravel(order="F"))


_lib = None
_ref = None
_mag = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "_build", "liboracle.so")
        if not os.path.exists(path):
            build(ref=False)
        _lib = C.CDLL(path)
        _lib.orc_pattern.restype = C.c_long
    return _lib


def have_ref() -> bool:
    return os.path.exists(os.path.join(_HERE, "_ref", "libfemeko_ref.so"))


def ref():
    global _ref
    if _ref is None:
        _ref = C.CDLL(os.path.join(_HERE, "_ref", "libfemeko_ref.so"))
        _ref.ref_stiffness_csc.restype = C.c_long
    return _ref


def ref_magnetostatics():
    global _mag
    if _mag is None:
        _mag = C.CDLL(os.path.join(_HERE, "_ref", "magnetostatics.so"))
    return _mag


# ----------------------------------------------------------------------------- restatement
def abcd(p, nodes, nd):
    r = np.zeros(4)
    lib().orc_abcd(_d(_flat(p, np.float64)), _i(_i32(nodes)), C.c_int(int(nd)), _d(r))
    return r


def local_stiffness(p, t, VE, f=None):
    nt = t.shape[1]
    Ak = np.zeros(16 * nt)
    lib().orc_local_stiffness(_d(_flat(p, np.float64)), _i(_flat(t, np.int32)), _d(_f64(VE)), _d(_f64(f)),
                              C.c_int(nt), _d(Ak))
    return Ak.reshape(nt, 16).T  # 16 x nt view, Julia layout


def local_mass(VE, rho=None):
    nt = len(VE)
    Mk = np.zeros(16 * nt)
    lib().orc_local_mass(_d(_f64(VE)), _d(_f64(rho)), C.c_int(nt), _d(Mk))
    return Mk.reshape(nt, 16).T


def local_tangential(p, t, VE, Hvec, dmu):
    nt = t.shape[1]
    Ak = np.zeros(16 * nt)
    lib().orc_local_tangential(_d(_flat(p, np.float64)), _i(_flat(t, np.int32)), _d(_f64(VE)),
                               _d(_flat(Hvec, np.float64)), _d(_f64(dmu)), C.c_int(nt), _d(Ak))
    return Ak.reshape(nt, 16).T


def pattern(t, nv):
    nt = t.shape[1]
    tf = _flat(t, np.int32)
    rowptr = np.zeros(nv + 1, dtype=np.int32)
    nnz = lib().orc_pattern(_i(tf), C.c_int(nv), C.c_int(nt), _i(rowptr), None)
    colind = np.zeros(nnz, dtype=np.int32)
    lib().orc_pattern(_i(tf), C.c_int(nv), C.c_int(nt), _i(rowptr), _i(colind))
    return rowptr, colind


def elem2nnz(t, rowptr, colind):
    nt = t.shape[1]
    m = np.zeros(16 * nt, dtype=np.int32)
    lib().orc_elem2nnz(_i(_flat(t, np.int32)), C.c_int(nt), _i(rowptr), _i(colind), _i(m))
    return m


def assemble_julia(e2nnz, nnz, Ak, scale=None, n_major_single=False):
    nt = Ak.shape[1]
    vals = np.zeros(nnz)
    lib().orc_assemble_julia(_i(e2nnz), C.c_int(nt), C.c_long(nnz), _d(_flat(Ak, np.float64)), _d(_f64(scale)),
                             C.c_int(int(n_major_single)), _d(vals))
    return vals


def assemble_eigen(e2nnz, nnz, Ak, mu=None):
    nt = Ak.shape[1]
    vals = np.zeros(nnz)
    lib().orc_assemble_eigen(_i(e2nnz), C.c_int(nt), C.c_long(nnz), _d(_flat(Ak, np.float64)), _d(_f64(mu)), _d(vals))
    return vals


def stiffness(mesh, f=None, order="julia"):
    """stiffnessMatrix(mesh, f) -> (rowptr, colind, vals) on the canonical pattern."""
    rowptr, colind = pattern(mesh.t, mesh.nv)
    m = elem2nnz(mesh.t, rowptr, colind)
    if order == "julia":
        Ak = local_stiffness(mesh.p, mesh.t, mesh.VE, f)
        vals = assemble_julia(m, len(colind), Ak)
    else:
        Ak = local_stiffness(mesh.p, mesh.t, mesh.VE, None)
        vals = assemble_eigen(m, len(colind), Ak, f)
    return rowptr, colind, vals


def mass(mesh, rho=None):
    rowptr, colind = pattern(mesh.t, mesh.nv)
    m = elem2nnz(mesh.t, rowptr, colind)
    return rowptr, colind, assemble_julia(m, len(colind), local_mass(mesh.VE, rho))


def tangential(mesh, Hvec, dmu):
    rowptr, colind = pattern(mesh.t, mesh.nv)
    m = elem2nnz(mesh.t, rowptr, colind)
    return rowptr, colind, assemble_julia(m, len(colind), local_tangential(mesh.p, mesh.t, mesh.VE, Hvec, dmu))


def lagrange(mesh):
    Cv = np.zeros(mesh.nv)
    lib().orc_lagrange(_i(_flat(mesh.t, np.int32)), _d(_f64(mesh.VE)), C.c_int(mesh.nv), C.c_int(mesh.nt), _d(Cv))
    return Cv


def boundary_integral(mesh, F, shell_id, form="julia"):
    RHS = np.zeros(mesh.nv)
    Fv = _f64(F)
    if form == "julia":
        ids = _i32(np.atleast_1d(shell_id))
        lib().orc_boundary_integral_julia(_i(_flat(mesh.surfaceT, np.int32)), _d(_flat(mesh.normal, np.float64)),
                                          _d(_f64(mesh.AE)), C.c_int(mesh.ns), _d(Fv), _i(ids), C.c_int(len(ids)),
                                          C.c_int(mesh.nv), _d(RHS))
    else:
        lib().orc_boundary_integral_cpp(_d(_flat(mesh.p, np.float64)), _i(_flat(mesh.surfaceT, np.int32)),
                                        _d(_flat(mesh.normal, np.float64)), C.c_int(mesh.ns), _d(Fv),
                                        C.c_int(int(np.atleast_1d(shell_id)[0])), C.c_int(mesh.nv), _d(RHS))
    return RHS


def recover_field(mesh, u, compat_bug=False):
    Hf = np.zeros(3 * mesh.nt)
    H = np.zeros(mesh.nt)
    lib().orc_recover_field(_d(_flat(mesh.p, np.float64)), _i(_flat(mesh.t, np.int32)), _d(_f64(u)), C.c_int(mesh.nt),
                            C.c_int(int(compat_bug)), _d(Hf), _d(H))
    return np.asfortranarray(Hf.reshape(mesh.nt, 3).T), H


def gradient(x, y):
    z = np.zeros(len(x))
    lib().orc_gradient(_d(_f64(x)), _d(_f64(y)), C.c_int(len(x)), _d(z))
    return z


def material_permeability(H, B):
    n = len(H)
    mu, dmu, nu = np.zeros(n), np.zeros(n), np.zeros(n)
    lib().orc_material_permeability(_d(_f64(H)), _d(_f64(B)), C.c_int(n), _d(mu), _d(dmu), _d(nu))
    return mu, dmu, nu


BC = {"extrapolate": 0, "zero": 1, "error": 2, "nearest": 3}


def splev(tk, c, k, x, bc="nearest"):
    tk, c = _f64(tk), _f64(c)
    x = np.atleast_1d(np.asarray(x, dtype=np.float64))
    out = np.zeros(x.size)
    err = 0
    v = C.c_double()
    for m, xv in enumerate(x.ravel()):
        err |= lib().orc_splev(_d(tk), C.c_int(len(tk)), _d(c), C.c_int(k), C.c_double(xv), C.c_int(BC[bc]), C.byref(v))
        out[m] = v.value
    if err:
        raise ValueError("spline evaluated outside its knots with bc='error'")
    return out.reshape(x.shape)


def spmv(rowptr, colind, vals, x):
    n = len(rowptr) - 1
    y = np.zeros(n)
    lib().orc_spmv(C.c_int(n), _i(rowptr), _i(colind), _d(_f64(vals)), _d(_f64(x)), _d(y))
    return y


def pcg(rowptr, colind, vals, b, tol=2.220446049250313e-16, max_iter=None):
    n = len(rowptr) - 1
    x = np.zeros(n)
    it, err = C.c_int(), C.c_double()
    info = lib().orc_pcg(C.c_int(n), _i(rowptr), _i(colind), _d(_f64(vals)), _d(_f64(b)), _d(x), C.c_double(tol),
                         C.c_int(2 * n if max_iter is None else max_iter), C.byref(it), C.byref(err))
    return x, it.value, err.value, info


# ----------------------------------------------------------------------------- compiled reference
def ref_abcd(p, nodes, nd):
    r = np.zeros(4)
    pf = _flat(p, np.float64)
    ref().ref_abcd(_d(pf), C.c_int(p.shape[1]), _i(_i32(nodes)), C.c_int(int(nd)), _d(r))
    return r


def ref_local_stiffness(mesh):
    Ak = np.zeros(16 * mesh.nt)
    ref().ref_local_stiffness(_d(_flat(mesh.p, np.float64)), _i(_flat(mesh.t, np.int32)), C.c_int(mesh.nv),
                              C.c_int(mesh.nt), _d(_f64(mesh.VE)), _d(Ak))
    return Ak.reshape(mesh.nt, 16).T


def ref_stiffness(mesh, mu=None, want_time=False):
    """Reference stiffnessMatrix (FEMc.cpp:102-167) -> CSC (outer, inner, values)."""
    pf, tf, ve, muv = _flat(mesh.p, np.float64), _flat(mesh.t, np.int32), _f64(mesh.VE), _f64(mu)
    outer = np.zeros(mesh.nv + 1, dtype=np.int32)
    secs = C.c_double()
    # two passes would assemble twice; over-allocate with the structural bound instead
    rowptr, _ = pattern(mesh.t, mesh.nv)
    cap = int(rowptr[-1])
    inner = np.zeros(cap, dtype=np.int32)
    vals = np.zeros(cap)
    nnz = ref().ref_stiffness_csc(_d(pf), _i(tf), C.c_int(mesh.nv), C.c_int(mesh.nt), _d(ve), _d(muv), _i(outer),
                                  _i(inner), _d(vals), C.byref(secs))
    out = (outer, inner[:nnz], vals[:nnz])
    return out + (secs.value,) if want_time else out


def ref_stiffness_seconds(mesh, mu):
    """Wall time of the reference assembly only (no copy-out)."""
    secs = C.c_double()
    ref().ref_stiffness_csc(_d(_flat(mesh.p, np.float64)), _i(_flat(mesh.t, np.int32)), C.c_int(mesh.nv),
                            C.c_int(mesh.nt), _d(_f64(mesh.VE)), _d(_f64(mu)), None, None, None, C.byref(secs))
    return secs.value


def ref_lagrange(mesh):
    Cv = np.zeros(mesh.nv)
    ref().ref_lagrange(_i(_flat(mesh.t, np.int32)), _d(_f64(mesh.VE)), C.c_int(mesh.nv), C.c_int(mesh.nt), _d(Cv))
    return Cv


def ref_boundary_integral(mesh, F, shell_id):
    RHS = np.zeros(mesh.nv)
    ref().ref_boundary_integral(_d(_flat(mesh.p, np.float64)), _i(_flat(mesh.surfaceT, np.int32)),
                                _d(_flat(mesh.normal, np.float64)), C.c_int(mesh.nv), C.c_int(mesh.ns), _d(_f64(F)),
                                C.c_int(int(shell_id)), _d(RHS))
    return RHS


def ref_cg(outer, inner, values, b, tol=0.0, max_iter=0):
    """Eigen ConjugateGradient (Lower, Jacobi) as magnetostatics.cpp:26-38; tol/max_iter <= 0 keep
    Eigen's defaults.  Returns x, iterations, error estimate, info, seconds."""
    n = len(outer) - 1
    x = np.zeros(n)
    it, err, secs = C.c_int(), C.c_double(), C.c_double()
    info = ref().ref_cg(C.c_int(n), _i(_i32(outer)), _i(_i32(inner)), _d(_f64(values)), _d(_f64(b)), _d(x),
                        C.c_double(tol), C.c_int(max_iter), C.byref(it), C.byref(err), C.byref(secs))
    return x, it.value, err.value, info, secs.value


def ref_scalar_potential(mesh, mu, Hext, shell_id):
    """The reference's own ccall entry point, called exactly as run.jl:66-79 does."""
    u = np.zeros(mesh.nv)
    ref_magnetostatics().scalarPotential(
        _d(u), _d(_flat(mesh.p, np.float64)), _i(_flat(mesh.t, np.int32)), _i(_flat(mesh.surfaceT, np.int32)),
        _d(_flat(mesh.normal, np.float64)), C.c_int(mesh.nv), C.c_int(mesh.nt), C.c_int(mesh.ns), _d(_f64(mesh.VE)),
        _d(_f64(mu)), _d(_f64(Hext)), C.c_int(int(shell_id)))
    return u


# ----------------------------------------------------------------------------- micromagnetics demag field
def node_volumes(mesh, inside):
    """Box-method node volumes, LL.cpp:29-36."""
    vol = np.zeros(mesh.nv)
    ins = _i32(inside)
    lib().orc_node_volumes(_i(_flat(mesh.t, np.int32)), _d(_f64(mesh.VE)), _i(ins), C.c_int(len(ins)), C.c_int(mesh.nv), _d(vol))
    return vol


def demag_rhs(mesh, inside, M):
    """Load vector of LL::magnetostaticField, LL.cpp:292-312.  M is 3 x nv."""
    rhs = np.zeros(mesh.nv)
    ins = _i32(inside)
    lib().orc_demag_rhs(_d(_flat(mesh.p, np.float64)), _i(_flat(mesh.t, np.int32)), _d(_f64(mesh.VE)), _i(ins), C.c_int(len(ins)),
                        C.c_int(mesh.nv), _d(_flat(M, np.float64)), _d(rhs))
    return rhs


def demag_box(mesh, inside, u, vol):
    """Element field -> nodes by the box method, LL.cpp:317-346 (NaN where vol == 0, like the reference)."""
    Hd = np.zeros(3 * mesh.nv)
    ins = _i32(inside)
    with np.errstate(all="ignore"):
        lib().orc_demag_box(_d(_flat(mesh.p, np.float64)), _i(_flat(mesh.t, np.int32)), _d(_f64(mesh.VE)), _i(ins), C.c_int(len(ins)),
                            C.c_int(mesh.nv), _d(_f64(u)), _d(_f64(vol)), _d(Hd))
    return np.asfortranarray(Hd.reshape(mesh.nv, 3).T)


def demag_field(mesh, inside, M, tol=2.220446049250313e-16):
    """LL::magnetostaticField restated end to end: RHS, Jacobi-PCG (Eigen defaults) on stiffnessMatrix(p,t,VE), box method."""
    rowptr, colind, vals = stiffness(mesh, None, order="eigen")
    rhs = demag_rhs(mesh, inside, M)
    u, it, _, _ = pcg(rowptr, colind, vals, rhs, tol)
    vol = node_volumes(mesh, inside)
    return demag_box(mesh, inside, u, vol), u, it, rhs, vol


_ref_ll = None


def have_ref_ll() -> bool:
    return os.path.exists(os.path.join(_HERE, "_ref", "libfemeko_ref_ll.so"))


def ref_demag_field(mesh, inside, inside_nodes, M):
    """The UNMODIFIED reference class LL (ref_ll_shim.cpp): returns (Hd 3 x nv, volumes)."""
    global _ref_ll
    if _ref_ll is None:
        _ref_ll = C.CDLL(os.path.join(_HERE, "_ref", "libfemeko_ref_ll.so"))
    Hd = np.zeros(3 * mesh.nv)
    vol = np.zeros(mesh.nv)
    pf, tf = _flat(mesh.p, np.float64), _flat(mesh.t, np.int32)
    ins, insn = _i32(inside), _i32(inside_nodes)
    Mf = _flat(M, np.float64).copy()
    VE = _f64(mesh.VE).copy()
    _ref_ll.ref_demag_field(_d(pf), _i(tf), _d(VE), _i(ins), _i(insn), C.c_int(mesh.nv), C.c_int(mesh.nt), C.c_int(len(ins)),
                            C.c_int(len(insn)), _d(Mf), _d(Hd), _d(vol))
    return np.asfortranarray(Hd.reshape(mesh.nv, 3).T), vol


# ----------------------------------------------------------------------------- nonlinear drivers (restated loops)
MU0 = 4e-7 * np.pi


def _csr(rowptr, colind, vals, n):
    import scipy.sparse as sp

    return sp.csr_matrix((vals, colind, rowptr), shape=(n, n))


def bordered_solve(A, Lag, b):
    """u = ([A Lag; Lag' 0] \\ [b; 0])[1:nv]  (magneticField.jl:107, manyBodyIntersecting.jl:155, NewtonRaphson.jl:166)."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla

    n = A.shape[0]
    K = sp.bmat([[A, sp.csr_matrix(Lag.reshape(n, 1))], [sp.csr_matrix(Lag.reshape(1, n)), None]], format="csc")
    return spla.spsolve(K, np.concatenate([b, [0.0]]))[:n]


def line_search(u, du, A, RHS, min_step=1e-4):
    """lineSearch of src/FEM3D.jl:123-152 restated; updates u in place, returns (residue, accepted alpha or 0)."""
    alf = 1.0
    residual_old = np.linalg.norm(RHS - A @ u)
    u_trial = u + du
    residue = np.linalg.norm(RHS - A @ u_trial)
    while residue > residual_old and alf > min_step:
        alf *= 0.9
        u_trial = u + alf * du
        residue = np.linalg.norm(RHS - A @ u_trial)
    if alf < min_step:
        return residue, 0.0
    u[:] = u_trial
    return residue, alf


def _solve(mesh, mu, b, Lag, solver, tol, extra_vals=None):
    """One linear solve of a nonlinear pass: 'direct' = the reference Julia's bordered sparse solve; 'pcg' = the C port
    of Eigen's Jacobi-CG on the oracle matrix; 'ref' = compiled reference stiffnessMatrix + Eigen ConjugateGradient
    (magnetostatics.cpp:26-38).  Returns (u, iterations)."""
    if solver == "ref":
        outer, inner, v = ref_stiffness(mesh, mu)
        x, it, _, _, _ = ref_cg(outer, inner, v, b, tol, 0)
        return x, it
    rp, ci, v = stiffness(mesh, mu, order="julia")
    if extra_vals is not None:
        v = v + extra_vals
    if solver == "direct":
        return bordered_solve(_csr(rp, ci, v, mesh.nv), Lag, b), 0
    x, it, _, _ = pcg(rp, ci, v, b, tol)
    return x, it


def picard_loop(mesh, Bext, tck, tol=1e-4, max_att=100, shell_id=0, solver="direct", cg_tol=1e-12, bc="nearest"):
    """Picard loop of Magnetostatics/manyBodyIntersecting.jl:141-210 restated (one material = the inside elements).
    Returns dict(u, Hfield, H, mu, att, div, history, cg_iters)."""
    RHS = boundary_integral(mesh, Bext, [shell_id], "julia")
    Lag = lagrange(mesh)
    mu = np.full(mesh.nt, MU0)
    H = np.zeros(mesh.nt)
    ins = np.asarray(mesh.InsideElements)
    att, div, hist, cg_it = 0, 1.0, [], 0
    u = np.zeros(mesh.nv)
    Hf = np.zeros((3, mesh.nt))
    while div > tol and att < max_att:
        att += 1
        Hold = H.copy()
        u, it = _solve(mesh, mu, -RHS, Lag, solver, cg_tol)
        cg_it += it
        Hf, H = recover_field(mesh, u)
        mu[ins] = splev(tck[0], tck[1], 3, H[ins], bc)
        if not np.all(np.isfinite(mu)):
            raise FloatingPointError("Nans/Infs in mu")
        div = MU0 * np.max(np.abs(H[ins] - Hold[ins]))
        hist.append((att, div, it))
    return dict(u=u, Hfield=Hf, H=H, mu=mu, att=att, div=div, history=hist, cg_iters=cg_it)


def newton_loop(mesh, Bext, tck, tck_dmu, tol=1e-4, max_att=100, shell_id=0, compat_bug=True, bc="nearest"):
    """Newton-Raphson loop of Magnetostatics/NewtonRaphson.jl:148-214 restated (bordered direct solves)."""
    RHS = -boundary_integral(mesh, Bext, [shell_id], "julia")
    Lag = lagrange(mesh)
    mu = np.full(mesh.nt, MU0)
    dmu = np.zeros(mesh.nt)
    H = np.zeros(mesh.nt)
    Hf = np.zeros((3, mesh.nt))
    u = np.zeros(mesh.nv)
    ins = np.asarray(mesh.InsideElements)
    att, div, hist = 0, np.inf, []
    while div > tol and att < max_att:
        att += 1
        Hold = H.copy()
        rp, ci, v = stiffness(mesh, mu, order="julia")
        A = _csr(rp, ci, v, mesh.nv)
        vt = tangential(mesh, Hf, dmu)[2] if att > 1 else np.zeros_like(v)
        du = bordered_solve(_csr(rp, ci, v + vt, mesh.nv), Lag, RHS - A @ u)
        residue, alf = line_search(u, du, A, RHS)
        Hf, H = recover_field(mesh, u, compat_bug=compat_bug)
        div = MU0 * np.max(np.abs(H[ins] - Hold[ins]))
        mu[ins] = splev(tck[0], tck[1], 3, H[ins], bc)
        dmu[ins] = splev(tck_dmu[0], tck_dmu[1], 3, H[ins], bc)
        if not (np.all(np.isfinite(mu)) and np.all(np.isfinite(dmu))):
            raise FloatingPointError("Nans/Infs in mu")
        hist.append((att, div, residue, alf))
    return dict(u=u, Hfield=Hf, H=H, mu=mu, dmu=dmu, att=att, div=div, history=hist)


# --------------------------------------------------------------------------------------------------
# 2-D P1 triangles (src/FEM2D.jl) -- numpy restatement, test infrastructure only
# --------------------------------------------------------------------------------------------------
def abc(p, nds, nd):
    """src/FEM2D.jl:5-25: coefficients (a, b, c) of the hat function of node nd on the triangle nds, from the 3 x 3
    solve M \\ [1, 0, 0] with the node itself in the first row."""
    others = [int(i) for i in nds if i != nd]
    M = np.array([[1.0, p[0, nd], p[1, nd]], [1.0, p[0, others[0]], p[1, others[0]]], [1.0, p[0, others[1]], p[1, others[1]]]])
    return np.linalg.solve(M, np.array([1.0, 0.0, 0.0]))


def pattern2d(t, nv):
    """Canonical CSR pattern of a triangle mesh: node adjacency + diagonal, ascending columns."""
    import scipy.sparse as sp

    r = np.concatenate([t[i] for i in range(3) for _ in range(3)]).astype(np.int64)
    c = np.concatenate([t[j] for _ in range(3) for j in range(3)]).astype(np.int64)
    A = sp.csr_matrix((np.ones(r.size), (r, c)), shape=(nv, nv))
    A.sort_indices()
    return A.indptr.astype(np.int32), A.indices.astype(np.int32)


def _assemble2d(mesh, Ak):
    """A = sum over the 9 local pairs (i outer, j inner) of sparse(t[i,:], t[j,:], Ak[n,:]), duplicates summed in element
    order, the 9 partial matrices folded left to right (src/FEM2D.jl:66-73); values on the canonical pattern."""
    rp, ci = pattern2d(mesh.t, mesh.nv)
    vals = np.zeros(len(ci))
    key = {}
    rows = np.repeat(np.arange(mesh.nv), np.diff(rp))
    for q, (r, c) in enumerate(zip(rows, ci)):
        key[(int(r), int(c))] = q
    n = 0
    for i in range(3):
        for j in range(3):
            part = np.zeros(len(ci))
            for k in range(mesh.nt):
                part[key[(int(mesh.t[i, k]), int(mesh.t[j, k]))]] += Ak[n, k]
            vals = vals + part
            n += 1
    return rp, ci, vals


def stiffness2d(mesh, mu=None):
    """stiffnessMatrix2D, src/FEM2D.jl:48-76: Ak[:, k] = vec(VE[k] mu[k] (b b' + c c'))."""
    mu = np.ones(mesh.nt) if mu is None else np.asarray(mu, dtype=np.float64)
    Ak = np.zeros((9, mesh.nt))
    for k in range(mesh.nt):
        nds = mesh.t[:, k]
        b, c = np.zeros(3), np.zeros(3)
        for i in range(3):
            _, b[i], c[i] = abc(mesh.p, nds, int(nds[i]))
        aux = mesh.VE[k] * mu[k] * (np.outer(b, b) + np.outer(c, c))
        Ak[:, k] = aux.ravel(order="F")
    return _assemble2d(mesh, Ak)


def mass2d(mesh):
    """massMatrix2D, src/FEM2D.jl:139-161."""
    Ml = (1.0 / 12.0) * np.array([[2.0, 1, 1], [1, 2, 1], [1, 1, 2]])
    Mk = np.outer(Ml.ravel(order="F"), mesh.VE)
    return _assemble2d(mesh, Mk)


# --------------------------------------------------------------------------------------------------
# P2 (10-node) tetrahedra (src/FEM3D.jl:21-37, 324-395, 494-561, 653-755) -- numpy restatement, test infrastructure only
# --------------------------------------------------------------------------------------------------
def quadratic_basis(p, nds, nd):
    """quadraticBasis, src/FEM3D.jl:21-37: coefficients S[0..9] of f = S0 + S1 x + S2 y + S3 z + S4 x^2 + S5 xy + S6 xz +
    S7 y^2 + S8 yz + S9 z^2 with f(nd) = 1 and f = 0 on the element's nine other nodes (10 x 10 solve)."""
    nodes = [int(nd)] + [int(i) for i in nds if i != nd]
    x, y, z = p[0, nodes], p[1, nodes], p[2, nodes]
    M = np.stack([np.ones(10), x, y, z, x * x, x * y, x * z, y * y, y * z, z * z], axis=1)
    rhs = np.zeros(10)
    rhs[0] = 1.0
    return np.linalg.solve(M, rhs)


def gauss_quadrature_3d_4():
    """GaussQuadrature3D(4), src/FEM3D.jl:711-745: 15 points on the reference tetrahedron, weights summing to 1/6."""
    pts = np.array([
        [0.2500000000000000, 0.2500000000000000, 0.2500000000000000], [0.0000000000000000, 0.3333333333333333, 0.3333333333333333],
        [0.3333333333333333, 0.3333333333333333, 0.3333333333333333], [0.3333333333333333, 0.3333333333333333, 0.0000000000000000],
        [0.3333333333333333, 0.0000000000000000, 0.3333333333333333], [0.7272727272727273, 0.0909090909090909, 0.0909090909090909],
        [0.0909090909090909, 0.0909090909090909, 0.0909090909090909], [0.0909090909090909, 0.0909090909090909, 0.7272727272727273],
        [0.0909090909090909, 0.7272727272727273, 0.0909090909090909], [0.4334498464263357, 0.0665501535736643, 0.0665501535736643],
        [0.0665501535736643, 0.4334498464263357, 0.0665501535736643], [0.0665501535736643, 0.0665501535736643, 0.4334498464263357],
        [0.0665501535736643, 0.4334498464263357, 0.4334498464263357], [0.4334498464263357, 0.0665501535736643, 0.4334498464263357],
        [0.4334498464263357, 0.4334498464263357, 0.0665501535736643]])
    w = np.array([0.1817020685825351] + [0.0361607142857143] * 4 + [0.0698714945161738] * 4 + [0.0656948493683187] * 6) / 6.0
    return w, pts


def quadratic_basis_all(mesh):
    """S[:, i, k] for every node i of every element k (what the reference drivers precompute once)."""
    S = np.zeros((10, 10, mesh.nt))
    for k in range(mesh.nt):
        nds = mesh.t[:, k]
        for i in range(10):
            S[:, i, k] = quadratic_basis(mesh.p, nds, nds[i])
    return S


def pattern_npe(t, nv):
    import scipy.sparse as sp

    npe = t.shape[0]
    r = np.concatenate([t[i] for i in range(npe) for _ in range(npe)]).astype(np.int64)
    c = np.concatenate([t[j] for _ in range(npe) for j in range(npe)]).astype(np.int64)
    A = sp.csr_matrix((np.ones(r.size), (r, c)), shape=(nv, nv))
    A.sort_indices()
    return A.indptr.astype(np.int32), A.indices.astype(np.int32)


def _assemble_npe(mesh, Ak):
    """A = sum over the npe^2 local pairs (i outer, j inner) of sparse(t[i,:], t[j,:], Ak[n,:]) (src/FEM3D.jl:384-391);
    values on the canonical pattern (duplicates summed by scipy in double precision)."""
    import scipy.sparse as sp

    npe = mesh.t.shape[0]
    rp, ci = pattern_npe(mesh.t, mesh.nv)
    A = sp.csr_matrix((np.zeros(len(ci)), ci, rp), shape=(mesh.nv, mesh.nv))
    n = 0
    for i in range(npe):
        for j in range(npe):
            A = A + sp.csr_matrix((Ak[n], (mesh.t[i].astype(np.int64), mesh.t[j].astype(np.int64))), shape=(mesh.nv, mesh.nv))
            n += 1
    P = sp.csr_matrix((np.ones(len(ci)), ci, rp), shape=(mesh.nv, mesh.nv))   # keep structural zeros
    A = (A + 0 * P).tocsr()
    A.sort_indices()
    out = np.zeros(len(ci))
    rows = np.repeat(np.arange(mesh.nv), np.diff(rp))
    key = {(int(r), int(c)): q for q, (r, c) in enumerate(zip(rows, ci))}
    Ar = np.repeat(np.arange(mesh.nv), np.diff(A.indptr))
    for r, c, v in zip(Ar, A.indices, A.data):
        out[key[(int(r), int(c))]] = v
    return rp, ci, out


def quadratic_local_stiffness(mesh, S, mu=None):
    """quadraticLocalStiffnessMatrix, src/FEM3D.jl:324-376 (the "10 node quadrature": mean over the element's nodes)."""
    mu = np.ones(mesh.nt) if mu is None else np.asarray(mu, dtype=np.float64)
    Ak = np.zeros((100, mesh.nt))
    for k in range(mesh.nt):
        nds = mesh.t[:, k]
        x, y, z = mesh.p[0, nds], mesh.p[1, nds], mesh.p[2, nds]
        Sk = S[:, :, k]
        dx = Sk[1][:, None] + 2 * Sk[4][:, None] * x + Sk[5][:, None] * y + Sk[6][:, None] * z     # [basis, node]
        dy = Sk[2][:, None] + 2 * Sk[7][:, None] * y + Sk[5][:, None] * x + Sk[8][:, None] * z
        dz = Sk[3][:, None] + 2 * Sk[9][:, None] * z + Sk[6][:, None] * x + Sk[8][:, None] * y
        temp = (dx @ dx.T + dy @ dy.T + dz @ dz.T) / 10 * mu[k] * mesh.VE[k]
        Ak[:, k] = temp.ravel(order="F")
    return Ak


def quadratic_stiffness(mesh, S=None, mu=None):
    S = quadratic_basis_all(mesh) if S is None else S
    return _assemble_npe(mesh, quadratic_local_stiffness(mesh, S, mu))


def quadratic_mass(mesh, S=None, F=None):
    """quadraticMassMatrix, src/FEM3D.jl:494-561 (weights sum to 1/6 and multiply VE, exactly as the reference does)."""
    S = quadratic_basis_all(mesh) if S is None else S
    F = np.ones(mesh.nt) if F is None else np.asarray(F, dtype=np.float64)
    w, pts = gauss_quadrature_3d_4()
    Ml = np.zeros((100, mesh.nt))
    for k in range(mesh.nt):
        v = mesh.p[:, mesh.t[:4, k]]
        J = (v[:, 1:] - v[:, :1])
        xyz = v[:, :1] + J @ pts.T                       # 3 x 15
        x, y, z = xyz
        Sk = S[:, :, k]
        mono = np.stack([np.ones_like(x), x, y, z, x * x, x * y, x * z, y * y, y * z, z * z])   # 10 x 15
        phi = Sk.T @ mono                                 # [basis, point]
        Mk = (phi * (w * mesh.VE[k])) @ phi.T * F[k]
        Ml[:, k] = Mk.ravel(order="F")
    return _assemble_npe(mesh, Ml)


def quadratic_convection(mesh, u, S=None, F=None, compat_bug=True):
    """quadraticConvectionMatrix, src/FEM3D.jl:563-648, literally: Ck[i, j] += w VE (u_avg . gradPhi_j) phi_i F at the 15 points,
    u_avg = mean of u over the four vertices; Clocal[:, k] = Ck[:] (COLUMN-major) and then sparse(t[i], t[j], Clocal[n]) with
    n = 10 i + j (row-major): the reference assembles the transpose of Ck.  compat_bug = False assembles Ck itself."""
    S = quadratic_basis_all(mesh) if S is None else S
    F = np.ones(mesh.nt) if F is None else np.asarray(F, dtype=np.float64)
    u = np.asarray(u, dtype=np.float64)
    w, pts = gauss_quadrature_3d_4()
    Cl = np.zeros((100, mesh.nt))
    for k in range(mesh.nt):
        v = mesh.p[:, mesh.t[:4, k]]
        J = (v[:, 1:] - v[:, :1])
        x, y, z = v[:, :1] + J @ pts.T
        Sk = S[:, :, k]
        mono = np.stack([np.ones_like(x), x, y, z, x * x, x * y, x * z, y * y, y * z, z * z])
        phi = Sk.T @ mono                                                  # [basis, point]
        gx = Sk[1][:, None] + 2 * Sk[4][:, None] * x + Sk[5][:, None] * y + Sk[6][:, None] * z
        gy = Sk[2][:, None] + 2 * Sk[7][:, None] * y + Sk[5][:, None] * x + Sk[8][:, None] * z
        gz = Sk[3][:, None] + 2 * Sk[9][:, None] * z + Sk[6][:, None] * x + Sk[8][:, None] * y
        ua = u[:, mesh.t[:4, k]].mean(axis=1)
        gj = ua[0] * gx + ua[1] * gy + ua[2] * gz                          # [basis j, point]
        Ck = (phi * (w * mesh.VE[k])) @ gj.T * F[k]                        # Ck[i, j]
        Cl[:, k] = Ck.ravel(order="F") if compat_bug else Ck.ravel(order="C")
    return _assemble_npe(mesh, Cl)


# --------------------------------------------------------------------------------------------------
# Nedelec (lowest-order edge) elements: src/FEM3D.jl:153-271, Magnetostatics/nedelec3D.jl:104-250 -- numpy restatement,
# test infrastructure only.  Julia-only code (no C++ twin, Julia absent here): parity unpinned against Julia itself.
# --------------------------------------------------------------------------------------------------
NED_EDGES = ((0, 1), (1, 2), (2, 0), (3, 0), (2, 3), (1, 3))   # NodesFromLocalEdge before the sort (src/mesh.jl:603-637)


def edge_numbering(t10):
    """mesh.edges / edge2localMap of src/mesh.jl:170-185: unique mid-node labels in order of first appearance (Julia's
    unique over the column-major 6 x nt block); returns (edges, map) with map[mid label] = 0-based edge label."""
    mids = np.asarray(t10[4:10]).reshape(-1, order="F")
    seen, edges = {}, []
    for m in mids.tolist():
        if m not in seen:
            seen[m] = len(edges)
            edges.append(m)
    mp = np.full(max(edges) + 1, -1, dtype=np.int64)
    mp[np.asarray(edges)] = np.arange(len(edges))
    return np.asarray(edges, dtype=np.int64), mp


def _hat_all(mesh):
    """abcd of the four vertices of every element (src/FEM3D.jl:5-17): hat[k, :, i] = (a, b, c, d) of vertex i."""
    t = np.asarray(mesh.t[:4], dtype=np.int64)
    P = np.asarray(mesh.p)[:, t]                                    # 3 x 4 x nt
    M = np.concatenate([np.ones((1, 4, t.shape[1])), P]).transpose(2, 1, 0)   # nt x vertex x (1, x, y, z)
    return np.linalg.solve(M, np.broadcast_to(np.eye(4), M.shape))


def nedelec_curls(mesh):
    """curl N_e = 2 L_e grad(lambda_i) x grad(lambda_j) of the six edges of every element, (i, j) = the edge's vertices
    sorted by global label (src/FEM3D.jl:169-181, src/mesh.jl:630-634): nt x 6 x 3."""
    t = np.asarray(mesh.t[:4], dtype=np.int64)
    nt = t.shape[1]
    hat = _hat_all(mesh)
    p = np.asarray(mesh.p)
    ks = np.arange(nt)
    out = np.zeros((nt, 6, 3))
    for ie, (a, b) in enumerate(NED_EDGES):
        swap = t[a] > t[b]
        i = np.where(swap, b, a)
        j = np.where(swap, a, b)
        L = np.linalg.norm(p[:, t[j, ks]] - p[:, t[i, ks]], axis=0)
        gi = hat[ks, 1:4, i]
        gj = hat[ks, 1:4, j]
        out[:, ie] = (2.0 * L)[:, None] * np.cross(gi, gj)
    return out


def nedelec_local_stiffness(mesh):
    """nedelecLocalStiffness (src/FEM3D.jl:153-194): Ak[6 ie + je, k] = VE[k] curlN_ie . curlN_je."""
    c = nedelec_curls(mesh)
    Ak = np.zeros((36, mesh.nt))
    for ie in range(6):
        for je in range(6):
            Ak[6 * ie + je] = mesh.VE * np.einsum("kd,kd->k", c[:, ie], c[:, je])
    return Ak


def _assemble_dofs(E, ndof, Ak):
    """A = sum over the local pairs (i outer, j inner) of sparse(E[i], E[j], Ak[n], ndof, ndof) on the canonical pattern
    (nedelec3D.jl:166-178, src/FEM3D.jl:257-268)."""
    class _M:
        pass
    m = _M()
    m.t, m.nv = np.asarray(E), int(ndof)
    return _assemble_npe(m, Ak)


def nedelec_stiffness(mesh, nu=None):
    """Global curl-curl matrix of the Newton loop (nedelec3D.jl:163-178): sum of sparse(edge1, edge2, Ak[n, :] .* nu)."""
    edges, mp = edge_numbering(mesh.t)
    E = mp[np.asarray(mesh.t[4:10])]
    Ak = nedelec_local_stiffness(mesh)
    if nu is not None:
        Ak = Ak * np.asarray(nu)[None, :]
    return _assemble_dofs(E, len(edges), Ak)


def nedelec_tangential(mesh, dnu, Bfield, B):
    """nedelecTangentialStiffness (src/FEM3D.jl:197-271): Alocal[i, j] = VE (curlN_i . B)(curlN_j . B) dnu / |B| for j >= i,
    mirrored below the diagonal."""
    edges, mp = edge_numbering(mesh.t)
    E = mp[np.asarray(mesh.t[4:10])]
    c = nedelec_curls(mesh)
    Bf = np.asarray(Bfield).T                                       # nt x 3
    d = np.einsum("ked,kd->ke", c, Bf)
    Ak = np.zeros((36, mesh.nt))
    with np.errstate(divide="ignore", invalid="ignore"):
        for i in range(6):
            for j in range(i, 6):
                v = mesh.VE * d[:, i] * d[:, j] * np.asarray(dnu) / np.asarray(B)
                Ak[6 * i + j] = v
                Ak[6 * j + i] = v
    return _assemble_dofs(E, len(edges), Ak)


def nedelec_load(mesh, Bext):
    """Load vector of nedelec3D.jl:108-146: RHS[e] += VE[k] Bext . curlN, elements outer, local edges inner."""
    edges, mp = edge_numbering(mesh.t)
    E = mp[np.asarray(mesh.t[4:10])]
    c = nedelec_curls(mesh)
    contrib = mesh.VE[:, None] * (c @ np.asarray(Bext, dtype=np.float64))   # nt x 6
    rhs = np.zeros(len(edges))
    np.add.at(rhs, E.T.reshape(-1), contrib.reshape(-1))             # k-major, the loop order of the reference
    return rhs


def nedelec_recover_B(mesh, u):
    """Bfield[:, k] = sum_i u[e_i] curlN_i, B = |Bfield| (nedelec3D.jl:196-236)."""
    edges, mp = edge_numbering(mesh.t)
    E = mp[np.asarray(mesh.t[4:10])]
    c = nedelec_curls(mesh)
    Bf = np.zeros((3, mesh.nt))
    for i in range(6):
        Bf += (np.asarray(u)[E[i]][:, None] * c[:, i]).T
    return Bf, np.sqrt(Bf[0] ** 2 + Bf[1] ** 2 + Bf[2] ** 2)


def nedelec_newton_loop(mesh, Bext, tck_nu, tck_dnu, tol=1e-4, max_att=30, cg_tol=1.4901161193847656e-08, bc="nearest"):
    """Newton-Raphson loop on the magnetic vector potential (nedelec3D.jl:148-250).  The reference solves with
    IterativeSolvers.cg on the singular curl-curl system at its default reltol = sqrt(eps); here scipy's CG at cg_tol (B = curl of
    u is gauge independent).  A much tighter cg_tol than the rounding noise of RHS - A u allows (its null-space component) makes
    CG run to its iteration limit and drift along the gradients near convergence: keep cg_tol >= ~1e-12 / (last div)."""
    import scipy.sparse.linalg as spl

    edges, _ = edge_numbering(mesh.t)
    ne = len(edges)
    RHS = nedelec_load(mesh, Bext)
    nu = np.ones(mesh.nt)
    dnu = np.zeros(mesh.nt)
    u = np.zeros(ne)
    Bf = np.zeros((3, mesh.nt))
    B = np.zeros(mesh.nt)
    ins = np.asarray(mesh.InsideElements)
    att, div, hist = 0, np.inf, []
    while div > tol and att < max_att:
        att += 1
        Bold = B.copy()
        rp, ci, v = nedelec_stiffness(mesh, nu)
        A = _csr(rp, ci, v, ne)
        vt = nedelec_tangential(mesh, dnu, Bf, B)[2] if att > 1 else np.zeros_like(v)
        du, info = spl.cg(_csr(rp, ci, v + vt, ne), RHS - A @ u, rtol=cg_tol, atol=0.0, maxiter=20 * ne)
        residue, alf = line_search(u, du, A, RHS)
        Bf, B = nedelec_recover_B(mesh, u)
        nu[ins] = splev(tck_nu[0], tck_nu[1], 3, B[ins], bc)
        dnu[ins] = splev(tck_dnu[0], tck_dnu[1], 3, B[ins], bc)
        if not (np.all(np.isfinite(nu)) and np.all(np.isfinite(dnu))):
            raise FloatingPointError("Nans in the interpolation")
        div = float(np.max(np.abs(B - Bold)))
        hist.append((att, div, residue, alf))
    return dict(u=u, Bfield=Bf, B=B, nu=nu, dnu=dnu, att=att, div=div, history=hist)
