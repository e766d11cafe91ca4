"""Nonlinear magnetostatics drivers on the GPU path -- the loops of the reference scripts with the same
names, stopping rules and update order, every numeric step a call into libfemeko_cuda.so:

    picard(...)          Magnetostatics/manyBodyIntersecting.jl:141-210  (fixed point on mu(|H|))
    newton_raphson(...)  Magnetostatics/NewtonRaphson.jl:148-214         (A + At, lineSearch, dmu/dH spline)

The reference solves the bordered system ``[A Lag; Lag' 0] \\ [b; 0]`` with a sparse direct solver; here it is the
Jacobi-PCG with ``project_nullspace`` and ``w = lagrange(mesh)``, which returns the same u (the multiplier only
makes b consistent and fixes the free constant, see include/femeko_cuda.h).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import lib as L
from .fem import Femeko, Spline1D
from .mesh import MESH, mu0


@dataclass
class NonlinearResult:
    u: np.ndarray
    Hfield: np.ndarray
    H: np.ndarray
    mu: np.ndarray
    att: int
    div: float
    history: list = field(default_factory=list)    # (att, div, pcg iterations[, residue, step])
    pcg_iters: int = 0


def _setup(fk: Femeko, mesh: MESH, spl: Spline1D, spl_dmu: Spline1D | None):
    fk.set_spline(0, spl)
    if spl_dmu is not None:
        fk.set_spline(L.FK_MAX_SPLINES // 2, spl_dmu)
    fk.set_inside(mesh.InsideElements)
    fk.set_mu(np.full(mesh.nt, mu0))                 # mu = zeros(nt) .+ mu0
    fk.reset_field()                                 # Hold = zeros(nt)
    Lag = fk.lagrange()
    fk.set_nullspace(Lag)
    return Lag


def picard(mesh: MESH, Bext, spl: Spline1D, picardDeviation: float = 1e-4, maxAtt: int = 100, shell_id=0,
           fk: Femeko | None = None, tol: float = 1e-12, recurrence: int = 0, verbose: bool = False) -> NonlinearResult:
    """manyBodyIntersecting.jl:141-210.  RHS = BoundaryIntegral(mesh, mu0*Hext, shell_id) (`Bext` = mu0*Hext);
    each pass: A = stiffnessMatrix(mesh, mu); u = [A Lag;Lag' 0] \\ [-RHS;0]; H = -grad u; mu[inside] = spl(|H|);
    div = mu0*max|H - Hold| over the inside elements.  On the GPU a pass is one PCG solve plus one fused sweep
    (fk_picard_step: H, div, mu and the stiffness matrix of the next pass, the mesh read once)."""
    fk = fk or Femeko(mesh)
    _setup(fk, mesh, spl, None)
    RHS = fk.boundary_integral(mesh.surfaceT, mesh.normal, mesh.AE, np.asarray(Bext, dtype=np.float64), shell_id)
    b = -RHS
    opts = dict(tol=tol, project_nullspace=True, recurrence=recurrence)
    fk.assemble_stiffness(fk.get_mu(), slot=0, want_vals=False)          # A(mu0); later passes: the sweep's matrix
    u = np.zeros(mesh.nv)
    res = NonlinearResult(u=u, Hfield=None, H=None, mu=None, att=0, div=1.0)
    while res.div > picardDeviation and res.att < maxAtt:
        res.att += 1
        u, it, _ = fk.pcg(b, slot=0, **opts)                              # u = [A Lag;Lag' 0] \ [-RHS;0]
        res.div = fk.picard_step(u, slot=0)                               # H, |H|, mu = spl(|H|), div, and A(mu) of the next pass
        res.pcg_iters += it
        res.history.append((res.att, res.div, it))
        if verbose:
            print(res.att, " | mu0 |H(n)-H(n-1)| = ", res.div)
    res.u = u
    res.Hfield, res.H = fk.get_field()
    res.mu = fk.get_mu()
    return res


def newton_raphson(mesh: MESH, Bext, spl: Spline1D, spl_dmu: Spline1D, maxDeviation: float = 1e-4, maxAtt: int = 100,
                   shell_id=0, fk: Femeko | None = None, tol: float = 1e-12, compat_bug: bool = True, recurrence: int = 0,
                   verbose: bool = False) -> NonlinearResult:
    """NewtonRaphson.jl:148-214: RHS = -BoundaryIntegral(mesh, Bext, shell_id); each pass A = A(mu), At = tangential
    (from the second pass on), du = [A+At Lag;Lag' 0] \\ [RHS - A*u; 0], lineSearch(u, du, A, RHS), H recovery
    (compat_bug reproduces the row-2 defect of :187-189), div, mu = spl(H), dmu = spl_dmu(H)."""
    fk = fk or Femeko(mesh)
    _setup(fk, mesh, spl, spl_dmu)
    RHS = -fk.boundary_integral(mesh.surfaceT, mesh.normal, mesh.AE, np.asarray(Bext, dtype=np.float64), shell_id)
    u = np.zeros(mesh.nv)
    res = NonlinearResult(u=u, Hfield=None, H=None, mu=None, att=0, div=np.inf)
    fk.assemble_stiffness(fk.get_mu(), slot=0, want_vals=False)          # A(mu0); later passes: the sweep's matrix
    while res.div > maxDeviation and res.att < maxAtt:
        res.att += 1
        slot = 0
        if res.att > 1:                                                   # At from the stored Hfield / dmu of the last sweep
            Hf, _ = fk.get_field()
            _, dmu = fk.get_mu(want_dmu=True)
            fk.assemble_tangential(Hf, dmu, slot=2, want_vals=False)
            fk.matrix_axpby(3, 1.0, 0, 1.0, 2)                           # A + At
            slot = 3
        r, _ = fk.residual(u, RHS, slot=0)                                # RHS - A*u
        du, it, _ = fk.pcg(r, slot=slot, tol=tol, project_nullspace=True, recurrence=recurrence)
        residue, step = fk.line_search(u, du, RHS, slot=0)
        res.div = fk.picard_step(u, slot=0, compat_bug=compat_bug)        # H, div, mu, dmu, A(mu) for the next pass
        res.pcg_iters += it
        res.history.append((res.att, res.div, it, residue, step))
        if verbose:
            print(res.att, " | mu0 |H(n)-H(n-1)| = ", res.div, " , |y-Ax| = ", residue)
    res.Hfield, res.H = fk.get_field()
    res.mu = fk.get_mu()
    return res


@dataclass
class NedelecResult:
    u: np.ndarray            # magnetic vector potential on the edges
    Bfield: np.ndarray       # 3 x nt
    B: np.ndarray            # nt
    nu: np.ndarray           # relative reluctance
    dnu: np.ndarray
    att: int
    div: float
    history: list = field(default_factory=list)
    pcg_iters: int = 0


def nedelec3D(mesh: MESH, Bext, spl_nu: Spline1D, spl_dnu: Spline1D, tol: float = 1e-4, maxAtt: int = 30, fk: Femeko | None = None,
              cg_tol: float = 1.4901161193847656e-08, cg_maxiter: int | None = None, verbose: bool = False) -> NedelecResult:
    """Newton-Raphson on the magnetic vector potential with Nedelec elements (Magnetostatics/nedelec3D.jl:104-250): load
    vector VE Bext . curlN (:108-146); each pass A = curl-curl matrix scaled by nu (:163-178), At = nedelecTangentialStiffness
    from the second pass on (:181-185), du = cg(A + At, RHS - A u) (:188), lineSearch (:191), B = curl u and |B| (:194-236),
    nu = spl(B), dnu = spl_dnu(B) on the inside elements (:239-240), div = max |B - Bold| (:248).  The matrices, the solve, the
    line search and the field recovery run on the GPU; the two spline evaluations over the inside elements are fk_spline_apply.
    The curl-curl system is singular (gradients); like the reference's cg call the solve is a plain Jacobi-CG on the
    consistent system at IterativeSolvers' default reltol = sqrt(eps) -- B does not depend on the gauge."""
    fk = fk or Femeko(mesh, nedelec=True)
    fk.set_spline(0, spl_nu)
    fk.set_spline(1, spl_dnu)
    ins = np.asarray(mesh.InsideElements, dtype=np.int64)
    RHS = fk.nedelec_load_vector(np.asarray(Bext, dtype=np.float64))
    nu = np.ones(mesh.nt)
    dnu = np.zeros(mesh.nt)
    u = np.zeros(fk.nv)
    Bf = np.zeros((3, mesh.nt), order="F")
    B = np.zeros(mesh.nt)
    res = NedelecResult(u=u, Bfield=Bf, B=B, nu=nu, dnu=dnu, att=0, div=np.inf)
    while res.div > tol and res.att < maxAtt:
        res.att += 1
        Bold = B.copy()
        fk.assemble_nedelec_stiffness(nu, slot=0, want_vals=False)
        slot = 0
        if res.att > 1:
            fk.assemble_nedelec_tangential(dnu, Bf, B, slot=2, want_vals=False)
            fk.matrix_axpby(3, 1.0, 0, 1.0, 2)                           # A + At
            slot = 3
        r, _ = fk.residual(u, RHS, slot=0)                                # RHS - A*u
        du, it, _ = fk.pcg(r, slot=slot, tol=cg_tol, max_iter=cg_maxiter or 0)
        residue, step = fk.line_search(u, du, RHS, slot=0)
        Bf, B = fk.nedelec_recover_field(u)
        if ins.size:
            nu[ins] = fk.spline_apply(0, B[ins])
            dnu[ins] = fk.spline_apply(1, B[ins])
        if not (np.all(np.isfinite(nu)) and np.all(np.isfinite(dnu))):
            raise FloatingPointError("Nans in the interpolation")          # nedelec3D.jl:243-245
        res.div = float(np.max(np.abs(B - Bold)))
        res.pcg_iters += it
        res.history.append((res.att, res.div, it, residue, step))
        if verbose:
            print(res.att, " , |B(n)-B(n-1)| = ", res.div, " , |y-Ax| = ", residue)
    res.Bfield, res.B = Bf, B
    return res
