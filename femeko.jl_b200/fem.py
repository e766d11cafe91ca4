"""Host-side mirror of the reference's FEM operator interface for the P1 tetrahedral hot path.

Same names, argument meaning and error behaviour as the Julia functions (src/FEM3D.jl) so that the
drivers read like the reference's; every numeric operation is a call into libfemeko_cuda.so:

    stiffnessMatrix(mesh, f)            src/FEM3D.jl:40-56      -> fk_assemble_stiffness
    localStiffnessMatrix(mesh, f)       src/FEM3D.jl:59-76      -> fk_local_stiffness
    massMatrix(mesh)                    src/FEM3D.jl:466-492    -> fk_assemble_mass
    tangentialStiffnessMatrix(...)      src/FEM3D.jl:79-119     -> fk_assemble_tangential
    lagrange(mesh)                      src/FEM3D.jl:274-281    -> fk_lagrange
    BoundaryIntegral(mesh, F, shell_id) src/FEM3D.jl:284-297    -> fk_boundary_integral
    cg(A, b)                            IterativeSolvers.cg call sites (run.jl:141, LL.jl:185) and
                                        Eigen's Jacobi-PCG (magnetostatics.cpp:26-38) -> fk_pcg
    lineSearch(u, du, A, RHS)           src/FEM3D.jl:123-152    -> fk_spmv
    recoverField(mesh, u, mu)           magneticField.jl:111-132 -> fk_recover_field
    Spline1D(x, y; bc)                  Dierckx.jl (NewtonRaphson.jl:57-63) -> fk_spline_fit/eval

Indices are 0-based (see mesh.py).  Matrices come back as :class:`SparseCSR` on the canonical
pattern (structural node adjacency, ascending columns; CSR == CSC by symmetry).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import lib as L
from .mesh import MESH


class Femeko:
    """Owner of one ``fk_handle``: the mesh resident on the GPU plus its symbolic artefacts."""

    def __init__(self, mesh: MESH, use_mesh_VE: bool = True, device: int | None = None, n_owned: int = 0, **opts):
        self.lib = L.load()
        self.mesh = mesh
        o = L.fk_create_opts()
        nedelec = bool(opts.pop("nedelec", False))
        for k, v in opts.items():
            setattr(o, k, int(v))
        if device is None:
            device = -1
        o.device = int(device)
        o.n_owned = int(n_owned)
        p = L.f64(mesh.p)
        t = L.i32(mesh.t)
        self.dim = 3
        if t.shape[0] == 3:   # 2-D P1 triangles ride the same engine as the elements (n0, n1, n2, n2) (fk_assemble_*2d)
            self.dim = 2
            t = np.vstack([t, t[2:3]])
            if p.shape[0] == 2:
                p = np.vstack([p, np.zeros((1, p.shape[1]))])
        self.edge_dofs = nedelec
        if t.shape[0] == 10 and not self.edge_dofs:  # P2 tetrahedra (fk_assemble_*_p2)
            self.dim, o.nodes_per_element = 3, 10
        assert p.shape[0] == 3 and t.shape[0] in (4, 10), "mesh.p must be 3 x nv and mesh.t 4 x nt (3 x nt triangles, 10 x nt P2)"
        p = np.asfortranarray(p)
        t = np.asfortranarray(t)
        VE = L.f64(mesh.VE) if (use_mesh_VE and mesh.VE is not None) else None
        h = C.c_void_p()
        if self.edge_dofs:    # Nedelec elements on an order-2 mesh: degrees of freedom = edges (Magnetostatics/nedelec3D.jl)
            from .mesh import edgeDofs
            assert t.shape[0] == 10, "edge elements need an order-2 mesh (10 x nt connectivity: the mid nodes label the edges)"
            self.edges6 = edgeDofs(mesh)
            t4 = np.asfortranarray(t[:4])
            L.check(self.lib.fk_create_edge(L.dptr(p), int(p.shape[1]), L.iptr(t4), L.iptr(self.edges6), int(mesh.ne), int(t.shape[1]),
                                            L.dptr(VE), C.byref(o), C.byref(h)))
        else:
            L.check(self.lib.fk_create(L.dptr(p), L.iptr(t), int(p.shape[1]), int(t.shape[1]), L.dptr(VE), C.byref(o), C.byref(h)))
        self.h = h
        nv, nt, nnz, ntl = C.c_int(), C.c_int(), C.c_longlong(), C.c_int()
        L.check(self.lib.fk_sizes(h, C.byref(nv), C.byref(nt), C.byref(nnz), C.byref(ntl)))
        self.nv, self.nt, self.nnz, self.ntiles = nv.value, nt.value, nnz.value, ntl.value
        self.n_rows = n_owned if n_owned > 0 else self.nv
        self._pattern = None

    # -- lifecycle ------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self.lib.fk_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, stream_ptr: int):
        L.check(self.lib.fk_set_stream(self.h, C.c_void_p(stream_ptr)))

    def tune(self, what: int, value: int):
        """Launch-shape tuning (fk_tune): what 0 = asm_cfg, 1 = spmv_variant, 2 = spmv_stages."""
        L.check(self.lib.fk_tune(self.h, int(what), int(value)))

    def comm_mode(self) -> int:
        """0 = single GPU, 1 = NCCL send/recv + all-reduce, 2 = peer memory over NVLink (fk_comm_mode)."""
        m = C.c_int()
        L.check(self.lib.fk_comm_mode(self.h, C.byref(m)))
        return m.value

    def synchronize(self):
        L.check(self.lib.fk_synchronize(self.h))

    def geo_stats(self):
        """(entry pairs, contributions, padded slots, device bytes) of the rescale-assembly plan (built if needed)."""
        v = [C.c_longlong() for _ in range(4)]
        L.check(self.lib.fk_geo_stats(self.h, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)

    def trace_read(self):
        """Timeline of the last solve (after tune(11, n)): array [iteration slot, kernel (0 SpMV, 1 update, 2 direction, 3),
        {start, end, longest cross-GPU wait, wait of CTA 0}] in ns; unused records have start = 2^64 - 1."""
        cap = 1 << 16
        buf = np.zeros(4 * cap, dtype=np.uint64)
        n = C.c_int()
        L.check(self.lib.fk_trace_read(self.h, buf.ctypes.data_as(C.POINTER(C.c_ulonglong)), cap, C.byref(n)))
        return buf[:4 * n.value].reshape(-1, 4, 4)

    def launch_count(self) -> int:
        n = C.c_longlong()
        L.check(self.lib.fk_launch_count(self.h, C.byref(n)))
        return n.value

    # -- symbolic artefacts ---------------------------------------------------------------------
    def pattern(self):
        if self._pattern is None:
            rowptr = np.zeros(self.nv + 1, dtype=np.int32)
            colind = np.zeros(self.nnz, dtype=np.int32)
            L.check(self.lib.fk_csr_pattern(self.h, L.iptr(rowptr), L.iptr(colind)))
            self._pattern = (rowptr, colind)
        return self._pattern

    def elem2nnz(self):
        m = np.zeros(16 * self.nt, dtype=np.int32)
        L.check(self.lib.fk_elem2nnz(self.h, L.iptr(m)))
        return m

    def volumes(self):
        v = np.zeros(self.nt)
        L.check(self.lib.fk_volumes(self.h, L.dptr(v)))
        return v

    def tile_stats(self):
        st, a, b, c = C.c_longlong(), C.c_int(), C.c_int(), C.c_int()
        L.check(self.lib.fk_tile_stats(self.h, C.byref(st), C.byref(a), C.byref(b), C.byref(c)))
        return {"tiles": self.ntiles, "staged_elems": st.value, "recompute_factor": st.value / max(1, self.nt),
                "max_elems": a.value, "max_nnz": b.value, "max_ents": c.value}

    def sell_stats(self):
        """(padded entries, slices, slices with 16-bit columns, bytes of the packed column stream)"""
        v = [C.c_longlong() for _ in range(4)]
        L.check(self.lib.fk_sell_stats(self.h, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)

    def device_ptr(self, which: int) -> int:
        p = C.c_void_p()
        L.check(self.lib.fk_device_ptr(self.h, which, C.byref(p)))
        return p.value

    # -- assembly (host pointers: the Julia call shape) -----------------------------------------
    def _vals(self, want):
        return np.zeros(self.nnz) if want else None

    def assemble_stiffness(self, f=None, slot=0, want_vals=True):
        v = self._vals(want_vals)
        f = self._coef(f)
        L.check(self.lib.fk_assemble_stiffness(self.h, slot, L.dptr(f), L.dptr(v)))
        return v

    def assemble_stiffness2d(self, f=None, slot=0, want_vals=True):
        assert self.dim == 2, "handle was not created from a triangle mesh"
        v = self._vals(want_vals)
        L.check(self.lib.fk_assemble_stiffness2d(self.h, slot, L.dptr(self._coef(f)), L.dptr(v)))
        return v

    def assemble_mass2d(self, rho=None, slot=1, want_vals=True):
        assert self.dim == 2, "handle was not created from a triangle mesh"
        v = self._vals(want_vals)
        L.check(self.lib.fk_assemble_mass2d(self.h, slot, L.dptr(self._coef(rho)), L.dptr(v)))
        return v

    def assemble_stiffness_p2(self, mu=None, slot=0, want_vals=True):
        v = self._vals(want_vals)
        L.check(self.lib.fk_assemble_stiffness_p2(self.h, slot, L.dptr(self._coef(mu)), L.dptr(v)))
        return v

    def assemble_convection_p2(self, u, F=None, slot=2, want_vals=True, compat_bug=True):
        """quadraticConvectionMatrix (src/FEM3D.jl:563-648): u = velocity, 3 x nv; the values are not symmetric.  compat_bug:
        assemble the local matrices transposed, as the reference does (:632-645)."""
        v = self._vals(want_vals)
        u = np.asfortranarray(np.asarray(u, dtype=np.float64))
        assert u.shape == (3, self.nv)
        L.check(self.lib.fk_assemble_convection_p2(self.h, slot, L.dptr(u), L.dptr(self._coef(F)), int(compat_bug), L.dptr(v)))
        return v

    def assemble_mass_p2(self, F=None, slot=1, want_vals=True):
        v = self._vals(want_vals)
        L.check(self.lib.fk_assemble_mass_p2(self.h, slot, L.dptr(self._coef(F)), L.dptr(v)))
        return v

    # -- Nedelec (edge) elements: handles made with Femeko(mesh, nedelec=True) ------------------------------------------
    def nedelec_local_stiffness(self):
        """nedelecLocalStiffness (src/FEM3D.jl:153-194): 36 x nt, Ak[6 ie + je, k]."""
        Ak = np.zeros((36, self.nt), order="F")
        L.check(self.lib.fk_nedelec_local_stiffness(self.h, L.dptr(Ak)))
        return Ak

    def assemble_nedelec_stiffness(self, nu=None, slot=0, want_vals=True):
        v = self._vals(want_vals)
        L.check(self.lib.fk_assemble_nedelec_stiffness(self.h, slot, L.dptr(self._coef(nu)), L.dptr(v)))
        return v

    def assemble_nedelec_tangential(self, dnu, Bfield, B, slot=2, want_vals=True):
        v = self._vals(want_vals)
        Bf = np.asfortranarray(np.asarray(Bfield, dtype=np.float64))
        assert Bf.shape == (3, self.nt)
        L.check(self.lib.fk_assemble_nedelec_tangential(self.h, slot, L.dptr(self._coef(dnu)), L.dptr(Bf), L.dptr(self._coef(B)), L.dptr(v)))
        return v

    def nedelec_load_vector(self, Bext):
        rhs = np.zeros(self.nv)
        L.check(self.lib.fk_nedelec_load_vector(self.h, L.dptr(np.ascontiguousarray(Bext, dtype=np.float64)), L.dptr(rhs)))
        return rhs

    def nedelec_recover_field(self, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        assert u.shape[0] == self.nv
        Bf = np.zeros((3, self.nt), order="F")
        B = np.zeros(self.nt)
        L.check(self.lib.fk_nedelec_recover_field(self.h, L.dptr(u), L.dptr(Bf), L.dptr(B)))
        return Bf, B

    def assemble_mass(self, rho=None, slot=1, want_vals=True):
        v = self._vals(want_vals)
        rho = self._coef(rho)
        L.check(self.lib.fk_assemble_mass(self.h, slot, L.dptr(rho), L.dptr(v)))
        return v

    def assemble_tangential(self, Hfield, dmu, slot=2, want_vals=True):
        v = self._vals(want_vals)
        Hf = np.asfortranarray(L.f64(Hfield))
        assert Hf.shape == (3, self.nt), "H_vec must be 3 x nt"
        L.check(self.lib.fk_assemble_tangential(self.h, slot, L.dptr(Hf), L.dptr(self._coef(dmu)), L.dptr(v)))
        return v

    def _coef(self, f):
        if f is None:
            return None
        f = np.ascontiguousarray(f, dtype=np.float64)
        if f.shape != (self.nt,):
            raise ValueError(f"coefficient vector must have length nt = {self.nt}")
        return f

    def local_stiffness(self, f=None):
        Ak = np.zeros((16, self.nt), order="F")
        L.check(self.lib.fk_local_stiffness(self.h, L.dptr(self._coef(f)), L.dptr(Ak)))
        return Ak

    def matrix_axpby(self, dst, alpha, x, beta=0.0, y=0):
        L.check(self.lib.fk_matrix_axpby(self.h, dst, float(alpha), x, float(beta), y))

    def matrix_get(self, slot):
        v = np.zeros(self.nnz)
        L.check(self.lib.fk_matrix_get(self.h, slot, L.dptr(v)))
        return v

    def matrix_set(self, slot, vals):
        v = np.ascontiguousarray(vals, dtype=np.float64)
        assert v.shape == (self.nnz,)
        L.check(self.lib.fk_matrix_set(self.h, slot, L.dptr(v)))

    # -- vectors --------------------------------------------------------------------------------
    def lagrange(self):
        Cv = np.zeros(self.nv)
        L.check(self.lib.fk_lagrange(self.h, L.dptr(Cv)))
        return Cv

    def boundary_integral(self, surfaceT, normal, AE, F, shell_id):
        sT = np.asfortranarray(L.i32(surfaceT))
        nrm = np.asfortranarray(L.f64(normal))
        ids = np.ascontiguousarray(np.atleast_1d(shell_id), dtype=np.int32)
        Fv = np.ascontiguousarray(F, dtype=np.float64)
        assert Fv.shape == (3,)
        RHS = np.zeros(self.nv)
        AEv = None if AE is None else np.ascontiguousarray(AE, dtype=np.float64)
        L.check(self.lib.fk_boundary_integral(self.h, L.iptr(sT), L.dptr(nrm), L.dptr(AEv), int(sT.shape[1]), L.dptr(Fv),
                                              L.iptr(ids), int(ids.size), L.dptr(RHS)))
        return RHS

    # -- SpMV / PCG -----------------------------------------------------------------------------
    def spmv(self, x, slot=0):
        x = np.ascontiguousarray(x, dtype=np.float64)
        assert x.shape == (self.nv,)
        y = np.zeros(self.nv)
        L.check(self.lib.fk_spmv(self.h, slot, L.dptr(x), L.dptr(y)))
        return y

    @staticmethod
    def pcg_opts(tol=0.0, max_iter=0, warm_start=False, check_every=0, tol_disable=False, no_precond=False, recurrence=0,
                 project_nullspace=False):
        return L.fk_pcg_opts(float(tol), int(max_iter), int(warm_start), int(check_every), int(tol_disable), int(no_precond),
                             int(recurrence), int(project_nullspace))

    # -- constraints ----------------------------------------------------------------------------
    def set_dirichlet(self, fixed, gD=None):
        """Fixed DOFs (0-based) and their values (Fluids/fluids3D.jl:142-162); an empty list clears the set."""
        ids = np.ascontiguousarray(fixed, dtype=np.int32)
        g = None if gD is None else np.ascontiguousarray(np.broadcast_to(np.asarray(gD, dtype=np.float64), ids.shape))
        L.check(self.lib.fk_set_dirichlet(self.h, L.iptr(ids), int(ids.size), L.dptr(g)))

    def apply_dirichlet(self, slot=0, b=None):
        """RHS lift + identity rows / columns on `slot` (fluids3D.jl:164-169); returns the lifted right-hand side."""
        bb = None if b is None else np.array(b, dtype=np.float64, copy=True)
        L.check(self.lib.fk_apply_dirichlet(self.h, slot, L.dptr(bb)))
        return bb

    def set_nullspace(self, w=None):
        """Weights of fk_pcg_opts.project_nullspace; lagrange(mesh) gives the bordered-system solution
        (magneticField.jl:99-107)."""
        wv = None if w is None else np.ascontiguousarray(w, dtype=np.float64)
        L.check(self.lib.fk_set_nullspace(self.h, L.dptr(wv)))

    def heat_step(self, T, slot_mass=1, slot_lhs=3, **kw):
        """T <- (M + dt A)^-1 (M T) (heat3D.jl:86): returns (T, iterations, relative residual)."""
        Tn = np.array(T, dtype=np.float64, copy=True)
        o = self.pcg_opts(**kw)
        it, rr = C.c_int(), C.c_double()
        L.check(self.lib.fk_heat_step(self.h, slot_mass, slot_lhs, L.dptr(Tn), C.byref(o), C.byref(it), C.byref(rr)))
        return Tn, it.value, rr.value

    def heat_step_dev(self, T_ptr: int, slot_mass=1, slot_lhs=3, **kw):
        o = self.pcg_opts(**kw)
        it, rr = C.c_int(), C.c_double()
        L.check(self.lib.fk_heat_step_dev(self.h, slot_mass, slot_lhs, C.c_void_p(T_ptr), C.byref(o), C.byref(it), C.byref(rr)))
        return it.value, rr.value

    def pcg(self, b, slot=0, x0=None, **kw):
        """Jacobi-PCG with Eigen's recurrence; returns (x, iterations, relative residual)."""
        b = np.ascontiguousarray(b, dtype=np.float64)
        assert b.shape[0] >= self.n_rows
        x = np.zeros(self.nv)
        if x0 is not None:
            x[:] = x0
            kw["warm_start"] = True
        o = self.pcg_opts(**kw)
        it, rr = C.c_int(), C.c_double()
        L.check(self.lib.fk_pcg(self.h, slot, L.dptr(b), L.dptr(x), C.byref(o), C.byref(it), C.byref(rr)))
        return x, it.value, rr.value

    def residual(self, x, b, slot=0, want_r=True):
        """r = b - A x and its 2-norm (the `RHS - A*u` of NewtonRaphson.jl:166)."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        assert x.shape[0] >= self.nv and b.shape[0] >= self.n_rows
        r = np.zeros(self.nv) if want_r else None
        nr = C.c_double()
        L.check(self.lib.fk_residual(self.h, slot, L.dptr(x), L.dptr(b), L.dptr(r), C.byref(nr)))
        return (r, nr.value) if want_r else nr.value

    def line_search(self, u, du, rhs, slot=0, min_step=1e-4):
        """lineSearch(u, du, A, RHS) of src/FEM3D.jl:123-152 on the device; updates u in place.
        Returns (residue, accepted step; 0 = u unchanged)."""
        assert u.dtype == np.float64 and u.flags.c_contiguous and u.shape[0] >= self.nv
        du = np.ascontiguousarray(du, dtype=np.float64)
        rhs = np.ascontiguousarray(rhs, dtype=np.float64)
        res, alf = C.c_double(), C.c_double()
        L.check(self.lib.fk_line_search(self.h, slot, L.dptr(u), L.dptr(du), L.dptr(rhs), float(min_step), C.byref(res), C.byref(alf)))
        return res.value, alf.value

    # -- field recovery / nonlinear -------------------------------------------------------------
    def recover_field(self, u, mu=None, compat_bug=False, want_B=True):
        u = np.ascontiguousarray(u, dtype=np.float64)
        assert u.shape[0] >= self.nv
        Hf = np.zeros((3, self.nt), order="F")
        Hn = np.zeros(self.nt)
        B = np.zeros(self.nt) if (mu is not None and want_B) else None
        L.check(self.lib.fk_recover_field(self.h, L.dptr(u), L.dptr(self._coef(mu)), int(compat_bug), L.dptr(Hf), L.dptr(Hn),
                                          L.dptr(B)))
        return Hf, Hn, B

    def set_spline(self, table_id, spl: "Spline1D"):
        L.check(self.lib.fk_set_spline(self.h, table_id, L.dptr(spl.t), L.dptr(spl.c), len(spl.t), spl.k, L.BC[spl.bc]))

    def spline_apply(self, table_id, x):
        """spl(x) evaluated by the device splev (what the Picard pre-pass evaluates per element)."""
        xv = np.ascontiguousarray(x, dtype=np.float64).ravel()
        y = np.zeros(xv.size)
        L.check(self.lib.fk_spline_apply(self.h, int(table_id), L.dptr(xv), int(xv.size), L.dptr(y)))
        return y.reshape(np.shape(x))

    def set_inside(self, ids):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        L.check(self.lib.fk_set_inside(self.h, L.iptr(ids), int(ids.size)))

    def set_material(self, matid):
        m = np.ascontiguousarray(matid, dtype=np.uint8)
        assert m.shape == (self.nt,)
        L.check(self.lib.fk_set_material(self.h, m.ctypes.data_as(L.c_ubyte_p)))

    def set_mu(self, mu):
        L.check(self.lib.fk_set_mu(self.h, L.dptr(self._coef(mu))))

    def get_mu(self, want_dmu=False):
        mu = np.zeros(self.nt)
        dmu = np.zeros(self.nt) if want_dmu else None
        L.check(self.lib.fk_get_mu(self.h, L.dptr(mu), L.dptr(dmu)))
        return (mu, dmu) if want_dmu else mu

    def get_field(self):
        Hf = np.zeros((3, self.nt), order="F")
        Hn = np.zeros(self.nt)
        L.check(self.lib.fk_get_field(self.h, L.dptr(Hf), L.dptr(Hn)))
        return Hf, Hn

    def reset_field(self):
        """|H|_old = 0 (the `Hold = zeros(nt)` a nonlinear solve starts from, NewtonRaphson.jl:139-140)."""
        L.check(self.lib.fk_reset_field(self.h))

    def picard_step(self, u, slot=0, compat_bug=False):
        """Fused sweep: H from u, div = mu0*max|dH|, mu = spl(|H|), stiffness(mu) -> slot.  Returns div.
        Raises FemekoError("Nans/Infs in mu") / out-of-range errors like the reference does."""
        u = np.ascontiguousarray(u, dtype=np.float64)
        assert u.shape[0] >= self.nv
        div = C.c_double()
        L.check(self.lib.fk_picard_step(self.h, slot, L.dptr(u), int(compat_bug), C.byref(div)))
        return div.value

    def picard_pcg(self, u, b, slot=0, x0=None, compat_bug=False, **kw):
        """One nonlinear iteration of the reference loop (NewtonRaphson.jl:148-214): picard_step(u) then pcg(b);
        one C call, the upload of b overlaps the sweep.  Returns (x, div, iterations, relative residual)."""
        u = np.ascontiguousarray(u, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        assert u.shape[0] >= self.nv and b.shape[0] >= self.n_rows
        x = np.zeros(self.nv)
        if x0 is not None:
            x[:] = x0
            kw["warm_start"] = True
        o = self.pcg_opts(**kw)
        div, it, rr = C.c_double(), C.c_int(), C.c_double()
        L.check(self.lib.fk_picard_pcg(self.h, slot, L.dptr(u), int(compat_bug), L.dptr(b), L.dptr(x), C.byref(o), C.byref(div),
                                       C.byref(it), C.byref(rr)))
        return x, div.value, it.value, rr.value

    # -- micromagnetics demagnetising field (LL.cpp:287-348) -------------------------------------
    def demag_setup(self, inside_elements, slot=0):
        """LL constructor work: box-method node volumes + stiffnessMatrix(p, t, VE) resident in `slot`."""
        ids = np.ascontiguousarray(inside_elements, dtype=np.int32)
        vol = np.zeros(self.nv)
        L.check(self.lib.fk_demag_setup(self.h, slot, L.iptr(ids), int(ids.size), L.dptr(vol)))
        self._demag_slot = slot
        return vol

    def demag_field(self, M, nan_outside=True, want_u=False, **kw):
        """One LL::magnetostaticField(): returns (Hd 3 x nv, iterations[, u])."""
        M = np.asfortranarray(M, dtype=np.float64)
        assert M.shape == (3, self.nv)
        Hd = np.zeros((3, self.nv), order="F")
        u = np.zeros(self.nv) if want_u else None
        o = self.pcg_opts(**kw)
        it, rr = C.c_int(), C.c_double()
        L.check(self.lib.fk_demag_field(self.h, self._demag_slot, L.dptr(M), C.byref(o), int(nan_outside), L.dptr(Hd), L.dptr(u),
                                        C.byref(it), C.byref(rr)))
        return (Hd, it.value, u) if want_u else (Hd, it.value)

    def demag_field_dev(self, M_ptr: int, Hd_ptr: int, u_ptr: int | None = None, nan_outside=True, **kw):
        o = self.pcg_opts(**kw)
        it, rr = C.c_int(), C.c_double()
        L.check(self.lib.fk_demag_field_dev(self.h, self._demag_slot, C.c_void_p(M_ptr), C.byref(o), int(nan_outside),
                                            C.c_void_p(Hd_ptr), C.c_void_p(u_ptr), C.byref(it), C.byref(rr)))
        return it.value, rr.value

    # -- device-pointer forms (zero-copy: torch tensors' data_ptr(), or fk_device_ptr arrays) ----
    def assemble_stiffness_dev(self, f_ptr: int | None, slot=0):
        L.check(self.lib.fk_assemble_stiffness_dev(self.h, slot, C.c_void_p(f_ptr)))

    def assemble_mass_dev(self, rho_ptr: int | None = None, slot=1):
        L.check(self.lib.fk_assemble_mass_dev(self.h, slot, C.c_void_p(rho_ptr)))

    def spmv_dev(self, x_ptr: int, y_ptr: int, slot=0):
        L.check(self.lib.fk_spmv_dev(self.h, slot, C.c_void_p(x_ptr), C.c_void_p(y_ptr)))

    def pcg_dev(self, b_ptr: int, x_ptr: int, slot=0, **kw):
        o = self.pcg_opts(**kw)
        it, rr = C.c_int(), C.c_double()
        L.check(self.lib.fk_pcg_dev(self.h, slot, C.c_void_p(b_ptr), C.c_void_p(x_ptr), C.byref(o), C.byref(it), C.byref(rr)))
        return it.value, rr.value

    def recover_field_dev(self, u_ptr: int, mu_ptr: int | None, Hfield_ptr: int | None, Hn_ptr: int | None,
                          B_ptr: int | None, compat_bug=False):
        L.check(self.lib.fk_recover_field_dev(self.h, C.c_void_p(u_ptr), C.c_void_p(mu_ptr), int(compat_bug),
                                              C.c_void_p(Hfield_ptr), C.c_void_p(Hn_ptr), C.c_void_p(B_ptr)))

    def picard_step_dev(self, u_ptr: int, slot=0, compat_bug=False, sync=True):
        """sync=False only enqueues the sweep (no host round trip); fetch div later with picard_result()."""
        div = C.c_double()
        L.check(self.lib.fk_picard_step_dev(self.h, slot, C.c_void_p(u_ptr), int(compat_bug), C.byref(div) if sync else None))
        return div.value if sync else None

    def picard_result(self):
        div = C.c_double()
        L.check(self.lib.fk_picard_result(self.h, C.byref(div)))
        return div.value


class FemekoMulti:
    """ONE process driving `ngpu` GPUs (fk_multi_*): the global mesh goes in, the library partitions it (RCB of the nodes,
    owned / halo rows, ghost elements), builds one handle per GPU wired over peer memory, and every call takes and
    returns global arrays in the mesh's own numbering -- what a single Julia process would @ccall."""

    def __init__(self, mesh: MESH, ngpu: int, use_mesh_VE: bool = True, devices=None, **opts):
        self.lib = L.load()
        self.mesh = mesh
        o = L.fk_create_opts()
        for k, v in opts.items():
            setattr(o, k, int(v))
        p = np.asfortranarray(L.f64(mesh.p))
        t = np.asfortranarray(L.i32(mesh.t))
        VE = L.f64(mesh.VE) if (use_mesh_VE and mesh.VE is not None) else None
        dv = None if devices is None else np.ascontiguousarray(devices, dtype=np.int32)
        h = C.c_void_p()
        L.check(self.lib.fk_multi_create(L.dptr(p), L.iptr(t), int(p.shape[1]), int(t.shape[1]), L.dptr(VE), int(ngpu), L.iptr(dv),
                                         C.byref(o), C.byref(h)))
        self.h, self.ngpu, self.nv, self.nt = h, int(ngpu), int(p.shape[1]), int(t.shape[1])

    def close(self):
        if getattr(self, "h", None) is not None and self.h.value:
            self.lib.fk_multi_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        """per GPU: owned rows, local nodes (owned + halo), local elements (with ghosts), neighbours"""
        n = C.c_int()
        a = [np.zeros(self.ngpu, dtype=np.int32) for _ in range(4)]
        L.check(self.lib.fk_multi_info(self.h, C.byref(n), *[L.iptr(x) for x in a]))
        return dict(n_owned=a[0], nv_local=a[1], nt_local=a[2], nneigh=a[3])

    def partition(self, rank):
        inf = self.info()
        l2g = np.zeros(int(inf["nv_local"][rank]), dtype=np.int32)
        e2g = np.zeros(int(inf["nt_local"][rank]), dtype=np.int32)
        L.check(self.lib.fk_multi_partition(self.h, int(rank), L.iptr(l2g), L.iptr(e2g)))
        return l2g, e2g

    def _elem(self, f):
        if f is None:
            return None
        f = np.ascontiguousarray(f, dtype=np.float64)
        assert f.shape == (self.nt,)
        return f

    def assemble_stiffness(self, f=None, slot=0):
        L.check(self.lib.fk_multi_assemble_stiffness(self.h, slot, L.dptr(self._elem(f))))

    def assemble_mass(self, rho=None, slot=1):
        L.check(self.lib.fk_multi_assemble_mass(self.h, slot, L.dptr(self._elem(rho))))

    def matrix_axpby(self, dst, alpha, x, beta=0.0, y=0):
        L.check(self.lib.fk_multi_matrix_axpby(self.h, dst, float(alpha), x, float(beta), y))

    def set_spline(self, idx, spl):
        L.check(self.lib.fk_multi_set_spline(self.h, int(idx), L.dptr(spl.t), L.dptr(spl.c), len(spl.t), spl.k, L.BC[spl.bc]))

    def set_inside(self, ids):
        ids = np.ascontiguousarray(ids, dtype=np.int32)
        L.check(self.lib.fk_multi_set_inside(self.h, L.iptr(ids), int(ids.size)))

    def set_mu(self, mu):
        L.check(self.lib.fk_multi_set_mu(self.h, L.dptr(self._elem(mu))))

    def get_mu(self):
        mu = np.zeros(self.nt)
        L.check(self.lib.fk_multi_get_mu(self.h, L.dptr(mu)))
        return mu

    def get_field(self):
        Hf = np.zeros((3, self.nt), order="F")
        Hn = np.zeros(self.nt)
        L.check(self.lib.fk_multi_get_field(self.h, L.dptr(Hf), L.dptr(Hn)))
        return Hf, Hn

    def reset_field(self):
        L.check(self.lib.fk_multi_reset_field(self.h))

    def spmv(self, x, slot=0):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.zeros(self.nv)
        L.check(self.lib.fk_multi_spmv(self.h, slot, L.dptr(x), L.dptr(y)))
        return y

    def pcg(self, b, slot=0, x0=None, **kw):
        b = np.ascontiguousarray(b, dtype=np.float64)
        x = np.zeros(self.nv) if x0 is None else np.array(x0, dtype=np.float64, copy=True)
        o = Femeko.pcg_opts(warm_start=x0 is not None, **kw)
        it, rr = C.c_int(), C.c_double()
        L.check(self.lib.fk_multi_pcg(self.h, slot, L.dptr(b), L.dptr(x), C.byref(o), C.byref(it), C.byref(rr)))
        return x, it.value, rr.value

    def picard_step(self, u, slot=0, compat_bug=False):
        u = np.ascontiguousarray(u, dtype=np.float64)
        div = C.c_double()
        L.check(self.lib.fk_multi_picard_step(self.h, slot, L.dptr(u), int(compat_bug), C.byref(div)))
        return div.value


    def demag_setup(self, inside_elements, slot=0):
        """fk_multi_demag_setup: inside_elements = GLOBAL element ids (ascending); returns the box-method node volumes."""
        ids = np.ascontiguousarray(inside_elements, dtype=np.int32)
        vol = np.zeros(self.nv)
        L.check(self.lib.fk_multi_demag_setup(self.h, slot, L.iptr(ids), int(ids.size), L.dptr(vol)))
        self._demag_slot = slot
        return vol

    def demag_field(self, M, nan_outside=True, want_u=False, **kw):
        """One LL::magnetostaticField() over all parts: M, Hd 3 x nv in the global numbering; returns (Hd, iterations[, u])."""
        M = np.asfortranarray(M, dtype=np.float64)
        assert M.shape == (3, self.nv)
        Hd = np.zeros((3, self.nv), order="F")
        u = np.zeros(self.nv) if want_u else None
        o = Femeko.pcg_opts(**kw)
        it, rr = C.c_int(), C.c_double()
        L.check(self.lib.fk_multi_demag_field(self.h, self._demag_slot, L.dptr(M), C.byref(o), int(nan_outside), L.dptr(Hd), L.dptr(u),
                                              C.byref(it), C.byref(rr)))
        return (Hd, it.value, u) if want_u else (Hd, it.value)


class SparseCSR:
    """Assembled matrix on the canonical pattern; keeps its handle + slot so solves stay on the GPU."""

    def __init__(self, fk: Femeko, slot: int, vals: np.ndarray):
        self.fk, self.slot, self.vals = fk, slot, vals
        self.rowptr, self.colind = fk.pattern()
        self.shape = (fk.nv, fk.nv)

    def toscipy(self):
        import scipy.sparse as sp

        return sp.csr_matrix((self.vals, self.colind, self.rowptr), shape=self.shape)

    def _sync(self):
        """Make the device slot hold this matrix (another assembly may have reused the slot)."""
        self.fk.matrix_set(self.slot, self.vals)

    def __matmul__(self, x):
        self._sync()
        return self.fk.spmv(x, self.slot)


def _handle(mesh: MESH) -> Femeko:
    fk = getattr(mesh, "_fk", None)
    if fk is None or not fk.h.value:
        fk = Femeko(mesh)
        mesh._fk = fk
    return fk


def stiffnessMatrix(mesh: MESH, f=None) -> SparseCSR:
    fk = _handle(mesh)
    return SparseCSR(fk, 0, fk.assemble_stiffness(f, slot=0))


def massMatrix(mesh: MESH) -> SparseCSR:
    fk = _handle(mesh)
    return SparseCSR(fk, 1, fk.assemble_mass(None, slot=1))


def stiffnessMatrix2D(mesh: MESH, mu=None) -> SparseCSR:
    """src/FEM2D.jl:48-76 on a triangle mesh (mesh.t is 3 x nt, mesh.VE the triangle areas)."""
    fk = _handle(mesh)
    return SparseCSR(fk, 0, fk.assemble_stiffness2d(mu, slot=0))


def massMatrix2D(mesh: MESH) -> SparseCSR:
    """src/FEM2D.jl:139-161."""
    fk = _handle(mesh)
    return SparseCSR(fk, 1, fk.assemble_mass2d(None, slot=1))


def quadraticStiffnessMatrix(mesh: MESH, S=None, mu=None) -> SparseCSR:
    """src/FEM3D.jl:378-395 on a 10-node mesh.  S (the reference's fitted basis coefficients) is accepted for call
    compatibility and ignored: on straight-sided elements the fits are the Lagrange P2 functions the kernel uses."""
    fk = _handle(mesh)
    return SparseCSR(fk, 0, fk.assemble_stiffness_p2(mu, slot=0))


def quadraticMassMatrix(mesh: MESH, S=None, F=None) -> SparseCSR:
    """src/FEM3D.jl:494-561."""
    fk = _handle(mesh)
    return SparseCSR(fk, 1, fk.assemble_mass_p2(F, slot=1))


def _edge_handle(mesh: MESH) -> Femeko:
    fk = getattr(mesh, "_fk_edge", None)
    if fk is None or not fk.h.value:
        fk = Femeko(mesh, nedelec=True)
        mesh._fk_edge = fk
    return fk


def nedelecLocalStiffness(mesh: MESH) -> np.ndarray:
    """src/FEM3D.jl:153-194 on an order-2 mesh: 36 x nt."""
    return _edge_handle(mesh).nedelec_local_stiffness()


def nedelecStiffnessMatrix(mesh: MESH, nu=None) -> SparseCSR:
    """The global curl-curl matrix the reference's Newton loop builds inline (Magnetostatics/nedelec3D.jl:163-178)."""
    fk = _edge_handle(mesh)
    return SparseCSR(fk, 0, fk.assemble_nedelec_stiffness(nu, slot=0))


def nedelecTangentialStiffness(mesh: MESH, dnu, Bfield, B) -> SparseCSR:
    """src/FEM3D.jl:197-271."""
    fk = _edge_handle(mesh)
    return SparseCSR(fk, 2, fk.assemble_nedelec_tangential(dnu, Bfield, B, slot=2))


def quadraticConvectionMatrix(mesh: MESH, S=None, u=None, F=None, compat_bug: bool = True) -> SparseCSR:
    """src/FEM3D.jl:563-648 (S accepted for call compatibility and ignored, as in quadraticStiffnessMatrix); by default the
    local matrices are assembled transposed, exactly as the reference does."""
    fk = _handle(mesh)
    return SparseCSR(fk, 2, fk.assemble_convection_p2(u, F, slot=2, compat_bug=compat_bug))


def tangentialStiffnessMatrix(mesh: MESH, H_vec, dmu) -> SparseCSR:
    fk = _handle(mesh)
    return SparseCSR(fk, 2, fk.assemble_tangential(H_vec, dmu, slot=2))


def localStiffnessMatrix(mesh: MESH, f=None) -> np.ndarray:
    return _handle(mesh).local_stiffness(f)


def lagrange(mesh: MESH) -> np.ndarray:
    return _handle(mesh).lagrange()


def BoundaryIntegral(mesh: MESH, F, shell_id) -> np.ndarray:
    return _handle(mesh).boundary_integral(mesh.surfaceT, mesh.normal, mesh.AE, F, shell_id)


def cg(A: SparseCSR, b, Pl="jacobi", reltol=None, maxiter=None, x0=None, log=False):
    """Conjugate gradients on the GPU.  ``Pl="jacobi"`` is Eigen's default solver of the C++ path
    (tol = eps, 2n iterations); ``Pl=None`` mirrors IterativeSolvers.cg(A, b) (no preconditioner,
    reltol = sqrt(eps), maxiter = n)."""
    n = A.shape[0]
    if Pl is None:
        reltol = np.sqrt(np.finfo(float).eps) if reltol is None else reltol
        maxiter = n if maxiter is None else maxiter
    A._sync()
    x, it, rr = A.fk.pcg(b, slot=A.slot, x0=x0, tol=reltol or 0.0, max_iter=maxiter or 0, no_precond=Pl is None)
    return (x, it, rr) if log else x


def lineSearch(u, du, A: SparseCSR, RHS, minStep: float = 1e-4):
    """Backtracking of src/FEM3D.jl:123-152 (fk_line_search: trial vectors, products and norms stay on the GPU);
    updates u in place, returns the residual norm."""
    A._sync()
    residue, alf = A.fk.line_search(u, du, RHS, slot=A.slot, min_step=minStep)
    if alf == 0.0:
        print("Warning: N-R could not decrease the residue further")
    return residue


def recoverField(mesh: MESH, u, mu=None, compat_bug: bool = False):
    """H_vectorField (3 x nt), |H| and B = mu.*H as in magneticField.jl:111-132."""
    return _handle(mesh).recover_field(u, mu, compat_bug)


NOALLOC_MAX_ROWS = 1 << 26   # include/femeko_cuda.h: FK_NOALLOC_MAX_ROWS


def residual_norm_noalloc(b_ptr: int, Ax_ptr: int, n: int, r_ptr: int | None = None, norm_dev_ptr: int | None = None,
                          want_host: bool = True, stream: int | None = None):
    """fk_residual_norm_noalloc (no counterpart in the reference): r = b - Ax and ||r|| on DEVICE pointers of the current
    device, with the reduction scratch in the library's static device array -- the call allocates nothing.  Returns the norm
    (synchronises `stream`) when want_host, else None.  n above NOALLOC_MAX_ROWS raises FemekoError."""
    lib = L.load()
    nr = C.c_double()
    L.check(lib.fk_residual_norm_noalloc(C.c_void_p(b_ptr), C.c_void_p(Ax_ptr), int(n), C.c_void_p(r_ptr), C.c_void_p(norm_dev_ptr),
                                         C.byref(nr) if want_host else None, C.c_void_p(stream)))
    return nr.value if want_host else None


class Spline1D:
    """Dierckx.jl ``Spline1D(x, y; k=3, s=0, bc)``: interpolating cubic B-spline with FITPACK's
    knot choice; fields ``t`` (knots), ``c`` (coefficients), ``k`` like the Julia object."""

    def __init__(self, x, y, bc: str = "nearest"):
        if bc not in L.BC:
            raise ValueError("bc must be one of " + ", ".join(L.BC))
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.float64)
        if x.shape != y.shape:
            raise ValueError("x and y must have the same length")
        m = x.size
        self.k, self.bc = 3, bc
        self.t = np.zeros(m + 4)
        self.c = np.zeros(m + 4)
        L.check(L.load().fk_spline_fit(L.dptr(x), L.dptr(y), m, L.dptr(self.t), L.dptr(self.c)))

    def __call__(self, x):
        xv = np.ascontiguousarray(np.atleast_1d(x), dtype=np.float64)
        out = np.zeros(xv.size)
        rc = L.load().fk_spline_eval(L.dptr(self.t), L.dptr(self.c), len(self.t), self.k, L.BC[self.bc], L.dptr(xv.ravel()),
                                     int(xv.size), L.dptr(out))
        L.check(rc)
        out = out.reshape(xv.shape)
        return out if np.ndim(x) else float(out[0])
