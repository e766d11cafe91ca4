#!/usr/bin/env python
"""Where does the host-pointer step (fk_picard_pcg) spend its time?  wall clock per call with the set-aside toggling on / off."""
import ctypes as C, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import femeko_jl_b200 as fk
from femeko_jl_b200 import lib as L
MU0 = 4e-7 * np.pi
m = fk.kuhn_cube(119, inside_edge=0.2)
g = np.load(os.path.join(ROOT, "tests", "golden", "gd_270K.npz"))
for mode in (2, 0):
    h = fk.Femeko(m, use_mesh_VE=False, device=0)
    h.tune(15, mode)
    h.set_spline(0, fk.Spline1D(g["HofM"], g["mu"], bc="nearest")); h.set_inside(m.InsideElements); h.set_mu(np.full(m.nt, MU0))
    u = torch.from_numpy(-(np.array([0.5, 0, 0]) / MU0) @ m.p).pin_memory()
    b = torch.randn(h.nv, dtype=torch.float64).pin_memory(); x = torch.zeros(h.nv, dtype=torch.float64).pin_memory()
    o = h.pcg_opts(max_iter=100, tol_disable=True, check_every=100)
    it, rr, dv = C.c_int(), C.c_double(), C.c_double()
    hp = lambda t: C.cast(t.data_ptr(), L.c_double_p)
    def step():
        L.check(h.lib.fk_picard_pcg(h.h, 0, hp(u), 0, hp(b), hp(x), C.byref(o), C.byref(dv), C.byref(it), C.byref(rr)))
    for _ in range(4): step()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(10): step()
    torch.cuda.synchronize(); print(f"mode {mode}: fk_picard_pcg wall {(time.perf_counter() - t0) / 10 * 1e3:.3f} ms per call", flush=True)
    h.close()
