#!/usr/bin/env python
"""Profiling driver (run under ncu): a few launches of each assembly kernel at configs[1] size."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import femeko_jl_b200 as fk
MU0 = 4e-7 * np.pi
n = int(sys.argv[1]) if len(sys.argv) > 1 else 119
opts = dict(a.split("=") for a in sys.argv[2:])
m = fk.kuhn_cube(n, inside_edge=0.2)
g = np.load(os.path.join(ROOT, "tests", "golden", "gd_270K.npz"))
h = fk.Femeko(m, use_mesh_VE=False, device=0, **{k: int(v) for k, v in opts.items()})
h.set_spline(0, fk.Spline1D(g["HofM"], g["mu"], bc="nearest"))
h.set_spline(fk.lib.FK_MAX_SPLINES // 2, fk.Spline1D(g["HofM"], g["dmu"], bc="nearest"))
h.set_inside(m.InsideElements)
h.set_mu(np.full(m.nt, MU0))
u = torch.from_numpy(-(np.array([0.5, 0, 0]) / MU0) @ m.p).cuda()
for _ in range(2):
    h.picard_step_dev(u.data_ptr(), slot=0)
    h.assemble_stiffness_dev(h.device_ptr(3), slot=1)
    h.assemble_mass_dev(None, slot=2)
b = torch.randn(h.nv, dtype=torch.float64, device="cuda")
x = torch.zeros(h.nv, dtype=torch.float64, device="cuda")
h.pcg_dev(b.data_ptr(), x.data_ptr(), slot=1, max_iter=4, tol_disable=True, check_every=4)
h.synchronize()
print("ok", h.ntiles, h.tile_stats())
h.close()
