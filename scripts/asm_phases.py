#!/usr/bin/env python
"""Profiling experiment (run under gpurun): kernel-only time of the assembly modes with phase 1 and / or
phase 2 skipped (fk_tune 3), to see which part of the tile pipeline bounds the kernel."""
import json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import femeko_jl_b200 as fk
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from sweep import timed, MU0
n = int(sys.argv[1]) if len(sys.argv) > 1 else 119
opts = dict(a.split("=") for a in sys.argv[2:])
dev = torch.device("cuda", 0); torch.cuda.set_device(0)
stream = torch.cuda.Stream(device=dev)
m = fk.kuhn_cube(n, inside_edge=0.2)
g = np.load(os.path.join(ROOT, "tests", "golden", "gd_270K.npz"))
h = fk.Femeko(m, use_mesh_VE=False, device=0, **{k: int(v) for k, v in opts.items()})
h.set_stream(stream.cuda_stream)
h.set_spline(0, fk.Spline1D(g["HofM"], g["mu"], bc="nearest"))
h.set_spline(fk.lib.FK_MAX_SPLINES // 2, fk.Spline1D(g["HofM"], g["dmu"], bc="nearest"))
h.set_inside(m.InsideElements)
h.set_mu(np.full(m.nt, MU0))
u = torch.from_numpy(-(np.array([0.5, 0, 0]) / MU0) @ m.p).to(dev)
for dbg in [int(x) for x in os.environ.get('FK_DBG', '0,1,2,3').split(',')]:
    h.tune(3, dbg)
    r = {"dbg": dbg, "opts": opts,
         "picard_ms": round(timed(stream, lambda: h.picard_step_dev(u.data_ptr(), slot=0), 10), 4),
         "stiff_ms": round(timed(stream, lambda: h.assemble_stiffness_dev(h.device_ptr(3), slot=1), 10), 4),
         "mass_ms": round(timed(stream, lambda: h.assemble_mass_dev(None, slot=2), 10), 4)}
    print(json.dumps(r), flush=True)
h.tune(3, 0)
h.close()
