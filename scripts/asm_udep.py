#!/usr/bin/env python
"""Profiling experiment: does the Picard sweep's time depend on the values of the potential?"""
import json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import femeko_jl_b200 as fk
from sweep import timed, MU0
dev = torch.device("cuda", 0); torch.cuda.set_device(0)
stream = torch.cuda.Stream(device=dev)
m = fk.kuhn_cube(119, inside_edge=0.2)
g = np.load(os.path.join(ROOT, "tests", "golden", "gd_270K.npz"))
h = fk.Femeko(m, use_mesh_VE=False, device=0)
h.set_stream(stream.cuda_stream)
h.set_spline(0, fk.Spline1D(g["HofM"], g["mu"], bc="nearest"))
h.set_inside(m.InsideElements)
h.set_mu(np.full(m.nt, MU0))
rng = np.random.default_rng(0)
lin = -(np.array([0.5, 0, 0]) / MU0) @ m.p
cases = {"linear": lin, "zeros": np.zeros(m.nv), "random": rng.standard_normal(m.nv),
         "sparse": np.where(rng.random(m.nv) < 0.02, 1.0, 0.0), "tiny": 1e-160 * rng.standard_normal(m.nv),
         "denorm": 1e-310 * rng.standard_normal(m.nv)}
for name, u in cases.items():
    ud = torch.from_numpy(np.ascontiguousarray(u)).to(dev)
    print(json.dumps({"u": name, "picard_ms": round(timed(stream, lambda: h.picard_step_dev(ud.data_ptr(), slot=0), 10), 4)}), flush=True)
h.close()
