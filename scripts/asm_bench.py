#!/usr/bin/env python
"""Assembly micro-benchmark (B200): CUDA-event times of the Picard sweep and of a plain stiffness assembly through the
tile kernel and through the rescale path (every launch shape of fk_tune 12), at one Kuhn size.  One JSON line per case.
    python scripts/asm_bench.py [N=119] [reps=20] [short]       short: two launches per path only (for ncu)"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import femeko_jl_b200 as fk  # noqa: E402

MU0 = 4e-7 * np.pi
n = int(sys.argv[1]) if len(sys.argv) > 1 else 119
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
short = len(sys.argv) > 3
m = fk.kuhn_cube(n, inside_edge=0.2)
g = np.load(os.path.join(ROOT, "tests", "golden", "gd_270K.npz"))
h = fk.Femeko(m, use_mesh_VE=False, device=0)
h.set_spline(0, fk.Spline1D(g["HofM"], g["mu"], bc="nearest"))
h.set_spline(fk.lib.FK_MAX_SPLINES // 2, fk.Spline1D(g["HofM"], g["dmu"], bc="nearest"))
h.set_inside(m.InsideElements)
h.set_mu(np.full(m.nt, MU0))
u = torch.from_numpy(-(np.array([0.5, 0, 0]) / MU0) @ m.p).cuda()
big = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")     # L2 flush between timed launches
nt, nv, nnz = h.nt, h.nv, h.nnz
picard_bytes = (16 + 8 + 64 + 24 + 8 + 8) * nt + (24 + 8) * nv + 8 * nnz
stiff_bytes = 88 * nt + 24 * nv + 8 * nnz


rnd = torch.randint(0, 255, (256 << 20,), dtype=torch.uint8, device="cuda") if os.environ.get("FK_FLUSH", "zero") == "random" else None


def timed(fn, reps):
    ts = []
    for _ in range(reps):
        if rnd is not None:
            big.copy_(rnd)      # flush with ordinary data: L2 is left full of dirty non-zero lines
        else:
            big.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def report(name, t, nbytes):
    med, best = t
    print(json.dumps({"case": name, "N": n, "ms_median": round(med, 4), "ms_best": round(best, 4),
                      "algorithmic_GBps": round(nbytes / med / 1e6, 1), "frac_of_6550": round(nbytes / med / 1e6 / 6550.7, 3),
                      "gtets_per_s": round(nt / med / 1e6, 2)}), flush=True)


sweep = lambda: h.picard_step_dev(u.data_ptr(), slot=0, sync=False)   # noqa: E731
stiff = lambda: h.assemble_stiffness_dev(h.device_ptr(3), slot=1)      # noqa: E731
r = 2 if short else reps
h.tune(10, 0)
sweep(); stiff(); h.synchronize()
report("picard_sweep_tile", timed(sweep, r), picard_bytes)
report("stiffness_tile", timed(stiff, r), stiff_bytes)
h.tune(10, 2)
sweep(); stiff(); h.synchronize()
print(json.dumps({"geo_stats(items, contributions, slots, bytes)": h.geo_stats()}), flush=True)
for win in ([4096] if short else [int(w) for w in os.environ.get('FK_WINDOWS', '4096,16384,65536,262144,1048576').split(',')]):
    h.tune(13, win)
    for v in [0]:
        h.tune(12, v)
        sweep(); stiff(); h.synchronize()
        gs = h.geo_stats()
        report(f"picard_sweep_rescale_v{v}_win{win}_pad{gs[2] / gs[1]:.3f}", timed(sweep, r), picard_bytes)
        report(f"stiffness_rescale_v{v}_win{win}", timed(stiff, r), stiff_bytes)
# the sweep as the bench step sees it: right after a PCG solve (its vectors pinned in L2) instead of after an L2 flush
b = torch.randn(h.nv, dtype=torch.float64, device="cuda")
x = torch.zeros(h.nv, dtype=torch.float64, device="cuda")
for persist in ([1] if short else [1, 0]):
    h.tune(2, persist)
    ts = []
    for _ in range(r):
        h.pcg_dev(b.data_ptr(), x.data_ptr(), slot=1, max_iter=20, tol_disable=True, check_every=20)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sweep()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    report(f"picard_sweep_rescale_after_pcg_l2persist{persist}", (ts[len(ts) // 2], ts[0]), picard_bytes)
# the two streaming kernels of the sweep side by side (fk_tune 17 = 16 a + b: sums on a CTAs per SM, field recovery on b), timed
# behind a PCG solve like the bench step
h.tune(2, 1)
h.tune(12, int(os.environ.get('FK_GEO_VARIANT', '0')))     # launch shape of the streamed sums for the split sweep
for split in ([0x32] if short else [int(v, 16) for v in os.environ.get('FK_SPLITS', '0,32,41,22,31,42,23,33').split(',')]):
    h.tune(17, split)
    sweep(); h.synchronize()
    ts = []
    for _ in range(r):
        h.pcg_dev(b.data_ptr(), x.data_ptr(), slot=1, max_iter=20, tol_disable=True, check_every=20)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sweep()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    report(f"picard_sweep_rescale_after_pcg_split{split >> 4}_{split & 15}", (ts[len(ts) // 2], ts[0]), picard_bytes)
h.tune(17, 0x41)
# critical path of the side-by-side sweep: leave one kernel out at a time (fk_tune 19; results are garbage)
if not short:
    for dbg, nm in ((0, "all"), (1, "no_field_recovery"), (2, "no_rowsum"), (4, "no_sums"), (3, "sums_only"), (6, "field_recovery_only")):
        h.tune(19, dbg)
        sweep(); h.synchronize()
        ts = []
        for _ in range(r):
            h.pcg_dev(b.data_ptr(), x.data_ptr(), slot=1, max_iter=20, tol_disable=True, check_every=20)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); sweep(); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        report(f"sweep_4_1_{nm}", (ts[len(ts) // 2], ts[0]), picard_bytes)
    h.tune(19, 0)
h.picard_result()
h.close()
