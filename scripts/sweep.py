#!/usr/bin/env python
"""Tuning sweep on one B200 (run under gpurun): assembly tile size x launch shape, SpMV variant x
stages x block size.  Prints one line per configuration; nothing here is a bench number."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import femeko_jl_b200 as fk  # noqa: E402

MU0 = 4e-7 * np.pi


def timed(stream, fn, reps):
    fn()
    stream.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        fn()
    e1.record(stream)
    stream.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=119)
    ap.add_argument("--tiles", default="512,1024,2048")
    ap.add_argument("--cfgs", default="0,1,2,3,4")
    ap.add_argument("--blocks", default="1,1024", help="SELL sigma values")
    ap.add_argument("--stages", default="0,4,6", help="SpMV CTAs per SM caps")
    ap.add_argument("--out", default="gpurun_out/sweep.jsonl")
    ap.add_argument("--no-spmv", action="store_true")
    ap.add_argument("--depths", default="2", help="assembly ring depths")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream(device=dev)
    m = fk.kuhn_cube(args.n, inside_edge=0.2)
    g = np.load(os.path.join(ROOT, "tests", "golden", "gd_270K.npz"))
    out = open(os.path.join(ROOT, args.out), "w")
    ref_vals = None

    def emit(d):
        print(json.dumps(d), flush=True)
        out.write(json.dumps(d) + "\n")
        out.flush()

    u = torch.from_numpy(-(np.array([0.5, 0, 0]) / MU0) @ m.p).to(dev)
    first = True
    for tile in [int(x) for x in args.tiles.split(",")]:
        blocks = [int(x) for x in args.blocks.split(",")] if first else [1024]
        for blk in blocks:
            t0 = time.time()
            try:
                h = fk.Femeko(m, use_mesh_VE=False, device=0, tile_max_elems=tile, tile_max_rows=4096, sell_sigma=blk)
            except Exception as e:  # noqa: BLE001
                emit({"tile": tile, "blk": blk, "error": str(e)})
                continue
            setup = time.time() - t0
            h.set_stream(stream.cuda_stream)
            h.set_spline(0, fk.Spline1D(g["HofM"], g["mu"], bc="nearest"))
            h.set_spline(fk.lib.FK_MAX_SPLINES // 2, fk.Spline1D(g["HofM"], g["dmu"], bc="nearest"))
            h.set_inside(m.InsideElements)
            h.set_mu(np.full(m.nt, MU0))
            st = h.tile_stats()
            if blk == blocks[0] or not first:
                for depth, cfg in [(d, c) for d in [int(x) for x in args.depths.split(",")] for c in [int(x) for x in args.cfgs.split(",")]]:
                    try:
                        h.tune(0, cfg)
                        ms_p = timed(stream, lambda: h.picard_step_dev(u.data_ptr(), slot=0), 10)
                        ms_s = timed(stream, lambda: h.assemble_stiffness_dev(h.device_ptr(3), slot=1), 10)
                        ms_m = timed(stream, lambda: h.assemble_mass_dev(None, slot=2), 10)
                        v = h.matrix_get(1)
                        if ref_vals is None:
                            ref_vals = v
                        same = bool(np.array_equal(v, ref_vals))
                        emit({"kind": "asm", "tile": tile, "cfg": cfg, "depth": depth, "picard_ms": round(ms_p, 4), "stiff_ms": round(ms_s, 4),
                              "mass_ms": round(ms_m, 4), "staged_ratio": round(st["staged_elems"] / m.nt, 3), "ntiles": h.ntiles,
                              "bitwise_same_as_first": same, "setup_s": round(setup, 2)})
                    except Exception as e:  # noqa: BLE001
                        emit({"kind": "asm", "tile": tile, "cfg": cfg, "depth": depth, "error": str(e)})
                h.tune(0, 0)
            if first and not args.no_spmv:
                h.assemble_stiffness_dev(h.device_ptr(3), slot=0)
                b = torch.from_numpy(np.random.default_rng(0).standard_normal(h.nv)).to(dev)
                x = torch.zeros(h.nv, dtype=torch.float64, device=dev)
                y = torch.zeros(h.nv, dtype=torch.float64, device=dev)
                y0 = x0 = None
                try:
                    from cuda.bindings import runtime as crt
                    emit({"kind": "l2", "max_persist": int(crt.cudaDeviceGetAttribute(crt.cudaDeviceAttr.cudaDevAttrMaxPersistingL2CacheSize, 0)[1]),
                          "max_window": int(crt.cudaDeviceGetAttribute(crt.cudaDeviceAttr.cudaDevAttrMaxAccessPolicyWindowSize, 0)[1]),
                          "l2": int(crt.cudaDeviceGetAttribute(crt.cudaDeviceAttr.cudaDevAttrL2CacheSize, 0)[1])})
                except Exception as e:  # noqa: BLE001
                    emit({"kind": "l2", "error": str(e)})
                for ctas in [int(v) for v in args.stages.split(",")]:
                    try:
                        h.tune(1, max(ctas, 0))
                        h.tune(2, 0 if ctas < 0 else 1)   # ctas = -1: default grid with the L2 window off
                        ms_spmv = timed(stream, lambda: h.spmv_dev(b.data_ptr(), y.data_ptr(), slot=0), 20)
                        yv = y.cpu().numpy().copy()
                        kw = dict(max_iter=100, tol_disable=True, check_every=100)
                        ms_pcg = timed(stream, lambda: h.pcg_dev(b.data_ptr(), x.data_ptr(), slot=0, **kw), 20)
                        xv = x.cpu().numpy().copy()
                        if y0 is None:
                            y0, x0 = yv, xv
                        emit({"kind": "spmv", "sigma": blk, "ctas_per_sm": ctas, "spmv_us": round(ms_spmv * 1e3, 2),
                              "pcg_iter_us": round(ms_pcg * 10, 2), "spmv_bitwise_same": bool(np.array_equal(yv, y0)),
                              "pcg_x_maxrel": float(np.max(np.abs(xv - x0)) / (np.max(np.abs(x0)) + 1e-300)),
                              "nnz": h.nnz})
                    except Exception as e:  # noqa: BLE001
                        emit({"kind": "spmv", "sigma": blk, "ctas_per_sm": ctas, "error": str(e)})
            h.close()
        first = False
    out.close()


if __name__ == "__main__":
    main()
