#!/usr/bin/env python
"""SpMV launch-shape sweep on one B200 (run under gpurun): variants of fk_tune(4, .) x wide / packed columns.
Prints one line per configuration; nothing here is a bench number."""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import femeko_jl_b200 as fk  # noqa: E402
from sweep import timed, MU0  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=119)
    ap.add_argument("--variants", default="0,1,2,3,4,5")
    ap.add_argument("--ctas", default="0")
    ap.add_argument("--wide", default="0,1")
    ap.add_argument("--out", default="gpurun_out/sweep_spmv.jsonl")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream(device=dev)
    m = fk.kuhn_cube(args.n, inside_edge=0.2)
    out = open(os.path.join(ROOT, args.out), "w")

    def emit(d):
        print(json.dumps(d), flush=True)
        out.write(json.dumps(d) + "\n")
        out.flush()

    mu = np.full(m.nt, MU0)
    mu[m.InsideElements] = 3 * MU0
    y0 = x0 = None
    for wide in [int(v) for v in args.wide.split(",")]:
        h = fk.Femeko(m, use_mesh_VE=False, device=0, sell_wide_cols=wide)
        h.set_stream(stream.cuda_stream)
        h.assemble_stiffness(mu, want_vals=False)
        b = torch.from_numpy(np.random.default_rng(0).standard_normal(h.nv)).to(dev)
        x = torch.zeros(h.nv, dtype=torch.float64, device=dev)
        y = torch.zeros(h.nv, dtype=torch.float64, device=dev)
        for var in [int(v) for v in args.variants.split(",")]:
            for ctas in [int(v) for v in args.ctas.split(",")]:
                try:
                    h.tune(4, var)
                    h.tune(1, ctas)
                    ms_spmv = timed(stream, lambda: h.spmv_dev(b.data_ptr(), y.data_ptr(), slot=0), 50)
                    yv = y.cpu().numpy().copy()
                    kw = dict(max_iter=100, tol_disable=True, check_every=100)
                    ms_pcg = timed(stream, lambda: h.pcg_dev(b.data_ptr(), x.data_ptr(), slot=0, **kw), 20)
                    xv = x.cpu().numpy().copy()
                    if y0 is None:
                        y0, x0 = yv, xv
                    emit({"kind": "spmv", "wide": wide, "variant": var, "ctas_per_sm": ctas, "spmv_us": round(ms_spmv * 1e3, 2),
                          "pcg_iter_us": round(ms_pcg * 10, 2), "spmv_bitwise_same": bool(np.array_equal(yv, y0)),
                          "pcg_bitwise_same": bool(np.array_equal(xv, x0)), "sell": h.sell_stats()})
                except Exception as e:  # noqa: BLE001
                    emit({"kind": "spmv", "wide": wide, "variant": var, "ctas_per_sm": ctas, "error": str(e)})
        h.close()
    out.close()


if __name__ == "__main__":
    main()
