#!/usr/bin/env python
"""Multi-GPU parity check (launch with torch.distributed.run, one rank per GPU): every rank holds one
slab of a Kuhn box, assembles its owned rows, and the ranks solve together (halo exchange + all-reduce
inside the PCG); rank 0 compares potential, iteration count and the Picard convergence metric with a
single-GPU solve of the whole box.  Prints one JSON line; exit code 1 on mismatch."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import femeko_jl_b200 as fk  # noqa: E402
from femeko_jl_b200 import partition as P  # noqa: E402
from femeko_jl_b200.mesh import kuhn_box  # noqa: E402

MU0 = 4e-7 * np.pi


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lm, pr = P.kuhn_slab(N, rank, world, inside_edge=0.4)
    h = fk.Femeko(lm, use_mesh_VE=False, device=local, n_owned=pr.n_owned)
    pr.attach(h, dist)
    mode = h.comm_mode()
    g = np.load(os.path.join(ROOT, "tests", "golden", "gd_270K.npz"))
    h.set_spline(0, fk.Spline1D(g["HofM"], g["mu"], bc="nearest"))
    h.set_inside(lm.InsideElements)
    mu = np.full(lm.nt, MU0)
    mu[lm.InsideElements] = 3 * MU0
    h.set_mu(mu)
    h.assemble_stiffness(mu, slot=0, want_vals=False)
    F = np.array([0.5, 0.1, -0.2])
    b = -h.boundary_integral(lm.surfaceT, lm.normal, lm.AE, F, 0)
    rec = int(os.environ.get("FK_RECURRENCE", "0"))
    if os.environ.get("FK_HALO_MODE"):
        h.tune(5, int(os.environ["FK_HALO_MODE"]))
    x, it, rr = h.pcg(b, slot=0, tol=1e-10, max_iter=5000, recurrence=rec)
    xb, itb, _ = h.pcg(b, slot=0, tol=1e-10, max_iter=5000, recurrence=rec)
    bitwise = bool(np.array_equal(x, xb) and it == itb)          # run-to-run determinism across the ranks
    # default options on the singular system: tol = eps is out of reach, every rank must stop at the same global 2n
    xd, itd, _ = h.pcg(b, slot=0, recurrence=rec, project_nullspace=True)
    div = h.picard_step(x * 2e5, slot=1)          # scaled so |H| sits on the nonlinear part of the table
    x2, it2, _ = h.pcg(b, slot=1, tol=1e-10, max_iter=5000, x0=x, recurrence=rec)
    res = torch.zeros((2, pr.nv_global), dtype=torch.float64, device=dev)
    idx = torch.from_numpy(pr.l2g[:pr.n_owned]).to(dev)
    res[0, idx] = torch.from_numpy(x[:pr.n_owned]).to(dev)
    res[1, idx] = torch.from_numpy(x2[:pr.n_owned]).to(dev)
    dist.all_reduce(res)
    ok = True
    if rank == 0:
        gm = kuhn_box((world * N, N, N), inside_edge=0.4)
        hs = fk.Femeko(gm, use_mesh_VE=False, device=local)
        hs.set_spline(0, fk.Spline1D(g["HofM"], g["mu"], bc="nearest"))
        hs.set_inside(gm.InsideElements)
        mug = np.full(gm.nt, MU0)
        mug[gm.InsideElements] = 3 * MU0
        hs.set_mu(mug)
        hs.assemble_stiffness(mug, slot=0, want_vals=False)
        bs = -hs.boundary_integral(gm.surfaceT, gm.normal, gm.AE, F, 0)
        xs, its, _ = hs.pcg(bs, slot=0, tol=1e-10, max_iter=5000)
        divs = hs.picard_step(xs * 2e5, slot=1)
        xs2, its2, _ = hs.pcg(bs, slot=1, tol=1e-10, max_iter=5000, x0=xs)
        r = res.cpu().numpy()
        e1 = float(np.linalg.norm(r[0] - xs) / np.linalg.norm(xs))
        e2 = float(np.linalg.norm(r[1] - xs2) / np.linalg.norm(xs2))
        ed = abs(div - divs) / max(abs(divs), 1e-300)
        ok = e1 < 1e-8 and e2 < 1e-8 and abs(it - its) <= 2 and abs(it2 - its2) <= 2 and ed < 1e-6 and bitwise
        print(json.dumps({"world": world, "N": N, "comm_mode": mode, "recurrence": rec, "bitwise": bitwise, "default_opts_iters": itd,
                          "iters": [it, its], "iters2": [it2, its2], "relerr_x": e1,
                          "relerr_x2": e2, "div": [div, divs], "ok": bool(ok)}), flush=True)
        hs.close()
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, src=0)
    h.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
