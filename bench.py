#!/usr/bin/env python
"""bench.py -- the hot path of BASELINE.json on B200: one nonlinear Picard step of 3-D P1
magnetostatics = fused [H recovery + mu(|H|) spline + stiffness assembly] sweep, then a Jacobi-PCG
solve of a fixed iteration count on the freshly assembled matrix.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (C ABI)
    python bench.py --impl reference --steps K --warmup W     # the reference's CPU path, same metric

Workload at N = 1: BASELINE.json configs[1] -- Kuhn cube N = 119 (10 110 954 tets, 1 728 000 nodes),
Gd mu(H) table at 270 K (NewtonRaphson.jl:41-63), B_ext = (0.5, 0, 0) T.  At N > 1 every rank holds
one N = 119 slab of a box that is `world` cubes long (weak scaling): owned + halo rows, ghost
elements, NCCL halo exchange + all-reduce inside the PCG.

Numbers printed (one JSON line, rank 0):
  value   Mtets/s through one step with every input resident in HBM (device-pointer entry points)
  e2e     the same through the host-pointer C ABI a Julia driver would @ccall: potential u and
          right-hand side b copied host->device, solution x and div copied back, every step
  roofline            the fused PCG iteration (the step's dominant kernel group) vs measured HBM peak
  roofline_assembly   the fused Picard assembly kernel vs the same peak
  cpu_baseline        the compiled reference C++ (oracle/_ref) on a bounded sample, host cores
Only the cpu_baseline / --impl reference legs touch oracle/; the GPU legs call libfemeko_cuda.so.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = json.load(open(os.path.join(ROOT, "BASELINE.json")))["metric"]
UNIT = "Mtets/s"   # tets swept by one Picard step (fused assembly + P PCG iterations) per second
MU0 = 4e-7 * np.pi


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def mesh_module():
    """femeko.jl_b200/mesh.py (pure numpy generators) loaded on its own: the CPU arm must not import the product
    package, so that the only native code it maps is oracle/."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("_fk_mesh_standalone", os.path.join(ROOT, "femeko.jl_b200", "mesh.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    spec.loader.exec_module(mod)
    return mod


def gd_table():
    g = np.load(os.path.join(ROOT, "tests", "golden", "gd_270K.npz"))
    return g["HofM"], g["mu"], g["dmu"]


def pcg_bytes(n, nnz):
    """Algorithmic bytes of one Jacobi-PCG iteration: 12 nnz + 108 n + 4."""
    return 12 * nnz + 108 * n + 4


def picard_bytes(nt, nv, nnz):
    """Algorithmic bytes of the fused Picard sweep: t 16 + mu 8 + map 64 per tet,
    compulsory p 24 nv and u 8 nv, CSR values 8 nnz, writes Hfield 24 + |H| 8 + mu 8 per tet."""
    return (16 + 8 + 64 + 24 + 8 + 8) * nt + (24 + 8) * nv + 8 * nnz


# --------------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md "clocks DURING the timed region")
# --------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                smax.append(float(c[2]))
                power.append(float(c[3]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(smax), reasons=sorted(reasons), samples=len(sm),
                       power_w_max=max(power))
        return out


# --------------------------------------------------------------------------------------------------
# the problem (one rank's share)
# --------------------------------------------------------------------------------------------------
def build_problem(N, rank, world):
    """Mesh partition of this rank + boundary load.  world == 1: the whole Kuhn cube."""
    import femeko_jl_b200 as fk

    if world == 1:
        m = fk.kuhn_cube(N, inside_edge=0.2)
        return m, None
    from femeko_jl_b200 import partition as P

    return P.kuhn_slab(N, rank, world, inside_edge=0.2)


def kernel_split(tr):
    """Per-kernel device time inside the replayed PCG graph from the %globaltimer records of fk_trace_read:
    span = latest CTA end - earliest CTA start of a launch; gap = start of the next kernel - end of this one."""
    names = {0: "spmv", 1: "update", 2: "direction"}
    tr = tr.astype(np.int64)
    used = tr[:, :, 0] >= 0          # unused records carry start = 2^64 - 1 (-1 as int64)
    out = {}
    for k, nm in names.items():
        sel = used[:, k] & (tr[:, k, 1] > 0)
        if sel.sum() < 4:
            continue
        out[nm + "_us"] = round(float(np.median(tr[sel, k, 1] - tr[sel, k, 0])) / 1e3, 3)
        out[nm + "_wait_us"] = round(float(np.median(tr[sel, k, 2])) / 1e3, 3)
    # fused halo exchange (record 3 of the SpMV): when the flags left and when the first / last warp began to wait for the peers'
    sel = used[:, 0] & (tr[:, 3, 0] >= 0) & (tr[:, 3, 1] > 0)
    if sel.sum() >= 4:
        t0 = tr[sel, 0, 0]
        out["halo_release_at_us"] = round(float(np.median(tr[sel, 3, 0] - t0)) / 1e3, 3)
        out["halo_first_wait_at_us"] = round(float(np.median((~tr[sel, 3, 2]) - t0)) / 1e3, 3)
        out["halo_last_wait_at_us"] = round(float(np.median(tr[sel, 3, 1] - t0)) / 1e3, 3)
    ks = [k for k in names if names[k] + "_us" in out]
    its = np.where(np.all(used[:, ks] & (tr[:, ks, 1] > 0), axis=1))[0]
    if len(its) > 4 and len(ks) >= 2:
        gaps = []
        for a, b in zip(ks[:-1], ks[1:]):
            gaps.append(tr[its, b, 0] - tr[its, a, 1])
        out["gap_us"] = round(float(np.median(np.sum(gaps, axis=0))) / 1e3, 3)
        starts = np.sort(tr[its, ks[0], 0])
        d = np.diff(starts)
        out["iteration_us"] = round(float(np.median(d[d < 10 * np.median(d)])) / 1e3, 3) if len(d) else None
    return out


class Ctx:
    """Process-wide state of one rank."""

    def __init__(self, args):
        import torch

        self.rank, self.world, self.local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
        if self.world != args.gpus and self.world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device -- this framework has no CPU path (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = None
        # N > 1: run this rank (and first-touch its pinned host buffers) on the cores next to its GPU -- eight ranks pushing
        # their host copies through one socket is what the e2e leg lost most to (FK_BENCH_AFFINITY=0 leaves the affinity alone)
        self.host_affinity = None
        if self.world > 1 and os.environ.get("FK_BENCH_AFFINITY", "1") != "0":
            try:
                import pynvml

                pynvml.nvmlInit()
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                idx = int(vis.split(",")[self.local]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else self.local
                mask = pynvml.nvmlDeviceGetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(idx), ((os.cpu_count() or 64) + 63) // 64)
                cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
                allowed = os.sched_getaffinity(0)
                near = cpus & allowed
                if near and near != allowed:
                    os.sched_setaffinity(0, near)
                    self.host_affinity = f"{len(near)} of {len(allowed)} cores (those next to GPU {idx})"
                else:
                    self.host_affinity = f"unchanged ({len(allowed)} cores allowed, {len(near)} of them next to GPU {idx})"
            except Exception as e:  # noqa: BLE001 -- affinity is an optimisation of the harness, never a requirement
                self.host_affinity = f"unchanged ({type(e).__name__})"
        if self.world > 1:
            import torch.distributed as dist

            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        self.stream = torch.cuda.Stream(device=self.dev)
        # torch.distributed.run pins OMP_NUM_THREADS = 1: give every rank its share of the host cores for the symbolic phase
        from femeko_jl_b200 import lib as L

        share = max(1, (os.cpu_count() or 1) // max(1, self.world))
        self.host_threads = L.load().fk_set_host_threads(max(1, min(share, len(os.sched_getaffinity(0)))))

    def barrier(self):
        import torch

        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, v):
        import torch

        if self.dist is None:
            return float(v)
        tt = torch.tensor([v], dtype=torch.float64, device=self.dev)
        self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX)
        return float(tt.item())

    def sum_over_ranks(self, vals):
        import torch

        if self.dist is None:
            return [int(v) for v in vals]
        tt = torch.tensor(list(vals), dtype=torch.float64, device=self.dev)
        self.dist.all_reduce(tt)
        return [int(v) for v in tt.tolist()]


def bench_problem(args, ctx, mesh, part, full):
    """Set one problem up on this rank's GPU and time the step on it.  full: also the end-to-end legs through the
    host-pointer C ABI and the untimed nonlinear solve (the headline workload); otherwise device-resident timing only
    (the strong-scaling block).  Returns a dict of raw figures (max over ranks where it is a time)."""
    import ctypes as C

    import torch

    import femeko_jl_b200 as fk
    from femeko_jl_b200 import lib as L

    rank, world, local, dist, dev, stream = ctx.rank, ctx.world, ctx.local, ctx.dist, ctx.dev, ctx.stream
    Piters = args.pcg_iters
    t_setup = time.time()
    n_owned = part.n_owned if part is not None else 0
    h = fk.Femeko(mesh, use_mesh_VE=False, device=local, n_owned=n_owned)
    for kv in [t for t in args.tune.split(",") if t]:
        h.tune(int(kv.split("=")[0]), int(kv.split("=")[1]))
    h.set_stream(stream.cuda_stream)
    if part is not None:
        part.attach(h, dist)
    Hk, muk, dmuk = gd_table()
    h.set_spline(0, fk.Spline1D(Hk, muk, bc="nearest"))
    h.set_spline(fk.lib.FK_MAX_SPLINES // 2, fk.Spline1D(Hk, dmuk, bc="nearest"))
    h.set_inside(mesh.InsideElements)
    h.set_mu(np.full(mesh.nt, MU0))
    Bext = np.array([0.5, 0.0, 0.0])                                  # NewtonRaphson.jl:32-34
    rhs = -h.boundary_integral(mesh.surfaceT, mesh.normal, mesh.AE, Bext, 0)   # RHS = -BoundaryIntegral, :109
    n_rows, nv, nt, nnz = h.n_rows, h.nv, h.nt, h.nnz
    nt_owned = part.nt_owned if part is not None else nt
    setup_s = time.time() - t_setup

    with torch.cuda.stream(stream):
        b_dev = torch.from_numpy(rhs).to(dev)
        u_dev = torch.zeros(nv, dtype=torch.float64, device=dev)
        x_dev = torch.zeros(nv, dtype=torch.float64, device=dev)
    stream.synchronize()
    # untimed: linear solve (mu = mu0 everywhere) for a physical potential to linearise about
    it0 = 0
    if args.presolve and full:
        h.assemble_stiffness_dev(h.device_ptr(3), slot=0)
        it0, rr0 = h.pcg_dev(b_dev.data_ptr(), u_dev.data_ptr(), slot=0, tol=1e-8, max_iter=20000)
    else:   # profiling runs / the strong-scaling block: the uniform applied field H = B_ext/mu0, u = -H.x
        with torch.cuda.stream(stream):
            u_dev.copy_(torch.from_numpy(-(Bext / MU0) @ mesh.p))
        stream.synchronize()

    opts = dict(max_iter=Piters, tol_disable=True, check_every=Piters + (Piters & 1), recurrence=args.recurrence)

    def step_dev(ev=None):
        if ev:
            ev[0].record(stream)
        h.picard_step_dev(u_dev.data_ptr(), slot=0, sync=False)      # enqueued; div comes back with the solve's poll
        if ev:
            ev[1].record(stream)
        h.pcg_dev(b_dev.data_ptr(), x_dev.data_ptr(), slot=0, **opts)
        if ev:
            ev[2].record(stream)
        return h.picard_result()

    def timed(fn, steps, warmup, with_events):
        for _ in range(warmup):
            fn() if not with_events else fn(None)
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)] if with_events else None
        ctx.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = h.launch_count()
        e0.record(stream)
        for s in range(steps):
            fn(evs[s]) if with_events else fn()
        e1.record(stream)
        ctx.barrier()
        ms = ctx.max_over_ranks(e0.elapsed_time(e1))
        return ms, h.launch_count() - l0, evs

    W, K = max(3, args.warmup), args.steps
    out = {"h": h, "mesh": mesh, "n_rows": n_rows, "nv": nv, "nt": nt, "nnz": nnz, "nt_owned": nt_owned, "setup_s": setup_s, "it0": it0,
           "b_dev": b_dev, "u_dev": u_dev}
    sampler = ClockSampler(local)
    if rank == 0 and full:
        sampler.start()
    out["ms_dev"], out["launches"], evs = timed(step_dev, K, W, True)
    out["asm_ms"] = ctx.max_over_ranks(statistics.mean(e[0].elapsed_time(e[1]) for e in evs))
    out["pcg_ms"] = ctx.max_over_ranks(statistics.mean(e[1].elapsed_time(e[2]) for e in evs))
    if full:
        # pinned host mirrors for the e2e leg (what a Julia driver would own)
        u_host = torch.empty(nv, dtype=torch.float64).pin_memory()
        b_host = torch.empty(nv, dtype=torch.float64).pin_memory()
        x_host = torch.empty(nv, dtype=torch.float64).pin_memory()
        u_host.copy_(u_dev.cpu())
        b_host.copy_(torch.from_numpy(rhs))

        def host_ptr(tn):
            return C.cast(tn.data_ptr(), L.c_double_p)

        o = h.pcg_opts(**opts)
        it_c, rr_c, div_c = C.c_int(), C.c_double(), C.c_double()

        def step_e2e():
            L.check(h.lib.fk_picard_pcg(h.h, 0, host_ptr(u_host), 0, host_ptr(b_host), host_ptr(x_host), C.byref(o), C.byref(div_c),
                                        C.byref(it_c), C.byref(rr_c)))

        out["ms_e2e"], _, _ = timed(step_e2e, K, W, False)
        # the same call with PAGEABLE host buffers (what a Julia Array is)
        u_pg, b_pg, x_pg = u_host.numpy().copy(), b_host.numpy().copy(), np.zeros(nv)

        def step_e2e_pageable():
            L.check(h.lib.fk_picard_pcg(h.h, 0, L.dptr(u_pg), 0, L.dptr(b_pg), L.dptr(x_pg), C.byref(o), C.byref(div_c),
                                        C.byref(it_c), C.byref(rr_c)))

        kp = max(2, K // 2)
        ms_pg, _, _ = timed(step_e2e_pageable, kp, 2, False)
        out["ms_e2e_pg"] = ms_pg * K / kp
        # ... and with those same arrays page-locked in place through the library (fk_host_register: what a Julia caller can do)
        for a in (u_pg, b_pg, x_pg):
            L.check(h.lib.fk_host_register(C.c_void_p(a.ctypes.data), a.nbytes))
        ms_rg, _, _ = timed(step_e2e_pageable, kp, 2, False)
        for a in (u_pg, b_pg, x_pg):
            L.check(h.lib.fk_host_unregister(C.c_void_p(a.ctypes.data)))
        out["ms_e2e_reg"] = ms_rg * K / kp
        out["clocks"] = sampler.stop() if rank == 0 else {}

    # outside the timed region: device-side timeline of one more solve (per-kernel spans inside the replayed graph)
    out["split"] = None
    if args.trace:
        h.tune(11, Piters)
        step_dev(None)
        h.synchronize()
        split = kernel_split(h.trace_read())
        h.tune(11, 0)
        if dist is not None:   # the slowest rank's view of every figure
            keys = sorted(k for k in split if split[k] is not None)
            tt = torch.tensor([split[k] for k in keys], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            split = {k: round(float(v), 3) for k, v in zip(keys, tt.tolist())}
        out["split"] = split
    out["step_dev"] = step_dev
    return out


def nonlinear_solve_block(args, ctx, P):
    """Outside the timed region: a complete nonlinear solve at the bench size (Picard to mu0*max|dH| < 1e-4,
    NewtonRaphson.jl:37-38,148; PCG to 1e-10 each pass, warm-started), checked by the true residual of the last pass."""
    import ctypes as C

    import torch

    from femeko_jl_b200 import lib as L

    h, mesh, dist, dev, stream = P["h"], P["mesh"], ctx.dist, ctx.dev, ctx.stream
    b_dev, n_rows = P["b_dev"], P["n_rows"]
    with torch.cuda.stream(stream):
        u2 = P["u_dev"].clone()
        r2 = torch.empty_like(u2)
    h.set_mu(np.full(mesh.nt, MU0))
    h.reset_field()
    h.assemble_stiffness_dev(h.device_ptr(3), slot=0)
    ctx.barrier()
    t0 = time.time()
    tot_it, att, div, rr = 0, 0, float("inf"), 0.0
    while div > 1e-4 and att < 100:
        it, rr = h.pcg_dev(b_dev.data_ptr(), u2.data_ptr(), slot=0, tol=1e-10, max_iter=20000, warm_start=att > 0,
                           recurrence=args.recurrence)
        div = h.picard_step_dev(u2.data_ptr(), slot=0)
        tot_it += it
        att += 1
    h.synchronize()
    secs = time.time() - t0
    nr = C.c_double()
    L.check(h.lib.fk_residual_dev(h.h, 0, C.c_void_p(u2.data_ptr()), C.c_void_p(b_dev.data_ptr()), C.c_void_p(r2.data_ptr()),
                                  C.byref(nr)))
    with torch.cuda.stream(stream):
        bn = float(torch.linalg.vector_norm(b_dev[:n_rows]).item())
    if dist is not None:
        tb = torch.tensor([bn * bn], dtype=torch.float64, device=dev)
        dist.all_reduce(tb)
        bn = float(tb.sqrt().item())
    return {"picard_iters": att, "pcg_iters_total": tot_it, "seconds": round(secs, 4), "div": div, "last_relres": rr,
            "converged": bool(div <= 1e-4),
            "check": {"what": "nonlinear residual ||b - A(mu(H(u))) u|| / ||b|| of the returned potential (matrix re-assembled from it)",
                      "value": nr.value / max(bn, 1e-300)}}


def multi_parity_block(args, ctx):
    """N > 1, outside every timed region: the partitioned path against a single-GPU solve of the same global problem
    (rank 0), on a small box cut into x-slabs (the weak-scaling layout) and on a cube cut into px x py x pz blocks (the
    strong-scaling layout, up to 7 neighbours): linear solve, one Picard sweep with the Gd table, second solve."""
    import torch

    import femeko_jl_b200 as fk
    from femeko_jl_b200 import partition as PT
    from femeko_jl_b200.mesh import kuhn_box

    rank, world, local, dist, dev = ctx.rank, ctx.world, ctx.local, ctx.dist, ctx.dev
    Hk, muk, _ = gd_table()
    spl = fk.Spline1D(Hk, muk, bc="nearest")
    F = np.array([0.5, 0.1, -0.2])
    out = {}
    for layout in ("slabs", "blocks"):
        N = args.parity_n
        if layout == "slabs":
            lm, pr = PT.kuhn_slab(N, rank, world, inside_edge=0.4)
        else:
            lm, pr = PT.kuhn_block(2 * N, rank, PT.GRIDS[world], inside_edge=0.4)

        def solve(mesh, n_owned, attach):
            h = fk.Femeko(mesh, use_mesh_VE=False, device=local, n_owned=n_owned)
            if attach is not None:
                attach.attach(h, dist)
            h.set_spline(0, spl)
            h.set_inside(mesh.InsideElements)
            mu = np.full(mesh.nt, MU0)
            mu[mesh.InsideElements] = 3 * MU0
            h.set_mu(mu)
            h.assemble_stiffness(mu, slot=0, want_vals=False)
            b = -h.boundary_integral(mesh.surfaceT, mesh.normal, mesh.AE, F, 0)
            x, it, _ = h.pcg(b, slot=0, tol=1e-10, max_iter=5000, recurrence=args.recurrence)
            div = h.picard_step(x * 2e5, slot=1)          # scaled so |H| sits on the nonlinear part of the table
            x2, it2, _ = h.pcg(b, slot=1, tol=1e-10, max_iter=5000, x0=x, recurrence=args.recurrence)
            mode = h.comm_mode()
            h.close()
            return x, it, x2, it2, div, mode

        x, it, x2, it2, div, mode = solve(lm, pr.n_owned, pr)
        res = torch.zeros((2, pr.nv_global), dtype=torch.float64, device=dev)
        idx = torch.from_numpy(pr.l2g[:pr.n_owned]).to(dev)
        res[0, idx] = torch.from_numpy(x[:pr.n_owned]).to(dev)
        res[1, idx] = torch.from_numpy(x2[:pr.n_owned]).to(dev)
        dist.all_reduce(res)
        if rank == 0:
            gm = kuhn_box((world * N, N, N), inside_edge=0.4) if layout == "slabs" else fk.kuhn_cube(2 * N, inside_edge=0.4)
            xs, its, xs2, its2, divs, _ = solve(gm, 0, None)
            r = res.cpu().numpy()
            e1 = float(np.linalg.norm(r[0] - xs) / np.linalg.norm(xs))
            e2 = float(np.linalg.norm(r[1] - xs2) / np.linalg.norm(xs2))
            out[layout] = {"global_tets": int(gm.nt), "neighbours_rank0": len(pr.neigh), "comm_mode": mode, "relerr_x": e1,
                           "relerr_x_after_sweep": e2, "iters": [it, its], "iters_after_sweep": [it2, its2],
                           "div_rel_diff": abs(div - divs) / max(abs(divs), 1e-300),
                           "ok": bool(e1 < 1e-8 and e2 < 1e-8 and abs(it - its) <= 2 and abs(it2 - its2) <= 2)}
        dist.barrier()
    if rank == 0:
        out["relerr_x"] = max(out[k]["relerr_x_after_sweep"] for k in ("slabs", "blocks"))
        out["ok"] = bool(all(out[k]["ok"] for k in ("slabs", "blocks")))
        out["against"] = "single-GPU solve of the same global mesh on rank 0 (tol 1e-10): x-slabs (weak layout) and node blocks (strong layout)"
    return out


def run_ours(args):
    import femeko_jl_b200 as fk

    ctx = Ctx(args)
    rank, world = ctx.rank, ctx.world
    N, Piters = args.n, args.pcg_iters
    mesh, part = build_problem(N, rank, world)
    P = bench_problem(args, ctx, mesh, part, full=True)
    h = P["h"]
    K, W = args.steps, max(3, args.warmup)
    n_rows, nv, nt, nnz, nt_owned = P["n_rows"], P["nv"], P["nt"], P["nnz"], P["nt_owned"]
    ms_dev, asm_ms, pcg_ms, ms_e2e, ms_e2e_pg = P["ms_dev"], P["asm_ms"], P["pcg_ms"], P["ms_e2e"], P["ms_e2e_pg"]
    nonlin = nonlinear_solve_block(args, ctx, P) if args.full_solve else None
    world_tets, world_rows, world_nnz = ctx.sum_over_ranks([nt_owned, n_rows, nnz])
    sell = h.sell_stats()
    comm_mode = h.comm_mode()
    geo = h.geo_stats() if "10=0" not in args.tune else None
    h.close()
    del P["h"], P["step_dev"], P["b_dev"], P["u_dev"]

    # N > 1: parity of the partitioned path and the strong-scaling workload, each outside the headline's timed region
    multi_parity = multi_parity_block(args, ctx) if world > 1 and args.parity_n > 0 else None
    strong = None
    if args.strong_n > 0 and (world > 1 or args.strong_single):
        from femeko_jl_b200 import partition as PT

        t0 = time.time()
        if world > 1:
            sm, sp = PT.kuhn_block(args.strong_n, rank, PT.GRIDS[world], inside_edge=0.2)
        else:
            sm, sp = fk.kuhn_cube(args.strong_n, inside_edge=0.2), None
        gen_s = time.time() - t0
        S = bench_problem(args, ctx, sm, sp, full=False)
        s_tets, s_rows, s_nnz = ctx.sum_over_ranks([S["nt_owned"], S["n_rows"], S["nnz"]])
        s_halo = ctx.sum_over_ranks([S["nv"] - S["n_rows"]])[0]
        ms = S["ms_dev"] / K
        strong = {"workload": f"configs[3]: ONE Kuhn cube N={args.strong_n} ({s_tets} tets, {s_rows} rows) cut into "
                              f"{'x'.join(str(g) for g in PT.GRIDS[world])} node blocks, same step as the headline "
                              f"(Picard sweep + {Piters} Jacobi-PCG iterations), device-resident",
                  "scaling": "strong", "grid": list(PT.GRIDS[world]), "tets_total": s_tets, "rows_total": s_rows, "halo_rows_total": s_halo,
                  "ms_per_step": round(ms, 4), "mtets_per_s": round(s_tets / (ms * 1e-3) / 1e6, 1),
                  "pcg_iteration_us": round(S["pcg_ms"] / Piters * 1e3, 2), "picard_assembly_ms": round(S["asm_ms"], 4),
                  "kernel_split_us": S["split"], "mesh_generation_s": round(gen_s, 1), "setup_s_untimed": round(S["setup_s"], 1),
                  "host_threads_per_rank": ctx.host_threads,
                  "note": "efficiency is for the driver to compute from the per-N lines: time per step should fall as 1/N"}
        S["h"].close()

    if rank == 0:
        peak, peak_src = measured_peaks()
        ms_step = ms_dev / K
        value = world_tets / (ms_step * 1e-3) / 1e6
        e2e_value = world_tets / (ms_e2e / K * 1e-3) / 1e6
        it_bytes = pcg_bytes(n_rows, nnz)                        # this rank's share, per iteration
        pcg_gbs = it_bytes * Piters / (pcg_ms * 1e-3) / 1e9
        asm_bytes = picard_bytes(nt, nv, nnz)
        asm_gbs = asm_bytes / (asm_ms * 1e-3) / 1e9
        traffic = {}
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):   # ncu captures are per configuration: only a matching one is quoted, else null
            key = f"n{N}_gpus{world}_rec{args.recurrence}"
            traffic = json.load(open(tpath)).get(key, {})
        # what cannot live in L2: the matrix stream (values + packed columns + slice metadata); the vectors are pinned
        floor_bytes = int(8 * sell[0] + sell[3] + 144 * sell[1])
        rec_name = ("single-reduction recurrence: spmv_sell_kernel<DOT> + cg_update_kernel, 2 kernels"
                    if args.recurrence == 1 else "Eigen recurrence: spmv_sell_kernel<DOT> + pcg_update_kernel + pcg_direction_kernel")
        sweep_name = ("Picard sweep, rescale path: material_update_kernel, then geo_rescale_kernel + diag_from_rowsum_kernel (4 CTAs "
                      "per SM) beside picard_element_kernel + max fold (1 CTA per SM, second stream)"
                      if geo else "Picard sweep: material_update_kernel + assemble_kernel<ASM_PICARD> + max fold")
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"configs[1]: nonlinear magnetostatics Picard step, Kuhn cube N={N} per GPU "
                                   f"({nt_owned} tets, {n_rows} rows, {nnz} nnz per GPU), Gd mu(H) table @270K, "
                                   f"{Piters} Jacobi-PCG iterations per step (fixed count, x0=0, "
                                   f"{'single-reduction' if args.recurrence == 1 else 'Eigen'} recurrence)",
                       "tets_total": world_tets, "rows_total": world_rows, "nnz_total": world_nnz, "pcg_iters_per_step": Piters,
                       "l2": "inputs larger than L2 (CSR values+indices %.0f MB per GPU vs 126 MB L2)" % (12 * nnz / 1e6),
                       "parallelism": "single GPU" if world == 1 else
                       f"node-partitioned slabs x{world}, " + ("halo rows + all-reduce over NVLink peer memory inside the PCG kernels"
                                                                if comm_mode == 2 else "NCCL send/recv halo + ncclAllReduce"),
                       "host_affinity": ctx.host_affinity},
            "components": {"picard_assembly_ms": round(asm_ms, 4), "picard_assembly_mtets_per_s": round(nt / asm_ms / 1e3, 1),
                           "pcg_ms": round(pcg_ms, 4), "pcg_iters_per_s": round(Piters / (pcg_ms * 1e-3), 1),
                           "setup_s_untimed": round(P["setup_s"], 2), "linear_presolve_iters": P["it0"]},
            "e2e": {"value": round(e2e_value, 3), "unit": UNIT, "ms_per_step": round(ms_e2e / K, 4),
                    "h2d_bytes_per_step": 16 * nv, "d2h_bytes_per_step": 8 * n_rows + 8 + 4 + 64,
                    "host_buffers": "pinned",
                    "pageable": {"value": round(world_tets / (ms_e2e_pg / K * 1e-3) / 1e6, 3), "ms_per_step": round(ms_e2e_pg / K, 4),
                                 "note": "same call with pageable host arrays (a Julia Array)"},
                    "registered": {"value": round(world_tets / (P["ms_e2e_reg"] / K * 1e-3) / 1e6, 3), "ms_per_step": round(P["ms_e2e_reg"] / K, 4),
                                   "note": "the same pageable arrays page-locked in place with fk_host_register"}},
            "gpu_launches": int(P["launches"]),
            "roofline": {"kernel": "fused Jacobi-PCG iteration (" + rec_name + ")",
                         "bound": "hbm", "achieved": round(pcg_gbs, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(pcg_gbs / peak, 4), "traffic": traffic.get("pcg_iteration_bytes"), "traffic_source": traffic.get("source"),
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": it_bytes,
                         "avg_launch_ms": round(pcg_ms / Piters, 5),
                         "note": "achieved uses SURVEY 8d's common denominator (12 B per CSR entry + 108 B per row); the "
                                 "SpMV actually streams %.2f B per entry (packed 16-bit column offsets, %d of %d slices) and "
                                 "the PCG vectors are pinned in L2, so frac can exceed 1" % (
                                     8 + sell[3] / max(sell[0], 1), sell[2], sell[1]),
                         "layout_bytes_per_launch": int(8 * sell[0] + sell[3] + 144 * sell[1] + 108 * n_rows),
                         "floor_bytes": floor_bytes,
                         "frac_of_floor": round(floor_bytes / (pcg_ms / Piters * 1e-3) / 1e9 / peak, 4),
                         "floor_note": "L2-aware floor: only the matrix stream (values, packed columns, slice metadata) must "
                                       "come from HBM each iteration; frac_of_floor = floor_bytes / t / peak"},
            "roofline_assembly": {"kernel": sweep_name + " (device time of the enqueued sweep; its scalars return with the solve's poll)",
                                  "bound": "hbm", "achieved": round(asm_gbs, 1),
                                  "peak": peak, "unit": "GB/s", "frac": round(asm_gbs / peak, 4),
                                  "traffic": traffic.get("picard_assembly_bytes") if geo else None, "algorithmic_bytes_per_launch": asm_bytes,
                                  "avg_launch_ms": round(asm_ms, 5)},
            "clocks": P["clocks"],
        }
        if geo:
            line["roofline_assembly"]["layout_bytes_per_launch"] = int(geo[3] + 8 * h_nnz_sell(sell) + 64 * nt)
            line["roofline_assembly"]["layout_note"] = ("what the rescale path streams per sweep: 12 B per contribution slot (%d slots for %d "
                                                        "contributions; 6 per tetrahedron when the diagonal comes from the row sums), 8 B per "
                                                        "entry pair, the values written once (the row-sum pass re-reads the off-diagonal values "
                                                        "it folds, mostly from L2), ~64 B per element in field recovery + material pass"
                                                        % (geo[2], geo[1]))
        if P["split"]:
            line["roofline"]["kernel_split_us"] = P["split"]
            line["roofline"]["kernel_split_note"] = ("device-side %globaltimer records inside the replayed graph (fk_tune 11), median over the "
                                                     "iterations of one extra untimed solve; wait = longest cross-GPU spin of any CTA")
        if nonlin:
            line["nonlinear_solve"] = nonlin
        if multi_parity:
            line["multi_parity"] = multi_parity
        if strong:
            line["strong"] = strong
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, budget_s=args.cpu_budget)
            line["parity"] = parity_block(args, ctx.local)
            if args.solve_n > 0:
                line["solve"] = solve_block(args, ctx.local)
        print(json.dumps(line), flush=True)
    if ctx.dist is not None:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()


def run_single_process(args):
    """bench.py --single-process --gpus N (NOT under torchrun): ONE process, fk_multi_* -- the global mesh goes in, the
    library partitions it and drives N GPUs with one host thread each; every step passes GLOBAL host arrays through the
    C ABI (what a single Julia process would do), so the figure is end to end by construction."""
    import femeko_jl_b200 as fk

    ngpu, N, Piters = args.gpus, args.n, args.pcg_iters
    t0 = time.time()
    mesh = fk.kuhn_cube(N, inside_edge=0.2)
    gen_s = time.time() - t0
    t0 = time.time()
    hm = fk.FemekoMulti(mesh, ngpu, use_mesh_VE=False)
    Hk, muk, _ = gd_table()
    hm.set_spline(0, fk.Spline1D(Hk, muk, bc="nearest"))
    hm.set_inside(mesh.InsideElements)
    hm.set_mu(np.full(mesh.nt, MU0))
    setup_s = time.time() - t0
    Bext = np.array([0.5, 0.0, 0.0])
    h0 = fk.Femeko(mesh, use_mesh_VE=False, device=0) if args.sp_check else None
    if h0 is not None:
        rhs = -h0.boundary_integral(mesh.surfaceT, mesh.normal, mesh.AE, Bext, 0)
    else:
        tmp = fk.Femeko(mesh, use_mesh_VE=False, device=0)
        rhs = -tmp.boundary_integral(mesh.surfaceT, mesh.normal, mesh.AE, Bext, 0)
        tmp.close()
    u = -(Bext / MU0) @ mesh.p
    opts = dict(max_iter=Piters, tol_disable=True, check_every=Piters + (Piters & 1), recurrence=args.recurrence)
    W, K = max(3, args.warmup), args.steps

    def step():
        div = hm.picard_step(u, slot=0)
        x, it, rr = hm.pcg(rhs, slot=0, **opts)
        return x, div

    for _ in range(W):
        step()
    t0 = time.time()
    for _ in range(K):
        x, div = step()
    ms = (time.time() - t0) / K * 1e3
    line = {"metric": METRIC, "value": round(mesh.nt / (ms * 1e-3) / 1e6, 3), "unit": UNIT, "n_gpus": ngpu, "steps": K, "warmup": W,
            "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"single process, fk_multi_*: ONE Kuhn cube N={N} ({mesh.nt} tets) partitioned by the library over {ngpu} "
                                   f"GPU(s); step = Picard sweep + {Piters} Jacobi-PCG iterations, GLOBAL host arrays in and out every call "
                                   "(wall clock, host scatter / gather of the vectors included)",
                       "parts": {k: [int(v) for v in vals] for k, vals in hm.info().items()}},
            "e2e": {"value": round(mesh.nt / (ms * 1e-3) / 1e6, 3), "unit": UNIT, "h2d_bytes_per_step": 16 * mesh.nv,
                    "d2h_bytes_per_step": 8 * mesh.nv, "host_buffers": "pageable"},
            "components": {"mesh_generation_s": round(gen_s, 1), "setup_s_untimed": round(setup_s, 1)}}
    if h0 is not None:   # same step on one GPU through the same host-pointer API
        h0.set_spline(0, fk.Spline1D(Hk, muk, bc="nearest"))
        h0.set_inside(mesh.InsideElements)
        h0.set_mu(np.full(mesh.nt, MU0))
        h0.picard_step(u, slot=0)
        x1, _, _ = h0.pcg(rhs, slot=0, **opts)
        line["parity"] = {"against": "the same step on one GPU (fk_picard_step + fk_pcg)", "x_rel_l2": float(np.linalg.norm(x - x1) / np.linalg.norm(x1))}
        h0.close()
    hm.close()
    print(json.dumps(line), flush=True)


def h_nnz_sell(sell):
    return int(sell[0])


# --------------------------------------------------------------------------------------------------
# configs[2]: Heat/heat3D.jl implicit time stepping (--workload heat)
# --------------------------------------------------------------------------------------------------
HEAT_DT = 1e-3      # heat3D.jl:10


class CpuHeat:
    """heat3D.jl:51-86 on the host: conductivity 1 / 2 inside, T0 = 0 / 10 on the inside nodes, A from the compiled
    reference stiffnessMatrix (FEMc.cpp:102-133; C port when it is absent), M from the C port of massMatrix
    (src/FEM3D.jl:466-492: Julia only), every step T <- (M + dt A)^-1 (M T) with Eigen's Jacobi-CG (x0 = 0)."""

    def __init__(self, N, tol):
        from oracle import oracle as orc

        self.orc, self.tol = orc, tol
        self.kind = "reference" if orc.have_ref() else "port"
        m = self.m = mesh_module().kuhn_cube(N, inside_edge=0.2)
        k = np.ones(m.nt)
        k[m.InsideElements] = 2.0
        self.rp, self.ci = orc.pattern(m.t, m.nv)
        if self.kind == "reference":
            _, _, A = orc.ref_stiffness(m, k)
        else:
            _, _, A = orc.stiffness(m, k, order="eigen")
        _, _, self.M = orc.mass(m)
        self.LHS = self.M + HEAT_DT * A
        self.T = np.zeros(m.nv)
        self.T[m.InsideNodes] = 10.0
        self.iters = 0

    def step(self):
        orc = self.orc
        b = orc.spmv(self.rp, self.ci, self.M, self.T)
        if self.kind == "reference":
            self.T, it = orc.ref_cg(self.rp, self.ci, self.LHS, b, self.tol, 0)[:2]
        else:
            self.T, it = orc.pcg(self.rp, self.ci, self.LHS, b, self.tol)[:2]
        self.iters += int(it)


def heat_problem(ctx, args, mesh, part):
    import torch

    import femeko_jl_b200 as fk

    h = fk.Femeko(mesh, use_mesh_VE=False, device=ctx.local, n_owned=part.n_owned if part is not None else 0)
    for kv in [t for t in args.tune.split(",") if t]:
        h.tune(int(kv.split("=")[0]), int(kv.split("=")[1]))
    h.set_stream(ctx.stream.cuda_stream)
    if part is not None:
        part.attach(h, ctx.dist)
    k = np.ones(mesh.nt)
    k[mesh.InsideElements] = 2.0
    h.assemble_stiffness(k, slot=0, want_vals=False)
    h.assemble_mass(slot=1, want_vals=False)
    h.matrix_axpby(3, 1.0, 1, HEAT_DT, 0)                    # M + dt A on the shared pattern (heat3D.jl:86)
    T0 = np.zeros(h.nv)
    T0[mesh.InsideNodes] = 10.0
    with torch.cuda.stream(ctx.stream):
        T_dev = torch.from_numpy(T0).to(ctx.dev)
    ctx.stream.synchronize()
    return h, T0, T_dev


def run_heat(args):
    import ctypes as C

    import torch

    import femeko_jl_b200 as fk
    from femeko_jl_b200 import lib as L

    ctx = Ctx(args)
    rank, world, stream = ctx.rank, ctx.world, ctx.stream
    N = args.n
    t0 = time.time()
    mesh, part = build_problem(N, rank, world)
    h, T0, T_dev = heat_problem(ctx, args, mesh, part)
    setup_s = time.time() - t0
    n_rows, nv, nt, nnz = h.n_rows, h.nv, h.nt, h.nnz
    nt_owned = part.nt_owned if part is not None else nt
    W, K = max(3, args.warmup), args.steps
    its = []

    def step_dev():
        it, _ = h.heat_step_dev(T_dev.data_ptr(), 1, 3, tol=args.heat_tol, recurrence=args.recurrence)
        its.append(it)

    T_host = torch.from_numpy(T0.copy()).pin_memory()
    o = h.pcg_opts(tol=args.heat_tol, recurrence=args.recurrence)
    it_c, rr_c = C.c_int(), C.c_double()

    def step_e2e():
        L.check(h.lib.fk_heat_step(h.h, 1, 3, C.cast(T_host.data_ptr(), L.c_double_p), C.byref(o), C.byref(it_c), C.byref(rr_c)))

    def timed(fn):
        for _ in range(W):
            fn()
        ctx.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = h.launch_count()
        e0.record(stream)
        for _ in range(K):
            fn()
        e1.record(stream)
        ctx.barrier()
        return ctx.max_over_ranks(e0.elapsed_time(e1)), h.launch_count() - l0

    sampler = ClockSampler(ctx.local)
    if rank == 0:
        sampler.start()
    ms_dev, launches = timed(step_dev)
    its_timed = its[-K:]
    ms_e2e, _ = timed(step_e2e)
    clocks = sampler.stop() if rank == 0 else {}
    world_tets, world_rows, world_nnz = ctx.sum_over_ranks([nt_owned, n_rows, nnz])
    sell = h.sell_stats()
    comm_mode = h.comm_mode()
    h.close()
    if rank == 0:
        peak, peak_src = measured_peaks()
        ms_step = ms_dev / K
        tot_it = int(sum(its_timed))
        # per step: the mass product (12 nnz + 16 n) and `it` PCG iterations + the initial residual product
        step_bytes = (12 * nnz + 16 * n_rows) * 2 * K + pcg_bytes(n_rows, nnz) * tot_it
        gbs = step_bytes / (ms_dev * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": round(world_tets / (ms_step * 1e-3) / 1e6, 3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"configs[2]: Heat/heat3D.jl implicit time stepping, Kuhn cube N={N} per GPU ({nt_owned} tets, {n_rows} rows per "
                                   f"GPU), conductivity 1 / 2 inside, T0 = 0 / 10, dt = {HEAT_DT}; a step = T <- (M + dt A)^-1 (M T): mass "
                                   f"SpMV + Jacobi-PCG to {args.heat_tol:g} warm-started from T (matrices assembled once, untimed); value = tets "
                                   "advanced one time step per second",
                       "tets_total": world_tets, "rows_total": world_rows, "pcg_iters_per_step": round(tot_it / K, 2),
                       "l2": "inputs larger than L2 (matrix stream %.0f MB per GPU vs 126 MB L2)" % (12 * nnz / 1e6),
                       "parallelism": "single GPU" if world == 1 else f"node-partitioned slabs x{world}, comm mode {comm_mode}"},
            "components": {"pcg_iters_total": tot_it, "pcg_iters_per_step": its_timed, "setup_s_untimed": round(setup_s, 2)},
            "e2e": {"value": round(world_tets / (ms_e2e / K * 1e-3) / 1e6, 3), "unit": UNIT, "ms_per_step": round(ms_e2e / K, 4),
                    "h2d_bytes_per_step": 8 * n_rows, "d2h_bytes_per_step": 8 * n_rows + 64, "host_buffers": "pinned"},
            "gpu_launches": int(launches),
            "roofline": {"kernel": "heat step = 2 SpMV (mass product, initial residual) + Jacobi-PCG iterations (spmv_sell_kernel<DOT> + "
                                   "pcg_update_kernel + pcg_direction_kernel)", "bound": "hbm", "achieved": round(gbs, 1), "peak": peak,
                         "unit": "GB/s", "frac": round(gbs / peak, 4), "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(step_bytes / K), "avg_launch_ms": round(ms_step, 5),
                         "note": "bytes per step by SURVEY 8d's figures (12 nnz + 108 n per iteration, 12 nnz + 16 n per product); with so few "
                                 "iterations per step the polls and the set-up kernels of every solve weigh in"},
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            cs = CpuHeat(args.cpu_n, args.heat_tol)
            cs.step()
            t1 = time.time()
            reps = 0
            while reps < 3 or (time.time() - t1 < args.cpu_budget and reps < 50):
                cs.step()
                reps += 1
            per = (time.time() - t1) / reps
            line["cpu_baseline"] = {"value": round(cs.m.nt / per / 1e6, 4), "unit": UNIT, "cores": 1, "host_cores": os.cpu_count(),
                                    "kind": cs.kind, "sample": f"Kuhn cube N={args.cpu_n} ({cs.m.nt} tets), {reps} time steps: mass product + "
                                                               f"Eigen Jacobi-CG to {args.heat_tol:g} on M + dt A (x0 = 0), single thread",
                                    "s_per_step": round(per, 4)}
            line["parity"] = heat_parity(args, ctx.local)
        print(json.dumps(line), flush=True)
    if ctx.dist is not None:
        ctx.dist.barrier()
        ctx.dist.destroy_process_group()


def heat_parity(args, device):
    """Outside the timed region: five time steps at Kuhn N = --heat-parity-n on the GPU against (a) a sparse direct solve of
    the oracle-assembled M + dt A (what heat3D.jl's `\\` does) and (b) the CPU arm."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla

    import femeko_jl_b200 as fk

    cs = CpuHeat(args.heat_parity_n, 1e-12)
    m = cs.m
    h = fk.Femeko(m, device=device)
    k = np.ones(m.nt)
    k[m.InsideElements] = 2.0
    h.assemble_stiffness(k, slot=0, want_vals=False)
    h.assemble_mass(slot=1, want_vals=False)
    h.matrix_axpby(3, 1.0, 1, HEAT_DT, 0)
    T = cs.T.copy()
    Td = cs.T.copy()
    Mm = sp.csr_matrix((cs.M, cs.ci, cs.rp), shape=(m.nv, m.nv))
    lu = spla.splu(sp.csr_matrix((cs.LHS, cs.ci, cs.rp), shape=(m.nv, m.nv)).tocsc())
    for _ in range(5):
        T, _, _ = h.heat_step(T, 1, 3, tol=1e-12, recurrence=args.recurrence)
        Td = lu.solve(Mm @ Td)
        cs.step()
    h.close()

    def rl2(a, b):
        return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))

    out = {"against": f"Kuhn N={args.heat_parity_n}, 5 steps: sparse direct solve of the oracle / reference matrices; CPU arm ({cs.kind})",
           "T_rel_l2_vs_direct": rl2(T, Td), "T_rel_l2_vs_cpu_arm": rl2(T, cs.T),
           "heat_conserved_rel": float(abs((Mm @ T).sum() - (Mm @ Td).sum()) / abs((Mm @ Td).sum()))}
    out["ok"] = bool(out["T_rel_l2_vs_direct"] < 1e-8 and out["T_rel_l2_vs_cpu_arm"] < 1e-8)
    return out


def run_reference_heat(args):
    if env_int("RANK", 0) != 0:
        return
    W, K = max(1, args.warmup), args.steps
    cs = CpuHeat(args.cpu_n, args.heat_tol)
    for _ in range(min(W, 2)):
        cs.step()
    t0 = time.time()
    for _ in range(K):
        cs.step()
    per = (time.time() - t0) / K
    value = round(cs.m.nt / per / 1e6, 4)
    sample = (f"each step = one implicit time step on a Kuhn cube N={args.cpu_n} ({cs.m.nt} tets): mass product + Eigen Jacobi-CG to "
              f"{args.heat_tol:g} on M + dt A (x0 = 0); single-threaded reference")
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": K, "warmup": W,
                      "ms_per_step": round(per * 1e3, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                      "data": "synthetic", "config": {"workload": f"configs[2] bounded sample: heat3D.jl time stepping, Kuhn cube N={args.cpu_n}"},
                      "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "host_cores": os.cpu_count(), "kind": cs.kind, "sample": sample},
                      "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


# --------------------------------------------------------------------------------------------------
# CPU arm: the compiled reference (oracle/_ref) -- or the C port when it is absent
# --------------------------------------------------------------------------------------------------
# ---------------------------------------------------------------------------------------------------------------------
# --workload demag: SURVEY 8 row f-1, the repeated solve of the Landau-Lifshitz loop (LL::magnetostaticField, LL.cpp:287-348):
# load vector from the nodal magnetisation, Jacobi-PCG on the constant stiffness matrix from x0 = 0 (CG.solve(RHS), LL.cpp:316),
# element field, box method back to the nodes.  One GPU (fk_demag_* does not take partitioned handles).
# ---------------------------------------------------------------------------------------------------------------------
def demag_magnetisations(mesh, count, seed=5):
    """A short cycle of nodal magnetisations (3 x nv, zero outside the magnet): unit vectors turning about z with a
    smooth spatial twist, the kind of state an LL time loop hands over step after step."""
    insn = np.unique(mesh.t[:, mesh.InsideElements])
    rng = np.random.default_rng(seed)
    ph0 = rng.uniform(0, 2 * np.pi)
    out = []
    for s in range(count):
        M = np.zeros((3, mesh.nv), order="F")
        ang = ph0 + 0.05 * s + 2.0 * mesh.p[2, insn]
        M[0, insn], M[1, insn], M[2, insn] = np.cos(ang), np.sin(ang), 0.2
        out.append(M)
    return out


class CpuDemag:
    """LL::magnetostaticField on the host: C port (oracle/) of the load vector, Eigen's Jacobi-CG and the box method, at the
    tolerance of the GPU arm (the compiled class LL runs Eigen's default epsilon tolerance and cannot be told otherwise)."""

    def __init__(self, N, tol):
        from oracle import oracle as orc

        self.orc, self.tol = orc, tol
        self.kind = "port"
        self.m = mesh_module().kuhn_cube(N, inside_edge=0.5)
        self.Ms = demag_magnetisations(self.m, 4)
        self.rp, self.ci, self.A = orc.stiffness(self.m, None, order="eigen")
        self.vol = orc.node_volumes(self.m, self.m.InsideElements)
        self.n, self.iters = 0, 0

    def step(self):
        orc, m = self.orc, self.m
        rhs = orc.demag_rhs(m, m.InsideElements, self.Ms[self.n % 4])
        u, it = orc.pcg(self.rp, self.ci, self.A, rhs, self.tol)[:2]
        self.Hd = orc.demag_box(m, m.InsideElements, u, self.vol)
        self.n += 1
        self.iters += int(it)


def demag_parity(args, device):
    """GPU demag field against the oracle on a small cube, same tolerance."""
    import femeko_jl_b200 as fk
    from oracle import oracle as orc

    m = mesh_module().kuhn_cube(args.heat_parity_n, inside_edge=0.5)
    M = demag_magnetisations(m, 1)[0]
    h = fk.Femeko(m, device=device)
    h.demag_setup(m.InsideElements.astype(np.int32), slot=0)
    Hd, it = h.demag_field(M, tol=args.demag_tol)
    h.close()
    Ho, _, ito, _, _ = orc.demag_field(m, m.InsideElements, M, tol=args.demag_tol)
    fin = np.isfinite(Ho)
    rel = float(np.linalg.norm(Hd[fin] - Ho[fin]) / np.linalg.norm(Ho[fin]))
    return {"against": f"oracle LL::magnetostaticField restatement, Kuhn N={args.heat_parity_n}, tol {args.demag_tol:g}",
            "Hd_rel_l2": rel, "iterations": [int(it), int(ito)], "nan_pattern_equal": bool(np.array_equal(fin, np.isfinite(Hd))),
            "ok": bool(rel < 1e-7 and abs(it - ito) <= 2 and np.array_equal(fin, np.isfinite(Hd)))}


def run_demag(args):
    import ctypes as C

    import torch

    import femeko_jl_b200 as fk
    from femeko_jl_b200 import lib as L

    if args.gpus != 1:
        raise SystemExit("bench.py --workload demag runs on one GPU (fk_demag_* does not take partitioned handles)")
    ctx = Ctx(args)
    stream = ctx.stream
    N = args.n
    t0 = time.time()
    mesh = mesh_module().kuhn_cube(N, inside_edge=0.5)
    h = fk.Femeko(mesh, use_mesh_VE=False, device=ctx.local)
    for kv in [t for t in args.tune.split(",") if t]:
        h.tune(int(kv.split("=")[0]), int(kv.split("=")[1]))
    h.set_stream(stream.cuda_stream)
    h.demag_setup(mesh.InsideElements.astype(np.int32), slot=0)
    Ms = demag_magnetisations(mesh, 4)
    with torch.cuda.stream(stream):
        M_dev = [torch.from_numpy(np.ascontiguousarray(M.T)).to(ctx.dev) for M in Ms]      # 3 x nv column-major
        Hd_dev = torch.empty_like(M_dev[0])
    M_pin = [torch.from_numpy(np.ascontiguousarray(M.T)).pin_memory() for M in Ms]
    Hd_pin = torch.empty_like(M_pin[0]).pin_memory()
    stream.synchronize()
    setup_s = time.time() - t0
    nv, nt, nnz = h.nv, h.nt, h.nnz
    W, K = max(3, args.warmup), args.steps
    its, n = [], [0]

    def step_dev(warm=False):
        it, _ = h.demag_field_dev(M_dev[n[0] % 4].data_ptr(), Hd_dev.data_ptr(), tol=args.demag_tol, warm_start=warm)
        n[0] += 1
        its.append(it)

    o = h.pcg_opts(tol=args.demag_tol)
    it_c, rr_c = C.c_int(), C.c_double()

    def step_e2e():
        L.check(h.lib.fk_demag_field(h.h, 0, C.cast(M_pin[n[0] % 4].data_ptr(), L.c_double_p), C.byref(o), 1,
                                     C.cast(Hd_pin.data_ptr(), L.c_double_p), None, C.byref(it_c), C.byref(rr_c)))
        n[0] += 1

    def timed(fn):
        for _ in range(W):
            fn()
        ctx.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = h.launch_count()
        e0.record(stream)
        for _ in range(K):
            fn()
        e1.record(stream)
        ctx.barrier()
        return e0.elapsed_time(e1), h.launch_count() - l0

    sampler = ClockSampler(ctx.local)
    sampler.start()
    ms_dev, launches = timed(step_dev)
    its_timed = its[-K:]
    ms_e2e, _ = timed(step_e2e)
    clocks = sampler.stop()
    ms_warm, _ = timed(lambda: step_dev(True))
    its_warm = its[-K:]
    h.close()
    peak, peak_src = measured_peaks()
    ms_step = ms_dev / K
    tot_it = int(sum(its_timed))
    step_bytes = pcg_bytes(nv, nnz) * tot_it + K * (12 * nnz + 16 * nv + 60 * nt)
    gbs = step_bytes / (ms_dev * 1e-3) / 1e9
    line = {
        "metric": METRIC, "value": round(nt / (ms_step * 1e-3) / 1e6, 3), "unit": UNIT, "n_gpus": 1, "steps": K, "warmup": W,
        "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"SURVEY 8 f-1: demagnetising field of the Landau-Lifshitz loop (LL::magnetostaticField, LL.cpp:287-348), Kuhn cube "
                               f"N={N} ({nt} tets, {nv} rows), magnet = the inner half cube ({mesh.nInside} tets); a step = load vector from the "
                               f"nodal magnetisation + Jacobi-PCG from x0 = 0 to {args.demag_tol:g} on the resident stiffness matrix + element field "
                               "+ box method to the nodes; value = tets of the mesh per second of one field evaluation",
                   "tets_total": nt, "rows_total": nv, "pcg_iters_per_step": round(tot_it / K, 1),
                   "l2": "inputs larger than L2 (matrix stream %.0f MB vs 126 MB L2)" % (12 * nnz / 1e6), "parallelism": "single GPU"},
        "components": {"pcg_iters_per_step": its_timed, "pcg_us_per_iteration": round(ms_dev * 1e3 / max(tot_it, 1), 2),
                       "warm_started": {"ms_per_step": round(ms_warm / K, 4), "pcg_iters_per_step": its_warm,
                                        "note": "same steps with the previous potential as initial guess (fk_pcg_opts.warm_start): what an "
                                                "LL loop on this library would use; the reference always starts from 0"},
                       "setup_s_untimed": round(setup_s, 2)},
        "e2e": {"value": round(nt / (ms_e2e / K * 1e-3) / 1e6, 3), "unit": UNIT, "ms_per_step": round(ms_e2e / K, 4),
                "h2d_bytes_per_step": 24 * nv, "d2h_bytes_per_step": 24 * nv + 12, "host_buffers": "pinned"},
        "gpu_launches": int(launches),
        "roofline": {"kernel": "Jacobi-PCG iterations of the demag solve (spmv_sell_kernel<DOT> + pcg_update_kernel + pcg_direction_kernel) + "
                               "demag_rhs / element-field / box-method kernels", "bound": "hbm", "achieved": round(gbs, 1), "peak": peak,
                     "unit": "GB/s", "frac": round(gbs / peak, 4), "traffic": None, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": int(step_bytes / K), "avg_launch_ms": round(ms_step, 5),
                     "note": "bytes per step by SURVEY 8d's figures (12 nnz + 108 n per iteration) + one residual product and the field kernels"},
        "clocks": clocks,
    }
    if not args.no_cpu_baseline:
        cs = CpuDemag(args.cpu_n, args.demag_tol)
        cs.step()
        t1 = time.time()
        reps = 0
        while reps < 2 or (time.time() - t1 < args.cpu_budget and reps < 50):
            cs.step()
            reps += 1
        per = (time.time() - t1) / reps
        line["cpu_baseline"] = {"value": round(cs.m.nt / per / 1e6, 4), "unit": UNIT, "cores": 1, "host_cores": os.cpu_count(), "kind": cs.kind,
                                "sample": f"Kuhn cube N={args.cpu_n} ({cs.m.nt} tets), {reps} field evaluations (C port of LL::magnetostaticField, "
                                          f"Eigen's Jacobi-CG from 0 to {args.demag_tol:g}, {cs.iters // (reps + 1)} iterations each)",
                                "s_per_step": round(per, 4)}
        line["parity"] = demag_parity(args, ctx.local)
    print(json.dumps(line), flush=True)


def run_reference_demag(args):
    if env_int("RANK", 0) != 0:
        return
    cs = CpuDemag(args.cpu_n, args.demag_tol)
    W, K = max(1, args.warmup), args.steps
    for _ in range(W):
        cs.step()
    t1 = time.time()
    for _ in range(K):
        cs.step()
    per = (time.time() - t1) / K
    v = round(cs.m.nt / per / 1e6, 4)
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": K, "warmup": W,
                      "ms_per_step": round(per * 1e3, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                      "data": "synthetic",
                      "config": {"workload": f"SURVEY 8 f-1 bounded sample: demagnetising field, Kuhn cube N={args.cpu_n}, Jacobi-PCG from 0 to {args.demag_tol:g}"},
                      "cpu_baseline": {"value": v, "unit": UNIT, "cores": 1, "host_cores": os.cpu_count(), "kind": cs.kind,
                                       "sample": f"Kuhn cube N={args.cpu_n} ({cs.m.nt} tets), {K} field evaluations"},
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}), flush=True)


def scipy_spline(x, y):
    """FITPACK curfit(k=3, s=0) through scipy (the Fortran Dierckx.jl wraps): (t, c) of Spline1D(x, y)."""
    from scipy.interpolate import splrep

    t, c, _ = splrep(x, y, k=3, s=0)
    return np.ascontiguousarray(t), np.ascontiguousarray(c)


class CpuStep:
    """One Picard step on the host with the reference's own code where it exists:
    field recovery + spline (C port of magneticField.jl:111-132 / FITPACK splev: the reference has
    them only in Julia), stiffnessMatrix (src/FEMc.cpp:102-133, compiled unchanged) and Eigen's
    Jacobi-PCG (magnetostatics.cpp:26-38) for a fixed iteration count.  Imports nothing from the product package
    (the mesh generator is loaded as a stand-alone numpy module, the spline is fitted by scipy)."""

    def __init__(self, N, piters):
        from oracle import oracle as orc

        self.orc = orc
        self.kind = "reference" if orc.have_ref() else "port"
        self.m = mesh_module().kuhn_cube(N, inside_edge=0.2)
        self.piters = piters
        Hk, muk, _ = gd_table()
        self.spl_t, self.spl_c = scipy_spline(Hk, muk)
        m = self.m
        self.rp, self.ci = orc.pattern(m.t, m.nv)
        self.b = -orc.boundary_integral(m, [0.5, 0.0, 0.0], [0], "julia")
        mu = np.full(m.nt, MU0)
        if self.kind == "reference":
            _, _, v = orc.ref_stiffness(m, mu)
            self.u = orc.ref_cg(self.rp, self.ci, v, self.b, 1e-8, 0)[0]
        else:
            _, _, v = orc.stiffness(m, mu, order="eigen")
            self.u = orc.pcg(self.rp, self.ci, v, self.b, 1e-8)[0]
        self.mu = mu
        self.out = None

    def step(self):
        orc, m = self.orc, self.m
        _, Hn = orc.recover_field(m, self.u)
        mu = self.mu.copy()
        mu[m.InsideElements] = orc.splev(self.spl_t, self.spl_c, 3, Hn[m.InsideElements], "nearest")
        if self.kind == "reference":
            outer, inner, v = orc.ref_stiffness(m, mu)
            x = orc.ref_cg(outer, inner, v, self.b, 1e-300, self.piters)[0]
        else:
            _, _, v = orc.stiffness(m, mu, order="eigen")
            x = orc.pcg(self.rp, self.ci, v, self.b, 0.0, self.piters)[0]
        self.out = {"Hn": Hn, "mu": mu, "vals": v, "x": x}


def parity_block(args, device):
    """Outside every timed region: the step bench.py times (sweep with the Gd table + fixed-count PCG) run by the GPU
    path and by the CPU arm on the SAME mesh and potential (Kuhn N = --cpu-n), outputs compared."""
    import femeko_jl_b200 as fk

    cs = CpuStep(args.cpu_n, args.pcg_iters)
    cs.step()
    m = cs.m
    Hk, muk, _ = gd_table()
    h = fk.Femeko(m, use_mesh_VE=cs.kind == "reference", device=device)   # the reference takes VE from its caller
    h.set_spline(0, fk.Spline1D(Hk, muk, bc="nearest"))
    h.set_inside(m.InsideElements)
    h.set_mu(np.full(m.nt, MU0))
    rp, ci = h.pattern()
    same_pattern = bool(np.array_equal(rp, cs.rp) and np.array_equal(ci, cs.ci))
    div = h.picard_step(cs.u, slot=0)
    v = h.matrix_get(0)
    x, it, _ = h.pcg(cs.b, slot=0, max_iter=args.pcg_iters, tol_disable=True, recurrence=args.recurrence)
    _, Hn = h.get_field()
    mu = h.get_mu()
    rows = np.repeat(np.arange(m.nv, dtype=np.int64), np.diff(rp))
    diag = np.zeros(m.nv)
    on = ci == rows
    diag[rows[on]] = np.abs(cs.out["vals"][on])

    def rl2(a, b):
        return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))

    out = {"against": f"CPU arm ({cs.kind}: FEMc.cpp stiffnessMatrix + Eigen Jacobi-CG x{args.pcg_iters}; C port of field recovery + splev) "
                      f"on Kuhn N={args.cpu_n}, same potential",
           "pattern_bit_exact": same_pattern,
           "matrix_max_abs_err_over_diag": float(np.max(np.abs(v - cs.out["vals"]) / diag[rows])),
           "mu_max_rel_err": float(np.max(np.abs(mu - cs.out["mu"]) / cs.out["mu"])),
           "Hnorm_rel_l2": rl2(Hn, cs.out["Hn"]),
           "x_rel_l2_after_fixed_iters": rl2(x, cs.out["x"]),
           "div": div}
    out["ok"] = bool(same_pattern and out["matrix_max_abs_err_over_diag"] < 1e-12 and out["mu_max_rel_err"] < 1e-12
                     and out["x_rel_l2_after_fixed_iters"] < 1e-8)
    h.close()
    return out


def cpu_solve(args):
    """The full nonlinear solve on the host: Picard loop of manyBodyIntersecting.jl:141-210 with the compiled reference
    stiffnessMatrix + Eigen Jacobi-CG (tol 1e-10, x0 = 0) inside, Kuhn N = --solve-n."""
    from oracle import oracle as orc

    m = mesh_module().kuhn_cube(args.solve_n, inside_edge=0.2)
    Hk, muk, _ = gd_table()
    tck = scipy_spline(Hk, muk)
    t0 = time.time()
    r = orc.picard_loop(m, [0.5, 0.0, 0.0], tck, tol=1e-4, solver="ref" if orc.have_ref() else "pcg", cg_tol=1e-10)
    return m, r, time.time() - t0


def solve_block(args, device, cpu=None):
    """Time to solution of the whole nonlinear problem, both arms on the SAME problem (same mesh, table, tolerances,
    x0 = 0 for every linear solve); the GPU arm runs through the host-pointer API like a Julia driver would."""
    import femeko_jl_b200 as fk

    m, r, cpu_s = cpu if cpu is not None else cpu_solve(args)
    Hk, muk, _ = gd_table()
    spl = fk.Spline1D(Hk, muk, bc="nearest")
    h = fk.Femeko(m, device=device)
    fk.picard(m, [0.5, 0.0, 0.0], spl, fk=h, tol=1e-10, recurrence=args.recurrence)      # warm-up (graphs, allocations)
    t0 = time.time()
    out = fk.picard(m, [0.5, 0.0, 0.0], spl, fk=h, tol=1e-10, recurrence=args.recurrence)
    gpu_s = time.time() - t0
    h.close()

    def rl2(a, b):
        return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))

    return {"workload": f"nonlinear magnetostatics to mu0*max|dH| < 1e-4 (Picard, manyBodyIntersecting.jl:141-210), Kuhn N={args.solve_n} "
                        f"({m.nt} tets), Gd table @270K, Jacobi-CG tol 1e-10, x0 = 0",
            "same_config": True, "gpu_seconds": round(gpu_s, 4), "cpu_seconds": round(cpu_s, 3), "cpu_cores": 1,
            "picard_passes": [out.att, r["att"]], "cg_iterations": [out.pcg_iters, r["cg_iters"]],
            "Hfield_rel_l2": rl2(out.Hfield, r["Hfield"]), "mu_rel_l2": rl2(out.mu, r["mu"]),
            "note": "gpu_seconds is wall clock through the host-pointer C ABI (vectors cross PCIe every call)"}


def cpu_baseline(args, budget_s=20.0):
    N = args.cpu_n
    cs = CpuStep(N, args.pcg_iters)
    t0 = time.time()
    cs.step()
    one = time.time() - t0
    reps = max(1, min(5, int(budget_s / max(one, 1e-3)) - 1))
    t0 = time.time()
    for _ in range(reps):
        cs.step()
    per = (time.time() - t0) / reps
    return {"value": round(cs.m.nt / per / 1e6, 4), "unit": UNIT, "cores": 1, "host_cores": os.cpu_count(), "kind": cs.kind,
            "sample": f"Kuhn cube N={N} ({cs.m.nt} tets), {reps} step(s) of [field recovery + spline (C port) + reference "
                      f"stiffnessMatrix (FEMc.cpp:102-133) + Eigen Jacobi-PCG x{args.pcg_iters} (magnetostatics.cpp:26-38)]; "
                      f"the reference path is single-threaded (omp pragma commented out FEMc.cpp:85; Eigen Lower-CG has no omp path)",
            "s_per_step": round(per, 4)}


def run_reference(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return
    W, K = max(1, args.warmup), args.steps
    cs = CpuStep(args.cpu_n, args.pcg_iters)
    for _ in range(min(W, 2)):
        cs.step()
    t0 = time.time()
    for _ in range(K):
        cs.step()
    per = (time.time() - t0) / K
    value = round(cs.m.nt / per / 1e6, 4)
    sample = (f"each step = one Picard step on a Kuhn cube N={args.cpu_n} ({cs.m.nt} tets): field recovery + spline (C port of "
              f"the Julia-only pieces) + reference stiffnessMatrix + Eigen Jacobi-PCG x{args.pcg_iters}; single-threaded reference")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": K, "warmup": W,
            "ms_per_step": round(per * 1e3, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"configs[1] bounded sample: nonlinear magnetostatics Picard step, Kuhn cube N={args.cpu_n}, "
                                   f"Gd mu(H) table @270K, {args.pcg_iters} Jacobi-PCG iterations per step (fixed count, x0=0)"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "host_cores": os.cpu_count(), "kind": cs.kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if args.solve_n > 0:
        m, r, secs = cpu_solve(args)
        line["solve"] = {"workload": f"nonlinear magnetostatics to mu0*max|dH| < 1e-4, Kuhn N={args.solve_n} ({m.nt} tets), Gd table "
                                     f"@270K, Jacobi-CG tol 1e-10, x0 = 0", "cpu_seconds": round(secs, 3), "cpu_cores": 1,
                         "picard_passes": r["att"], "cg_iterations": r["cg_iters"], "div": r["div"]}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", "--cells", dest="n", type=int, default=119, help="Kuhn cells per edge per GPU (119 = configs[1], 10.1 M tets)")
    ap.add_argument("--pcg-iters", type=int, default=100)
    ap.add_argument("--cpu-n", type=int, default=48, help="Kuhn cells per edge of the bounded CPU sample")
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    ap.add_argument("--solve-n", type=int, default=32, help="Kuhn cells per edge of the full nonlinear solve timed on both arms (0: skip)")
    ap.add_argument("--strong-n", type=int, default=256,
                    help="N > 1: also time the step on ONE global Kuhn cube of this many cells per edge cut into blocks (0: skip)")
    ap.add_argument("--strong-single", action="store_true", help="run the --strong-n cube on one GPU too (the 1-GPU point of the strong curve)")
    ap.add_argument("--parity-n", type=int, default=12, help="N > 1: cells per edge and rank of the multi-GPU parity check (0: skip)")
    ap.add_argument("--trace", type=int, default=1, help="1: add the per-kernel split of the PCG iteration (one extra untimed solve)")
    ap.add_argument("--recurrence", type=int, default=0, choices=[0, 1],
                    help="PCG recurrence: 1 = single reduction (2 kernels / iteration), 0 = Eigen's (3 kernels)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--tune", default="", help="experiments: comma-separated what=value pairs for fk_tune (e.g. 6=0)")
    ap.add_argument("--presolve", type=int, default=1, help="0: skip the untimed linear presolve (profiling runs)")
    ap.add_argument("--full-solve", type=int, default=1, help="also record one complete nonlinear solve (untimed)")
    ap.add_argument("--workload", default="picard", choices=["picard", "heat", "demag"],
                    help="picard = configs[1] (the headline); heat = configs[2], Heat/heat3D.jl time stepping (use --n 150); "
                         "demag = SURVEY 8 f-1, the repeated demagnetising-field solve of the LL loop (one GPU)")
    ap.add_argument("--demag-tol", type=float, default=1e-10)
    ap.add_argument("--heat-tol", type=float, default=1e-10)
    ap.add_argument("--heat-parity-n", type=int, default=24)
    ap.add_argument("--single-process", action="store_true",
                    help="drive --gpus N GPUs from THIS process through fk_multi_* (do not launch under torchrun); --n = global cube")
    ap.add_argument("--sp-check", type=int, default=1, help="--single-process: also run the step on one GPU and compare")
    args = ap.parse_args()
    if args.single_process:
        return run_single_process(args)
    if args.workload == "heat":
        run_reference_heat(args) if args.impl == "reference" else run_heat(args)
    elif args.workload == "demag":
        run_reference_demag(args) if args.impl == "reference" else run_demag(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
