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


def run_ours(args):
    import torch

    import femeko_jl_b200 as fk

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- this framework has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)

    N, Piters = args.n, args.pcg_iters
    t_setup = time.time()
    mesh, part = build_problem(N, rank, world)
    n_owned = part.n_owned if part is not None else 0
    h = fk.Femeko(mesh, use_mesh_VE=False, device=local, n_owned=n_owned)
    for kv in [t for t in args.tune.split(",") if t]:
        h.tune(int(kv.split("=")[0]), int(kv.split("=")[1]))
    stream = torch.cuda.Stream(device=dev)
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
    if args.presolve:
        h.assemble_stiffness_dev(h.device_ptr(3), slot=0)
        it0, rr0 = h.pcg_dev(b_dev.data_ptr(), u_dev.data_ptr(), slot=0, tol=1e-8, max_iter=20000)
    else:   # profiling runs: the uniform applied field H = B_ext/mu0, u = -H.x
        with torch.cuda.stream(stream):
            u_dev.copy_(torch.from_numpy(-(Bext / MU0) @ mesh.p))
        stream.synchronize()

    # pinned host mirrors for the e2e leg (what a Julia driver would own)
    u_host = torch.empty(nv, dtype=torch.float64).pin_memory()
    b_host = torch.empty(nv, dtype=torch.float64).pin_memory()
    x_host = torch.empty(nv, dtype=torch.float64).pin_memory()
    u_host.copy_(u_dev.cpu())
    b_host.copy_(torch.from_numpy(rhs))
    import ctypes as C

    from femeko_jl_b200 import lib as L

    def host_ptr(tn):
        return C.cast(tn.data_ptr(), L.c_double_p)

    opts = dict(max_iter=Piters, tol_disable=True, check_every=Piters + (Piters & 1))

    def step_dev(ev=None):
        if ev:
            ev[0].record(stream)
        div = h.picard_step_dev(u_dev.data_ptr(), slot=0)
        if ev:
            ev[1].record(stream)
        h.pcg_dev(b_dev.data_ptr(), x_dev.data_ptr(), slot=0, **opts)
        if ev:
            ev[2].record(stream)
        return div

    o = h.pcg_opts(**opts)
    it_c, rr_c, div_c = C.c_int(), C.c_double(), C.c_double()

    def step_e2e():
        L.check(h.lib.fk_picard_pcg(h.h, 0, host_ptr(u_host), 0, host_ptr(b_host), host_ptr(x_host), C.byref(o), C.byref(div_c),
                                    C.byref(it_c), C.byref(rr_c)))
        return div_c.value

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, with_events):
        for _ in range(warmup):
            fn() if not with_events else fn(None)
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)] if with_events else None
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = h.launch_count()
        e0.record(stream)
        for s in range(steps):
            fn(evs[s]) if with_events else fn()
        e1.record(stream)
        barrier()
        ms = e0.elapsed_time(e1)
        launches = h.launch_count() - l0
        if dist is not None:
            tt = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
        return ms, launches, evs

    W, K = max(3, args.warmup), args.steps
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_dev, launches, evs = timed(step_dev, K, W, True)
    ms_e2e, _, _ = timed(step_e2e, K, W, False)
    clocks = sampler.stop() if rank == 0 else {}

    asm_ms = statistics.mean(e[0].elapsed_time(e[1]) for e in evs)
    pcg_ms = statistics.mean(e[1].elapsed_time(e[2]) for e in evs)

    # untimed record: a complete nonlinear solve (Picard to mu0*max|dH| < 1e-4, PCG to 1e-10 each step)
    nonlin = None
    if args.full_solve:
        with torch.cuda.stream(stream):
            u2 = u_dev.clone()
        h.set_mu(np.full(mesh.nt, MU0))
        h.reset_field()
        t0 = time.time()
        tot_it, att, div = 0, 0, float("inf")
        while div > 1e-4 and att < 100:                      # NewtonRaphson.jl:37-38,148
            div = h.picard_step_dev(u2.data_ptr(), slot=0)
            it, rr = h.pcg_dev(b_dev.data_ptr(), u2.data_ptr(), slot=0, tol=1e-10, max_iter=20000, warm_start=True)
            tot_it += it
            att += 1
        h.synchronize()
        nonlin = {"picard_iters": att, "pcg_iters_total": tot_it, "seconds": round(time.time() - t0, 4), "div": div,
                  "last_relres": rr}

    world_tets = nt_owned
    world_rows = n_rows
    world_nnz = nnz
    if dist is not None:
        tt = torch.tensor([nt_owned, n_rows, nnz], dtype=torch.float64, device=dev)
        dist.all_reduce(tt)
        world_tets, world_rows, world_nnz = (int(v) for v in tt.tolist())

    if rank == 0:
        peak, peak_src = measured_peaks()
        ms_step = ms_dev / K
        value = world_tets / (ms_step * 1e-3) / 1e6
        e2e_value = world_tets / (ms_e2e / K * 1e-3) / 1e6
        sell = h.sell_stats()
        it_bytes = pcg_bytes(n_rows, nnz)                        # this rank's share, per iteration
        pcg_gbs = it_bytes * Piters / (pcg_ms * 1e-3) / 1e9
        asm_bytes = picard_bytes(nt, nv, nnz)
        asm_gbs = asm_bytes / (asm_ms * 1e-3) / 1e9
        traffic = {}
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            traffic = json.load(open(tpath))
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"configs[1]: nonlinear magnetostatics Picard step, Kuhn cube N={N} per GPU "
                                   f"({nt_owned} tets, {n_rows} rows, {nnz} nnz per GPU), Gd mu(H) table @270K, "
                                   f"{Piters} Jacobi-PCG iterations per step (fixed count, x0=0)",
                       "tets_total": world_tets, "rows_total": world_rows, "nnz_total": world_nnz, "pcg_iters_per_step": Piters,
                       "l2": "inputs larger than L2 (CSR values+indices %.0f MB per GPU vs 126 MB L2)" % (12 * nnz / 1e6),
                       "parallelism": "single GPU" if world == 1 else
                       f"node-partitioned slabs x{world}, " + ("halo rows + all-reduce over NVLink peer memory inside the PCG kernels"
                                                                if h.comm_mode() == 2 else "NCCL send/recv halo + ncclAllReduce")},
            "components": {"picard_assembly_ms": round(asm_ms, 4), "picard_assembly_mtets_per_s": round(nt / asm_ms / 1e3, 1),
                           "pcg_ms": round(pcg_ms, 4), "pcg_iters_per_s": round(Piters / (pcg_ms * 1e-3), 1),
                           "setup_s_untimed": round(setup_s, 2), "linear_presolve_iters": it0},
            "e2e": {"value": round(e2e_value, 3), "unit": UNIT, "ms_per_step": round(ms_e2e / K, 4),
                    "h2d_bytes_per_step": 16 * nv, "d2h_bytes_per_step": 8 * n_rows + 8 + 4 + 64},
            "gpu_launches": int(launches),
            "roofline": {"kernel": "fused Jacobi-PCG iteration (spmv_sell_kernel<DOT> + pcg_update_kernel + pcg_direction_kernel)",
                         "bound": "hbm", "achieved": round(pcg_gbs, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(pcg_gbs / peak, 4), "traffic": traffic.get("pcg_iteration_bytes"),
                         "peak_source": peak_src, "algorithmic_bytes_per_launch": it_bytes,
                         "avg_launch_ms": round(pcg_ms / Piters, 5),
                         "note": "achieved uses SURVEY 8d's common denominator (12 B per CSR entry + 108 B per row); the "
                                 "SpMV actually streams %.2f B per entry (packed 16-bit column offsets, %d of %d slices) and "
                                 "the PCG vectors are pinned in L2, so frac can exceed 1" % (
                                     8 + sell[3] / max(sell[0], 1), sell[2], sell[1]),
                         "layout_bytes_per_launch": int(8 * sell[0] + sell[3] + 144 * sell[1] + 108 * n_rows)},
            "roofline_assembly": {"kernel": "Picard sweep: material_update_kernel + assemble_kernel<ASM_PICARD> + max fold (fk_picard_step_dev, one host poll)", "bound": "hbm", "achieved": round(asm_gbs, 1),
                                  "peak": peak, "unit": "GB/s", "frac": round(asm_gbs / peak, 4),
                                  "traffic": traffic.get("picard_assembly_bytes"), "algorithmic_bytes_per_launch": asm_bytes,
                                  "avg_launch_ms": round(asm_ms, 5)},
            "clocks": clocks,
        }
        if nonlin:
            line["nonlinear_solve_untimed"] = nonlin
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, budget_s=args.cpu_budget)
        print(json.dumps(line), flush=True)
    h.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------------
# CPU arm: the compiled reference (oracle/_ref) -- or the C port when it is absent
# --------------------------------------------------------------------------------------------------
class CpuStep:
    """One Picard step on the host with the reference's own code where it exists:
    field recovery + spline (C port of magneticField.jl:111-132 / FITPACK splev: the reference has
    them only in Julia), stiffnessMatrix (src/FEMc.cpp:102-133, compiled unchanged) and Eigen's
    Jacobi-PCG (magnetostatics.cpp:26-38) for a fixed iteration count."""

    def __init__(self, N, piters):
        import femeko_jl_b200 as fk
        from oracle import oracle as orc

        self.orc, self.fk = orc, fk
        self.kind = "reference" if orc.have_ref() else "port"
        self.m = fk.kuhn_cube(N, inside_edge=0.2)
        self.piters = piters
        Hk, muk, _ = gd_table()
        self.spl = fk.Spline1D(Hk, muk, bc="nearest")
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

    def step(self):
        orc, m = self.orc, self.m
        _, Hn = orc.recover_field(m, self.u)
        mu = self.mu.copy()
        mu[m.InsideElements] = orc.splev(self.spl.t, self.spl.c, 3, Hn[m.InsideElements], "nearest")
        if self.kind == "reference":
            outer, inner, v = orc.ref_stiffness(m, mu)
            orc.ref_cg(outer, inner, v, self.b, 1e-300, self.piters)
        else:
            _, _, v = orc.stiffness(m, mu, order="eigen")
            orc.pcg(self.rp, self.ci, v, self.b, 0.0, self.piters)


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
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=119, help="Kuhn cells per edge per GPU (119 = configs[1], 10.1 M tets)")
    ap.add_argument("--pcg-iters", type=int, default=100)
    ap.add_argument("--cpu-n", type=int, default=48, help="Kuhn cells per edge of the bounded CPU sample")
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--tune", default="", help="experiments: comma-separated what=value pairs for fk_tune (e.g. 6=0)")
    ap.add_argument("--presolve", type=int, default=1, help="0: skip the untimed linear presolve (profiling runs)")
    ap.add_argument("--full-solve", type=int, default=1, help="also record one complete nonlinear solve (untimed)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
