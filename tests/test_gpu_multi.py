"""GPU (>= 2 devices): the partitioned solve -- halo exchange and all-reduce over peer memory (and over
NCCL with FK_NO_P2P=1) -- against a single-GPU solve of the same box.  Skipped on one-GPU boxes."""
import json
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:  # noqa: BLE001
        return 0


def _port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.gpu
@pytest.mark.parametrize("no_p2p,rec,halo", [("", 0, "2"), ("", 0, "1"), ("", 1, ""), ("1", 0, "")])
def test_partitioned_solve_matches_single_gpu(no_p2p, rec, halo):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2 if n < 4 else 4
    env = dict(os.environ, FK_NO_P2P=no_p2p, FK_RECURRENCE=str(rec), FK_HALO_MODE=halo)
    if not no_p2p:
        env.pop("FK_NO_P2P")
    if not halo:
        env.pop("FK_HALO_MODE")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(_port()), os.path.join(ROOT, "scripts", "mgpu_check.py"), "16"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert r.returncode == 0 and lines, r.stdout[-2000:] + r.stderr[-4000:]
    out = json.loads(lines[-1])
    assert out["ok"] and out["comm_mode"] == (1 if no_p2p else 2), out


@pytest.mark.gpu
@pytest.mark.parametrize("mesh", ["kuhn", "delaunay"])
def test_single_process_multi_gpu_matches_single_gpu(mesh):
    """fk_multi_* (ONE process, one host thread per GPU, peer memory, no NCCL / IPC): the global mesh goes in, the library
    partitions it; linear solve, Picard sweep with the real Gd table, second solve and the recovered field against a
    single-GPU handle on the same mesh."""
    import os

    import numpy as np

    import femeko_jl_b200 as fk

    n = _ngpu()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    ngpu = 2 if n < 4 else 4
    MU0 = 4e-7 * np.pi
    m = fk.kuhn_cube(14, inside_edge=0.4) if mesh == "kuhn" else fk.delaunay_ball(3000, seed=21)
    g = np.load(os.path.join(ROOT, "tests", "golden", "gd_270K.npz"))
    spl = fk.Spline1D(g["HofM"], g["mu"], bc="nearest")
    mu = np.full(m.nt, MU0)
    mu[m.InsideElements] = 3 * MU0

    hs = fk.Femeko(m, device=0)
    hs.set_spline(0, spl)
    hs.set_inside(m.InsideElements)
    hs.set_mu(mu)
    hs.assemble_stiffness(mu, slot=0, want_vals=False)
    b = -hs.boundary_integral(m.surfaceT, m.normal, m.AE, [0.5, 0.1, -0.2], 0)
    xs, its, _ = hs.pcg(b, slot=0, tol=1e-10, max_iter=5000)
    divs = hs.picard_step(xs * 2e5, slot=1)
    xs2, its2, _ = hs.pcg(b, slot=1, tol=1e-10, max_iter=5000, x0=xs)
    Hfs, Hns = hs.get_field()
    mus = hs.get_mu()
    ys = hs.spmv(xs, slot=1)
    hs.close()

    hm = fk.FemekoMulti(m, ngpu)
    inf = hm.info()
    assert int(inf["n_owned"].sum()) == m.nv and (inf["nneigh"] >= 1).all()
    hm.set_spline(0, spl)
    hm.set_inside(m.InsideElements)
    hm.set_mu(mu)
    hm.assemble_stiffness(mu, slot=0)
    x, it, _ = hm.pcg(b, slot=0, tol=1e-10, max_iter=5000)
    xb, itb, _ = hm.pcg(b, slot=0, tol=1e-10, max_iter=5000)
    assert np.array_equal(x, xb) and it == itb                      # run-to-run determinism across the GPUs
    div = hm.picard_step(xs * 2e5, slot=1)
    x2, it2, _ = hm.pcg(b, slot=1, tol=1e-10, max_iter=5000, x0=xs)
    Hf, Hn = hm.get_field()
    y = hm.spmv(xs, slot=1)
    rl2 = lambda a, c: float(np.linalg.norm(a - c) / np.linalg.norm(c))   # noqa: E731
    assert rl2(x, xs) < 1e-8 and abs(it - its) <= 2
    assert rl2(x2, xs2) < 1e-8 and abs(it2 - its2) <= 2
    assert abs(div - divs) <= 1e-9 * abs(divs)
    assert np.array_equal(hm.get_mu(), mus) and np.array_equal(Hn, Hns) and np.array_equal(Hf, Hfs)
    assert rl2(y, ys) < 1e-14
    hm.close()


@pytest.mark.gpu
@pytest.mark.parametrize("mesh", ["kuhn", "delaunay"])
def test_single_process_multi_gpu_demag_field(mesh):
    """fk_multi_demag_* (the LL loop's repeated solve, LL.cpp:287-348, on partitioned handles): node volumes bit for bit,
    demagnetising field and potential against a single-GPU handle and against the oracle on the same mesh."""
    import numpy as np

    import femeko_jl_b200 as fk
    from oracle import oracle as orc

    n = _ngpu()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    ngpu = 2 if n < 4 else 4
    m = fk.kuhn_cube(14, inside_edge=0.5) if mesh == "kuhn" else fk.delaunay_ball(3000, seed=22)
    ins = np.sort(m.InsideElements).astype(np.int32)
    insn = np.unique(m.t[:, ins])
    rng = np.random.default_rng(4)
    M = np.zeros((3, m.nv), order="F")
    M[:, insn] = rng.normal(size=(3, len(insn)))

    hs = fk.Femeko(m, device=0)
    vol_s = hs.demag_setup(ins, slot=0)
    Hs, its, us = hs.demag_field(M, want_u=True, tol=1e-12)
    hs.close()

    hm = fk.FemekoMulti(m, ngpu)
    vol = hm.demag_setup(ins, slot=0)
    assert np.array_equal(vol, vol_s) and np.array_equal(vol, orc.node_volumes(m, ins))
    Hd, it, u = hm.demag_field(M, want_u=True, tol=1e-12)
    Hd2, it2 = hm.demag_field(M, tol=1e-12)
    assert np.array_equal(Hd, Hd2, equal_nan=True) and it == it2        # run to run
    fin = np.isfinite(Hs)
    assert np.array_equal(fin, np.isfinite(Hd)) and abs(it - its) <= 2
    assert np.linalg.norm(Hd[fin] - Hs[fin]) < 1e-8 * np.linalg.norm(Hs[fin])
    assert np.linalg.norm((u - u.mean()) - (us - us.mean())) < 1e-8 * np.linalg.norm(us - us.mean())
    Ho = orc.demag_box(m, ins, u, vol)                                   # the box method of the oracle on the multi-GPU potential
    assert np.linalg.norm(Hd[fin] - Ho[fin]) < 1e-12 * np.linalg.norm(Ho[fin])
    Hz, _ = hm.demag_field(M, nan_outside=False, tol=1e-12)
    assert np.all(Hz[~fin] == 0.0)
    # warm start from the converged potential
    _, itw = hm.demag_field(M, tol=1e-10, warm_start=True)
    assert itw <= 2
    hm.close()
