"""CPU: the multi-GPU host logic (femeko.jl_b200/partition.py) -- RCB, owned/halo renumbering, ghost
elements, halo plan -- including a world_size-2 and -4 run over `gloo` in which every rank assembles
its owned rows with the oracle, exchanges halo rows with Partition.exchange_host and all-reduces the
PCG dot products; the distributed solution must equal the serial oracle solve."""
import os
import socket

import numpy as np
import pytest

import femeko_jl_b200 as fk
from conftest import rel_l2
from femeko_jl_b200 import partition as P
from femeko_jl_b200.mesh import kuhn_box
from oracle import oracle as orc


def test_rcb_is_balanced_and_deterministic():
    m = fk.delaunay_ball(1500, seed=6)
    for nparts in (2, 3, 4, 8):
        part = P.rcb(m.p, nparts)
        cnt = np.bincount(part, minlength=nparts)
        assert cnt.min() >= m.nv // nparts - 1 and cnt.max() <= m.nv // nparts + nparts
        assert np.array_equal(part, P.rcb(m.p, nparts))
    # two parts are separated by a plane orthogonal to the longest axis
    b = kuhn_box((8, 4, 4))
    part = P.rcb(b.p, 2)
    assert b.p[0, part == 0].max() <= b.p[0, part == 1].min()


@pytest.mark.parametrize("world", [1, 2, 3, 4])
def test_kuhn_slab_equals_partitioned_global_box(world):
    N = 4
    g = kuhn_box((world * N, N, N))
    part = P.slab_part(N, world)
    assert part.size == g.nv and set(np.unique(part)) == set(range(world))
    tot = 0
    for r in range(world):
        la, pa = P.partition_mesh(g, part, r, world)
        lb, pb = P.kuhn_slab(N, r, world)
        assert np.array_equal(pa.l2g, pb.l2g) and pa.n_owned == pb.n_owned and pa.nt_owned == pb.nt_owned
        assert np.array_equal(la.t, lb.t) and np.allclose(la.p, lb.p, atol=1e-15)
        assert np.array_equal(la.surfaceT, lb.surfaceT) and np.array_equal(la.InsideElements, lb.InsideElements)
        assert np.allclose(la.normal, lb.normal) and np.allclose(la.AE, lb.AE) and np.allclose(la.VE, lb.VE)
        assert pa.neigh == pb.neigh and np.array_equal(pa.send_idx, pb.send_idx) and np.array_equal(pa.send_ptr, pb.send_ptr)
        assert np.array_equal(pa.recv_start, pb.recv_start) and np.array_equal(pa.recv_count, pb.recv_count)
        tot += pa.nt_owned
    assert tot == g.nt


@pytest.mark.parametrize("N,grid", [(6, (2, 1, 1)), (7, (2, 2, 1)), (6, (2, 2, 2)), (5, (1, 1, 1))])
def test_kuhn_block_equals_partitioned_global_cube(N, grid):
    """The strong-scaling generator (one global cube, px x py x pz node blocks, every rank's share built directly)
    == partition_mesh of the materialised cube: node / element order, surface, inside set, halo plan."""
    g = fk.kuhn_cube(N, inside_edge=0.4)
    part = P.block_part(N, grid)
    world = grid[0] * grid[1] * grid[2]
    assert set(np.unique(part)) == set(range(world))
    tot = 0
    for r in range(world):
        la, pa = P.partition_mesh(g, part, r, world)
        lb, pb = P.kuhn_block(N, r, grid, inside_edge=0.4)
        assert np.array_equal(pa.l2g, pb.l2g) and pa.n_owned == pb.n_owned and pa.nt_owned == pb.nt_owned
        assert np.array_equal(la.t, lb.t) and np.array_equal(la.p, lb.p)
        assert np.array_equal(la.surfaceT, lb.surfaceT) and np.array_equal(la.InsideElements, lb.InsideElements)
        assert np.allclose(la.normal, lb.normal) and np.allclose(la.AE, lb.AE) and np.allclose(la.VE, lb.VE)
        assert list(pa.neigh) == list(pb.neigh) and np.array_equal(pa.send_idx, pb.send_idx) and np.array_equal(pa.send_ptr, pb.send_ptr)
        assert np.array_equal(pa.recv_start, pb.recv_start) and np.array_equal(pa.recv_count, pb.recv_count)
        tot += pa.nt_owned
    assert tot == g.nt
    if world == 8:   # Kuhn adjacency is directional: the corner block (0,0,0) talks to all seven others
        assert len(P.kuhn_block(N, 0, grid)[1].neigh) == 7


@pytest.mark.parametrize("mesh,nparts", [("delaunay", 4), ("kuhn", 8)])
def test_halo_plan_is_consistent_across_ranks(mesh, nparts):
    m = fk.delaunay_ball(1200, seed=9) if mesh == "delaunay" else fk.kuhn_cube(6)
    part = P.rcb(m.p, nparts)
    locs = [P.partition_mesh(m, part, r, nparts) for r in range(nparts)]
    owned_total = 0
    for r, (lm, pr) in enumerate(locs):
        owned_total += pr.n_owned
        assert np.all(part[pr.l2g[:pr.n_owned]] == r) and np.all(part[pr.l2g[pr.n_owned:]] != r)
        # every element touching an owned node is present, exactly once
        assert np.array_equal(pr.elem_l2g, np.flatnonzero((part[m.t] == r).any(axis=0)))
        for q, nb in enumerate(pr.neigh):
            other = locs[nb][1]
            qq = other.neigh.index(r)
            sent = pr.l2g[pr.send_idx[pr.send_ptr[q]:pr.send_ptr[q + 1]]]
            recv = other.l2g[other.recv_start[qq]:other.recv_start[qq] + other.recv_count[qq]]
            assert np.array_equal(sent, recv)                  # same nodes, same order, on both sides
        # halo rows are covered exactly by the receive ranges
        cover = np.zeros(lm.nv - pr.n_owned, dtype=int)
        for s, c in zip(pr.recv_start, pr.recv_count):
            cover[s - pr.n_owned:s - pr.n_owned + c] += 1
        assert np.all(cover == 1)
    assert owned_total == m.nv


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _dist_worker(rank, world, port, kind, out):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        if kind == "slab":
            lm, pr = P.kuhn_slab(5, rank, world, inside_edge=0.4)
        else:
            g = fk.delaunay_ball(900, seed=12)
            lm, pr = P.partition_mesh(g, P.rcb(g.p, world), rank, world)
        mu = np.full(lm.nt, fk.mu0)
        mu[lm.InsideElements] = 3 * fk.mu0
        rp, ci, v = orc.stiffness(lm, mu, order="eigen")          # local rows; only the owned ones are complete
        n = pr.n_owned
        b = -orc.boundary_integral(lm, [1.0, 0.2, -0.3], [0], "julia")[:n]
        rows = np.repeat(np.arange(lm.nv), np.diff(rp))
        own = rows < n
        import scipy.sparse as sp

        A = sp.csr_matrix((v[own], (rows[own], ci[own])), shape=(n, lm.nv))
        dinv = 1.0 / A.diagonal()

        def allsum(x):
            t = torch.tensor([x], dtype=torch.float64)
            dist.all_reduce(t)
            return float(t[0])

        # Eigen's recurrence (ConjugateGradient.h:59-82) with the two exchanges of SURVEY 8e
        x = np.zeros(n)
        r = b.copy()
        p = np.zeros(lm.nv)
        p[:n] = dinv * r
        rhs2 = allsum(b @ b)
        absNew = allsum(r @ p[:n])
        it = 0
        while it < 2000:
            pr.exchange_host(p, dist)
            Ap = A @ p
            alpha = absNew / allsum(p[:n] @ Ap)
            x += alpha * p[:n]
            r -= alpha * Ap
            it += 1
            if allsum(r @ r) < (1e-10 ** 2) * rhs2:
                break
            z = dinv * r
            absOld, absNew = absNew, allsum(r @ z)
            p[:n] = z + (absNew / absOld) * p[:n]
        res = torch.zeros(pr.nv_global, dtype=torch.float64)
        res[torch.from_numpy(pr.l2g[:n])] = torch.from_numpy(x)
        dist.all_reduce(res)
        if rank == 0:
            np.save(out, np.concatenate([[it], res.numpy()]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind,world", [("slab", 2), ("rcb", 2), ("rcb", 4)])
def test_distributed_pcg_over_gloo_matches_serial_oracle(kind, world, tmp_path):
    import torch.multiprocessing as mp

    out = str(tmp_path / "x.npy")
    mp.spawn(_dist_worker, args=(world, _free_port(), kind, out), nprocs=world, join=True)
    got = np.load(out)
    it, x = int(got[0]), got[1:]
    g = kuhn_box((world * 5, 5, 5), inside_edge=0.4) if kind == "slab" else fk.delaunay_ball(900, seed=12)
    mu = np.full(g.nt, fk.mu0)
    mu[g.InsideElements] = 3 * fk.mu0
    rp, ci, v = orc.stiffness(g, mu, order="eigen")
    b = -orc.boundary_integral(g, [1.0, 0.2, -0.3], [0], "julia")
    xs, its, _, info = orc.pcg(rp, ci, v, b, 1e-10)
    assert info == 0 and abs(it - its) <= 2
    assert rel_l2(x, xs) < 1e-8


@pytest.mark.parametrize("mesh,nparts", [("delaunay", 2), ("delaunay", 4), ("kuhn", 8), ("kuhn", 3)])
def test_cxx_partitioner_equals_python(mesh, nparts):
    """The C++ decomposition behind fk_multi_create (csrc/multi.cu: fk_rcb, fk_partition_local) == partition.py:
    node partition, local numbering (owned ascending, halo by owner then id), ghost elements, halo plan."""
    import ctypes as C

    from femeko_jl_b200 import lib as L

    lib = L.load()
    m = fk.delaunay_ball(900, seed=4) if mesh == "delaunay" else fk.kuhn_cube(5)
    p = np.asfortranarray(m.p, dtype=np.float64)
    t = np.asfortranarray(m.t, dtype=np.int32)
    part = np.zeros(m.nv, dtype=np.int32)
    L.check(lib.fk_rcb(L.dptr(p), m.nv, nparts, L.iptr(part)))
    assert np.array_equal(part, P.rcb(m.p, nparts))
    for r in range(nparts):
        lm, pr = P.partition_mesh(m, part, r, nparts)
        sizes = np.zeros(5, dtype=np.int32)
        none = [None] * 8
        L.check(lib.fk_partition_local(L.dptr(p), L.iptr(t), m.nv, m.nt, L.iptr(part), r, L.iptr(sizes), *none))
        n_owned, nvl, ntl, nn, nsend = (int(v) for v in sizes)
        assert (n_owned, nvl, ntl, nn) == (pr.n_owned, lm.nv, lm.nt, len(pr.neigh))
        l2g, e2g, tl = np.zeros(nvl, np.int32), np.zeros(ntl, np.int32), np.zeros((4, ntl), np.int32, order="F")
        neigh, sp, si = np.zeros(nn, np.int32), np.zeros(nn + 1, np.int32), np.zeros(nsend, np.int32)
        rs, rc = np.zeros(nn, np.int32), np.zeros(nn, np.int32)
        L.check(lib.fk_partition_local(L.dptr(p), L.iptr(t), m.nv, m.nt, L.iptr(part), r, L.iptr(sizes), L.iptr(l2g), L.iptr(e2g),
                                       L.iptr(tl), L.iptr(neigh), L.iptr(sp), L.iptr(si), L.iptr(rs), L.iptr(rc)))
        assert np.array_equal(l2g, pr.l2g) and np.array_equal(e2g, pr.elem_l2g) and np.array_equal(tl, lm.t)
        assert list(neigh) == [int(q) for q in pr.neigh] and np.array_equal(sp, pr.send_ptr) and np.array_equal(si, pr.send_idx)
        assert np.array_equal(rs, pr.recv_start) and np.array_equal(rc, pr.recv_count)
