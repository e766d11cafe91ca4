"""Host-side domain decomposition for 2/4/8 GPUs (one process per GPU).

The reference is a single process; this module is what the north star adds around the
hot path: nodes are split by recursive coordinate bisection (RCB), every rank renumbers its nodes
*owned first, then halo* (halo grouped by owning rank, ascending global id inside a group) and keeps
every element that touches an owned node (ghost elements are duplicated), so

  * assembly of the owned rows needs no communication and no atomics,
  * field recovery / mu(H) of the local elements is local once the halo entries of u are filled,
  * the only exchanges are the halo rows of the SpMV input and the PCG dot-product all-reduce
    (fk_set_halo / fk_comm_init in include/femeko_cuda.h).

`rcb` + `partition_mesh` work on any mesh held on the host; `kuhn_slab` generates one rank's share
of the weak-scaling benchmark box directly (no global mesh is ever materialised).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import os

import numpy as np

from .mesh import MESH, _box_surface, _finish, _orient_outward, box_inside_mask, kuhn_cells


def rcb(p: np.ndarray, nparts: int) -> np.ndarray:
    """Recursive coordinate bisection of the nodes (3 x nv coordinates) into `nparts` parts of
    (nearly) equal size: split the longest axis of the current bounding box at the weighted median,
    recurse.  Ties are broken by node id, so the result is deterministic.  Returns part id per node."""
    nv = p.shape[1]
    part = np.zeros(nv, dtype=np.int32)

    def split(idx, first, count):
        if count == 1:
            part[idx] = first
            return
        left_parts = count // 2
        ext = p[:, idx].max(axis=1) - p[:, idx].min(axis=1)
        ax = int(np.argmax(ext))
        order = np.lexsort((idx, p[ax, idx]))
        nleft = (idx.size * left_parts) // count
        split(idx[order[:nleft]], first, left_parts)
        split(idx[order[nleft:]], first + left_parts, count - left_parts)

    split(np.arange(nv, dtype=np.int64), 0, int(nparts))
    return part


@dataclass
class Partition:
    """One rank's view of a partitioned mesh (local ids: owned rows [0, n_owned), then halo rows)."""
    rank: int = 0
    world: int = 1
    n_owned: int = 0
    nt_owned: int = 0                 # local elements counted for this rank (owner = part of node t[0,k])
    l2g: np.ndarray = None            # local node -> global node id
    elem_l2g: np.ndarray = None       # local element -> global element id (None for generated slabs)
    neigh: list = field(default_factory=list)
    send_ptr: np.ndarray = None       # len(neigh)+1
    send_idx: np.ndarray = None       # local owned rows to send, per neighbour ascending global id
    recv_start: np.ndarray = None     # first local halo row filled by each neighbour
    recv_count: np.ndarray = None
    nv_global: int = 0

    # ---- device side -------------------------------------------------------------------------
    def attach(self, h, dist=None, unique_id: bytes | None = None):
        """Create the NCCL communicator of handle `h` and install the halo plan.  The 128-byte NCCL
        id comes from rank 0 through torch.distributed (any backend) unless given."""
        import ctypes as C

        from . import lib as L

        lib = L.load()
        if unique_id is None:
            buf = (C.c_ubyte * 128)()
            if self.rank == 0:
                L.check(lib.fk_nccl_unique_id(buf))
            import torch

            t = torch.tensor(list(buf), dtype=torch.uint8)
            if dist.get_backend() == "nccl":
                t = t.cuda()
            dist.broadcast(t, src=0)
            unique_id = bytes(t.cpu().tolist())
        idb = (C.c_ubyte * 128).from_buffer_copy(unique_id)
        L.check(lib.fk_comm_init(h.h, idb, self.rank, self.world))
        ne = np.ascontiguousarray(self.neigh, dtype=np.int32)
        sp = np.ascontiguousarray(self.send_ptr, dtype=np.int32)
        si = np.ascontiguousarray(self.send_idx, dtype=np.int32)
        rs = np.ascontiguousarray(self.recv_start, dtype=np.int32)
        rc = np.ascontiguousarray(self.recv_count, dtype=np.int32)
        L.check(lib.fk_set_halo(h.h, self.n_owned, len(self.neigh), L.iptr(ne), L.iptr(sp), L.iptr(si), L.iptr(rs), L.iptr(rc)))
        if self.world > 1 and dist is not None and not os.environ.get("FK_NO_P2P"):
            self._attach_p2p(h, dist)

    def _attach_p2p(self, h, dist):
        """Peer-memory path: all-gather the CUDA-IPC handles and every rank's receive plan, then map the
        peers (fk_comm_p2p_attach).  Set FK_NO_P2P=1 to stay on the NCCL send/recv + all-reduce path."""
        import ctypes as C

        import torch

        from . import lib as L

        lib = L.load()
        hb = (C.c_ubyte * 128)()
        L.check(lib.fk_comm_p2p_export(h.h, hb))
        mine = torch.tensor(list(hb), dtype=torch.uint8)
        if dist.get_backend() == "nccl":
            mine = mine.cuda()
        allh = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(allh, mine)
        plans = [None] * self.world
        dist.all_gather_object(plans, {"neigh": [int(v) for v in self.neigh], "recv_start": [int(v) for v in self.recv_start]})
        rstart, rslot = [], []
        for nb in self.neigh:
            qq = plans[int(nb)]["neigh"].index(self.rank)
            rslot.append(qq)
            rstart.append(plans[int(nb)]["recv_start"][qq])
        blob = np.concatenate([t.cpu().numpy() for t in allh]).astype(np.uint8)
        hs = (C.c_ubyte * blob.size).from_buffer_copy(blob.tobytes())
        a = np.ascontiguousarray(rstart, dtype=np.int32)
        b = np.ascontiguousarray(rslot, dtype=np.int32)
        L.check(lib.fk_comm_p2p_attach(h.h, hs, L.iptr(a), L.iptr(b)))
        dist.barrier()

    # ---- host side (any torch.distributed backend; used for set-up data and the gloo tests) ----
    def exchange_host(self, v: np.ndarray, dist) -> None:
        """Fill the halo rows of the local vector `v` (length >= n_owned + halo) from their owners."""
        import torch

        reqs, bufs = [], []
        for q, nb in enumerate(self.neigh):
            if self.recv_count[q] > 0:
                rb = torch.empty(int(self.recv_count[q]), dtype=torch.float64)
                bufs.append((q, rb))
                reqs.append(dist.irecv(rb, src=int(nb)))
            ids = self.send_idx[self.send_ptr[q]:self.send_ptr[q + 1]]
            if ids.size:
                reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(v[ids])), dst=int(nb)))
        for r in reqs:
            r.wait()
        for q, rb in bufs:
            s = int(self.recv_start[q])
            v[s:s + int(self.recv_count[q])] = rb.numpy()

    def to_global(self, v_local: np.ndarray, out: np.ndarray) -> None:
        out[self.l2g[:self.n_owned]] = v_local[:self.n_owned]


def _halo_plan(part_of_local: np.ndarray, l2g: np.ndarray, t_local: np.ndarray, rank: int, n_owned: int):
    """Neighbour lists from the local mesh alone.  Node a (owned here) is a halo node of rank q iff
    some element holds a together with a node owned by q; every such element is local here."""
    owner = part_of_local  # per local node
    halo_owner = owner[n_owned:]
    neigh = sorted(set(int(q) for q in np.unique(halo_owner)) | _send_neighbours(owner, t_local, rank))
    send_ptr, send_idx, recv_start, recv_count = [0], [], [], []
    eo = owner[t_local]                                    # 4 x nt_local
    for q in neigh:
        sel = (eo == q).any(axis=0)
        nodes = np.unique(t_local[:, sel])
        nodes = nodes[owner[nodes] == rank]
        nodes = nodes[np.argsort(l2g[nodes], kind="stable")]
        send_idx.append(nodes.astype(np.int32))
        send_ptr.append(send_ptr[-1] + nodes.size)
        pos = np.flatnonzero(halo_owner == q)
        recv_count.append(pos.size)
        recv_start.append(n_owned + (int(pos[0]) if pos.size else 0))
        if pos.size:
            assert pos[-1] - pos[0] + 1 == pos.size, "halo rows of one owner must be contiguous"
    send_idx = np.concatenate(send_idx) if send_idx else np.zeros(0, dtype=np.int32)
    return neigh, np.asarray(send_ptr, dtype=np.int32), send_idx, np.asarray(recv_start, dtype=np.int32), np.asarray(recv_count, dtype=np.int32)


def _send_neighbours(owner, t_local, rank):
    eo = owner[t_local]
    return set(int(q) for q in np.unique(eo) if q != rank)


def partition_mesh(mesh: MESH, part: np.ndarray, rank: int, world: int | None = None):
    """Local MESH + Partition of `rank` for the node partition `part` (part id per global node)."""
    world = int(part.max()) + 1 if world is None else world
    t = mesh.t
    owned = np.flatnonzero(part == rank)
    touch = (part[t] == rank).any(axis=0)
    elems = np.flatnonzero(touch)
    tl = t[:, elems]
    nodes = np.unique(tl)
    halo = nodes[part[nodes] != rank]
    halo = halo[np.lexsort((halo, part[halo]))]
    l2g = np.concatenate([owned, halo]).astype(np.int64)
    g2l = -np.ones(mesh.nv, dtype=np.int64)
    g2l[l2g] = np.arange(l2g.size)
    loc = MESH()
    loc.p = np.asfortranarray(mesh.p[:, l2g])
    loc.t = np.asfortranarray(g2l[tl].astype(np.int32))
    sT = mesh.surfaceT
    keep = (part[sT[:3]] == rank).any(axis=0)
    s = sT[:, keep].astype(np.int64)
    s[:3] = g2l[s[:3]]
    assert (s[:3] >= 0).all()
    loc.surfaceT = np.asfortranarray(s.astype(np.int32))
    loc.shell_id = getattr(mesh, "shell_id", [0])
    inside = np.zeros(mesh.nt, dtype=bool)
    inside[mesh.InsideElements] = True
    _finish(loc, inside[elems])
    loc.VE = mesh.VE[elems]
    loc.AE = mesh.AE[keep]
    loc.normal = np.asfortranarray(mesh.normal[:, keep])
    P = Partition(rank=rank, world=world, n_owned=int(owned.size), l2g=l2g, elem_l2g=elems, nv_global=mesh.nv)
    P.nt_owned = int((part[t[0, elems]] == rank).sum())
    P.neigh, P.send_ptr, P.send_idx, P.recv_start, P.recv_count = _halo_plan(part[l2g], l2g, loc.t, rank, P.n_owned)
    return loc, P


def slab_planes(N: int, world: int):
    """Node planes [a_r, a_{r+1}) of the (world*N + 1)-plane box owned by each rank."""
    total = world * N + 1
    return [(r * total) // world for r in range(world + 1)]


def slab_part(N: int, world: int) -> np.ndarray:
    """Node partition of kuhn_box((world*N, N, N)) into x-slabs (what RCB yields for that box)."""
    a = slab_planes(N, world)
    plane = np.repeat(np.arange(world * N + 1), (N + 1) * (N + 1))
    return (np.searchsorted(np.asarray(a[1:]), plane, side="right")).astype(np.int32)


def kuhn_slab(N: int, rank: int, world: int, inside_edge: float = 0.2):
    """Rank `rank`'s share of kuhn_box((world*N, N, N)) partitioned by slab_part, generated directly:
    identical (node order, element order, halo plan) to partition_mesh(kuhn_box(...), slab_part(...),
    rank) -- tests/test_partition.py checks that at small N."""
    n1 = N + 1
    pl = n1 * n1
    a = slab_planes(N, world)
    a0, a1 = a[rank], a[rank + 1]
    NX = world * N
    c0, c1 = max(a0 - 1, 0), min(a1, NX)              # cells touching an owned plane
    has_l, has_r = a0 > 0, a1 <= NX
    # global ids, then remap planes: owned planes first, halo planes after (left owner < right owner)
    tg = kuhn_cells(NX, c0, c1, Ny=N, Nz=N).astype(np.int64)
    n_owned = (a1 - a0) * pl
    plane = tg // pl
    rest = tg - plane * pl
    lp = plane - a0
    if has_l:
        lp = np.where(plane == a0 - 1, (a1 - a0), lp)
    if has_r:
        lp = np.where(plane == a1, (a1 - a0) + (1 if has_l else 0), lp)
    loc = MESH()
    loc.t = np.asfortranarray((lp * pl + rest).astype(np.int32))
    planes = list(range(a0, a1)) + ([a0 - 1] if has_l else []) + ([a1] if has_r else [])
    h = 1.0 / N
    g = (np.arange(n1, dtype=np.float64) - N / 2.0) * h
    gx = (np.asarray(planes, dtype=np.float64) - NX / 2.0) * h
    X, Y, Z = np.meshgrid(gx, g, g, indexing="ij")
    loc.p = np.asfortranarray(np.stack([X.ravel(), Y.ravel(), Z.ravel()]))
    l2g = (np.asarray(planes, dtype=np.int64)[:, None] * pl + np.arange(pl, dtype=np.int64)[None, :]).ravel()
    # outer surface: y/z faces over the local cells, x faces on the first / last rank
    faces = [2, 3, 4, 5] + ([0] if a0 == 0 else []) + ([1] if a1 == NX + 1 else [])
    sg = _box_surface((NX, N, N), True, x_range=(c0, c1), faces=tuple(sorted(faces)))
    centre_shift = np.zeros(3)
    pg_plane = sg[:3] // pl
    owned_any = ((pg_plane >= a0) & (pg_plane < a1)).any(axis=0)
    sg = sg[:, owned_any]
    sl = sg.copy()
    pp = sg[:3] // pl
    rr = sg[:3] - pp * pl
    lpp = pp - a0
    if has_l:
        lpp = np.where(pp == a0 - 1, (a1 - a0), lpp)
    if has_r:
        lpp = np.where(pp == a1, (a1 - a0) + (1 if has_l else 0), lpp)
    sl[:3] = lpp * pl + rr
    sl = _orient_outward(loc.p, sl, centre_shift)   # the box is convex and centred at the origin
    loc.surfaceT = np.asfortranarray(sl.astype(np.int32))
    loc.shell_id = [0]
    cen = (loc.p[:, loc.t[0]] + loc.p[:, loc.t[1]] + loc.p[:, loc.t[2]] + loc.p[:, loc.t[3]]) / 4
    _finish(loc, box_inside_mask(cen, NX * h, N * h, inside_edge))
    P = Partition(rank=rank, world=world, n_owned=int(n_owned), l2g=l2g, nv_global=(NX + 1) * pl)
    P.nt_owned = int((loc.t[0] < n_owned).sum())   # node t[0,k] is the cell's low corner: its owner counts the element
    owner = np.full(l2g.size, rank, dtype=np.int32)
    if has_l:
        owner[n_owned:n_owned + pl] = rank - 1
    if has_r:
        owner[n_owned + (pl if has_l else 0):] = rank + 1
    P.neigh, P.send_ptr, P.send_idx, P.recv_start, P.recv_count = _halo_plan(owner, l2g, loc.t, rank, P.n_owned)
    return loc, P


# --------------------------------------------------------------------------------------------------
# strong scaling: ONE global Kuhn cube cut into px x py x pz blocks of nodes (what RCB yields for a cube),
# every rank's share generated directly -- the global mesh (100 M tets at N = 256) is never materialised
# --------------------------------------------------------------------------------------------------
GRIDS = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}


def block_bounds(N: int, parts: int):
    """Node planes [a_r, a_{r+1}) of the N + 1 planes along one axis owned by block r."""
    return [(r * (N + 1)) // parts for r in range(parts + 1)]


def block_part(N: int, grid) -> np.ndarray:
    """Node partition of kuhn_cube(N) into grid = (px, py, pz) blocks; rank = (bx * py + by) * pz + bz."""
    n1 = N + 1
    b = [np.searchsorted(np.asarray(block_bounds(N, g)[1:]), np.arange(n1), side="right") for g in grid]
    part = (b[0][:, None, None] * grid[1] + b[1][None, :, None]) * grid[2] + b[2][None, None, :]
    return part.ravel().astype(np.int32)


def kuhn_block(N: int, rank: int, grid, inside_edge: float = 0.2):
    """Rank `rank`'s share of kuhn_cube(N) partitioned by block_part(N, grid), generated directly: identical (node
    order, element order, surface triangles, halo plan) to partition_mesh(kuhn_cube(N), block_part(N, grid), rank)
    -- tests/test_partition.py checks that at small N.  Work and memory are proportional to the rank's share."""
    from itertools import permutations

    px, py, pz = (int(g) for g in grid)
    world = px * py * pz
    n1 = N + 1
    bidx = (rank // (py * pz), (rank // pz) % py, rank % pz)
    bounds = [block_bounds(N, g) for g in (px, py, pz)]
    a0 = [bounds[d][bidx[d]] for d in range(3)]
    a1 = [bounds[d][bidx[d] + 1] for d in range(3)]
    c0 = [max(a0[d] - 1, 0) for d in range(3)]            # cells with a corner on an owned plane, per axis
    c1 = [min(a1[d], N) for d in range(3)]
    E = [c1[d] - c0[d] + 1 for d in range(3)]             # node box spanned by those cells
    ii, jj, kk = np.meshgrid(*[np.arange(c0[d], c1[d], dtype=np.int64) for d in range(3)], indexing="ij")
    ii, jj, kk = ii.ravel(), jj.ravel(), kk.ravel()
    ncell = ii.size
    # six tets per cell as (i, j, k) lattice paths, in the element order of kuhn_cells (cells lexicographic, 6 per cell)
    box = np.empty((ncell, 6, 4), dtype=np.int64)         # flat index inside the node box
    own = np.zeros((ncell, 6), dtype=bool)
    step = np.eye(3, dtype=np.int64)
    for q, perm in enumerate(permutations(range(3))):
        v = [ii.copy(), jj.copy(), kk.copy()]
        for m in range(4):
            if m > 0:
                ax = perm[m - 1]
                v[ax] = v[ax] + step[ax, ax]
            box[:, q, m] = ((v[0] - c0[0]) * E[1] + (v[1] - c0[1])) * E[2] + (v[2] - c0[2])
            own[:, q] |= ((v[0] >= a0[0]) & (v[0] < a1[0]) & (v[1] >= a0[1]) & (v[1] < a1[1]) & (v[2] >= a0[2]) & (v[2] < a1[2]))
    del ii, jj, kk
    box = box.reshape(ncell * 6, 4)[own.ravel()]          # elements touching an owned node, ascending global id
    nbox = E[0] * E[1] * E[2]
    used = np.zeros(nbox, dtype=bool)
    used[box.ravel()] = True
    bi, bj, bk = np.unravel_index(np.arange(nbox, dtype=np.int64), E)
    gi, gj, gk = bi + c0[0], bj + c0[1], bk + c0[2]
    owned_box = (gi >= a0[0]) & (gi < a1[0]) & (gj >= a0[1]) & (gj < a1[1]) & (gk >= a0[2]) & (gk < a1[2])
    assert used[owned_box].all()
    gid = (gi * n1 + gj) * n1 + gk
    blk = [np.searchsorted(np.asarray(bounds[d][1:]), g, side="right") for d, g in enumerate((gi, gj, gk))]
    owner_box = ((blk[0] * py + blk[1]) * pz + blk[2]).astype(np.int32)
    owned_idx = np.flatnonzero(owned_box)                  # box order == ascending global id
    halo_idx = np.flatnonzero(used & ~owned_box)
    halo_idx = halo_idx[np.lexsort((gid[halo_idx], owner_box[halo_idx]))]
    order = np.concatenate([owned_idx, halo_idx])
    lid = -np.ones(nbox, dtype=np.int64)
    lid[order] = np.arange(order.size)
    n_owned = int(owned_idx.size)
    l2g = gid[order]
    loc = MESH()
    loc.t = np.asfortranarray(lid[box].T.astype(np.int32))
    del box
    g = (np.arange(n1, dtype=np.float64) / N - 0.5) * 1.0
    loc.p = np.asfortranarray(np.stack([g[gi[order]], g[gj[order]], g[gk[order]]]))
    # outer surface: the global triangles (restricted to this rank's x range of cells) that hold an owned node
    sg = _box_surface((N, N, N), True, x_range=(c0[0], c1[0]))
    si, sj, sk = sg[:3] // (n1 * n1), (sg[:3] // n1) % n1, sg[:3] % n1
    s_own = ((si >= a0[0]) & (si < a1[0]) & (sj >= a0[1]) & (sj < a1[1]) & (sk >= a0[2]) & (sk < a1[2])).any(axis=0)
    sg, si, sj, sk = sg[:, s_own], si[:, s_own], sj[:, s_own], sk[:, s_own]
    sl = sg.copy()
    sl[:3] = lid[((si - c0[0]) * E[1] + (sj - c0[1])) * E[2] + (sk - c0[2])]
    assert (sl[:3] >= 0).all()
    sl = _orient_outward(loc.p, sl, np.zeros(3))
    loc.surfaceT = np.asfortranarray(sl.astype(np.int32))
    loc.shell_id = [0]
    cen = (loc.p[:, loc.t[0]] + loc.p[:, loc.t[1]] + loc.p[:, loc.t[2]] + loc.p[:, loc.t[3]]) / 4
    half = 0.5 * inside_edge
    _finish(loc, np.all(np.abs(cen) <= half, axis=0))
    P = Partition(rank=rank, world=world, n_owned=n_owned, l2g=l2g, nv_global=n1 ** 3)
    P.nt_owned = int((loc.t[0] < n_owned).sum())
    P.neigh, P.send_ptr, P.send_idx, P.recv_start, P.recv_count = _halo_plan(owner_box[order], l2g, loc.t, rank, n_owned)
    return loc, P
