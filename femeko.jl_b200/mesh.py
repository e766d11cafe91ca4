"""Host-side mirror of the reference mesh container and synthetic mesh generators.

`MESH` mirrors `mutable struct MESH` (reference src/mesh.jl:9-59): same field names, same memory
layout (Julia column-major ``3 x nv`` / ``4 x nt`` matrices are numpy Fortran-ordered arrays of the
same shape, so ``mesh.p[:, n]`` reads like the Julia code and ``mesh.p.ctypes`` is the pointer the
reference passes through ``@ccall``, Femeko_jl_C++/Magnetostatics/run.jl:66-79).

One deliberate difference: node / element indices are **0-based** here, i.e. already in the form
the reference hands to C (``t = mesh.t .- 1``, run.jl:50-52).

Gmsh is not available on the box, so meshes come from generators instead of
``Mesh(cells, meshSize, localSize)`` (src/mesh.jl:62-245):

* :func:`kuhn_cube`      structured cube, 6 tets per cell (Kuhn / Freudenthal subdivision);
* :func:`delaunay_ball`  unstructured "Gmsh-like" stand-in (scipy.spatial.Delaunay, seeded).

Geometry helpers restate src/mesh.jl: elementVolume :484-505, normalSurface :436-452,
areaTriangle :474-481.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from itertools import permutations

import numpy as np

mu0 = np.pi * 4e-7  # vacuum permeability, as in every reference driver (magneticField.jl:38)


@dataclass
class MESH:
    order: int = 1
    p: np.ndarray = None            # 3 x nv float64, Fortran order
    t: np.ndarray = None            # 4 x nt int32, Fortran order, 0-based node ids
    surfaceT: np.ndarray = None     # 4 x ns int32: 3 nodes (0-based) + boundary-surface id
    elementID: np.ndarray = None    # nt int32, parent cell id
    InsideElements: np.ndarray = None   # int32 list (0-based)
    InsideNodes: np.ndarray = None      # int32 list (0-based)
    VE: np.ndarray = None           # nt float64
    AE: np.ndarray = None           # ns float64
    normal: np.ndarray = None       # 3 x ns float64
    nv: int = 0
    nt: int = 0
    ns: int = 0
    ne: int = 0
    nInside: int = 0
    nInsideNodes: int = 0
    shell_id: list = field(default_factory=list)
    edges: np.ndarray = None        # order-2 meshes: mid-node label of every edge, in order of first appearance (src/mesh.jl:171)
    edge2localMap: np.ndarray = None    # mid-node label -> edge label (0-based; -1 where the label is not a mid node)


def elementVolume(p: np.ndarray, t: np.ndarray) -> np.ndarray:
    """|AB . (AC x AD)| / 6 for every element (src/mesh.jl:484-505), same operation order."""
    A = p[:, t[0]]
    AB = p[:, t[1]] - A
    AC = p[:, t[2]] - A
    AD = p[:, t[3]] - A
    c0 = AC[1] * AD[2] - AC[2] * AD[1]
    c1 = AC[2] * AD[0] - AC[0] * AD[2]
    c2 = AC[0] * AD[1] - AC[1] * AD[0]
    triple = AB[0] * c0 + AB[1] * c1 + AB[2] * c2
    return np.abs(triple) / 6.0


def _cross(v1, v2):
    return np.stack([v1[1] * v2[2] - v1[2] * v2[1],
                     v1[2] * v2[0] - v1[0] * v2[2],
                     v1[0] * v2[1] - v1[1] * v2[0]])


def normalSurface(p: np.ndarray, tri: np.ndarray) -> np.ndarray:
    """Unit normal (p2-p1) x (p3-p1) of every surface triangle (src/mesh.jl:436-452)."""
    p1 = p[:, tri[0]]
    n = _cross(p[:, tri[1]] - p1, p[:, tri[2]] - p1)
    return n / np.sqrt(n[0] ** 2 + n[1] ** 2 + n[2] ** 2)


def areaTriangle(p: np.ndarray, tri: np.ndarray) -> np.ndarray:
    """0.5*sqrt(det([x;y;1])^2 + det([y;z;1])^2 + det([z;x;1])^2) (src/mesh.jl:474-481)."""
    x, y, z = (p[d][tri[:3]] for d in range(3))

    def det3(a, b):  # det([a'; b'; 1 1 1])
        return (a[0] * (b[1] - b[2]) - a[1] * (b[0] - b[2]) + a[2] * (b[0] - b[1]))

    return 0.5 * np.sqrt(det3(x, y) ** 2 + det3(y, z) ** 2 + det3(z, x) ** 2)


def _finish(mesh: MESH, inside_mask: np.ndarray) -> MESH:
    """Fill the derived fields exactly as Mesh() does (src/mesh.jl:188-230)."""
    mesh.nv = int(mesh.p.shape[1])
    mesh.nt = int(mesh.t.shape[1])
    mesh.ns = int(mesh.surfaceT.shape[1])
    mesh.elementID = np.where(inside_mask, 1, 2).astype(np.int32)   # cell 1 = material, 2 = container
    mesh.InsideElements = np.flatnonzero(inside_mask).astype(np.int32)
    mesh.nInside = int(mesh.InsideElements.size)
    mesh.InsideNodes = np.unique(mesh.t[:, mesh.InsideElements]).astype(np.int32)
    mesh.nInsideNodes = int(mesh.InsideNodes.size)
    mesh.VE = elementVolume(mesh.p, mesh.t)
    mesh.AE = areaTriangle(mesh.p, mesh.surfaceT)
    mesh.normal = np.asfortranarray(normalSurface(mesh.p, mesh.surfaceT))
    return mesh


def kuhn_cells(N: int, i0: int = 0, i1: int | None = None, stride: tuple | None = None, Ny: int | None = None,
               Nz: int | None = None) -> np.ndarray:
    """Connectivity (4 x 6*ncells, int32, F-order) of the Kuhn subdivision of the cells with x-index
    in [i0, i1) of a lattice with N x Ny x Nz cells (Ny, Nz default to N).  Node id of lattice point
    (i, j, k) is ``(i*(Ny+1) + j)*(Nz+1) + k`` unless `stride` (si, sj, sk) overrides it.  The six
    tets of a cell are the six monotone lattice paths from corner (i,j,k) to (i+1,j+1,k+1), one per
    permutation of the axis order."""
    if i1 is None:
        i1 = N
    Ny = N if Ny is None else Ny
    Nz = N if Nz is None else Nz
    si, sj, sk = stride if stride is not None else ((Ny + 1) * (Nz + 1), Nz + 1, 1)
    ii, jj, kk = np.meshgrid(np.arange(i0, i1, dtype=np.int64), np.arange(Ny, dtype=np.int64),
                             np.arange(Nz, dtype=np.int64), indexing="ij")
    base = (ii * si + jj * sj + kk * sk).ravel()
    steps = (si, sj, sk)
    ncell = base.size
    t = np.empty((ncell, 6, 4), dtype=np.int32)
    for q, perm in enumerate(permutations(range(3))):
        v = base.copy()
        t[:, q, 0] = v
        for m, ax in enumerate(perm):
            v = v + steps[ax]
            t[:, q, m + 1] = v
    return np.asfortranarray(t.reshape(ncell * 6, 4).T)


def _box_surface(dims, single_id: bool, x_range=None, faces=(0, 1, 2, 3, 4, 5)) -> np.ndarray:
    """Boundary triangles (global lattice ids) of the Kuhn box with dims = (Nx, Ny, Nz) cells: every
    face lattice square is cut along the diagonal from its low corner to its high corner (the trace
    of the Kuhn subdivision).  `x_range` = (c0, c1) restricts the y/z faces to cells c0 <= i < c1;
    `faces` selects which of the six faces (-x, +x, -y, +y, -z, +z) to emit."""
    n = [d + 1 for d in dims]
    out = []
    sid = 0
    for axis in range(3):
        o = [d for d in range(3) if d != axis]
        r0 = (0, dims[o[0]])
        if x_range is not None and o[0] == 0:
            r0 = x_range
        a, b = np.meshgrid(np.arange(r0[0], r0[1], dtype=np.int64), np.arange(dims[o[1]], dtype=np.int64), indexing="ij")
        a = a.ravel()
        b = b.ravel()
        for side in (0, dims[axis]):
            if sid in faces:
                def nid(u, v):
                    idx = [None, None, None]
                    idx[axis] = np.full_like(u, side)
                    idx[o[0]] = u
                    idx[o[1]] = v
                    return (idx[0] * n[1] + idx[1]) * n[2] + idx[2]
                c00, c10, c01, c11 = nid(a, b), nid(a + 1, b), nid(a, b + 1), nid(a + 1, b + 1)
                ident = np.full_like(c00, 0 if single_id else sid)
                out.append(np.stack([c00, c10, c11, ident]))
                out.append(np.stack([c00, c11, c01, ident]))
            sid += 1
    if not out:
        return np.zeros((4, 0), dtype=np.int64)
    return np.concatenate(out, axis=1)


def _cube_surface(N: int, single_id: bool) -> np.ndarray:
    return _box_surface((N, N, N), single_id)


def _orient_outward(p: np.ndarray, surfaceT: np.ndarray, centre: np.ndarray) -> np.ndarray:
    """Swap two nodes where the triangle normal points towards `centre` (Gmsh boundary triangles
    of a closed volume are outward oriented; normalSurface relies on that)."""
    tri = surfaceT[:3]
    n = _cross(p[:, tri[1]] - p[:, tri[0]], p[:, tri[2]] - p[:, tri[0]])
    c = (p[:, tri[0]] + p[:, tri[1]] + p[:, tri[2]]) / 3.0 - centre[:, None]
    flip = (n * c).sum(axis=0) < 0
    s = surfaceT.copy()
    s[1, flip], s[2, flip] = surfaceT[2, flip], surfaceT[1, flip]
    return s


def kuhn_cube(N: int, L: float = 1.0, inside_edge: float = 0.2, inside_box=None,
              single_shell_id: bool = True) -> MESH:
    """Cube [-L/2, L/2]^3 with N cells per edge, 6N^3 tets, (N+1)^3 nodes.

    Inside region ("magnet"): elements whose centroid lies in the centred box of edge
    `inside_edge`*L (or half-extents `inside_box`), mirroring the material cube inside the
    container of the reference drivers (NewtonRaphson.jl:75-79, magneticField.jl:57-58)."""
    n1 = N + 1
    g = (np.arange(n1, dtype=np.float64) / N - 0.5) * L
    X, Y, Z = np.meshgrid(g, g, g, indexing="ij")
    mesh = MESH()
    mesh.p = np.asfortranarray(np.stack([X.ravel(), Y.ravel(), Z.ravel()]))
    mesh.t = kuhn_cells(N)
    s = _cube_surface(N, single_shell_id)
    s = _orient_outward(mesh.p, s, np.zeros(3))
    mesh.surfaceT = np.asfortranarray(s.astype(np.int32))
    mesh.shell_id = [0] if single_shell_id else list(range(6))
    cen = (mesh.p[:, mesh.t[0]] + mesh.p[:, mesh.t[1]] + mesh.p[:, mesh.t[2]] + mesh.p[:, mesh.t[3]]) / 4
    half = np.asarray(inside_box if inside_box is not None else [0.5 * inside_edge * L] * 3)
    inside = np.all(np.abs(cen) <= half[:, None], axis=0)
    return _finish(mesh, inside)


def kuhn_box(dims, h: float | None = None, inside_edge: float = 0.2, single_shell_id: bool = True) -> MESH:
    """Box of dims = (Nx, Ny, Nz) Kuhn cells of edge h (default 1/Ny), centred at the origin: the
    multi-GPU benchmark domain is `world` unit cubes side by side along x.  Inside region: in every
    unit-cube period along x, the elements whose centroid lies in the centred cube of edge
    `inside_edge` (one "magnet" per period, so partitions carry equal nonlinear work)."""
    Nx, Ny, Nz = (int(d) for d in dims)
    h = 1.0 / Ny if h is None else h
    gx = (np.arange(Nx + 1, dtype=np.float64) - Nx / 2.0) * h
    gy = (np.arange(Ny + 1, dtype=np.float64) - Ny / 2.0) * h
    gz = (np.arange(Nz + 1, dtype=np.float64) - Nz / 2.0) * h
    X, Y, Z = np.meshgrid(gx, gy, gz, indexing="ij")
    mesh = MESH()
    mesh.p = np.asfortranarray(np.stack([X.ravel(), Y.ravel(), Z.ravel()]))
    mesh.t = kuhn_cells(Nx, Ny=Ny, Nz=Nz)
    s = _box_surface((Nx, Ny, Nz), single_shell_id)
    s = _orient_outward(mesh.p, s, np.zeros(3))
    mesh.surfaceT = np.asfortranarray(s.astype(np.int32))
    mesh.shell_id = [0] if single_shell_id else list(range(6))
    cen = (mesh.p[:, mesh.t[0]] + mesh.p[:, mesh.t[1]] + mesh.p[:, mesh.t[2]] + mesh.p[:, mesh.t[3]]) / 4
    return _finish(mesh, box_inside_mask(cen, Nx * h, Ny * h, inside_edge))


def box_inside_mask(cen: np.ndarray, Lx: float, Ly: float, inside_edge: float) -> np.ndarray:
    """Periodic magnets along x: period Ly (one per unit cube), centred cube of edge inside_edge*Ly."""
    xs = cen[0] + Lx / 2.0
    xr = xs - np.floor(xs / Ly) * Ly - Ly / 2.0
    half = 0.5 * inside_edge * Ly
    return (np.abs(xr) <= half) & (np.abs(cen[1]) <= half) & (np.abs(cen[2]) <= half)


def delaunay_ball(npts: int = 4000, seed: int = 1234, R: float = 5.0, inside_half: float = 0.5,
                  dense_fraction: float = 0.4, min_quality: float = 1e-3) -> MESH:
    """Unstructured stand-in for the Gmsh "unit cube in a radius-5 sphere" meshes
    (NewtonRaphson.jl:75-79): Delaunay tetrahedralisation of seeded random points in the ball with a
    denser cloud around the material cube.  Slivers (V / Lmax^3 < min_quality; a regular tet has
    0.118) are dropped, as Gmsh's optimiser would: on a sliver the reference's own 4x4 solve
    (abcd) loses cond*eps digits, so no 1e-12 comparison is meaningful there.  Every boundary face
    (outer hull and the walls of the cavities left by dropped slivers) carries surface id 0 = the
    shell."""
    from scipy.spatial import Delaunay

    rng = np.random.default_rng(seed)
    n_dense = int(npts * dense_fraction)
    n_shell = max(64, int(npts * 0.15))
    n_bulk = npts - n_dense - n_shell
    d = rng.normal(size=(n_bulk, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    bulk = d * (R * rng.random(n_bulk) ** (1 / 3))[:, None]
    dense = (rng.random((n_dense, 3)) - 0.5) * (4 * inside_half)
    d = rng.normal(size=(n_shell, 3))
    shell = R * d / np.linalg.norm(d, axis=1)[:, None]
    pts = np.concatenate([dense, bulk, shell])
    tri = Delaunay(pts)
    t = tri.simplices.T.astype(np.int32)
    mesh = MESH()
    mesh.p = np.asfortranarray(pts.T.copy())
    V = elementVolume(mesh.p, t)
    Lmax = np.zeros(t.shape[1])
    for a in range(4):
        for b in range(a + 1, 4):
            Lmax = np.maximum(Lmax, np.linalg.norm(mesh.p[:, t[a]] - mesh.p[:, t[b]], axis=0))
    keep = V > min_quality * Lmax ** 3
    t = t[:, keep]
    # drop nodes no kept element references, keep ids contiguous (src/mesh.jl:120-121 assumes 1..nv)
    used = np.unique(t)
    remap = -np.ones(pts.shape[0], dtype=np.int64)
    remap[used] = np.arange(used.size)
    mesh.p = np.asfortranarray(mesh.p[:, used])
    mesh.t = np.asfortranarray(remap[t].astype(np.int32))
    # boundary faces = faces referenced by exactly one kept element; oriented away from that
    # element's opposite vertex, so the boundary (outer hull + cavity walls) is a union of closed,
    # outward-oriented surfaces and the Neumann right-hand side stays consistent
    nt = mesh.t.shape[1]
    faces = np.concatenate([mesh.t[[1, 2, 3]], mesh.t[[0, 2, 3]], mesh.t[[0, 1, 3]], mesh.t[[0, 1, 2]]], axis=1)
    opp = np.concatenate([mesh.t[0], mesh.t[1], mesh.t[2], mesh.t[3]])
    fs = np.sort(faces, axis=0)
    _, idx, cnt = np.unique(fs, axis=1, return_index=True, return_counts=True)
    sel = idx[cnt == 1]
    bf = faces[:, sel].astype(np.int64)
    nrm = _cross(mesh.p[:, bf[1]] - mesh.p[:, bf[0]], mesh.p[:, bf[2]] - mesh.p[:, bf[0]])
    inward = (nrm * (mesh.p[:, opp[sel]] - mesh.p[:, bf[0]])).sum(axis=0) > 0
    bf[1, inward], bf[2, inward] = bf[2, inward].copy(), bf[1, inward].copy()
    s = np.concatenate([bf, np.zeros((1, bf.shape[1]), dtype=np.int64)])
    mesh.surfaceT = np.asfortranarray(s.astype(np.int32))
    mesh.shell_id = [0]
    cen = (mesh.p[:, mesh.t[0]] + mesh.p[:, mesh.t[1]] + mesh.p[:, mesh.t[2]] + mesh.p[:, mesh.t[3]]) / 4
    inside = np.all(np.abs(cen) <= inside_half, axis=0)
    return _finish(mesh, inside)


# --------------------------------------------------------------------------------------------------
# 2-D triangle meshes (src/FEM2D.jl operates on MESH with t = 3 x nt, VE = triangle areas)
# --------------------------------------------------------------------------------------------------
def _finish2d(mesh: MESH, inside_mask: np.ndarray) -> MESH:
    p, t = mesh.p, mesh.t
    a, b, c = p[:2, t[0]], p[:2, t[1]], p[:2, t[2]]
    det = (b[0] - a[0]) * (c[1] - a[1]) - (c[0] - a[0]) * (b[1] - a[1])
    mesh.VE = np.abs(det) / 2.0
    mesh.nv, mesh.nt = int(p.shape[1]), int(t.shape[1])
    mesh.InsideElements = np.flatnonzero(inside_mask).astype(np.int32)
    mesh.nInside = int(mesh.InsideElements.size)
    mesh.InsideNodes = np.unique(t[:, mesh.InsideElements]).astype(np.int32)
    mesh.nInsideNodes = int(mesh.InsideNodes.size)
    mesh.surfaceT = np.zeros((4, 0), dtype=np.int32, order="F")
    mesh.ns = 0
    return mesh


def tri_grid(N: int, L: float = 1.0, inside_edge: float = 0.4) -> MESH:
    """Square [-L/2, L/2]^2 with N cells per edge, each cut along its low-to-high diagonal: 2 N^2 triangles,
    (N+1)^2 nodes, node id i (N+1) + j; p is 3 x nv with z = 0 (Gmsh hands 2-D meshes over that way)."""
    n1 = N + 1
    g = (np.arange(n1, dtype=np.float64) / N - 0.5) * L
    X, Y = np.meshgrid(g, g, indexing="ij")
    mesh = MESH()
    mesh.p = np.asfortranarray(np.stack([X.ravel(), Y.ravel(), np.zeros(n1 * n1)]))
    ii, jj = np.meshgrid(np.arange(N, dtype=np.int64), np.arange(N, dtype=np.int64), indexing="ij")
    c00 = (ii * n1 + jj).ravel()
    c10, c01, c11 = c00 + n1, c00 + 1, c00 + n1 + 1
    t = np.empty((c00.size, 2, 3), dtype=np.int32)
    t[:, 0] = np.stack([c00, c10, c11], axis=1)
    t[:, 1] = np.stack([c00, c11, c01], axis=1)
    mesh.t = np.asfortranarray(t.reshape(-1, 3).T)
    cen = (mesh.p[:2, mesh.t[0]] + mesh.p[:2, mesh.t[1]] + mesh.p[:2, mesh.t[2]]) / 3
    return _finish2d(mesh, np.all(np.abs(cen) <= 0.5 * inside_edge * L, axis=0))


def delaunay_disc(npts: int = 2000, seed: int = 7, R: float = 1.0, inside_half: float = 0.3) -> MESH:
    """Unstructured stand-in for a Gmsh 2-D mesh: Delaunay triangulation of random points in a disc (slivers removed)."""
    from scipy.spatial import Delaunay

    rng = np.random.default_rng(seed)
    r = R * np.sqrt(rng.uniform(0, 1, npts))
    th = rng.uniform(0, 2 * np.pi, npts)
    pts = np.stack([r * np.cos(th), r * np.sin(th)], axis=1)
    tri = Delaunay(pts).simplices.astype(np.int32)
    a, b, c = pts[tri[:, 0]], pts[tri[:, 1]], pts[tri[:, 2]]
    area = 0.5 * np.abs((b[:, 0] - a[:, 0]) * (c[:, 1] - a[:, 1]) - (c[:, 0] - a[:, 0]) * (b[:, 1] - a[:, 1]))
    tri = tri[area > 1e-10 * area.mean()]
    used = np.unique(tri)
    remap = -np.ones(npts, dtype=np.int64)
    remap[used] = np.arange(used.size)
    mesh = MESH()
    mesh.p = np.asfortranarray(np.vstack([pts[used].T, np.zeros((1, used.size))]))
    mesh.t = np.asfortranarray(remap[tri].T.astype(np.int32))
    cen = (mesh.p[:2, mesh.t[0]] + mesh.p[:2, mesh.t[1]] + mesh.p[:2, mesh.t[2]]) / 3
    return _finish2d(mesh, np.all(np.abs(cen) <= inside_half, axis=0))


# --------------------------------------------------------------------------------------------------
# P2 (10-node) tetrahedra: mesh.t is 10 x nt, Gmsh order (vertices, then mid nodes of edges 0-1 1-2 0-2 0-3 2-3 1-3)
# --------------------------------------------------------------------------------------------------
P2_EDGES = ((0, 1), (1, 2), (0, 2), (0, 3), (2, 3), (1, 3))


def to_p2(mesh: MESH) -> MESH:
    """Order-2 version of a P1 tetrahedral MESH (what Gmsh's setOrder(2) hands the reference, src/mesh.jl:20-21): one mid
    node per edge appended after the vertices (edge numbering = sorted unique (low, high) vertex pairs), straight edges."""
    t = np.asarray(mesh.t[:4], dtype=np.int64)
    nv = int(mesh.p.shape[1])
    pairs = np.concatenate([np.stack([np.minimum(t[a], t[b]), np.maximum(t[a], t[b])]) for a, b in P2_EDGES], axis=1)
    key = pairs[0] * nv + pairs[1]
    uniq, inv = np.unique(key, return_inverse=True)
    lo, hi = uniq // nv, uniq % nv
    out = MESH()
    out.p = np.asfortranarray(np.concatenate([mesh.p, 0.5 * (mesh.p[:, lo] + mesh.p[:, hi])], axis=1))
    mids = (nv + inv).reshape(6, t.shape[1])
    out.t = np.asfortranarray(np.vstack([t, mids]).astype(np.int32))
    out.nv, out.nt = int(out.p.shape[1]), int(t.shape[1])
    out.VE = np.array(mesh.VE, copy=True)
    out.InsideElements = np.array(mesh.InsideElements, copy=True)
    out.nInside = int(out.InsideElements.size)
    out.InsideNodes = np.unique(out.t[:, out.InsideElements]).astype(np.int32)
    out.nInsideNodes = int(out.InsideNodes.size)
    out.surfaceT = np.zeros((4, 0), dtype=np.int32, order="F")
    out.ns = 0
    out.order = 2
    return out


# local edge -> local vertex pair before the orientation swap (NodesFromLocalEdge, src/mesh.jl:603-637; Gmsh's edge order)
NEDELEC_EDGES = ((0, 1), (1, 2), (2, 0), (3, 0), (2, 3), (1, 3))


def edgeNumbering(mesh: MESH) -> MESH:
    """Edge labels of an order-2 mesh as Mesh() makes them (src/mesh.jl:170-185): mesh.edges = unique(t[5:10, :]) -- Julia's
    unique walks the 6 x nt block column by column and keeps the order of FIRST APPEARANCE --, mesh.ne, and
    mesh.edge2localMap[mid-node label] = edge label (0-based here).  Returns the mesh."""
    assert mesh.t.shape[0] == 10, "edge numbering needs an order-2 mesh (10 x nt connectivity)"
    mids = np.asarray(mesh.t[4:10]).reshape(-1, order="F")
    uniq, first = np.unique(mids, return_index=True)
    mesh.edges = uniq[np.argsort(first, kind="stable")].astype(np.int32)
    mesh.ne = int(mesh.edges.size)
    mesh.edge2localMap = np.full(int(mesh.edges.max()) + 1, -1, dtype=np.int32)
    mesh.edge2localMap[mesh.edges] = np.arange(mesh.ne, dtype=np.int32)
    return mesh


def edgeDofs(mesh: MESH) -> np.ndarray:
    """6 x nt edge labels of every element: edge2localMap[t[5:10, :]] (Magnetostatics/nedelec3D.jl:170)."""
    if mesh.edge2localMap is None:
        edgeNumbering(mesh)
    return np.asfortranarray(mesh.edge2localMap[np.asarray(mesh.t[4:10])].astype(np.int32))


def kuhn_cube_p2(N: int, L: float = 1.0, inside_edge: float = 0.2) -> MESH:
    return to_p2(kuhn_cube(N, L=L, inside_edge=inside_edge))
