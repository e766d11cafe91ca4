"""Mesh ingest / on-disk formats: meshes produced elsewhere feed the GPU path without
Gmsh on the box.

* :func:`save2file`, :func:`saveMesh`, :func:`loadMesh` -- the reference's text dumps
  (``save2file`` src/geometry.jl:236-243; the eight files written by ``Mesh(...; saveMesh=true)``
  src/mesh.jl:233-242): one matrix row per line, entries joined by ``" , "``; a vector is one value
  per line.  On disk the indices are Julia's (1-based: node ids in ``t`` / ``surfaceT[1:3]`` /
  ``InsideNodes``, element ids in ``InsideElements``, boundary-surface ids in ``surfaceT[4]``);
  in memory they are 0-based like everything the reference hands to C (run.jl:50-52).
* :func:`read_msh` -- Gmsh ``.msh`` ASCII reader (format 2.2 and 4.1) -> :class:`MESH`, restating
  what ``Mesh()`` pulls through the Gmsh API (src/mesh.jl:119-230): nodes in tag order renumbered
  contiguously, 4-node tetrahedra with their volume-entity tag as ``elementID``, 3-node boundary
  triangles with their surface-entity tag in row 4 of ``surfaceT``, inside elements = the tets of
  the chosen cell ids, then VE / AE / normal.

Host-side only; nothing here touches the GPU.
"""
from __future__ import annotations

import os

import numpy as np

from .mesh import MESH, areaTriangle, elementVolume, normalSurface

_ONE_BASED_ROWS = {"t": slice(None), "surfaceT": slice(None), "InsideNodes": slice(None), "InsideElements": slice(None)}
_FILES = ("t", "p", "InsideNodes", "surfaceT", "InsideElements", "VE", "AE", "normal")   # src/mesh.jl:234-241


def _julia_number(x) -> str:
    """Julia's ``print`` of an Int32 / Float64 (shortest round-trip digits; fixed notation for
    1e-4 <= |x| < 1e21, ``d.ddde±x`` otherwise; integral floats keep a trailing ``.0``)."""
    if isinstance(x, (int, np.integer)):
        return str(int(x))
    x = float(x)
    if x != x:
        return "NaN"
    if x in (float("inf"), float("-inf")):
        return "Inf" if x > 0 else "-Inf"
    if x == 0.0:
        return "-0.0" if np.signbit(x) else "0.0"
    r = repr(x)
    mant, _, ex = r.partition("e")
    a = abs(x)
    if 1e-4 <= a < 1e21:
        if ex:   # python switched to exponent form (>= 1e16 or < 1e-4): expand
            r = np.format_float_positional(x, unique=True, trim="0")
        return r if "." in r else r + ".0"
    if not ex:
        r = np.format_float_scientific(x, unique=True, trim="0")
        mant, _, ex = r.partition("e")
    if "." not in mant:
        mant += ".0"
    return f"{mant}e{int(ex)}"


def save2file(fileName: str, input) -> None:   # noqa: A002  (reference argument name)
    """``save2file(fileName, input)`` (src/geometry.jl:236-243): one row per line, joined by ' , '."""
    a = np.asarray(input)
    rows = a.reshape(-1, 1) if a.ndim == 1 else a
    with open(fileName, "w") as io:
        for row in rows:
            io.write(" , ".join(_julia_number(v) for v in row.tolist()) + "\n")


def _read_matrix(path: str, dtype) -> np.ndarray:
    with open(path) as f:
        rows = [[tok for tok in line.replace(",", " ").split()] for line in f if line.strip()]
    if not rows:
        return np.zeros((0,), dtype)
    w = len(rows[0])
    if any(len(r) != w for r in rows):
        raise ValueError(f"{path}: ragged rows")
    a = np.array(rows, dtype=np.float64)
    if np.issubdtype(dtype, np.integer):
        if not np.array_equal(a, np.round(a)):
            raise ValueError(f"{path}: non-integer index")
        a = a.astype(dtype)
    return a


def saveMesh(mesh: MESH, folder: str = ".") -> None:
    """Write the eight text files of ``Mesh(...; saveMesh=true)`` (src/mesh.jl:233-242) into ``folder``."""
    os.makedirs(folder, exist_ok=True)
    for name in _FILES:
        a = np.asarray(getattr(mesh, name))
        if name in _ONE_BASED_ROWS:
            a = a.astype(np.int64) + 1
        save2file(os.path.join(folder, name + ".txt"), a)


def loadMesh(folder: str = ".") -> MESH:
    """Read a mesh written by :func:`saveMesh` or by the reference.  Only ``t.txt`` and ``p.txt`` are
    required; missing derived files are recomputed the way ``Mesh()`` does (src/mesh.jl:216-230)."""
    def path(n):
        return os.path.join(folder, n + ".txt")

    def have(n):
        return os.path.exists(path(n))

    p = np.asfortranarray(_read_matrix(path("p"), np.float64))
    t = _read_matrix(path("t"), np.int64) - 1
    if p.ndim != 2 or p.shape[0] != 3:
        raise ValueError("p.txt must hold 3 rows (x, y, z) of nv columns")
    if t.ndim != 2 or t.shape[0] not in (4, 10):
        raise ValueError("t.txt must hold 4 (or 10) rows of nt columns")
    nv = p.shape[1]
    if t.size and (t.min() < 0 or t.max() >= nv):
        raise ValueError("t.txt: node id out of range")
    m = MESH(order=1 if t.shape[0] == 4 else 2, p=p, t=np.asfortranarray(t.astype(np.int32)))
    m.nv, m.nt = nv, int(t.shape[1])
    if have("surfaceT"):
        s = _read_matrix(path("surfaceT"), np.int64) - 1
        if s.size == 0:
            s = np.zeros((4, 0), np.int64)
        if s.ndim != 2 or s.shape[0] != 4:
            raise ValueError("surfaceT.txt must hold 4 rows (3 nodes + boundary id)")
        if s.size and (s[:3].min() < 0 or s[:3].max() >= nv):
            raise ValueError("surfaceT.txt: node id out of range")
    else:
        s = np.zeros((4, 0), np.int64)
    m.surfaceT = np.asfortranarray(s.astype(np.int32))
    m.ns = int(s.shape[1])
    ie = (_read_matrix(path("InsideElements"), np.int64).reshape(-1) - 1) if have("InsideElements") else np.zeros(0, np.int64)
    if ie.size and (ie.min() < 0 or ie.max() >= m.nt):
        raise ValueError("InsideElements.txt: element id out of range")
    m.InsideElements = ie.astype(np.int32)
    m.nInside = int(ie.size)
    if have("InsideNodes"):
        m.InsideNodes = (_read_matrix(path("InsideNodes"), np.int64).reshape(-1) - 1).astype(np.int32)
    else:
        m.InsideNodes = np.unique(m.t[:4, m.InsideElements]).astype(np.int32)
    m.nInsideNodes = int(m.InsideNodes.size)
    t4 = m.t[:4]
    m.VE = _read_matrix(path("VE"), np.float64).reshape(-1) if have("VE") else elementVolume(m.p, t4)
    m.AE = _read_matrix(path("AE"), np.float64).reshape(-1) if have("AE") else areaTriangle(m.p, m.surfaceT)
    m.normal = np.asfortranarray(_read_matrix(path("normal"), np.float64).reshape(3, -1)) if have("normal") and m.ns else \
        np.asfortranarray(normalSurface(m.p, m.surfaceT))
    if m.VE.size != m.nt or m.AE.size != m.ns or m.normal.shape != (3, m.ns):
        raise ValueError("VE / AE / normal do not match t / surfaceT")
    el = np.full(m.nt, 2, np.int32)
    el[m.InsideElements] = 1
    m.elementID = el
    return m


# ------------------------------------------------------------------------------------------------
# Gmsh .msh (ASCII) -> MESH
# ------------------------------------------------------------------------------------------------
_GMSH_NODES = {1: 2, 2: 3, 4: 4, 9: 6, 11: 10, 15: 1, 8: 3}   # element type -> node count (the ones a tet mesh holds)


def _sections(lines):
    """{name: list of lines} of every $Name ... $EndName block"""
    out, i, n = {}, 0, len(lines)
    while i < n:
        s = lines[i].strip()
        if s.startswith("$") and not s.startswith("$End"):
            name, j = s[1:], i + 1
            while j < n and lines[j].strip() != "$End" + name:
                j += 1
            if j >= n:
                raise ValueError(f"msh: section ${name} is not closed")
            out[name] = lines[i + 1: j]
            i = j
        i += 1
    return out


def _msh2(sec):
    nl = sec["Nodes"]
    nn = int(nl[0])
    tags = np.empty(nn, np.int64)
    xyz = np.empty((nn, 3))
    for i in range(nn):
        f = nl[1 + i].split()
        tags[i] = int(f[0])
        xyz[i] = [float(f[1]), float(f[2]), float(f[3])]
    el = sec["Elements"]
    elems = {}   # type -> list of (entity tag, node tags)
    for i in range(int(el[0])):
        f = [int(v) for v in el[1 + i].split()]
        etype, ntags = f[1], f[2]
        ent = f[4] if ntags >= 2 else (f[3] if ntags == 1 else 0)   # tags = physical, elementary entity, ...
        elems.setdefault(etype, []).append((ent, f[3 + ntags:]))
    return tags, xyz, elems


def _msh4(sec):
    nl = sec["Nodes"]
    nblocks, nn = (int(v) for v in nl[0].split()[:2])
    tags = np.empty(nn, np.int64)
    xyz = np.empty((nn, 3))
    i, k = 1, 0
    for _ in range(nblocks):
        _, _, parametric, nb = (int(v) for v in nl[i].split())
        if parametric:
            raise ValueError("msh: parametric node blocks are not supported")
        for j in range(nb):
            tags[k + j] = int(nl[i + 1 + j])
            xyz[k + j] = [float(v) for v in nl[i + 1 + nb + j].split()[:3]]
        i += 1 + 2 * nb
        k += nb
    el = sec["Elements"]
    nblocks = int(el[0].split()[0])
    elems = {}
    i = 1
    for _ in range(nblocks):
        _, ent, etype, nb = (int(v) for v in el[i].split())
        for j in range(nb):
            f = [int(v) for v in el[i + 1 + j].split()]
            elems.setdefault(etype, []).append((ent, f[1:]))
        i += 1 + nb
    return tags, xyz, elems


def read_msh(path: str, cells=(), order: int = 1) -> MESH:
    """Read a Gmsh ASCII ``.msh`` (format 2.2 or 4.1) into a :class:`MESH` the way ``Mesh()`` fills it
    from the Gmsh API (src/mesh.jl:119-230).  ``cells`` = volume-entity ids of the bodies inside the
    container (the reference's ``cells`` argument, ``(3, tag)`` pairs or bare tags): their tetrahedra
    become ``InsideElements``.  Node tags are renumbered to 0..nv-1 in tag order; row 4 of
    ``surfaceT`` holds the surface-entity id minus one (the 0-based form of run.jl:52)."""
    if order != 1:
        raise ValueError("read_msh: only first-order (4-node) tetrahedra are on the GPU path")
    with open(path) as f:
        lines = f.read().splitlines()
    sec = _sections(lines)
    if "MeshFormat" not in sec or "Nodes" not in sec or "Elements" not in sec:
        raise ValueError("msh: $MeshFormat / $Nodes / $Elements missing")
    fmt = sec["MeshFormat"][0].split()
    if int(fmt[1]) != 0:
        raise ValueError("msh: binary files are not supported (export with Mesh.Binary = 0)")
    ver = float(fmt[0])
    if 2.0 <= ver < 3.0:
        tags, xyz, elems = _msh2(sec)
    elif 4.0 <= ver < 5.0:
        tags, xyz, elems = _msh4(sec)
    else:
        raise ValueError(f"msh: unsupported format version {fmt[0]}")
    if 4 not in elems:
        raise ValueError("msh: no 4-node tetrahedra (element type 4) in the file")
    order_ = np.argsort(tags, kind="stable")
    tags, xyz = tags[order_], xyz[order_]
    if np.any(np.diff(tags) == 0):
        raise ValueError("msh: duplicate node tag")

    def renumber(a):
        idx = np.searchsorted(tags, a)
        if np.any(idx >= tags.size) or np.any(tags[np.minimum(idx, tags.size - 1)] != a):
            raise ValueError("msh: element refers to an unknown node tag")
        return idx

    tet = elems[4]
    t = renumber(np.array([nd for _, nd in tet], np.int64)).T
    m = MESH(p=np.asfortranarray(xyz.T.copy()), t=np.asfortranarray(t.astype(np.int32)))
    m.nv, m.nt = int(tags.size), int(t.shape[1])
    m.elementID = np.array([ent for ent, _ in tet], np.int32)
    tri = elems.get(2, [])
    s = np.zeros((4, len(tri)), np.int64)
    if tri:
        s[:3] = renumber(np.array([nd for _, nd in tri], np.int64)).T
        s[3] = np.array([ent for ent, _ in tri], np.int64) - 1
    m.surfaceT = np.asfortranarray(s.astype(np.int32))
    m.ns = len(tri)
    vol = {int(c[1]) if isinstance(c, (tuple, list)) else int(c) for c in cells}
    m.InsideElements = np.flatnonzero(np.isin(m.elementID, sorted(vol))).astype(np.int32) if vol else np.zeros(0, np.int32)
    m.nInside = int(m.InsideElements.size)
    inside = m.t[:, m.InsideElements].T.reshape(-1)            # unique(vec(t[:, InsideElements])): first-seen order
    _, first = np.unique(inside, return_index=True)
    m.InsideNodes = inside[np.sort(first)].astype(np.int32)
    m.nInsideNodes = int(m.InsideNodes.size)
    m.VE = elementVolume(m.p, m.t)
    m.AE = areaTriangle(m.p, m.surfaceT)
    m.normal = np.asfortranarray(normalSurface(m.p, m.surfaceT))
    return m
