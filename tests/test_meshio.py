"""Mesh ingest: the reference's save2file text dumps and Gmsh .msh files -> MESH.  CPU only."""
import os

import numpy as np
import pytest

import femeko_jl_b200 as fk
from femeko_jl_b200.meshio import _julia_number

FIELDS = ["p", "t", "surfaceT", "InsideElements", "InsideNodes", "VE", "AE", "normal"]


def test_save2file_format_matches_reference(tmp_path):
    """src/geometry.jl:236-243: rows joined by ' , '; vectors one value per line; ids 1-based on disk."""
    m = fk.kuhn_cube(2)
    fk.saveMesh(m, str(tmp_path))
    assert sorted(os.listdir(tmp_path)) == sorted(f + ".txt" for f in FIELDS)        # src/mesh.jl:234-241
    t_lines = open(tmp_path / "t.txt").read().splitlines()
    assert len(t_lines) == 4 and all(len(ln.split(" , ")) == m.nt for ln in t_lines)
    assert [int(v) for v in t_lines[0].split(" , ")] == (m.t[0] + 1).tolist()
    p_lines = open(tmp_path / "p.txt").read().splitlines()
    assert len(p_lines) == 3 and p_lines[0].split(" , ")[0] == "-0.5"
    ve = open(tmp_path / "VE.txt").read().splitlines()
    assert len(ve) == m.nt and float(ve[3]) == m.VE[3]
    s_lines = open(tmp_path / "surfaceT.txt").read().splitlines()
    assert len(s_lines) == 4 and int(s_lines[3].split(" , ")[0]) == m.surfaceT[3, 0] + 1
    # Julia number formatting (shortest round trip, trailing .0, d.de-x outside [1e-4, 1e21))
    assert [_julia_number(v) for v in (1.0, 0.1, 0.0001, 1e-6, 1.5e22, -2.5e-7, 3, 1e16)] == \
        ["1.0", "0.1", "0.0001", "1.0e-6", "1.5e22", "-2.5e-7", "3", "10000000000000000.0"]


@pytest.mark.parametrize("gen", ["kuhn", "delaunay"])
def test_text_roundtrip_is_exact(tmp_path, gen):
    m = fk.kuhn_cube(4) if gen == "kuhn" else fk.delaunay_ball(npts=300)
    fk.saveMesh(m, str(tmp_path))
    m2 = fk.loadMesh(str(tmp_path))
    for f in FIELDS:
        a, b = getattr(m, f), getattr(m2, f)
        assert a.shape == b.shape and a.dtype == b.dtype and np.array_equal(a, b), f
    assert (m2.nv, m2.nt, m2.ns, m2.nInside, m2.nInsideNodes) == (m.nv, m.nt, m.ns, m.nInside, m.nInsideNodes)
    # only p / t present: the derived fields are recomputed as Mesh() does
    for f in FIELDS[2:]:
        os.remove(tmp_path / (f + ".txt"))
    m3 = fk.loadMesh(str(tmp_path))
    assert np.array_equal(m3.VE, m.VE) and m3.ns == 0 and m3.nInside == 0


def test_text_reader_rejects_bad_files(tmp_path):
    m = fk.kuhn_cube(2)
    fk.saveMesh(m, str(tmp_path))
    bad = m.t.astype(np.int64) + 1
    bad[0, 0] = m.nv + 5
    fk.save2file(str(tmp_path / "t.txt"), bad)
    with pytest.raises(ValueError, match="out of range"):
        fk.loadMesh(str(tmp_path))
    open(tmp_path / "t.txt", "w").write("1 , 2 , 3\n1 , 2\n")
    with pytest.raises(ValueError, match="ragged"):
        fk.loadMesh(str(tmp_path))


def _write_msh(path, m, version, node_tag_gap=0):
    """A Gmsh ASCII file of mesh m: volume entities 1 (inside) / 2 (container), surface entity = id + 1; node tags
    1-based with an optional gap (Gmsh tags need not be contiguous)."""
    tag = np.arange(1, m.nv + 1) + node_tag_gap * (np.arange(m.nv) >= m.nv // 2)
    ent = np.full(m.nt, 2)
    ent[m.InsideElements] = 1
    with open(path, "w") as f:
        if version == 2:
            f.write("$MeshFormat\n2.2 0 8\n$EndMeshFormat\n$Nodes\n%d\n" % m.nv)
            for i in range(m.nv):
                f.write("%d %r %r %r\n" % (tag[i], *m.p[:, i].tolist()))
            f.write("$EndNodes\n$Elements\n%d\n" % (m.ns + m.nt))
            e = 1
            for i in range(m.ns):
                f.write("%d 2 2 0 %d %d %d %d\n" % (e, m.surfaceT[3, i] + 1, *tag[m.surfaceT[:3, i]]))
                e += 1
            for k in range(m.nt):
                f.write("%d 4 2 0 %d %d %d %d %d\n" % (e, ent[k], *tag[m.t[:, k]]))
                e += 1
            f.write("$EndElements\n")
        else:
            f.write("$MeshFormat\n4.1 0 8\n$EndMeshFormat\n$Nodes\n")
            half = m.nv // 2
            f.write("2 %d %d %d\n" % (m.nv, tag.min(), tag.max()))
            for lo, hi, e_ in ((0, half, 1), (half, m.nv, 2)):
                f.write("3 %d 0 %d\n" % (e_, hi - lo))
                f.write("".join("%d\n" % tag[i] for i in range(lo, hi)))
                f.write("".join("%r %r %r\n" % tuple(m.p[:, i].tolist()) for i in range(lo, hi)))
            f.write("$EndNodes\n$Elements\n")
            sids = np.unique(m.surfaceT[3])
            blocks = [(2, int(s) + 1, 2, np.flatnonzero(m.surfaceT[3] == s)) for s in sids] + \
                     [(3, v, 4, np.flatnonzero(ent == v)) for v in (1, 2) if np.any(ent == v)]
            # Mesh() keeps Gmsh's element order (by tag); write volume 2 before volume 1 to exercise entity blocks
            f.write("%d %d 1 %d\n" % (len(blocks), m.ns + m.nt, m.ns + m.nt))
            e = 1
            for dim, etag, etype, ids in blocks:
                f.write("%d %d %d %d\n" % (dim, etag, etype, ids.size))
                for i in ids:
                    nodes = tag[m.surfaceT[:3, i]] if etype == 2 else tag[m.t[:, i]]
                    f.write("%d %s\n" % (e, " ".join(str(v) for v in nodes)))
                    e += 1
            f.write("$EndElements\n")


@pytest.mark.parametrize("version,gap", [(2, 0), (2, 1000), (4, 0), (4, 7)])
def test_read_msh(tmp_path, version, gap):
    m = fk.kuhn_cube(3, single_shell_id=False) if version == 4 else fk.kuhn_cube(3)
    path = str(tmp_path / "mesh.msh")
    _write_msh(path, m, version, gap)
    r = fk.read_msh(path, cells=[(3, 1)])
    assert (r.nv, r.nt, r.ns) == (m.nv, m.nt, m.ns)
    assert np.array_equal(r.p, m.p)
    if version == 2:                         # element order preserved -> everything identical
        for f in FIELDS:
            a, b = getattr(m, f), getattr(r, f)
            if f == "InsideNodes":
                a, b = np.sort(a), np.sort(b)
            assert np.array_equal(a, b), f
    else:                                    # 4.1 groups elements by entity: same sets
        key = lambda t: set(map(tuple, t.T.tolist()))   # noqa: E731
        assert key(r.t) == key(m.t) and key(r.surfaceT) == key(m.surfaceT)
        assert set(r.InsideNodes.tolist()) == set(m.InsideNodes.tolist()) and r.nInside == m.nInside
    assert abs(r.VE.sum() - 1.0) < 1e-12 and abs(r.AE.sum() - 6.0) < 1e-12
    assert set(np.unique(r.elementID).tolist()) == {1, 2}
    assert fk.read_msh(path).nInside == 0   # no cells: no inside elements (src/mesh.jl:205-208)


def test_read_msh_rejects_bad_files(tmp_path):
    p = tmp_path / "bad.msh"
    p.write_text("$MeshFormat\n4.1 1 8\n$EndMeshFormat\n$Nodes\n$EndNodes\n$Elements\n$EndElements\n")
    with pytest.raises(ValueError, match="binary"):
        fk.read_msh(str(p))
    p.write_text("$MeshFormat\n2.2 0 8\n$EndMeshFormat\n$Nodes\n1\n1 0 0 0\n$EndNodes\n$Elements\n1\n1 15 2 0 1 1\n$EndElements\n")
    with pytest.raises(ValueError, match="tetrahedra"):
        fk.read_msh(str(p))
    p.write_text("$MeshFormat\n2.2 0 8\n$EndMeshFormat\n$Nodes\n1\n1 0 0 0\n$EndNodes\n$Elements\n1\n1 4 2 0 1 1 2 3 4\n$EndElements\n")
    with pytest.raises(ValueError, match="unknown node"):
        fk.read_msh(str(p))
