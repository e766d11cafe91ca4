"""Golden fixtures of the micromagnetics demagnetising field, made by RUNNING THE COMPILED
REFERENCE class LL (oracle/_ref/libfemeko_ref_ll.so = unmodified Femeko_jl_C++/Micromagnetics/LL.cpp behind
oracle/ref_ll_shim.cpp): one LL::magnetostaticField() for a seeded magnetisation on the meshes of the main
fixtures.  Run in the dev container only:

    python tests/golden/make_demag_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from conftest import load_golden  # noqa: E402
from oracle import oracle as orc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def make(name, seed):
    mesh, g = load_golden(name)
    inside = np.asarray(g["InsideElements"], dtype=np.int32)
    inside_nodes = np.unique(mesh.t[:, inside]).astype(np.int32)        # mesh.InsideNodes, src/mesh.jl:203-213
    rng = np.random.default_rng(seed)
    M = np.zeros((3, mesh.nv), order="F")
    m = rng.normal(size=(3, len(inside_nodes)))
    M[:, inside_nodes] = m / np.linalg.norm(m, axis=0)                  # unit vectors on the magnetic nodes (permalloy.jl:65-66 starts from M = Ms x)
    Hd, vol = orc.ref_demag_field(mesh, inside, inside_nodes, M)
    np.savez_compressed(os.path.join(HERE, "demag_" + name + ".npz"), InsideNodes=inside_nodes, M=M, Hd=Hd, volumes=vol)
    print(name, "magnetic elements", len(inside), "nodes", len(inside_nodes), "finite Hd columns", int(np.isfinite(Hd[0]).sum()))


if __name__ == "__main__":
    assert orc.have_ref_ll(), "build oracle/_ref first (make -C oracle ref)"
    make("kuhn6", 21)
    make("delaunay300", 22)
