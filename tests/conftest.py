import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for pth in (ROOT, os.path.join(ROOT, "tests")):
    if pth not in sys.path:
        sys.path.insert(0, pth)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _build_native():
    """Build the oracle (gcc) and the CUDA library (nvcc cross-compiles without a GPU)."""
    from oracle import oracle as orc
    from femeko_jl_b200 import lib as L

    orc.build(ref=os.path.isdir("/root/reference/src"))
    L.build()


GOLDEN = os.path.join(ROOT, "tests", "golden")


def load_golden(name):
    """Golden fixture -> (MESH, dict of reference outputs)."""
    import femeko_jl_b200 as fk
    from femeko_jl_b200.mesh import _finish

    g = dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    m = fk.MESH()
    m.p = np.asfortranarray(g["p"])
    m.t = np.asfortranarray(g["t"].astype(np.int32))
    m.surfaceT = np.asfortranarray(g["surfaceT"].astype(np.int32))
    m.shell_id = [0]
    inside = np.zeros(m.t.shape[1], dtype=bool)
    inside[g["InsideElements"]] = True
    _finish(m, inside)
    # the fixture's own geometry arrays are authoritative (they were the reference's inputs)
    m.VE, m.AE, m.normal = g["VE"], g["AE"], np.asfortranarray(g["normal"])
    return m, g


def rel_err_matrix(vals, ref, scale=None):
    """Matrix-value tolerance metric: max |a - ref| / max(|ref|, scale).

    `scale` (abs_sum_scale) is sum_k V_k f_k |grad l_i||grad l_j| = sum_k sqrt(A_k[i,i] A_k[j,j]),
    the Cauchy-Schwarz magnitude of what each entry sums.  On uniform Kuhn meshes about half of
    the structural entries are dot products of orthogonal gradients (exactly 0 here, ~1e-23 of
    rounding noise in the reference's QR-solved abcd), where a plain relative error is
    meaningless; measured against this scale the 1e-12 bar of the north star is well defined for
    every entry."""
    den = np.abs(ref) if scale is None else np.maximum(np.abs(ref), scale)
    den = np.where(den == 0, 1.0, den)
    return float(np.max(np.abs(vals - ref) / den))


def abs_sum_scale(mesh, Ak, f=None):
    """sum_k |f_k| sqrt(A_k[i,i] A_k[j,j]) per CSR entry (oracle, test infrastructure)."""
    from oracle import oracle as orc

    rp, ci = orc.pattern(mesh.t, mesh.nv)
    e2n = orc.elem2nnz(mesh.t, rp, ci)
    d = np.abs(np.stack([Ak[5 * i] for i in range(4)]))          # local diagonals, 4 x nt
    S = np.zeros_like(Ak)
    for i in range(4):
        for j in range(4):
            S[4 * j + i] = np.sqrt(d[i] * d[j])
    return orc.assemble_eigen(e2n, len(ci), S, None if f is None else np.abs(f))


def rel_l2(a, b):
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1e-300))
