"""GPU: fk_residual_norm_noalloc (include/femeko_cuda.h; csrc/noalloc.cu) -- the allocation-free residual norm whose scratch is a
static __device__ array -- against the existing entry point fk_residual_dev, through the C ABI, on two calls in a row.
No counterpart in the reference.  Collected last on
purpose."""
import ctypes as C

import numpy as np
import pytest

import femeko_jl_b200 as fk
from femeko_jl_b200 import lib as L

pytestmark = pytest.mark.gpu


def _existing(h, x_d, b_d, r_d):
    nr = C.c_double()
    L.check(h.lib.fk_residual_dev(h.h, 0, C.c_void_p(x_d.data_ptr()), C.c_void_p(b_d.data_ptr()), C.c_void_p(r_d.data_ptr()), C.byref(nr)))
    return nr.value


def test_noalloc_residual_matches_existing_entry_point_twice():
    import torch

    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    m = fk.kuhn_cube(20, inside_edge=0.5)                 # 9261 rows: three CTAs of 4096 rows, the last one ragged
    h = fk.Femeko(m)
    rng = np.random.default_rng(3)
    h.assemble_stiffness(rng.uniform(0.5, 2.0, m.nt), slot=0)
    n = m.nv
    norm_d = torch.zeros(1, dtype=torch.float64, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    assert fk.residual_norm_noalloc(None, None, 0) == 0.0   # per-device set-up (address of the static array, kernel load)
    seen = []
    for call in range(2):                                 # second call: the scratch holds the first call's partials and result
        x_d = torch.from_numpy(rng.normal(size=n) * 10.0 ** call).to(dev)
        b_d = torch.from_numpy(rng.normal(size=n)).to(dev)
        r_ref = torch.empty(n, dtype=torch.float64, device=dev)
        nr_ref = _existing(h, x_d, b_d, r_ref)
        Ax_d = torch.empty(n, dtype=torch.float64, device=dev)
        h.spmv_dev(x_d.data_ptr(), Ax_d.data_ptr(), slot=0)
        h.synchronize()
        r_new = torch.full((n,), float("nan"), dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        mem0 = torch.cuda.mem_get_info()[0]
        nr_new = fk.residual_norm_noalloc(b_d.data_ptr(), Ax_d.data_ptr(), n, r_new.data_ptr(), norm_d.data_ptr(), stream=stream)
        assert torch.cuda.mem_get_info()[0] >= mem0, "the call took device memory"
        assert torch.equal(r_new, r_ref)                  # same subtraction, element by element: bit for bit
        assert abs(nr_new - nr_ref) <= 1e-13 * nr_ref     # different (both deterministic) summation orders
        assert float(norm_d.item()) == nr_new
        assert abs(nr_new - float(np.linalg.norm(r_ref.cpu().numpy()))) <= 1e-13 * nr_ref
        # norm only (r = NULL, no host copy), twice: identical bits from the shared scratch
        norm_d.zero_()
        assert fk.residual_norm_noalloc(b_d.data_ptr(), Ax_d.data_ptr(), n, None, norm_d.data_ptr(), want_host=False, stream=stream) is None
        torch.cuda.synchronize()
        assert float(norm_d.item()) == nr_new
        seen.append(nr_new)
    assert seen[0] != seen[1]
    # edge cases: an empty vector, a single ragged CTA
    assert fk.residual_norm_noalloc(None, None, 0) == 0.0
    assert fk.residual_norm_noalloc(b_d.data_ptr(), Ax_d.data_ptr(), 5) == pytest.approx(
        float(torch.linalg.norm(b_d[:5] - Ax_d[:5]).item()), rel=1e-14)
    h.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch

    torch.cuda.set_device(0)
    z = torch.zeros(8, dtype=torch.float64, device="cuda:0")
    with pytest.raises(L.FemekoError, match="exceeds FK_NOALLOC_MAX_ROWS"):
        fk.residual_norm_noalloc(z.data_ptr(), z.data_ptr(), fk.NOALLOC_MAX_ROWS + 1)
    with pytest.raises(L.FemekoError, match="n < 0"):
        fk.residual_norm_noalloc(z.data_ptr(), z.data_ptr(), -1)
    # the limit itself is a property of the header and of the binding alike
    assert fk.NOALLOC_MAX_ROWS == 1 << 26
