"""Test helper: export the symbolic phase of libfemeko_cuda.so (host only, no GPU) and execute the
assembly gather plan in numpy, so the plan itself (tiles, slots, contributor lists) is validated on
the CPU against the oracle.  Test infrastructure only."""
import ctypes as C

import numpy as np

from femeko_jl_b200 import lib as L

SYM = np.array([[0, 1, 2, 3], [1, 4, 5, 6], [2, 5, 7, 8], [3, 6, 8, 9]])


def export_symbolic(mesh, **opts):
    lib = L.load()
    o = L.fk_create_opts()
    want_rescale = opts.pop("_rescale", True)
    for k, v in opts.items():
        setattr(o, k, int(v))
    p = np.asfortranarray(mesh.p, dtype=np.float64)
    t = np.asfortranarray(mesh.t, dtype=np.int32)
    s = C.c_void_p()
    L.check(lib.fk_symbolic_create(L.dptr(p), L.iptr(t), mesh.nv, mesh.nt, C.byref(o), C.byref(s)))
    nnz, ntiles, nte, nnzt, nent, nsl, nsell = (C.c_longlong(), C.c_int(), C.c_longlong(), C.c_longlong(), C.c_longlong(), C.c_int(),
                                                C.c_longlong())
    L.check(lib.fk_symbolic_sizes(s, C.byref(nnz), C.byref(ntiles), C.byref(nte), C.byref(nnzt), C.byref(nent), C.byref(nsl),
                                  C.byref(nsell)))
    T = ntiles.value
    n_rows = o.n_owned if o.n_owned > 0 else mesh.nv
    out = dict(
        rowptr=np.zeros(mesh.nv + 1, np.int32), colind=np.zeros(nnz.value, np.int32), diag=np.zeros(n_rows, np.int32),
        e2nnz=np.zeros(int(opts.get("nodes_per_element", 4) or 4) ** 2 * mesh.nt, np.int32), tile_elem_ptr=np.zeros(T + 1, np.int64), tile_elems=np.zeros(nte.value, np.uint32),
        tile_nnz_ptr=np.zeros(T + 1, np.int64), tile_gidx=np.zeros(nnzt.value, np.int32),
        tile_coff=np.zeros(nnzt.value + T, np.uint16), tile_ent_ptr=np.zeros(T + 1, np.int64),
        tile_ents=np.zeros(nent.value, np.uint16), slice_ptr=np.zeros(nsl.value + 1, np.int32),
        srow=np.zeros(32 * nsl.value, np.int32), sell_col=np.zeros(nsell.value, np.int32), csr2sell=np.zeros(nnz.value, np.int32))
    order = ["rowptr", "colind", "diag", "e2nnz", "tile_elem_ptr", "tile_elems", "tile_nnz_ptr", "tile_gidx", "tile_coff",
             "tile_ent_ptr", "tile_ents", "slice_ptr", "srow", "sell_col", "csr2sell"]
    L.check(lib.fk_symbolic_export(s, *[C.c_void_p(out[k].ctypes.data) for k in order]))
    ncz, n16 = C.c_longlong(), C.c_longlong()
    L.check(lib.fk_symbolic_colz(s, C.byref(ncz), C.byref(n16), None, None))
    out["slice_meta"] = np.zeros(4 * nsl.value, np.int32)
    out["colz"] = np.zeros(ncz.value, np.uint16)
    L.check(lib.fk_symbolic_colz(s, None, None, C.c_void_p(out["slice_meta"].ctypes.data), C.c_void_p(out["colz"].ctypes.data)))
    out["n_slices16"] = n16.value
    if want_rescale:
        ni, nsr, nslots = C.c_longlong(), C.c_int(), C.c_longlong()
        L.check(lib.fk_symbolic_rescale(s, C.byref(ni), C.byref(nsr), C.byref(nslots), None, None, None, None))
        out["r_slice_ptr"] = np.zeros(nsr.value + 1, np.uint32)
        out["r_dst1"] = np.zeros(32 * nsr.value, np.int32)
        out["r_dst2"] = np.zeros(32 * nsr.value, np.int32)
        out["r_cref"] = np.zeros(nslots.value, np.uint32)
        L.check(lib.fk_symbolic_rescale(s, None, None, None, *[C.c_void_p(out[k].ctypes.data) for k in ("r_slice_ptr", "r_dst1", "r_dst2", "r_cref")]))
        out["r_items"], out["r_nslices"], out["r_slots"] = ni.value, nsr.value, nslots.value
    lib.fk_symbolic_destroy(s)
    out["ntiles"] = T
    out["nnz"] = nnz.value
    out["nslices"] = nsl.value
    out["nnz_sell"] = nsell.value
    return out


def run_plan(plan, Ak):
    """Execute the gather plan on dense local matrices Ak (16 x nt, Julia layout): returns CSR values."""
    nt = Ak.shape[1]
    A10 = np.zeros((10, nt))
    for i in range(4):
        for j in range(i, 4):
            A10[SYM[i, j]] = Ak[4 * j + i]
    vals = np.full(plan["nnz"], np.nan)
    written = np.zeros(plan["nnz"], dtype=np.int32)
    for T in range(plan["ntiles"]):
        e0, e1 = plan["tile_elem_ptr"][T], plan["tile_elem_ptr"][T + 1]
        elems = (plan["tile_elems"][e0:e1] & 0x7FFFFFFF).astype(np.int64)
        n0, n1 = plan["tile_nnz_ptr"][T], plan["tile_nnz_ptr"][T + 1]
        coff = plan["tile_coff"][n0 + T: n1 + T + 1].astype(np.int64)
        ents = plan["tile_ents"][plan["tile_ent_ptr"][T]: plan["tile_ent_ptr"][T + 1]].astype(np.int64)
        assert coff[0] == 0 and coff[-1] == ents.size
        contrib = A10[ents & 15, elems[ents >> 4]]
        csum = np.concatenate([[0.0], np.cumsum(contrib)])
        # exact sequential sums per segment (cumsum differences would lose bits): loop when small
        g = plan["tile_gidx"][n0:n1]
        for q in range(n1 - n0):
            vals[g[q]] = contrib[coff[q]:coff[q + 1]].sum() if coff[q + 1] - coff[q] > 8 else sum(contrib[coff[q]:coff[q + 1]], 0.0)
        written[g] += 1
    return vals, written


def run_rescale_plan(plan, G10, s0, shift=28):
    """Execute the rescale plan (symbolic.hpp: RescalePlan) the way geo_rescale_kernel walks it: one lane per item,
    contributions left to right, s0[element] * G10[pair, element] rounded separately, the sum written to the item's
    entry and to its mirror.  G10: 10 x nt unscaled geometric factors (sym_index order).  Returns CSR values and the
    number of times each stored entry was written."""
    vs = np.full(plan["nnz_sell"], np.nan)
    written = np.zeros(plan["nnz_sell"], np.int32)
    sp, cref = plan["r_slice_ptr"].astype(np.int64), plan["r_cref"]
    for s in range(plan["r_nslices"]):
        w = int(sp[s + 1] - sp[s])
        block = cref[32 * sp[s]: 32 * sp[s + 1]].reshape(w, 32)
        acc = np.zeros(32)
        for e in range(w):
            c = block[e]
            live = c != 0xFFFFFFFF
            k = (c & ((1 << shift) - 1)).astype(np.int64)
            q = (c >> shift).astype(np.int64)
            prod = np.where(live, s0[np.where(live, k, 0)] * G10[np.where(live, q, 0), np.where(live, k, 0)], 0.0)
            acc = np.where(live, acc + prod, acc)
            if e > 0:   # a padding slot is never followed by a live one, and elements ascend inside a list
                prev = block[e - 1]
                assert not np.any(live & (prev == 0xFFFFFFFF))
                assert np.all((prev[live] & ((1 << shift) - 1)) <= (c[live] & ((1 << shift) - 1)))
        d1, d2 = plan["r_dst1"][32 * s: 32 * s + 32], plan["r_dst2"][32 * s: 32 * s + 32]
        vs[d1[d1 >= 0]] = acc[d1 >= 0]
        vs[d2[d2 >= 0]] = acc[d2 >= 0]
        np.add.at(written, d1[d1 >= 0], 1)
        np.add.at(written, d2[d2 >= 0], 1)
    return vs[plan["csr2sell"]], written[plan["csr2sell"]]


def sell_spmv(plan, vals_csr, x):
    """Execute the SELL-32-sigma layout in numpy the way spmv_sell_kernel walks it (one lane per row, left to
    right, separately rounded multiply and add; columns from the packed stream: 16-bit offsets from the slice's
    column base, or 32-bit columns where the base is -1): returns y over the assembled rows."""
    vs = np.zeros(plan["nnz_sell"])
    vs[plan["csr2sell"]] = vals_csr
    y = np.zeros(len(plan["diag"]))
    srow, meta, colz = plan["srow"], plan["slice_meta"].reshape(-1, 4), plan["colz"]
    assert np.array_equal(meta[:, 0], plan["slice_ptr"][:-1])
    live_all = np.zeros(plan["nnz_sell"], bool)
    live_all[plan["csr2sell"]] = True
    for s in range(plan["nslices"]):
        base, w, off, cbase = (int(v) for v in meta[s])
        assert w == (plan["slice_ptr"][s + 1] - base) // 32
        if cbase >= 0:   # groups of four steps per lane: word ((k // 4) * 32 + l) * 4 + k % 4
            g4 = (w + 3) // 4
            z = colz[32 * off: 32 * (off + 4 * g4)].astype(np.int64).reshape(g4, 32, 4)
            col = cbase + z.transpose(0, 2, 1).reshape(4 * g4, 32)[:w].reshape(-1)
        else:
            col = colz[32 * off: 32 * (off + 2 * w)].view(np.int32).astype(np.int64)
        r = srow[32 * s: 32 * s + 32]
        # the packed columns are the plain SELL columns wherever a value can be non-zero
        ref = plan["sell_col"][base: base + 32 * w]
        live = live_all[base: base + 32 * w]
        assert np.array_equal(col[live], ref[live])
        acc = np.zeros(32)
        for k in range(w):
            i = 32 * k + np.arange(32)
            acc = acc + vs[base + i] * x[col[i]]
        y[r[r >= 0]] = acc[r >= 0]
    return y
