// K1r: RESCALE assembly (included inside namespace fk by kernels.cuh; plan in symbolic.hpp: RescalePlan).
//
// Between two Picard steps only the element coefficient s_k = VE[k] mu[k] changes: A = sum_k s_k G_k, G_k the
// geometric part (grad lambda_i . grad lambda_j) of the local matrix.  The reference's nonlinear loop keeps the local
// matrices and rescales them by mu every iteration for exactly that reason (NewtonRaphson.jl:114-127,153-161).
// Here the geometric factors are computed ONCE per handle (geo_build_kernel), stored in the order the sums consume
// them, and a sweep is three streaming kernels with no shared memory and no block-level synchronisation:
//   material_update_kernel   mu(|H|) of the elements that carry a B-H curve (kernels.cuh, unchanged)
//   picard_element_kernel    H = -grad u, |H| of every element, s_k = VE[k] mu[k]
//   geo_rescale_kernel       one lane per stored entry PAIR {(r,c), (c,r)}: sum of s_k g in ascending element order
//                            (the order of the tile kernel and of Eigen's setFromTriplets), written to both entries
// Every product is rounded as the tile kernel rounds it (s_k first: VE[k]*f[k]*(...), src/FEM3D.jl:71), so the two
// paths agree bit for bit; the tile kernel (assemble.cuh) stays the from-scratch path.
// Bytes per sweep and tet on a Kuhn mesh: factors 8 + element id 4 per contribution (10.4 incl. padding), 8 per pair
// for the two value indices, 20 for the values -> ~165 B; element kernel ~70 B.

// the ten local pairs in sym_index order (symbolic.hpp)
__device__ __forceinline__ void pair_ij(const int q, int& i, int& j) {
  i = q < 4 ? 0 : (q < 7 ? 1 : (q < 9 ? 2 : 3));
  j = q - (i == 0 ? 0 : (i == 1 ? 3 : (i == 2 ? 5 : 6)));
}

// geometric factor of every contribution slot; strips the pair bits of the slot's element reference
static __global__ void __launch_bounds__(256) geo_build_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4,
                                                               uint32_t* __restrict__ cref, long long nslots,
                                                               double* __restrict__ g) {
  for (long long s = blockIdx.x * (long long)blockDim.x + threadIdx.x; s < nslots; s += (long long)gridDim.x * blockDim.x) {
    const uint32_t c = cref[s];
    if (c == 0xffffffffu) { g[s] = 0.0; continue; }
    const int k = (int)(c & 0x0fffffffu);
    int pi, pj;
    pair_ij((int)(c >> 28), pi, pj);
    TetGeom G;
    tet_geometry(p4, __ldg(t4 + k), G);
    double a0 = 0, a1 = 0, a2 = 0, b0 = 0, b1 = 0, b2 = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) {
      if (q == pi) { a0 = G.g[q][0]; a1 = G.g[q][1]; a2 = G.g[q][2]; }
      if (q == pj) { b0 = G.g[q][0]; b1 = G.g[q][1]; b2 = G.g[q][2]; }
    }
    g[s] = grad_dot3(a0, a1, a2, b0, b1, b2);
    cref[s] = (uint32_t)k;
  }
}

// s_k = VE[k] f[k] of a plain stiffness assembly (f nullable = 1)
static __global__ void __launch_bounds__(256) geo_scale_kernel(const double* __restrict__ VE, const double* __restrict__ f, int nt,
                                                               double* __restrict__ s0) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nt) return;
  s0[k] = f ? __dmul_rn(VE[k], f[k]) : __dmul_rn(VE[k], 1.0);
}

// Picard sweep, element part: field recovery of EVERY element (NewtonRaphson.jl:178-197) and its scale s_k = VE[k] mu[k];
// mu of the material elements was refreshed by material_update_kernel just before.  VE == nullptr: |det| / 6.
// The kernel is bound by the L1 tag stage (one 128-byte line per cycle and SM), not by HBM: a node is therefore fetched
// with ONE 256-bit load (load_node: LDG.E.256) instead of two 128-bit ones, and the three components
// of H leave through a per-warp shared-memory transpose so that a warp writes 768 contiguous bytes with three stores.
constexpr int kElemThreads = 256;
// Persistent grid (fk: picard_step_dev launches one wave of resident CTAs): the connectivity and the coefficient of the
// chunk a CTA will work on NEXT are requested before the gathers of the current one, so the two dependent memory
// round trips of an element (connectivity -> node / potential gathers) overlap across chunks.
// S0 = false: field recovery only -- the scales are kept current by geo_scale_kernel / material_update_kernel, so this
// kernel and geo_rescale_kernel are independent and run side by side on two streams (api.cu: picard_step_dev).
template <bool S0>
static __global__ void __launch_bounds__(kElemThreads) picard_element_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4,
                                                                             const double* __restrict__ u, const double* __restrict__ mu,
                                                                             const double* __restrict__ VE, int nt, int compat,
                                                                             double* __restrict__ Hn, double* __restrict__ Hfield,
                                                                             double* __restrict__ s0) {
  __shared__ double hs[kElemThreads / 32][96];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const long long stride = (long long)gridDim.x * kElemThreads;
  long long k = (long long)blockIdx.x * kElemThreads + threadIdx.x;
  int4 nd = make_int4(0, 0, 0, 0);
  double f = 0.0, V = 0.0;
  if (k < nt) { nd = __ldg(t4 + k); if (S0) { f = __ldcs(mu + k); if (VE) V = __ldcs(VE + k); } }
  for (long long k0 = (long long)blockIdx.x * kElemThreads; k0 < nt; k0 += stride, k += stride) {
    const long long kn = k + stride;
    int4 ndn = make_int4(0, 0, 0, 0);
    double fn = 0.0, Vn = 0.0;
    if (kn < nt) { ndn = __ldg(t4 + kn); if (S0) { fn = __ldcs(mu + kn); if (VE) Vn = __ldcs(VE + kn); } }
    double hx = 0.0, hy = 0.0, hz = 0.0;
    if (k < nt) {
      const double u0 = __ldg(u + nd.x), u1 = __ldg(u + nd.y), u2 = __ldg(u + nd.z), u3 = __ldg(u + nd.w);
      TetGeom G;
      tet_geometry(p4, nd, G);
      hx = -(u0 * G.g[0][0]) - u1 * G.g[1][0] - u2 * G.g[2][0] - u3 * G.g[3][0];
      hy = -(u0 * G.g[0][1]) - u1 * G.g[1][1] - u2 * G.g[2][1] - u3 * G.g[3][1];
      hz = -(u0 * G.g[0][2]) - u1 * G.g[1][2] - u2 * G.g[2][2] - u3 * G.g[3][2];
      if (compat) { hy += hz; hz = 0.0; }  // NewtonRaphson.jl:187-189: d-term lands in row 2
      __stcs(Hn + k, norm3(hx, hy, hz));
      if (S0) s0[k] = __dmul_rn(VE ? V : G.vol, f);
    }
    if (Hfield) {
      double* w = hs[wib];
      __syncwarp();
      w[3 * lane] = hx; w[3 * lane + 1] = hy; w[3 * lane + 2] = hz;
      __syncwarp();
      const long long kw = k - lane;                       // first element of this warp
      const int n3 = kw < nt ? 3 * (int)min((long long)32, nt - kw) : 0;
#pragma unroll
      for (int q = 0; q < 3; q++)
        if (32 * q + lane < n3) __stcs(Hfield + 3 * kw + 32 * q + lane, w[32 * q + lane]);
    }
    nd = ndn; f = fn; V = Vn;
  }
}

// One warp per slice of 32 items; a lane works through its item's contributions in batches of B: the B element ids and
// factors (coalesced, streaming) and then B independent gathers of s are in flight before the ordered fold starts.
// (A software pipeline over the slices of a warp -- next slice's first batch requested during the gathers of this
// one -- measured no faster at 85 registers and 2.3x slower where it spilled: profiles/r02g_*.)
constexpr int kGeoThreads = 256;
template <int B, int MINB>
static __global__ void __launch_bounds__(kGeoThreads, MINB) geo_rescale_kernel(const uint32_t* __restrict__ slice_ptr,
                                                                               const int* __restrict__ dst1, const int* __restrict__ dst2,
                                                                               const uint32_t* __restrict__ cref, const double* __restrict__ g,
                                                                               const double* __restrict__ s0, int nslices,
                                                                               double* __restrict__ vals) {
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  constexpr int WPB = kGeoThreads / 32;
  for (int s = blockIdx.x * WPB + wib; s < nslices; s += gridDim.x * WPB) {
    const uint32_t b0 = __ldg(slice_ptr + s);
    const int w = (int)(__ldg(slice_ptr + s + 1) - b0);
    const int d1 = __ldcs(dst1 + (size_t)s * 32 + lane), d2 = __ldcs(dst2 + (size_t)s * 32 + lane);
    const uint32_t* __restrict__ rp = cref + (size_t)b0 * 32 + lane;
    const double* __restrict__ gp = g + (size_t)b0 * 32 + lane;
    double acc = 0.0;
    for (int k = 0; k < w; k += B) {
      uint32_t rr[B];
      double gg[B], ss[B];
#pragma unroll
      for (int q = 0; q < B; q++) rr[q] = (k + q < w) ? __ldcs(rp + (size_t)(k + q) * 32) : 0xffffffffu;
#pragma unroll
      for (int q = 0; q < B; q++) gg[q] = (k + q < w) ? __ldcs(gp + (size_t)(k + q) * 32) : 0.0;
#pragma unroll
      for (int q = 0; q < B; q++) ss[q] = rr[q] != 0xffffffffu ? __ldg(s0 + rr[q]) : 0.0;
#pragma unroll
      for (int q = 0; q < B; q++)
        if (rr[q] != 0xffffffffu) acc = __dadd_rn(acc, __dmul_rn(ss[q], gg[q]));
    }
    if (d1 >= 0) {
      vals[d1] = acc;
      if (d2 >= 0) vals[d2] = acc;
    }
  }
}

// The same sums for a NON-SYMMETRIC matrix on the symmetric pattern (P2 convection matrix, p2.cuh): every contribution slot
// carries two factors, g1 for the item's canonical entry (dst1: the one with the lower row) and g2 for its mirror (dst2).
static __global__ void __launch_bounds__(kGeoThreads, 3) geo_rescale2_kernel(const uint32_t* __restrict__ slice_ptr,
                                                                             const int* __restrict__ dst1, const int* __restrict__ dst2,
                                                                             const uint32_t* __restrict__ cref, const double* __restrict__ g1,
                                                                             const double* __restrict__ g2, const double* __restrict__ s0,
                                                                             int nslices, double* __restrict__ vals) {
  constexpr int B = 4;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  constexpr int WPB = kGeoThreads / 32;
  for (int s = blockIdx.x * WPB + wib; s < nslices; s += gridDim.x * WPB) {
    const uint32_t b0 = __ldg(slice_ptr + s);
    const int w = (int)(__ldg(slice_ptr + s + 1) - b0);
    const int d1 = __ldcs(dst1 + (size_t)s * 32 + lane), d2 = __ldcs(dst2 + (size_t)s * 32 + lane);
    const size_t base = (size_t)b0 * 32 + lane;
    double acc1 = 0.0, acc2 = 0.0;
    for (int k = 0; k < w; k += B) {
      uint32_t rr[B];
      double ga[B], gb[B], ss[B];
#pragma unroll
      for (int q = 0; q < B; q++) rr[q] = (k + q < w) ? __ldcs(cref + base + (size_t)(k + q) * 32) : 0xffffffffu;
#pragma unroll
      for (int q = 0; q < B; q++) {
        ga[q] = (k + q < w) ? __ldcs(g1 + base + (size_t)(k + q) * 32) : 0.0;
        gb[q] = (k + q < w) ? __ldcs(g2 + base + (size_t)(k + q) * 32) : 0.0;
      }
#pragma unroll
      for (int q = 0; q < B; q++) ss[q] = rr[q] != 0xffffffffu ? __ldg(s0 + rr[q]) : 0.0;
#pragma unroll
      for (int q = 0; q < B; q++)
        if (rr[q] != 0xffffffffu) {
          acc1 = __dadd_rn(acc1, __dmul_rn(ss[q], ga[q]));
          acc2 = __dadd_rn(acc2, __dmul_rn(ss[q], gb[q]));
        }
    }
    if (d1 >= 0) {
      vals[d1] = acc1;
      if (d2 >= 0) vals[d2] = acc2;
    }
  }
}

// Diagonal of a matrix whose local matrices have zero row sums (stiffness: sum_j grad l_i . grad l_j = 0), from the
// off-diagonal entries the streamed sums just wrote: a_rr = -(sum_{c != r} a_rc), folded left to right in column order.
// One lane per row over the SELL slices, like the SpMV.  Replaces the 4 diagonal contributions of every element (40 % of
// the contribution stream); the result differs from the summed diagonal by rounding only (|delta| <~ 1e-15 sum |a_rc|).
static __global__ void __launch_bounds__(256) diag_from_rowsum_kernel(const int4* __restrict__ slice_meta, const int* __restrict__ srow,
                                                                      const int* __restrict__ sell_diag, int nslices, int n_rows,
                                                                      double* __restrict__ vals) {
  const int lane = threadIdx.x & 31;
  // last slices first: the streamed sums wrote the values in ascending order, so the tail is what L2 still holds
  const int sl = nslices - 1 - (int)(blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5));
  if (sl < 0) return;
  const int4 m = __ldg(slice_meta + sl);   // {value base, width, ...}
  const int r = __ldg(srow + sl * 32 + lane);
  const int dg = (r >= 0 && r < n_rows) ? __ldg(sell_diag + r) : -1;
  const double* __restrict__ v = vals + m.x + lane;
  double acc = 0.0;
  for (int k = 0; k < m.y; k += 4) {
    double a[4];
#pragma unroll
    for (int q = 0; q < 4; q++) a[q] = (k + q < m.y && m.x + (k + q) * 32 + lane != dg) ? __ldcg(v + (size_t)(k + q) * 32) : 0.0;
#pragma unroll
    for (int q = 0; q < 4; q++) acc = __dadd_rn(acc, a[q]);
  }
  if (dg >= 0) vals[dg] = -acc;
}
