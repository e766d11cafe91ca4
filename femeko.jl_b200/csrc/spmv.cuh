// K2a: SpMV on the SELL-32-sigma layout (included inside namespace fk; layout in symbolic.hpp).
//
// One thread owns one row, one warp one slice.  In step k the warp reads 32 consecutive values (256 B)
// and 32 consecutive columns (128 B) -- the matrix stream, 12 B per entry, is read exactly once and
// fully coalesced with streaming (evict-first) loads so it does not displace the PCG vectors from L2 --
// and gathers x through the read-only path; on meshes numbered with locality neighbouring rows have
// neighbouring columns, so the gather of a step touches 2-3 lines instead of 32.  Each row is folded
// left to right with separately rounded multiply and add: the summation order of Eigen's column-major
// product on a symmetric pattern (and of the CSR loop), bitwise reproducible, no shared memory, no
// block-level synchronisation.  DOT fuses p.Ap: per-CTA partial sums, folded in index order by the last
// CTA to finish (and all-reduced over peer memory on a partitioned problem): grid_reduce_finish.
//
// Measured alternatives (profiles/r01b_spmv_variants.md): a CSR-stream kernel staging products in shared
// memory (85 us at 25.6 M entries) and the same with the stream moved by TMA bulk copies (90-300 us):
// both were bound by the per-entry x gather in L1 (32 distinct lines per warp request), not by HBM.

constexpr int kSellC = 32;
constexpr int kSellUnroll = 5;

template <bool DOT>
static __global__ void __launch_bounds__(kSpmvThreads) spmv_sell_kernel(const int* __restrict__ slice_ptr, const int* __restrict__ srow,
                                                                  const int* __restrict__ col, const double* __restrict__ vals,
                                                                  const double* __restrict__ x, double* __restrict__ y, int nslices,
                                                                  double* __restrict__ part, const RedArgs red, const PcgCtl* ctl,
                                                                  int j) {
  __shared__ double red_s[33];
  if (pcg_inactive(ctl, j)) return;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  constexpr int WPB = kSpmvThreads / 32;
  double dot = 0.0;
  for (int sl = blockIdx.x * WPB + wib; sl < nslices; sl += gridDim.x * WPB) {
    const int base = __ldg(slice_ptr + sl), w = (__ldg(slice_ptr + sl + 1) - base) >> 5;
    const int r = __ldg(srow + sl * kSellC + lane);
    const double* __restrict__ v = vals + base + lane;
    const int* __restrict__ c = col + base + lane;
    double acc = 0.0;
    int k = 0;
    for (; k + kSellUnroll <= w; k += kSellUnroll) {
      double vv[kSellUnroll];
      int cc[kSellUnroll];
#pragma unroll
      for (int u = 0; u < kSellUnroll; u++) {
        cc[u] = __ldcs(c + (k + u) * kSellC);
        vv[u] = __ldcs(v + (k + u) * kSellC);
      }
      double xx[kSellUnroll];
#pragma unroll
      for (int u = 0; u < kSellUnroll; u++) xx[u] = __ldg(x + cc[u]);
#pragma unroll
      for (int u = 0; u < kSellUnroll; u++) acc = __dadd_rn(acc, __dmul_rn(vv[u], xx[u]));
    }
    for (; k < w; k++) acc = __dadd_rn(acc, __dmul_rn(__ldcs(v + k * kSellC), __ldg(x + __ldcs(c + k * kSellC))));
    if (r >= 0) {
      y[r] = acc;
      if (DOT) dot += acc * x[r];
    }
  }
  if (DOT) {
    const double v[1] = {block_sum(dot, red_s)};
    grid_reduce_finish<1>(v, part, gridDim.x, red, red_s);   // -> out[0] = p.Ap over all ranks
  }
}

// host I/O in CSR order: out[q] = vals[csr2sell[q]]  /  vals[csr2sell[q]] = in[q]
static __global__ void sell_to_csr_kernel(const double* __restrict__ vals, const int* __restrict__ csr2sell, long long nnz,
                                          double* __restrict__ out) {
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < nnz; q += (long long)gridDim.x * blockDim.x)
    out[q] = vals[csr2sell[q]];
}
static __global__ void csr_to_sell_kernel(const double* __restrict__ in, const int* __restrict__ csr2sell, long long nnz,
                                          double* __restrict__ vals) {
  for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < nnz; q += (long long)gridDim.x * blockDim.x)
    vals[csr2sell[q]] = in[q];
}
