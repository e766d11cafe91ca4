// K2a: SpMV on the SELL-32-sigma layout (included inside namespace fk; layout in symbolic.hpp).
//
// One thread owns one row, one warp one slice.  In step k the warp reads 32 consecutive values (256 B)
// and its columns (16-bit offsets from the slice's smallest column, four steps per 8-byte word, 256 B per warp
// load; 32-bit columns in a slice whose columns span >= 65536: packed stream, symbolic.hpp) -- the matrix stream, 10 B per entry on a mesh numbered
// with locality, is read exactly once and fully coalesced with streaming (evict-first) loads so it does not displace the PCG vectors from L2 --
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

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// a slice with 32-bit columns (columns span >= 65536: halo columns of a partition, meshes numbered without
// locality): same batches, one column load per entry
// COH: gather x with ld.global.cg (L2 only).  Used for the slices that read halo columns when the halo rows are
// written by peer GPUs while this kernel runs (fused exchange): the non-coherent path (ld.global.nc / L1) could serve
// a line fetched before the peer's store landed.
template <bool COH>
static __device__ __forceinline__ double gather_x(const double* p) { return COH ? __ldcg(p) : __ldg(p); }

template <int G, bool COH = false>
static __device__ __forceinline__ double sell_slice_row32(const double* __restrict__ v, const int* __restrict__ c, int w,
                                                          const double* __restrict__ x) {
  double acc = 0.0;
  for (int k = 0; k < w; k += 4 * G) {
    int cc[4 * G];
    double vv[4 * G], xx[4 * G];
#pragma unroll
    for (int u = 0; u < 4 * G; u++) cc[u] = (k + u < w) ? __ldcs(c + (k + u) * kSellC) : 0;
#pragma unroll
    for (int u = 0; u < 4 * G; u++) vv[u] = (k + u < w) ? __ldcs(v + (k + u) * kSellC) : 0.0;
#pragma unroll
    for (int u = 0; u < 4 * G; u++) xx[u] = (k + u < w) ? gather_x<COH>(x + cc[u]) : 0.0;
#pragma unroll
    for (int u = 0; u < 4 * G; u++)
      if (k + u < w) acc = __dadd_rn(acc, __dmul_rn(vv[u], xx[u]));
  }
  return acc;
}

// a slice with 16-bit column offsets: one 8-byte word per lane holds the columns of four consecutive steps, so
// a batch of 4 G entries issues G column loads + 4 G value loads, then 4 G independent gathers of x, and only
// then folds them left to right -- the memory-level parallelism comes from the batch, not from occupancy alone.
template <int G, bool COH = false>
static __device__ __forceinline__ double sell_slice_row16(const double* __restrict__ v, const uint2* __restrict__ c, int cbase, int w,
                                                          const double* __restrict__ x) {
  double acc = 0.0;
  const double* __restrict__ xb = x + cbase;
  for (int k = 0; k < w; k += 4 * G) {
    uint2 cw[G];
    double vv[4 * G], xx[4 * G];
#pragma unroll
    for (int g = 0; g < G; g++) cw[g] = (k + 4 * g < w) ? __ldcs(c + ((k >> 2) + g) * kSellC) : make_uint2(0u, 0u);
#pragma unroll
    for (int u = 0; u < 4 * G; u++) vv[u] = (k + u < w) ? __ldcs(v + (k + u) * kSellC) : 0.0;
#pragma unroll
    for (int u = 0; u < 4 * G; u++) {
      const unsigned word = (u & 2) ? cw[u >> 2].y : cw[u >> 2].x;
      const unsigned off = (u & 1) ? (word >> 16) : (word & 0xffffu);
      xx[u] = (k + u < w) ? gather_x<COH>(xb + off) : 0.0;
    }
#pragma unroll
    for (int u = 0; u < 4 * G; u++)
      if (k + u < w) acc = __dadd_rn(acc, __dmul_rn(vv[u], xx[u]));
  }
  return acc;
}

// COH: gather x through L2 only (the halo-coupled rows of the split exchange: their kernel runs beside the interior one,
// whose CTAs may hold the 128-byte line that straddles the last owned and the first halo row in L1 from before the
// peer's store landed -- the read-only path promises nothing for a line that is written during the kernel's lifetime)
template <bool DOT, int G, int MINB, bool COH = false>
static __global__ void __launch_bounds__(kSpmvThreads, MINB) spmv_sell_kernel(const int4* __restrict__ slice_meta, const int* __restrict__ srow,
                                                                  const unsigned short* __restrict__ colz,
                                                                  const double* __restrict__ vals, const double* __restrict__ x,
                                                                  double* __restrict__ y, int nslices, double* __restrict__ part,
                                                                  const RedArgs red, const PcgCtl* ctl, int j,
                                                                  const int* __restrict__ slice_list) {
  __shared__ double red_s[33];
  // the matrix stream does not depend on the previous kernel: pull the head of this warp's first slice towards L2
  // while the predecessor drains, then wait for its results (x, the control block)
  {
    const int q0 = blockIdx.x * (kSpmvThreads / 32) + (threadIdx.x >> 5);
    if (q0 < nslices) {
      const int sl0 = slice_list ? __ldg(slice_list + q0) : q0;
      const int4 m0 = __ldg(slice_meta + sl0);
      prefetch_l2(vals + m0.x + (threadIdx.x & 31));
      prefetch_l2(colz + (size_t)m0.z * kSellC + 4 * (threadIdx.x & 31));
    }
  }
  griddep_wait();
  griddep_launch();
  if (pcg_inactive(ctl, j)) return;
  unsigned long long* const tr = trace_rec(ctl, j, 0);
  trace_begin(tr);
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  constexpr int WPB = kSpmvThreads / 32;
  double dot = 0.0;
  // slice_list != nullptr: the kernel covers the nslices slices it names (partitions: interior / halo-coupled rows)
  for (int q = blockIdx.x * WPB + wib; q < nslices; q += gridDim.x * WPB) {
    const int sl = slice_list ? __ldg(slice_list + q) : q;
    const int4 m = __ldg(slice_meta + sl);   // {value base, width, column offset in 64-byte units, column base | -1}
    const int r = __ldg(srow + sl * kSellC + lane);
    const double* __restrict__ v = vals + m.x + lane;
    const unsigned short* __restrict__ cz = colz + (size_t)m.z * kSellC;
    const double acc = m.w >= 0 ? sell_slice_row16<G, COH>(v, reinterpret_cast<const uint2*>(cz) + lane, m.w, m.y, x)
                                : sell_slice_row32<(G > 2 ? 2 : G), COH>(v, reinterpret_cast<const int*>(cz) + lane, m.y, x);
    if (r >= 0) {
      y[r] = acc;
      if (DOT) dot += acc * x[r];
    }
  }
  if (DOT) {
    const double v[1] = {block_sum(dot, red_s)};
    grid_reduce_finish<1>(v, part, gridDim.x, red, red_s);   // -> out[0] = p.Ap over all ranks
  }
  trace_end(tr);
}


// ------------------------------------------------------------------------------------------------
// Partitioned problems, single-reduction recurrence: SpMV fused with the halo exchange of its input over NVLink
// peer memory -- ONE kernel, no separate push / wait launches, no second stream.
//   * the first hf.npush CTAs first store this rank's boundary rows of x straight into the neighbours' halo rows
//     (x was completed by the previous kernel); the last pusher to finish releases a system-scope sequence flag
//     to every neighbour;
//   * every CTA then walks slice_list: the slices that read no halo column first (the bulk: the peers' rows arrive
//     while they stream), then the halo-coupled ones, each warp waiting once for the neighbours' flags and gathering
//     x through L2 (ld.global.cg) there.
// The whole grid is resident (persistent launch), so the pushers always run; a rank only ever waits for stores that
// its peers issue at the START of the matching kernel, so there is no cyclic wait.
struct HaloFused {
  const int* send_idx;
  int send_ptr[kMaxNeigh + 1];
  double* dst[kMaxNeigh];                // neighbour q's copy of x at the rows it receives from me
  unsigned long long* flag[kMaxNeigh];   // neighbour q's flag word for me (its control block, kOffHalo2 + my slot there)
  unsigned long long* my_ctrl;
  unsigned int* ticket;                  // zero between kernels
  int nneigh, npush, n_int;              // slices [0, n_int) of slice_list read no halo column
};

template <int G, int MINB>
static __global__ void __launch_bounds__(kSpmvThreads, MINB) spmv_sell_fused_kernel(const int4* __restrict__ slice_meta, const int* __restrict__ srow,
                                                                        const unsigned short* __restrict__ colz,
                                                                        const double* __restrict__ vals, const double* __restrict__ x,
                                                                        double* __restrict__ y, int nslices, double* __restrict__ part,
                                                                        const PcgCtl* ctl, int j, const int* __restrict__ slice_list,
                                                                        const __grid_constant__ HaloFused hf) {
  __shared__ double red_s[33];
  __shared__ int last_pusher;
  if (pcg_inactive(ctl, j)) return;
  unsigned long long* const tr = trace_rec(ctl, j, 0);
  trace_begin(tr);
  const unsigned long long seq = cg_seq(ctl, ctl->base_it + j);
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  constexpr int WPB = kSpmvThreads / 32;
  // the LAST hf.npush CTAs push: the slice schedule below gives them one slice per warp less than the others, which pays
  // for the push (stores over NVLink, one system fence, the ticket), so they do not end the kernel late
  const int first_pusher = (int)gridDim.x - hf.npush;
  if ((int)blockIdx.x >= first_pusher) {
    const int pidx = (int)blockIdx.x - first_pusher;
    for (int q = 0; q < hf.nneigh; q++) {
      const int b = hf.send_ptr[q], n = hf.send_ptr[q + 1] - b;
      double* __restrict__ dst = hf.dst[q];
      for (int i = pidx * kSpmvThreads + threadIdx.x; i < n; i += hf.npush * kSpmvThreads) dst[i] = x[__ldg(hf.send_idx + b + i)];
    }
    __syncthreads();
    if (threadIdx.x == 0) {   // one fence per CTA: the barrier orders every thread's stores before it
      __threadfence_system();
      last_pusher = (atomicAdd(hf.ticket, 1u) == (unsigned)hf.npush - 1u) ? 1 : 0;
    }
    __syncthreads();
    if (last_pusher) {
      if ((int)threadIdx.x < hf.nneigh) {
        __threadfence_system();
        st_release_sys(hf.flag[threadIdx.x], seq);
      }
      if (threadIdx.x == 0) {
        *hf.ticket = 0u;
        if (tr) atomicMin(trace_rec(ctl, j, 3), gtime_ns());   // timeline record 3: {flags released, latest / ~earliest wait begin}
      }
    }
  }
  double dot = 0.0;
  bool waited = false;
  // static schedule: pass 0 covers the first Wn slices with the non-pushing warps only; every later pass is a stride of
  // all W warps.  Fixed per launch shape, so the per-CTA partial sums are reproducible.
  const int W = (int)gridDim.x * WPB, Wn = first_pusher * WPB, wid = (int)blockIdx.x * WPB + wib;
  for (int q = wid < Wn ? wid : Wn + wid; q < nslices; q = q < Wn ? Wn + wid : q + W) {
    const int sl = __ldg(slice_list + q);
    const int4 m = __ldg(slice_meta + sl);
    const int r = __ldg(srow + sl * kSellC + lane);
    const double* __restrict__ v = vals + m.x + lane;
    const unsigned short* __restrict__ cz = colz + (size_t)m.z * kSellC;
    double acc;
    if (q < hf.n_int) {
      acc = m.w >= 0 ? sell_slice_row16<G>(v, reinterpret_cast<const uint2*>(cz) + lane, m.w, m.y, x)
                     : sell_slice_row32<(G > 2 ? 2 : G)>(v, reinterpret_cast<const int*>(cz) + lane, m.y, x);
    } else {
      if (!waited) {   // once per warp: the neighbours' rows of this exchange have landed
        const unsigned long long tw0 = (tr && wib == 0) ? gtime_ns() : 0ULL;
        if (tr && threadIdx.x == 0) {
          atomicMax(trace_rec(ctl, j, 3) + 1, tw0);
          atomicMax(trace_rec(ctl, j, 3) + 2, ~tw0);
        }
        if (lane < hf.nneigh) spin_until_relaxed(hf.my_ctrl + kOffHalo2 + lane, seq, hf.my_ctrl + kOffErr);
        __syncwarp();
        waited = true;
        if (wib == 0) trace_wait(tr, tw0);
      }
      acc = m.w >= 0 ? sell_slice_row16<G, true>(v, reinterpret_cast<const uint2*>(cz) + lane, m.w, m.y, x)
                     : sell_slice_row32<(G > 2 ? 2 : G), true>(v, reinterpret_cast<const int*>(cz) + lane, m.y, x);
    }
    if (r >= 0) {
      y[r] = acc;
      dot += acc * x[r];
    }
  }
  const double bs = block_sum(dot, red_s);
  if (threadIdx.x == 0) part[blockIdx.x] = bs;
  trace_end(tr);
}

// launch shapes {entries per batch = 4 G, min CTAs per SM}; selected with fk_tune(h, 4, variant)
constexpr int kNumSpmvVariants = 6;
template <typename F>
static void spmv_variant_dispatch(int variant, F&& f) {
  switch (variant) {
    case 0: f(std::integral_constant<int, 1>{}, std::integral_constant<int, 8>{}); break;
    case 1: f(std::integral_constant<int, 1>{}, std::integral_constant<int, 6>{}); break;
    case 2: f(std::integral_constant<int, 2>{}, std::integral_constant<int, 5>{}); break;
    case 3: f(std::integral_constant<int, 2>{}, std::integral_constant<int, 4>{}); break;
    case 4: f(std::integral_constant<int, 4>{}, std::integral_constant<int, 3>{}); break;
    default: f(std::integral_constant<int, 4>{}, std::integral_constant<int, 2>{}); break;
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
