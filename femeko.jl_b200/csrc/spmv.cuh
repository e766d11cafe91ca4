// K2a: CSR SpMV with the matrix stream staged by TMA bulk copies (included inside namespace fk).
//
// The matrix (values 8 B + column 4 B per entry) is read exactly once per product and is by far the
// largest stream of a PCG iteration, so it is moved by the copy engine, not by the warps: each
// persistent CTA owns a ring of STAGES shared-memory stages; thread 0 issues three cp.async.bulk
// copies per row block (values, columns, row pointers) that complete on the stage's mbarrier, always
// STAGES blocks ahead of the consumers.  The warps only (1) multiply the staged values with gathered
// x (L1/L2 resident) in place and (2) fold each row's segment left to right -- the summation order of
// Eigen's column-major product on a symmetric pattern (SparseDenseProduct) -- and fuse p.Ap.
// Row blocks come from the symbolic phase: blk_info[b] = {r0, r1, q0 = rowptr[r0], q1 = rowptr[r1]}.
// Copies start at 16-byte aligned addresses at or below q0 / r0, so vals / colind / rowptr carry a few
// elements of slack at the end (allocated by the handle).

struct SpmvTmaArgs {
  const int* rowptr;
  const int* colind;
  const double* vals;
  const double* x;
  double* y;
  const int4* blk_info;
  int nblk;
  int cap;       // max entries per staged block
  int rows_cap;  // max rows per staged block
  int stages;
  double* part;
  const PcgCtl* ctl;
  int j;
};

__host__ __device__ inline int spmv_stage_vals_bytes(int cap) { return pad16((cap + 2) * 8); }
__host__ __device__ inline int spmv_stage_col_bytes(int cap) { return pad16((cap + 4) * 4); }
__host__ __device__ inline int spmv_stage_rp_bytes(int rows_cap) { return pad16((rows_cap + 1 + 4) * 4); }
__host__ __device__ inline int spmv_stage_bytes(int cap, int rows_cap) {
  return spmv_stage_vals_bytes(cap) + spmv_stage_col_bytes(cap) + spmv_stage_rp_bytes(rows_cap);
}

template <bool DOT>
static __global__ void __launch_bounds__(kSpmvThreads) spmv_tma_kernel(const SpmvTmaArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t full[8];
  __shared__ double red[33];
  if (pcg_inactive(a.ctl, a.j)) return;
  const int tid = threadIdx.x;
  const int S = a.stages;
  const int sbytes = spmv_stage_bytes(a.cap, a.rows_cap);
  const int vb = spmv_stage_vals_bytes(a.cap), cb = spmv_stage_col_bytes(a.cap);
  const int mine = (a.nblk - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // blocks of this CTA

  auto issue = [&](int it) {  // thread 0 only
    const int st = it % S;
    const int4 bi = __ldg(a.blk_info + blockIdx.x + (size_t)it * gridDim.x);
    unsigned char* base = smem_raw + (size_t)st * sbytes;
    if (bi.w - bi.z > a.cap) {  // long-row block: streamed straight from global by the consumers
      mbar_expect_tx(&full[st], 0);
      return;
    }
    const int qa = bi.z & ~1, qc = bi.z & ~3, ra = bi.x & ~3;
    const uint32_t nv_b = (uint32_t)pad16((bi.w - qa) * 8), nc_b = (uint32_t)pad16((bi.w - qc) * 4),
                   nr_b = (uint32_t)pad16((bi.y + 1 - ra) * 4);
    mbar_expect_tx(&full[st], nv_b + nc_b + nr_b);
    bulk_g2s(base, a.vals + qa, nv_b, &full[st]);
    bulk_g2s(base + vb, a.colind + qc, nc_b, &full[st]);
    bulk_g2s(base + vb + cb, a.rowptr + ra, nr_b, &full[st]);
  };

  if (tid == 0) {
    for (int s = 0; s < S; s++) mbar_init(&full[s], 1);
    mbar_init_fence();
    for (int it = 0; it < S && it < mine; it++) issue(it);
  }
  __syncthreads();

  double acc = 0.0;
  for (int it = 0; it < mine; it++) {
    const int st = it % S;
    const int4 bi = __ldg(a.blk_info + blockIdx.x + (size_t)it * gridDim.x);
    const int r0 = bi.x, r1 = bi.y, q0 = bi.z, q1 = bi.w;
    mbar_wait(&full[st], (uint32_t)((it / S) & 1));
    if (q1 - q0 <= a.cap) {
      unsigned char* base = smem_raw + (size_t)st * sbytes;
      double* v = reinterpret_cast<double*>(base) + (q0 & 1);
      const int* col = reinterpret_cast<const int*>(base + vb) + (q0 & 3);
      const int* rp = reinterpret_cast<const int*>(base + vb + cb) + (r0 & 3);
      const int cnt = q1 - q0;
      for (int q = tid; q < cnt; q += kSpmvThreads) v[q] *= __ldg(a.x + col[q]);
      __syncthreads();
      for (int r = tid; r < r1 - r0; r += kSpmvThreads) {
        const int b = rp[r] - q0, e = rp[r + 1] - q0;
        double s = 0.0;
        for (int q = b; q < e; q++) s += v[q];
        a.y[r0 + r] = s;
        if (DOT) acc += s * a.x[r0 + r];
      }
    } else {  // a row too long to stage (very high-degree node): whole CTA per row
      for (int r = r0; r < r1; r++) {
        double s = 0.0;
        for (int q = a.rowptr[r] + tid; q < a.rowptr[r + 1]; q += kSpmvThreads) s += a.vals[q] * __ldg(a.x + a.colind[q]);
        s = block_sum(s, red);
        if (tid == 0) {
          a.y[r] = s;
          if (DOT) acc += s * a.x[r];
        }
      }
    }
    __syncthreads();  // every reader of stage st is done
    if (tid == 0 && it + S < mine) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic writes (products) before the async refill
      issue(it + S);
    }
  }
  if (DOT) {
    acc = block_sum(acc, red);
    if (tid == 0) a.part[blockIdx.x] = acc;
  }
}
