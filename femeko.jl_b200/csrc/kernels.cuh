// Hand-written sm_100a fp64 kernels of the Femeko hot path.  All three kernel families are
// HBM-bound gather / stream kernels (no dense contraction -> no tensor cores):
//   K1  assemble_kernel<MODE>   P1 tet local matrices -> CSR values, tile-gather, no atomics
//   K2  spmv_sell_kernel / pcg_*  SELL-32 SpMV fused with p.Ap, fused x/r update + 2 dots, direction update
//   K3  recover_field_kernel    H = -grad(u), |H|, B = mu |H|
// Reference math: src/FEM3D.jl:5-17 (abcd), :59-76 (local stiffness), :466-476 (mass), :79-105
// (tangential), magneticField.jl:111-132 (field), ConjugateGradient.h:29-85 (PCG recurrence).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>

#include "../../include/femeko_cuda.h"
#include "symbolic.hpp"

namespace fk {

constexpr int kAsmThreads = 256;
constexpr int kSpmvThreads = 256;
constexpr int kVecThreads = 256;

// ------------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Programmatic dependent launch (the PCG iteration of one GPU): a kernel launched with the programmatic-serialization
// attribute may become resident while its predecessor still runs; it must not touch the predecessor's results before
// griddep_wait() (which returns once the predecessor has completed and its writes are visible; a no-op for a plain
// launch).  griddep_launch() lets the successor's CTAs take SM slots as soon as this kernel's CTAs free them.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Deterministic block reduction (fixed shuffle tree, warps folded in index order).  Every thread
// returns the total.  `scratch` holds >= 33 doubles.  Safe to call repeatedly.
__device__ __forceinline__ double block_sum(double v, double* scratch) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) scratch[w] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int i = 0; i < nw; i++) s += scratch[i];
    scratch[32] = s;
  }
  __syncthreads();
  return scratch[32];
}
__device__ __forceinline__ double block_max(double v, double* scratch) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) scratch[w] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = scratch[0];
    for (int i = 1; i < nw; i++) s = fmax(s, scratch[i]);
    scratch[32] = s;
  }
  __syncthreads();
  return scratch[32];
}
// Sum of a short global array by the whole block in a fixed order (every block gets the same bits).
__device__ __forceinline__ double block_sum_array(const double* __restrict__ a, int n, double* scratch) {
  double v = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) v += a[i];
  return block_sum(v, scratch);
}

// two sums in one pass (same tree and fold order as block_sum for each); scratch holds >= 66 doubles
__device__ __forceinline__ void block_sum2(double& a, double& b, double* scratch) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  a = warp_sum(a);
  b = warp_sum(b);
  __syncthreads();
  if (lane == 0) { scratch[w] = a; scratch[33 + w] = b; }
  __syncthreads();
  if (threadIdx.x < 2) {
    double s = 0.0;
    for (int i = 0; i < nw; i++) s += scratch[33 * threadIdx.x + i];
    scratch[33 * threadIdx.x + 32] = s;
  }
  __syncthreads();
  a = scratch[32];
  b = scratch[65];
}
__device__ __forceinline__ void block_sum2_array(const double* __restrict__ pa, const double* __restrict__ pb, int n, double* scratch,
                                                 double& a, double& b) {
  a = 0.0;
  b = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { a += pa[i]; b += pb[i]; }
  block_sum2(a, b, scratch);
}

// ------------------------------------------------------------------------------------------------
// Grid-wide reductions finished INSIDE the producing kernel, and across GPUs over peer memory.
// Every CTA publishes its partial sums; the CTA that takes the last ticket folds all partials in index
// order (so the result does not depend on which CTA was last), and -- on a partitioned problem -- stores
// the operands into every rank's mailbox over NVLink, waits for every rank's operands in its own mailbox
// and combines them in rank order: identical bits on every rank, no extra kernel, no NCCL call.
// Control block layout (64-bit words) of a rank: halo flags | mailbox [2 slots][kMaxRanks][kMailWords] |
// error | count of executed reductions | count of executed halo exchanges.
// ------------------------------------------------------------------------------------------------
constexpr int kMaxRanks = 8;       // one NVSwitch node
constexpr int kMaxNeigh = 32;
constexpr int kMailWords = 4;      // per (slot, rank): 3 values + sequence flag
constexpr int kOffMail = kMaxNeigh;
constexpr int kOffErr = kOffMail + 2 * kMaxRanks * kMailWords;
constexpr int kOffCnt = kOffErr + 1;
constexpr int kOffHaloCnt = kOffCnt + 1;
constexpr int kMail2Words = 4;     // consumer-side fold of the PCG iteration: 2 values x 2 LL words per (slot, rank)
constexpr int kOffMail2 = kOffHaloCnt + 1;
// single-reduction recurrence (cg_update_kernel / fused SpMV): halo flags of the in-kernel exchange and a mailbox of
// 3 values + sequence flag per (slot, rank)
constexpr int kOffHalo2 = kOffMail2 + 2 * kMaxRanks * kMail2Words;
constexpr int kMail3Words = 8;   // 4 values x 2 LL words
constexpr int kOffMail3 = kOffHalo2 + kMaxNeigh;
constexpr int kOffLimit = kOffMail3 + 4 * kMaxRanks * kMail3Words;   // spin limit in clock64 ticks (0 = default, ~2 s): fk_tune 16
constexpr int kCtrlWords = 512;
static_assert(kOffLimit + 1 <= kCtrlWords, "control block too small");
__device__ __forceinline__ long long spin_limit(const unsigned long long* err) {   // err = ctrl + kOffErr
  const unsigned long long v = err[kOffLimit - kOffErr];
  return v ? (long long)v : 4000000000LL;
}
constexpr int kSlabVecs = 8;       // x | r | p | Ap (w) | dinv | s | z | potential staging -- then the control block

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
// same contract as spin_until, polling with relaxed loads and ONE acquire load after success (the polls of ~600 CTAs
// then cost an L2 read each instead of an acquire).  Not a fence: under the load of a running SpMV a system-scope
// fence (MEMBAR.SYS drains the SM's outstanding writes) was measured at ~8 us per waiting warp (fk_tune 11 timeline).
__device__ __forceinline__ bool spin_until_relaxed(const unsigned long long* flag, unsigned long long seq, unsigned long long* err) {
  const long long t0 = clock64(), lim = spin_limit(err);
  bool ok = true;
  while (ld_relaxed_sys(flag) < seq) {
    if (clock64() - t0 > lim) {
      *err = 1ULL;
      ok = false;
      break;
    }
  }
  (void)ld_acquire_sys(flag);
  return ok;
}
// spin until *flag >= seq; gives up after ~2 s (a lost peer must not hang the GPU) and records the failure
__device__ __forceinline__ bool spin_until(const unsigned long long* flag, unsigned long long seq, unsigned long long* err) {
  const long long t0 = clock64(), lim = spin_limit(err);
  while (ld_acquire_sys(flag) < seq) {
    if (clock64() - t0 > lim) {
      *err = 1ULL;
      return false;
    }
  }
  return true;
}

struct RedArgs {
  unsigned int* ticket;                        // zero between kernels
  double* out;                                 // reduced scalars
  int op;                                      // 0 = sum, 1 = max
  int rank, nranks;                            // nranks <= 1 or my_ctrl == nullptr: this GPU only
  unsigned long long* const* peer_ctrl;        // control block of every rank (peer memory)
  unsigned long long* my_ctrl;
};

// exchange `nvals` (<= 3) scalars held in val[] (shared memory) with every rank; result back in val[].
// Called by a whole CTA (>= kMaxRanks threads).  The sequence number is a device-side count of EXECUTED
// reductions (every rank executes the same ones), so consecutive reductions alternate mailbox slots and a
// rank can run at most one reduction ahead of the slowest reader.
__device__ __forceinline__ void peer_allreduce(double* val, int nvals, const RedArgs& ra) {
  const unsigned long long seq = ra.my_ctrl[kOffCnt] + 1ULL;
  __syncthreads();
  const int slot = (int)(seq & 1ULL);
  if ((int)threadIdx.x < ra.nranks) {
    unsigned long long* box = ra.peer_ctrl[threadIdx.x] + kOffMail + ((size_t)slot * kMaxRanks + ra.rank) * kMailWords;
    for (int v = 0; v < nvals; v++) reinterpret_cast<double*>(box)[v] = val[v];
    __threadfence_system();
    st_release_sys(box + 3, seq);
    spin_until(ra.my_ctrl + kOffMail + ((size_t)slot * kMaxRanks + threadIdx.x) * kMailWords + 3, seq, ra.my_ctrl + kOffErr);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int v = 0; v < nvals; v++) {
      double s = 0.0;
      for (int r = 0; r < ra.nranks; r++) {
        const double x = reinterpret_cast<volatile double*>(ra.my_ctrl + kOffMail + ((size_t)slot * kMaxRanks + r) * kMailWords)[v];
        s = (ra.op == 0) ? s + x : fmax(s, x);
      }
      val[v] = s;
    }
    ra.my_ctrl[kOffCnt] = seq;
  }
  __syncthreads();
}

// v[k] (valid in thread 0): this CTA's partial of value k.  part: NV x stride scratch, stride >= gridDim.x.
// ra.ticket == nullptr: only publish the partials (measured on one B200: letting every CTA of the consumer
// fold the ~1200 partials in parallel is 7 us per iteration faster than the serial tail of a last CTA).
template <int NV>
__device__ __forceinline__ void grid_reduce_finish(const double (&v)[NV], double* __restrict__ part, int stride, const RedArgs& ra,
                                                   double* scratch /* >= 33 doubles */) {
  __shared__ int last_flag;
  __shared__ double val[3];
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) part[(size_t)k * stride + blockIdx.x] = v[k];
    if (ra.ticket == nullptr) {
      last_flag = 0;   // single GPU: the consuming kernel folds the partials itself (every CTA in parallel, no serial tail)
    } else {
      __threadfence();
      last_flag = (atomicAdd(ra.ticket, 1u) == gridDim.x - 1) ? 1 : 0;
    }
  }
  __syncthreads();
  if (!last_flag) return;
  __threadfence();
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double s = (ra.op == 0) ? 0.0 : -DBL_MAX;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) {
      const double x = __ldcg(part + (size_t)k * stride + i);
      s = (ra.op == 0) ? s + x : fmax(s, x);
    }
    s = (ra.op == 0) ? block_sum(s, scratch) : block_max(s, scratch);
    if (threadIdx.x == 0) val[k] = s;
  }
  __syncthreads();
  if (ra.nranks > 1 && ra.my_ctrl != nullptr) peer_allreduce(val, NV, ra);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) ra.out[k] = val[k];
    *ra.ticket = 0u;
  }
}

// Dot products of a PCG iteration on a partitioned problem, folded on the CONSUMER side: the producing kernel only
// publishes its per-CTA partials (no last-CTA tail); every CTA of the consuming kernel folds them (identical bits
// in every CTA), CTA 0 stores this GPU's sums into every peer's mailbox over NVLink, and every CTA waits for the
// peers' sums in its own GPU's mailbox and adds them in rank order -> identical bits on every rank.
// The mailbox speaks NCCL's low-latency (LL) protocol: a double travels as two 8-byte words, each {32 data bits | 32-bit
// tag}; an aligned 8-byte store is single-copy atomic, so a word whose tag matches carries its data -- no fence on
// either side, one NVLink crossing per reduction (measured with fk_tune 11: the release / acquire version with two
// system fences on the sender cost 8.8 us per reduction on two B200s).  Tags count the reductions of a handle
// (PcgCtl::tag_base + position inside the solve, identical on every rank), so a slot never sees the same tag twice in
// a row; consecutive reductions alternate slots, and a rank can only reach the next use of a slot after the peer's
// previous readers of it have finished (it needs the peer's contribution to the reduction in between).
struct PeerRed {
  int rank, nranks;                            // nranks <= 1: this GPU only
  unsigned long long* const* peer_ctrl;
  unsigned long long* my_ctrl;
};
__device__ __forceinline__ void st_relaxed_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void ll_send(unsigned long long* box, const double v, const unsigned tag) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  st_relaxed_sys(box, (b << 32) | (unsigned long long)tag);
  st_relaxed_sys(box + 1, (b & 0xffffffff00000000ULL) | (unsigned long long)tag);
}
__device__ __forceinline__ double ll_recv(const unsigned long long* box, const unsigned tag, unsigned long long* err) {
  const long long t0 = clock64(), lim = spin_limit(err);
  unsigned long long w0, w1;
  while (true) {
    w0 = ld_relaxed_sys(box);
    w1 = ld_relaxed_sys(box + 1);
    if ((unsigned)w0 == tag && (unsigned)w1 == tag) break;
    if (clock64() - t0 > lim) { *err = 1ULL; break; }   // a lost peer must not hang the GPU
  }
  return __longlong_as_double((long long)((w0 >> 32) | (w1 & 0xffffffff00000000ULL)));
}
// NV values per (slot, rank), two words each, at ctrl[base + (slot * kMaxRanks + rank) * 2 NV]
template <int NV>
__device__ __forceinline__ void peer_fold_ll(double (&v)[NV], const PeerRed& pr, const unsigned tag, const int slot, const int base) {
  if (pr.nranks <= 1) return;
  __shared__ double sh[NV * kMaxRanks];
  const int t = threadIdx.x;
  if (t < pr.nranks * NV) {   // one thread per (rank, value): the stores and the polls of a reduction run side by side
    const int r = t / NV, k = t - r * NV;
    if (r == pr.rank) sh[k * kMaxRanks + r] = v[k];
    else {
      if (blockIdx.x == 0) ll_send(pr.peer_ctrl[r] + base + ((size_t)slot * kMaxRanks + pr.rank) * (2 * NV) + 2 * k, v[k], tag);
      sh[k * kMaxRanks + r] = ll_recv(pr.my_ctrl + base + ((size_t)slot * kMaxRanks + r) * (2 * NV) + 2 * k, tag, pr.my_ctrl + kOffErr);
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double s = 0.0;
    for (int r = 0; r < pr.nranks; r++) s += sh[k * kMaxRanks + r];
    v[k] = s;
  }
  __syncthreads();
}
// the two reductions of Eigen's recurrence (p.Ap | r.r, r.z): slots alternate with the parity of seq
template <int NV>
__device__ __forceinline__ void peer_fold(double (&v)[NV], const PeerRed& pr, unsigned long long seq, unsigned tag_base) {
  static_assert(2 * NV <= kMail2Words, "mailbox holds 2 values");
  // both reductions share the slots (one 4-word record per rank): a 1-value reduction uses the first two words
  if (pr.nranks <= 1) return;
  __shared__ double sh[NV * kMaxRanks];
  const int t = threadIdx.x, slot = (int)(seq & 1ULL);
  const unsigned tag = tag_base + (unsigned)seq;
  if (t < pr.nranks * NV) {
    const int r = t / NV, k = t - r * NV;
    if (r == pr.rank) sh[k * kMaxRanks + r] = v[k];
    else {
      if (blockIdx.x == 0) ll_send(pr.peer_ctrl[r] + kOffMail2 + ((size_t)slot * kMaxRanks + pr.rank) * kMail2Words + 2 * k, v[k], tag);
      sh[k * kMaxRanks + r] = ll_recv(pr.my_ctrl + kOffMail2 + ((size_t)slot * kMaxRanks + r) * kMail2Words + 2 * k, tag, pr.my_ctrl + kOffErr);
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double s = 0.0;
    for (int r = 0; r < pr.nranks; r++) s += sh[k * kMaxRanks + r];
    v[k] = s;
  }
  __syncthreads();
}

// Four values (gamma, r.r, delta, sum r: the single fused reduction of the single-reduction recurrence), same protocol on
// its own mailbox; consecutive reductions alternate slots by the parity of seq, consecutive solves by the parity of the
// epoch (a rank two hops away may start its next solve while a slow CTA here still reads the last reduction of this one).
__device__ __forceinline__ void peer_fold3(double (&v)[4], const PeerRed& pr, unsigned long long seq, int epoch_parity, unsigned tag_base) {
  peer_fold_ll<4>(v, pr, tag_base + (unsigned)seq, (int)(seq & 1ULL) + 2 * epoch_parity, kOffMail3);
}

// ------------------------------------------------------------------------------------------------
// P1 geometry: gradients of the barycentric shape functions of one tetrahedron.
// Replaces the four 4x4 solves of abcd (src/FEM3D.jl:5-17 / FEMc.cpp:18-42): with
// E = [x1-x0 | x2-x0 | x3-x0], rows of inv(E) are grad(lambda_1..3), grad(lambda_0) = -(sum).
// Nodes are padded to 32 B (x,y,z,0) so one node = one aligned sector = two 128-bit loads.
// ------------------------------------------------------------------------------------------------
// |h|: exact zeros (potential still flat, e.g. the first Picard step) skip the fp64 sqrt, whose zero / denormal
// inputs take the slow path of the software sequence; sqrt(0) = 0, so the value is the same bit for bit
__device__ __forceinline__ double norm3(double hx, double hy, double hz) {
  const double h2 = hx * hx + hy * hy + hz * hz;
  return h2 > 0.0 ? sqrt(h2) : h2;
}

struct TetGeom {
  double g[4][3];  // grad(lambda_i)
  double vol;      // |det E| / 6  (elementVolume, src/mesh.jl:484-505)
};

// one node = one aligned 32-byte sector = ONE 256-bit load (LDG.E.256 on sm_100a): the gather kernels are bound by the
// L1 tag stage (one line per cycle), so half the load instructions is half the tag traffic of the coordinate gathers
__device__ __forceinline__ void load_node(const double4* __restrict__ p4, int n, double& x, double& y, double& z) {
  [[maybe_unused]] double w;
  asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(x), "=d"(y), "=d"(z), "=d"(w) : "l"(p4 + n));
}

__device__ __forceinline__ void tet_geometry_xyz(const double x0, const double y0, const double z0, const double x1, const double y1,
                                                 const double z1, const double x2, const double y2, const double z2, const double x3,
                                                 const double y3, const double z3, TetGeom& G) {
  const double ax = x1 - x0, ay = y1 - y0, az = z1 - z0;  // AB
  const double bx = x2 - x0, by = y2 - y0, bz = z2 - z0;  // AC
  const double cx = x3 - x0, cy = y3 - y0, cz = z3 - z0;  // AD
  // cofactors
  const double c1x = by * cz - bz * cy, c1y = bz * cx - bx * cz, c1z = bx * cy - by * cx;  // AC x AD
  const double c2x = cy * az - cz * ay, c2y = cz * ax - cx * az, c2z = cx * ay - cy * ax;  // AD x AB
  const double c3x = ay * bz - az * by, c3y = az * bx - ax * bz, c3z = ax * by - ay * bx;  // AB x AC
  const double det = ax * c1x + ay * c1y + az * c1z;                                      // AB . (AC x AD)
  const double inv = 1.0 / det;
  G.g[1][0] = c1x * inv; G.g[1][1] = c1y * inv; G.g[1][2] = c1z * inv;
  G.g[2][0] = c2x * inv; G.g[2][1] = c2y * inv; G.g[2][2] = c2z * inv;
  G.g[3][0] = c3x * inv; G.g[3][1] = c3y * inv; G.g[3][2] = c3z * inv;
  G.g[0][0] = -(G.g[1][0] + G.g[2][0] + G.g[3][0]);
  G.g[0][1] = -(G.g[1][1] + G.g[2][1] + G.g[3][1]);
  G.g[0][2] = -(G.g[1][2] + G.g[2][2] + G.g[3][2]);
  G.vol = fabs(det) / 6.0;
}

// grad(lambda_i) . grad(lambda_j) with the roundings spelled out (one product, two fused multiply-adds), so that every
// kernel that forms it -- the tile kernel, the dense local matrices, the one-off factors of the rescale path -- gets
// the same bits whatever the compiler would contract
__device__ __forceinline__ double grad_dot3(const double a0, const double a1, const double a2, const double b0, const double b1,
                                            const double b2) {
  return __fma_rn(a2, b2, __fma_rn(a1, b1, __dmul_rn(a0, b0)));
}

__device__ __forceinline__ void tet_geometry(const double4* __restrict__ p4, const int4 nd, TetGeom& G) {
  double x0, y0, z0, x1, y1, z1, x2, y2, z2, x3, y3, z3;
  load_node(p4, nd.x, x0, y0, z0);
  load_node(p4, nd.y, x1, y1, z1);
  load_node(p4, nd.z, x2, y2, z2);
  load_node(p4, nd.w, x3, y3, z3);
  tet_geometry_xyz(x0, y0, z0, x1, y1, z1, x2, y2, z2, x3, y3, z3, G);
}

// ------------------------------------------------------------------------------------------------
// FITPACK splev / fpbspl on the device (the Dierckx Spline1D call of NewtonRaphson.jl:203-207).
// ------------------------------------------------------------------------------------------------
struct SplineDev {
  const double* t;  // n knots
  const double* c;  // n coefficients (n-k-1 used)
  int n, k, bc;
};
struct SplineSet {
  SplineDev s[FK_MAX_SPLINES];
};

__device__ __forceinline__ double splev_dev(const SplineDev& S, double x, int& err) {
  const double* __restrict__ t = S.t;
  const int n = S.n, k = S.k;
  const double tb = __ldg(t + k), te = __ldg(t + n - k - 1);
  if (x < tb || x > te || !(x == x)) {
    if (S.bc == FK_BC_ZERO) return 0.0;
    if (S.bc == FK_BC_ERROR || !(x == x)) { err = 1; return x; }
    if (S.bc == FK_BC_NEAREST) x = (x < tb) ? tb : te;
  }
  // largest l in [k, n-k-2] with t[l] <= x (binary search; FITPACK walks linearly)
  int lo = k, hi = n - k - 2;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (__ldg(t + mid) <= x) lo = mid; else hi = mid - 1;
  }
  const int l = lo;
  double h[6], hh[5];
  h[0] = 1.0;
  for (int j = 1; j <= k; j++) {
    for (int i = 0; i < j; i++) hh[i] = h[i];
    h[0] = 0.0;
    for (int i = 0; i < j; i++) {
      const double tli = __ldg(t + l + i + 1), tlj = __ldg(t + l + i + 1 - j);
      if (tli == tlj) { h[i + 1] = 0.0; continue; }
      const double f = hh[i] / (tli - tlj);
      h[i] = h[i] + f * (tli - x);
      h[i + 1] = f * (x - tlj);
    }
  }
  double sp = 0.0;
  for (int j = 0; j <= k; j++) sp += __ldg(S.c + l - k + j) * h[j];
  return sp;
}

#include "assemble.cuh"
#include "rescale.cuh"
#include "p2.cuh"
#include "nedelec.cuh"
#include "demag.cuh"

// mu(|H|), dmu/dH of the material elements from the B-H spline (NewtonRaphson.jl:178-207; manyBodyIntersecting.jl:189-197),
// run just before the fused sweep of a Picard step: one thread per listed element recomputes H from the potential,
// evaluates the spline(s), stores mu / dmu and folds |  |H| - |H|old  | into a per-CTA maximum (the convergence
// metric of NewtonRaphson.jl:200).  |H|old is read here, before the sweep overwrites it.  Linear elements are never
// touched, so the sweep itself carries no spline code.
static __global__ void __launch_bounds__(256) material_update_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4,
                                                              const int* __restrict__ list, int n,
                                                              const unsigned char* __restrict__ matid,
                                                              const double* __restrict__ u, const SplineSet spl, int compat,
                                                              const double* __restrict__ Hn_old, double* __restrict__ mu,
                                                              double* __restrict__ dmu, double* __restrict__ part_max,
                                                              int* __restrict__ errflag, const double* __restrict__ VE = nullptr,
                                                              double* __restrict__ s0 = nullptr) {
  __shared__ double red[33];
  double dmax = 0.0;
  int err = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int k = __ldg(list + i);
    const int mat = (int)__ldg(matid + k);
    if (mat <= 0) continue;
    const int4 nd = __ldg(t4 + k);
    TetGeom G;
    tet_geometry(p4, nd, G);
    const double u0 = __ldg(u + nd.x), u1 = __ldg(u + nd.y), u2 = __ldg(u + nd.z), u3 = __ldg(u + nd.w);
    double hx = -(u0 * G.g[0][0]) - u1 * G.g[1][0] - u2 * G.g[2][0] - u3 * G.g[3][0];
    double hy = -(u0 * G.g[0][1]) - u1 * G.g[1][1] - u2 * G.g[2][1] - u3 * G.g[3][1];
    double hz = -(u0 * G.g[0][2]) - u1 * G.g[1][2] - u2 * G.g[2][2] - u3 * G.g[3][2];
    if (compat) { hy += hz; hz = 0.0; }  // NewtonRaphson.jl:187-189: d-term lands in row 2
    const double Hn = norm3(hx, hy, hz);
    const double f = splev_dev(spl.s[mat - 1], Hn, err);
    double df = 0.0;
    const SplineDev& D = spl.s[FK_MAX_SPLINES / 2 + mat - 1];
    if (dmu && D.n > 0) df = splev_dev(D, Hn, err);
    if (!(f == f) || fabs(f) > DBL_MAX || !(df == df) || fabs(df) > DBL_MAX) err |= 2;  // Nans/Infs in mu
    dmax = fmax(dmax, fabs(Hn - Hn_old[k]));
    mu[k] = f;
    if (dmu) dmu[k] = df;
    if (s0) s0[k] = __dmul_rn(VE ? VE[k] : G.vol, f);   // rescale path: the element scale follows its mu (rescale.cuh)
  }
  if (err) atomicOr(errflag, err);
  dmax = block_max(dmax, red);
  if (threadIdx.x == 0) part_max[blockIdx.x] = dmax;
}

// y = spl(x) on the device (fk_spline_apply): the same splev_dev the Picard pre-pass calls
static __global__ void splev_kernel(const SplineDev S, const double* __restrict__ x, int n, double* __restrict__ y, int* __restrict__ errflag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int err = 0;
  y[i] = splev_dev(S, x[i], err);
  if (err) atomicOr(errflag, err);
}

// dense 16 x nt local stiffness (localStiffnessMatrix, src/FEM3D.jl:59-76) -- API completeness
static __global__ void local_stiffness_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4,
                                       const double* __restrict__ VE, const double* __restrict__ f, int nt,
                                       double* __restrict__ Ak) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nt) return;
  TetGeom G;
  tet_geometry(p4, __ldg(t4 + k), G);
  const double V = VE ? VE[k] : G.vol;
  const double s0 = f ? V * f[k] : V;
  for (int j = 0; j < 4; j++)
    for (int i = 0; i < 4; i++)
      Ak[16 * (size_t)k + 4 * j + i] = __dmul_rn(s0, grad_dot3(G.g[i][0], G.g[i][1], G.g[i][2], G.g[j][0], G.g[j][1], G.g[j][2]));
}

static __global__ void volumes_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4, int nt, double* __restrict__ VE) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nt) return;
  TetGeom G;
  tet_geometry(p4, __ldg(t4 + k), G);
  VE[k] = G.vol;
}

// ------------------------------------------------------------------------------------------------
// K3: field recovery  H_k = -sum_i u[t_i] grad(lambda_i)  (magneticField.jl:111-132)
// ------------------------------------------------------------------------------------------------
static __global__ void __launch_bounds__(256) recover_field_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4,
                                                            const double* __restrict__ u, const double* __restrict__ mu,
                                                            int nt, int compat, double* __restrict__ Hfield,
                                                            double* __restrict__ Hn, double* __restrict__ B) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nt) return;
  const int4 nd = __ldg(t4 + k);
  TetGeom G;
  tet_geometry(p4, nd, G);
  const double u0 = __ldg(u + nd.x), u1 = __ldg(u + nd.y), u2 = __ldg(u + nd.z), u3 = __ldg(u + nd.w);
  double hx = -(u0 * G.g[0][0]) - u1 * G.g[1][0] - u2 * G.g[2][0] - u3 * G.g[3][0];
  double hy = -(u0 * G.g[0][1]) - u1 * G.g[1][1] - u2 * G.g[2][1] - u3 * G.g[3][1];
  double hz = -(u0 * G.g[0][2]) - u1 * G.g[1][2] - u2 * G.g[2][2] - u3 * G.g[3][2];
  if (compat) { hy += hz; hz = 0.0; }
  const double H = norm3(hx, hy, hz);
  if (Hfield) { Hfield[3 * (size_t)k] = hx; Hfield[3 * (size_t)k + 1] = hy; Hfield[3 * (size_t)k + 2] = hz; }
  if (Hn) Hn[k] = H;
  if (B) B[k] = __ldg(mu + k) * H;
}

// ------------------------------------------------------------------------------------------------
// K2: SpMV + Jacobi-PCG
// ------------------------------------------------------------------------------------------------
struct PcgCtl {
  double absNew[2];  // r.z of the current / next iteration (ping-pong by iteration parity)
  double rhs2, thr, r2;
  int base_it, max_iter, done, iters;
  unsigned long long epoch;   // solve counter (same on every rank): high half of the peer_fold sequence numbers
  double gamma[2], alpha[2];  // single-reduction recurrence: (r,u) and alpha of the previous / current iteration (ping-pong)
  // fk_pcg_opts.project_nullspace: the mean of r (rounding drift out of range(A) on a singular Neumann system) measured by
  // one iteration is subtracted by the next one, so the recursive residual can fall to the requested tolerance
  double rmean[2], inv_n;
  int proj;
  // fk_tune(h, 11, iterations): device-side timeline of the iteration's kernels (nullptr = off).  Record of iteration it, kernel k
  // at trace[((it % trace_cap) * kTraceKernels + k) * 4 ..]: {min CTA start, max CTA end, longest cross-GPU wait, waits of CTA 0} in
  // ns of %globaltimer
  unsigned long long* trace;
  int trace_cap;
  unsigned tag_base;   // LL mailbox tags of this solve start here (host: running bound on the reductions of earlier solves)
};
constexpr int kTraceKernels = 4;   // 0 SpMV, 1 update (cg_update), 2 direction, 3 halo push / wait kernels

__device__ __forceinline__ unsigned long long gtime_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned long long* trace_rec(const PcgCtl* c, int j, int kernel) {
  if (c == nullptr || c->trace == nullptr) return nullptr;
  return c->trace + ((size_t)((c->base_it + j) % c->trace_cap) * kTraceKernels + kernel) * 4;
}
__device__ __forceinline__ void trace_begin(unsigned long long* tr) {
  if (tr != nullptr && threadIdx.x == 0) atomicMin(tr, gtime_ns());
}
__device__ __forceinline__ void trace_end(unsigned long long* tr) {
  if (tr != nullptr && threadIdx.x == 0) atomicMax(tr + 1, gtime_ns());
}
__device__ __forceinline__ void trace_wait(unsigned long long* tr, unsigned long long t0) {
  if (tr != nullptr && threadIdx.x == 0) {
    const unsigned long long dt = gtime_ns() - t0;
    atomicMax(tr + 2, dt);
    if (blockIdx.x == 0) tr[3] = dt;
  }
}

__device__ __forceinline__ unsigned long long pcg_seq(const PcgCtl* c, int j, int which) {
  return (c->epoch << 32) + 2ULL * (unsigned long long)(c->base_it + j) + (unsigned long long)which + 1ULL;
}

// sequence numbers of the single-reduction recurrence: one reduction (V_it) and one halo exchange (S_it) per iteration
__device__ __forceinline__ unsigned long long cg_seq(const PcgCtl* c, int it) {
  return (c->epoch << 32) + (unsigned long long)it + 1ULL;
}

__device__ __forceinline__ bool pcg_inactive(const PcgCtl* c, int j) {
  return c != nullptr && (c->done != 0 || c->base_it + j >= c->max_iter);
}

#include "spmv.cuh"

// dinv = 1/diag(A), 1 where the diagonal is zero or missing (BasicPreconditioners.h:62-75)
static __global__ void jacobi_kernel(const double* __restrict__ vals, const int* __restrict__ diag, int n, int identity,
                              double* __restrict__ dinv) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  double d = 1.0;
  if (!identity && diag[r] >= 0) {
    const double v = vals[diag[r]];
    if (v != 0.0) d = 1.0 / v;
  }
  dinv[r] = d;
}

// r = b - Ax0 (Ax0 nullable), p = dinv r, partials of b.b, r.r, r.p   (ConjugateGradient.h:38-59)
static __global__ void __launch_bounds__(kVecThreads) pcg_init_kernel(const double* __restrict__ b, const double* __restrict__ Ax0,
                                                               const double* __restrict__ dinv, int n, double* __restrict__ r,
                                                               double* __restrict__ p, double* __restrict__ part3, const RedArgs red) {
  __shared__ double red_s[33];
  double s_bb = 0.0, s_rr = 0.0, s_rp = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double bi = b[i];
    const double ri = Ax0 ? bi - Ax0[i] : bi;
    const double pi = dinv[i] * ri;
    r[i] = ri;
    p[i] = pi;
    s_bb += bi * bi;
    s_rr += ri * ri;
    s_rp += ri * pi;
  }
  s_bb = block_sum(s_bb, red_s);
  s_rr = block_sum(s_rr, red_s);
  s_rp = block_sum(s_rp, red_s);
  const double v[3] = {s_bb, s_rr, s_rp};
  grid_reduce_finish<3>(v, part3, gridDim.x, red, red_s);   // -> out[0..2] = (b.b, r.r, r.p) over all ranks
}

// single block: fold the partials of (b.b, r.r, r.p) (npart == 1: already reduced scalars) -> thresholds and
// the early exits of ConjugateGradient.h:43-56
static __global__ void pcg_init_fin_kernel(const double* __restrict__ pbb, const double* __restrict__ prr, const double* __restrict__ prp,
                                    int npart, double tol, int max_iter, unsigned long long epoch, int proj, double inv_n,
                                    PcgCtl* ctl, unsigned long long* trace = nullptr, int trace_cap = 0, unsigned tag_base = 0u) {
  __shared__ double red[33];
  const double bb = block_sum_array(pbb, npart, red);
  const double rr = block_sum_array(prr, npart, red);
  const double rp = block_sum_array(prp, npart, red);
  if (threadIdx.x == 0) {
    ctl->rhs2 = bb;
    ctl->thr = fmax(tol * tol * bb, DBL_MIN);
    ctl->r2 = rr;
    ctl->absNew[0] = rp;
    ctl->absNew[1] = 0.0;
    ctl->base_it = 0;
    ctl->max_iter = max_iter;
    ctl->iters = 0;
    ctl->epoch = epoch;
    ctl->proj = proj;
    ctl->inv_n = inv_n;
    ctl->rmean[0] = ctl->rmean[1] = 0.0;
    ctl->trace = trace_cap > 0 ? trace : nullptr;
    ctl->trace_cap = trace_cap;
    ctl->tag_base = tag_base;
    ctl->done = (bb == 0.0 || rr < ctl->thr) ? 1 : 0;
  }
}

// The two vector kernels of an iteration work on pairs of rows (128-bit loads; every vector of the PCG slab is
// 256-byte aligned) and keep kVecUnroll pairs per thread in flight: all loads of a pass are issued before the first
// use.  Launched with kPcgVecCtas CTAs per SM (fk_handle_s::grid_pcg) so the whole grid is resident and the fold
// of the producer's partial sums in the prologue runs once per CTA, not once per wave.
constexpr int kVecUnroll = 2;
constexpr int kPcgVecCtas = 4;

// x += alpha p ; r -= alpha Ap ; partials of r.r and r.(dinv r)        (ConjugateGradient.h:62-76)
static __global__ void __launch_bounds__(kVecThreads, kPcgVecCtas) pcg_update_kernel(double* __restrict__ x, double* __restrict__ r,
                                                                 const double* __restrict__ p, const double* __restrict__ Ap,
                                                                 const double* __restrict__ dinv, int n,
                                                                 const double* __restrict__ part_pAp, int npart,
                                                                 double* __restrict__ part2, const RedArgs red, const PcgCtl* ctl,
                                                                 int j, const PeerRed pr) {
  __shared__ double red_s[66];
  griddep_wait();
  griddep_launch();
  if (pcg_inactive(ctl, j)) return;
  unsigned long long* const tr = trace_rec(ctl, j, 1);
  trace_begin(tr);
  const int n2 = n >> 1, T = gridDim.x * blockDim.x;
  double2* __restrict__ x2 = reinterpret_cast<double2*>(x);
  double2* __restrict__ r2 = reinterpret_cast<double2*>(r);
  const double2* __restrict__ p2 = reinterpret_cast<const double2*>(p);
  const double2* __restrict__ A2 = reinterpret_cast<const double2*>(Ap);
  const double2* __restrict__ d2 = reinterpret_cast<const double2*>(dinv);
  double2 xv[kVecUnroll], rv[kVecUnroll], pv[kVecUnroll], av[kVecUnroll], dv[kVecUnroll];
  auto load = [&](int i0) {
#pragma unroll
    for (int u = 0; u < kVecUnroll; u++) {
      const int i = i0 + u * T;
      if (i < n2) { pv[u] = p2[i]; av[u] = A2[i]; rv[u] = r2[i]; xv[u] = x2[i]; dv[u] = d2[i]; }
    }
  };
  int i0 = blockIdx.x * blockDim.x + threadIdx.x;
  load(i0);   // the first batch is in flight while the producer's partial sums are folded
  double pAp_v[1] = {block_sum_array(part_pAp, npart, red_s)};   // npart == 1: the all-reduced scalar (NCCL path)
  const unsigned long long tw0 = tr ? gtime_ns() : 0ULL;
  peer_fold<1>(pAp_v, pr, pcg_seq(ctl, j, 0), ctl->tag_base);
  trace_wait(tr, tw0);
  const double alpha = ctl->absNew[j & 1] / pAp_v[0];
  const bool proj = ctl->proj != 0;
  const double m_prev = proj ? ctl->rmean[j & 1] : 0.0;   // mean of r after the previous update
  double s_rr = 0.0, s_rz = 0.0, s_r = 0.0;
  while (i0 < n2) {
#pragma unroll
    for (int u = 0; u < kVecUnroll; u++) {
      const int i = i0 + u * T;
      if (i < n2) {
        xv[u].x += alpha * pv[u].x;
        xv[u].y += alpha * pv[u].y;
        rv[u].x -= alpha * av[u].x;
        rv[u].y -= alpha * av[u].y;
        if (proj) { rv[u].x -= m_prev; rv[u].y -= m_prev; s_r += rv[u].x + rv[u].y; }
        x2[i] = xv[u];
        r2[i] = rv[u];
        s_rr += rv[u].x * rv[u].x;
        s_rz += rv[u].x * (dv[u].x * rv[u].x);
        s_rr += rv[u].y * rv[u].y;
        s_rz += rv[u].y * (dv[u].y * rv[u].y);
      }
    }
    i0 += kVecUnroll * T;
    if (i0 < n2) load(i0);
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {   // odd row count: the last row
    const int i = n - 1;
    x[i] += alpha * p[i];
    const double ri = r[i] - alpha * Ap[i] - m_prev;
    r[i] = ri;
    s_rr += ri * ri;
    s_rz += ri * (dinv[i] * ri);
    s_r += ri;
  }
  if (proj) {   // third partial: sum r (single GPU / publish-only layouts: row 2 of part2)
    s_r = block_sum(s_r, red_s);
    if (threadIdx.x == 0) part2[2 * (size_t)gridDim.x + blockIdx.x] = s_r;
  }
  block_sum2(s_rr, s_rz, red_s);
  const double v[2] = {s_rr, s_rz};
  grid_reduce_finish<2>(v, part2, gridDim.x, red, red_s);   // -> out[0..1] = (r.r, r.z) over all ranks
  trace_end(tr);
}

// convergence test, beta, p = dinv r + beta p                           (ConjugateGradient.h:72-81)
static __global__ void __launch_bounds__(kVecThreads, kPcgVecCtas) pcg_direction_kernel(double* __restrict__ p, const double* __restrict__ r,
                                                                    const double* __restrict__ dinv, int n,
                                                                    const double* __restrict__ part2, int npart, int stride,
                                                                    PcgCtl* ctl, int j, const PeerRed pr) {
  __shared__ double red[66];
  griddep_wait();
  griddep_launch();
  if (pcg_inactive(ctl, j)) return;
  unsigned long long* const tr = trace_rec(ctl, j, 2);
  trace_begin(tr);
  const int n2 = n >> 1, T = gridDim.x * blockDim.x;
  double2* __restrict__ p2 = reinterpret_cast<double2*>(p);
  const double2* __restrict__ r2 = reinterpret_cast<const double2*>(r);
  const double2* __restrict__ d2 = reinterpret_cast<const double2*>(dinv);
  double2 rv[kVecUnroll], pv[kVecUnroll], dv[kVecUnroll];
  auto load = [&](int i0) {
#pragma unroll
    for (int u = 0; u < kVecUnroll; u++) {
      const int i = i0 + u * T;
      if (i < n2) { dv[u] = d2[i]; rv[u] = r2[i]; pv[u] = p2[i]; }
    }
  };
  int i0 = blockIdx.x * blockDim.x + threadIdx.x;
  load(i0);   // in flight during the fold of (r.r, r.z)
  double d2v[2];
  block_sum2_array(part2, part2 + stride, npart, red, d2v[0], d2v[1]);
  const unsigned long long tw0 = tr ? gtime_ns() : 0ULL;
  peer_fold<2>(d2v, pr, pcg_seq(ctl, j, 1), ctl->tag_base);
  trace_wait(tr, tw0);
  const double rr = d2v[0], rz = d2v[1];
  const double absOld = ctl->absNew[j & 1];
  const bool conv = rr < ctl->thr;
  if (ctl->proj) {   // uniform: every CTA takes the same branch
    const double sr = block_sum_array(part2 + 2 * (size_t)stride, npart, red);
    if (blockIdx.x == 0 && threadIdx.x == 0) ctl->rmean[(j + 1) & 1] = sr * ctl->inv_n;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    ctl->r2 = rr;
    if (conv) { ctl->iters = ctl->base_it + j; ctl->done = 1; }
    else ctl->absNew[(j + 1) & 1] = rz;
  }
  if (conv) { trace_end(tr); return; }
  const double beta = rz / absOld;
  while (i0 < n2) {
#pragma unroll
    for (int u = 0; u < kVecUnroll; u++) {
      const int i = i0 + u * T;
      if (i < n2) {
        pv[u].x = dv[u].x * rv[u].x + beta * pv[u].x;
        pv[u].y = dv[u].y * rv[u].y + beta * pv[u].y;
        p2[i] = pv[u];
      }
    }
    i0 += kVecUnroll * T;
    if (i0 < n2) load(i0);
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) p[n - 1] = dinv[n - 1] * r[n - 1] + beta * p[n - 1];
  trace_end(tr);
}


// ------------------------------------------------------------------------------------------------
// Single-reduction (Chronopoulos-Gear) Jacobi-PCG: fk_pcg_opts.recurrence = 1.
//   u = M^-1 r, w = A u, gamma = (r,u), delta = (w,u), beta = gamma/gamma_old,
//   alpha = gamma / (delta - beta gamma / alpha_old);  p = u + beta p;  s = w + beta s;  x += alpha p;  r -= alpha s.
// Same iterates as Eigen's loop (ConjugateGradient.h:59-82) in exact arithmetic, but all three dot products of an
// iteration are consumed at ONE point (the prologue of the vector kernel), so an iteration is two kernels --
// S_it: w = A z fused with (w,z) [spmv_sell_kernel<DOT>]; V_it: everything else -- and, on a partitioned problem, one
// cross-GPU reduction instead of two.  V_it folds (gamma, r.r, delta) of the CURRENT residual, tests convergence
// (Eigen's test: ||r||^2 < tol^2 ||b||^2, iterations reported as Eigen counts them), then makes update it+1.
// ------------------------------------------------------------------------------------------------
// r = b - A x0 (Ax0 nullable), z = dinv r; per-CTA partials of (r.z, r.r, b.b) at part[0 | stride | 2 stride]
static __global__ void __launch_bounds__(kVecThreads) cg_init_kernel(const double* __restrict__ b, const double* __restrict__ Ax0,
                                                              const double* __restrict__ dinv, int n, double* __restrict__ r,
                                                              double* __restrict__ z, double* __restrict__ part, int stride) {
  __shared__ double red_s[33];
  double s_bb = 0.0, s_rr = 0.0, s_rz = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double bi = b[i];
    const double ri = Ax0 ? bi - Ax0[i] : bi;
    const double zi = dinv[i] * ri;
    r[i] = ri;
    z[i] = zi;
    s_bb += bi * bi;
    s_rr += ri * ri;
    s_rz += ri * zi;
  }
  s_bb = block_sum(s_bb, red_s);
  s_rr = block_sum(s_rr, red_s);
  s_rz = block_sum(s_rz, red_s);
  if (threadIdx.x == 0) {
    part[blockIdx.x] = s_rz;
    part[stride + blockIdx.x] = s_rr;
    part[2 * stride + blockIdx.x] = s_bb;
  }
}

// V_it.  partA / partB: (gamma, r.r) partials, rows `stride` apart; V_it reads the buffer of parity it & 1 (written by
// V_it-1, or by cg_init_kernel for it = 0) and writes the other one -- a CTA may finish its pass while another still
// folds the old partials.  part_s: delta partials of the SpMV, nps.
static __global__ void __launch_bounds__(kVecThreads, kPcgVecCtas) cg_update_kernel(double* __restrict__ x, double* __restrict__ r,
                                                                double* __restrict__ p, double* __restrict__ sv,
                                                                const double* __restrict__ w, const double* __restrict__ dinv,
                                                                double* __restrict__ z, int n, double* __restrict__ partA,
                                                                double* __restrict__ partB, int npv, int stride,
                                                                const double* __restrict__ part_s, int nps, PcgCtl* ctl, int j,
                                                                const PeerRed pr) {
  __shared__ double red_s[66];
  if (pcg_inactive(ctl, j)) return;
  unsigned long long* const tr = trace_rec(ctl, j, 1);
  trace_begin(tr);
  const int it = ctl->base_it + j;
  const double* __restrict__ part_in = (it & 1) ? partB : partA;
  double* __restrict__ part_out = (it & 1) ? partA : partB;
  const int n2 = n >> 1, T = gridDim.x * blockDim.x;
  double2* __restrict__ x2 = reinterpret_cast<double2*>(x);
  double2* __restrict__ r2 = reinterpret_cast<double2*>(r);
  double2* __restrict__ p2 = reinterpret_cast<double2*>(p);
  double2* __restrict__ s2 = reinterpret_cast<double2*>(sv);
  double2* __restrict__ z2 = reinterpret_cast<double2*>(z);
  const double2* __restrict__ w2 = reinterpret_cast<const double2*>(w);
  const double2* __restrict__ d2 = reinterpret_cast<const double2*>(dinv);
  const bool first = it == 0;   // p and s hold nothing yet: p = u, s = w
  double2 xv = make_double2(0, 0), rv = xv, pv = xv, sv2 = xv, wv = xv, dv = xv;
  auto load = [&](int i) {
    if (i < n2) {
      rv = r2[i]; wv = w2[i]; dv = d2[i]; xv = x2[i];
      if (!first) { pv = p2[i]; sv2 = s2[i]; }
    }
  };
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  load(i);   // in flight during the fold
  const bool proj = ctl->proj != 0;
  double v3[4];
  block_sum2_array(part_in, part_in + stride, npv, red_s, v3[0], v3[1]);
  v3[2] = block_sum_array(part_s, nps, red_s);
  v3[3] = (proj && !first) ? block_sum_array(part_in + 2 * (size_t)stride, npv, red_s) : 0.0;   // sum r of the current residual
  const unsigned long long tw0 = tr ? gtime_ns() : 0ULL;
  peer_fold3(v3, pr, cg_seq(ctl, it), (int)(ctl->epoch & 1ULL), ctl->tag_base);
  trace_wait(tr, tw0);
  const double gamma = v3[0], rr = v3[1], delta = v3[2];
  const double m_prev = v3[3] * ctl->inv_n;   // mean of the current r: taken out by this update
  const bool conv = rr < ctl->thr;
  double alpha, beta = 0.0;
  if (first) alpha = gamma / delta;
  else {
    beta = gamma / ctl->gamma[(it - 1) & 1];
    alpha = gamma / (delta - beta * gamma / ctl->alpha[(it - 1) & 1]);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    ctl->r2 = rr;
    if (conv) { ctl->iters = it > 0 ? it - 1 : 0; ctl->done = 1; }
    else { ctl->gamma[it & 1] = gamma; ctl->alpha[it & 1] = alpha; }
  }
  if (conv) { trace_end(tr); return; }
  double s_rr = 0.0, s_rz = 0.0, s_r = 0.0;
  while (i < n2) {
    const double zx = dv.x * rv.x, zy = dv.y * rv.y;   // u = M^-1 r of the current residual (== z[i])
    if (first) { pv.x = zx; pv.y = zy; sv2 = wv; }
    else {
      pv.x = zx + beta * pv.x; pv.y = zy + beta * pv.y;
      sv2.x = wv.x + beta * sv2.x; sv2.y = wv.y + beta * sv2.y;
    }
    xv.x += alpha * pv.x; xv.y += alpha * pv.y;
    rv.x -= alpha * sv2.x; rv.y -= alpha * sv2.y;
    if (proj) { rv.x -= m_prev; rv.y -= m_prev; s_r += rv.x + rv.y; }
    const double2 zn = make_double2(dv.x * rv.x, dv.y * rv.y);
    p2[i] = pv; s2[i] = sv2; x2[i] = xv; r2[i] = rv; z2[i] = zn;
    s_rr += rv.x * rv.x; s_rz += rv.x * zn.x;
    s_rr += rv.y * rv.y; s_rz += rv.y * zn.y;
    i += T;
    load(i);
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {   // odd row count: the last row
    const int q = n - 1;
    const double zq = dinv[q] * r[q];
    const double pq = first ? zq : zq + beta * p[q];
    const double sq = first ? w[q] : w[q] + beta * sv[q];
    p[q] = pq; sv[q] = sq;
    x[q] += alpha * pq;
    const double rq = r[q] - alpha * sq - (proj ? m_prev : 0.0);
    r[q] = rq;
    const double zn = dinv[q] * rq;
    z[q] = zn;
    s_rr += rq * rq; s_rz += rq * zn;
    s_r += rq;
  }
  if (proj) {
    s_r = block_sum(s_r, red_s);
    if (threadIdx.x == 0) part_out[2 * (size_t)stride + blockIdx.x] = s_r;
  }
  block_sum2(s_rz, s_rr, red_s);
  if (threadIdx.x == 0) { part_out[blockIdx.x] = s_rz; part_out[stride + blockIdx.x] = s_rr; }
  trace_end(tr);
}

// End of a chunk.  If no further V kernel will run (max_iter reached), fold and test the residual of the last update
// here (Eigen's last loop trip, ConjugateGradient.h:72-73).  On a partitioned problem that reduction is also what keeps
// a rank from starting its next solve (and pushing new halo rows) while a peer's last SpMV still reads the old ones.
static __global__ void __launch_bounds__(256) cg_chunk_end_kernel(const double* __restrict__ partA, const double* __restrict__ partB,
                                                            int npv, int stride, PcgCtl* ctl, int chunk, const PeerRed pr) {
  __shared__ double red_s[66];
  const bool done = ctl->done != 0;
  const bool tail = !done && ctl->base_it + chunk >= ctl->max_iter;
  __syncthreads();
  if (tail) {
    const int itd = ctl->max_iter;   // updates made
    const double* part_in = (itd & 1) ? partB : partA;
    double v3[4];
    block_sum2_array(part_in, part_in + stride, npv, red_s, v3[0], v3[1]);
    v3[2] = 0.0;
    v3[3] = 0.0;
    peer_fold3(v3, pr, cg_seq(ctl, itd), (int)(ctl->epoch & 1ULL), ctl->tag_base);   // V_itd never runs: the sequence number is free
    if (threadIdx.x == 0) {
      ctl->r2 = v3[1];
      if (v3[1] < ctl->thr) { ctl->iters = itd - 1; ctl->done = 1; }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0 && !ctl->done) ctl->base_it += chunk;
}

// records of the device-side timeline: {min start = max u64, max end = 0, longest wait = 0, wait of CTA 0 = 0}
static __global__ void trace_reset_kernel(unsigned long long* tr, int nrec) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nrec) { tr[4 * i] = ~0ULL; tr[4 * i + 1] = 0ULL; tr[4 * i + 2] = 0ULL; tr[4 * i + 3] = 0ULL; }
}

static __global__ void pcg_advance_kernel(PcgCtl* ctl, int chunk) {
  if (!ctl->done) ctl->base_it += chunk;
}

// fold per-CTA partials into scalars (multi-GPU: feeds the NCCL all-reduce)
static __global__ void fold_partials_kernel(const double* __restrict__ part, int npart, int stride, int nvals, double* __restrict__ out) {
  __shared__ double red[33];
  for (int v = 0; v < nvals; v++) {
    const double s = block_sum_array(part + (size_t)v * stride, npart, red);
    if (threadIdx.x == 0) out[v] = s;
  }
}

// r = b - Ax (r nullable), per-CTA partials of ||b - Ax||^2  (lineSearch / Newton right-hand side, src/FEM3D.jl:128-139)
static __global__ void __launch_bounds__(kVecThreads) residual_kernel(const double* __restrict__ b, const double* __restrict__ Ax, int n,
                                                               double* __restrict__ r, double* __restrict__ part) {
  __shared__ double red_s[33];
  double s = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double v = b[i] - Ax[i];
    if (r) r[i] = v;
    s += v * v;
  }
  s = block_sum(s, red_s);
  if (threadIdx.x == 0) part[blockIdx.x] = s;
}

// ------------------------------------------------------------------------------------------------
// constraints: null-space projection (fk_pcg_opts.project_nullspace) and Dirichlet elimination
// ------------------------------------------------------------------------------------------------
// per-CTA partials of (sum a [mode 0] | sum a w [mode 1],  sum w) at part[0 | stride]; w == nullptr: ones
static __global__ void __launch_bounds__(kVecThreads) ns_partials_kernel(const double* __restrict__ a, const double* __restrict__ w, int n,
                                                                  int mode, double* __restrict__ part, int stride) {
  __shared__ double red_s[66];
  double s0 = 0.0, s1 = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double wi = w ? w[i] : 1.0;
    s0 += mode ? a[i] * wi : a[i];
    s1 += wi;
  }
  block_sum2(s0, s1, red_s);
  if (threadIdx.x == 0) { part[blockIdx.x] = s0; part[stride + blockIdx.x] = s1; }
}
// mode 0: out = a - (S0 / S1) w   (right-hand side made consistent: the multiplier term of the bordered system)
// mode 1: out = a - (S0 / S1)     (solution shifted so that w . x = 0)
static __global__ void __launch_bounds__(kVecThreads) ns_apply_kernel(const double* __restrict__ a, const double* __restrict__ w, int n,
                                                               int mode, const double* __restrict__ S, double* __restrict__ out) {
  const double c = S[1] != 0.0 ? S[0] / S[1] : 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    out[i] = a[i] - (mode ? c : c * (w ? w[i] : 1.0));
}

// Dirichlet lift (Fluids/fluids3D.jl:164-169): b[free] -= (A g)[free], b[fixed] = g, g = gD on the fixed rows, 0 elsewhere
static __global__ void dirichlet_lift_kernel(double* __restrict__ b, const double* __restrict__ Ag, const double* __restrict__ g,
                                      const unsigned char* __restrict__ mask, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) b[i] = mask[i] ? g[i] : b[i] - Ag[i];
}
// fixed rows and columns of a matrix in the SELL layout -> identity: one lane per row, like the SpMV
static __global__ void __launch_bounds__(256) dirichlet_mask_kernel(const int4* __restrict__ slice_meta, const int* __restrict__ srow,
                                                              const unsigned short* __restrict__ colz, const int* __restrict__ sell_diag,
                                                              const unsigned char* __restrict__ mask, int nslices, int n_rows,
                                                              double* __restrict__ vals) {
  const int lane = threadIdx.x & 31;
  const int sl = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (sl >= nslices) return;
  const int4 m = slice_meta[sl];
  const int r = srow[sl * 32 + lane];
  if (r < 0 || r >= n_rows) return;
  const bool rfix = mask[r] != 0;
  const int dg = sell_diag[r];
  const unsigned short* cz = colz + (size_t)m.z * 32;
  for (int k = 0; k < m.y; k++) {
    int c;
    if (m.w >= 0) c = m.w + (int)cz[((size_t)(k >> 2) * 32 + lane) * 4 + (k & 3)];
    else c = reinterpret_cast<const int*>(cz)[(size_t)k * 32 + lane];
    const int idx = m.x + k * 32 + lane;
    if (rfix) vals[idx] = (idx == dg) ? 1.0 : 0.0;
    else if (mask[c]) vals[idx] = 0.0;
  }
}
static __global__ void scatter_kernel(double* __restrict__ dst, const int* __restrict__ ids, const double* __restrict__ v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[ids[i]] = v ? v[i] : 0.0;
}
// dst[ids[i]] = full[ids[i]] (fixed DOFs of an initial guess <- the scattered Dirichlet values)
static __global__ void copy_rows_kernel(double* __restrict__ dst, const int* __restrict__ ids, const double* __restrict__ full, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[ids[i]] = full[ids[i]];
}
static __global__ void scatter_flag_kernel(unsigned char* __restrict__ dst, const int* __restrict__ ids, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[ids[i]] = 1;
}

// ------------------------------------------------------------------------------------------------
// small vector / matrix-value kernels
// ------------------------------------------------------------------------------------------------
static __global__ void axpby_kernel(double* __restrict__ dst, double alpha, const double* __restrict__ x, double beta,
                             const double* __restrict__ y, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dst[i] = (beta != 0.0) ? alpha * x[i] + beta * y[i] : alpha * x[i];
}
static __global__ void fill_kernel(double* __restrict__ dst, double v, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) dst[i] = v;
}
static __global__ void pad_nodes_kernel(const double* __restrict__ p, int nv, double4* __restrict__ p4) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n < nv) p4[n] = make_double4(p[3 * (size_t)n], p[3 * (size_t)n + 1], p[3 * (size_t)n + 2], 0.0);
}
static __global__ void gather_diag_kernel(const double* __restrict__ vals, const int* __restrict__ diag, int n, double* __restrict__ out) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) out[r] = diag[r] >= 0 ? vals[diag[r]] : 0.0;
}
static __global__ void max_array_kernel(const double* __restrict__ a, int n, double* __restrict__ out) {
  __shared__ double red[33];
  double v = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) v = fmax(v, a[i]);
  v = block_max(v, red);
  if (threadIdx.x == 0) out[0] = v;
}
static __global__ void set_material_kernel(unsigned char* __restrict__ matid, const int* __restrict__ ids, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) matid[ids[i]] = 1;
}
// BoundaryIntegral (src/FEM3D.jl:284-297 / FEMc.cpp:170-200) as a per-node gather over the node's
// surface triangles in ascending triangle order (= the order the reference loop adds them).
static __global__ void boundary_integral_kernel(const double4* __restrict__ p4, const int* __restrict__ surfaceT,
                                         const double* __restrict__ normal, const double* __restrict__ AE,
                                         const int* __restrict__ n2s_ptr, const int* __restrict__ n2s, int nv, double Fx,
                                         double Fy, double Fz, double* __restrict__ RHS) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= nv) return;
  double acc = 0.0;
  for (int q = n2s_ptr[n]; q < n2s_ptr[n + 1]; q++) {
    const int s = n2s[q];
    double area;
    if (AE) area = AE[s];
    else {  // areaTriangle of FEMc.cpp:45-54
      double x0, y0, z0, x1, y1, z1, x2, y2, z2;
      load_node(p4, surfaceT[4 * (size_t)s], x0, y0, z0);
      load_node(p4, surfaceT[4 * (size_t)s + 1], x1, y1, z1);
      load_node(p4, surfaceT[4 * (size_t)s + 2], x2, y2, z2);
      const double ax = x1 - x0, ay = y1 - y0, az = z1 - z0, bx = x2 - x0, by = y2 - y0, bz = z2 - z0;
      const double cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
      area = 0.5 * sqrt(cx * cx + cy * cy + cz * cz);
    }
    const double dotv = normal[3 * (size_t)s] * Fx + normal[3 * (size_t)s + 1] * Fy + normal[3 * (size_t)s + 2] * Fz;
    acc += dotv * area / 3;
  }
  RHS[n] = acc;
}

// halo exchange packing: sendbuf[i] = v[idx[i]]
static __global__ void pack_kernel(const double* __restrict__ v, const int* __restrict__ idx, int n, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = v[idx[i]];
}

}  // namespace fk
