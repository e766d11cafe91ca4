// fk_residual_norm_noalloc: r = b - Ax and ||r||_2 WITHOUT any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).  It is the allocation-free variant of residual_kernel + the fold of its per-CTA
// partials (kernels.cuh, api.cu: residual_dev), whose scratch -- the partials, the result and the ticket of the last CTA --
// normally lives in buffers the handle allocates.  Here the scratch is ONE file-scope __device__ array of a fixed size:
// every GPU that loads the library holds kScratchBytes (131 328 B) of it until the process exits.
//
// * The array's name is a host-side stand-in in host code, not a device address: the device address is taken once per
//   device, with that device current, by cudaGetSymbolAddress (scratch_of_current_device) and handed to the kernel.
// * Every call on a device shares the array and finds the previous call's contents in it: the ticket is cleared on the
//   caller's stream before each launch (and again by the folding CTA), partials and results are overwritten before they
//   are read.  Calls that overlap on one GPU (two streams, two host threads) RACE on it: serialise them.
// * No cudaMalloc / cudaFree / cudaMallocHost / library routine that allocates is called; the only runtime calls are
//   cudaGetDevice, cudaGetSymbolAddress (first call per device), cudaMemsetAsync (4 bytes), the launch, and -- only when the
//   caller asks for the norm on the host -- cudaMemcpyAsync + cudaStreamSynchronize.
// This translation unit is self-contained on purpose (no handle, no torch, nothing from kernels.cuh).
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <mutex>
#include <stdexcept>
#include <string>

#include "femeko_cuda.h"

namespace fk {
void set_error(const std::string& s);   // api.cu: the thread's fk_last_error text

namespace noalloc {
constexpr int kThreads = 256;
constexpr int kRowsPerThread = 16;
constexpr int kRowsPerCta = kThreads * kRowsPerThread;                   // 4096: fixed, so the summation order depends on n alone
constexpr int kMaxParts = (int)(FK_NOALLOC_MAX_ROWS / kRowsPerCta);      // 16384 per-CTA partials
constexpr int kHead = 32;                                                // doubles: [0] ||r||^2, [1] ||r||, [2] ticket (as unsigned)
constexpr size_t kScratchBytes = sizeof(double) * (size_t)(kHead + kMaxParts);
constexpr int kMaxDevices = 64;
static_assert(kScratchBytes == 131328, "the size stated in include/femeko_cuda.h");

static __device__ __align__(256) double g_scratch[kHead + kMaxParts];

// fixed shuffle tree, warps folded in index order; only thread 0 holds the total
__device__ __forceinline__ double cta_sum(double v, double* sh /* 8 */) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < kThreads / 32; w++) s += sh[w];
  return s;
}

// CTA c owns rows [4096 c, 4096 (c + 1)): r = b - Ax (r nullable), partial of ||r||^2 to scratch; the CTA that draws the last
// ticket folds the partials in index order (deterministic: the result does not depend on which CTA finishes last)
static __global__ void __launch_bounds__(kThreads) residual_norm_kernel(const double* __restrict__ b, const double* __restrict__ Ax,
                                                                        long long n, double* __restrict__ r,
                                                                        double* __restrict__ scratch, double* __restrict__ norm_out) {
  __shared__ double sh[kThreads / 32];
  __shared__ bool last;
  double* part = scratch + kHead;
  unsigned* ticket = reinterpret_cast<unsigned*>(scratch + 2);
  const long long base = (long long)blockIdx.x * kRowsPerCta;
  double s = 0.0;
#pragma unroll 4
  for (int q = 0; q < kRowsPerThread; q++) {
    const long long i = base + (long long)q * kThreads + threadIdx.x;
    if (i < n) {
      const double v = b[i] - Ax[i];
      if (r) r[i] = v;
      s += v * v;
    }
  }
  s = cta_sum(s, sh);
  if (threadIdx.x == 0) {
    part[blockIdx.x] = s;
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  double t = 0.0;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += kThreads) t += __ldcg(part + i);
  t = cta_sum(t, sh);
  if (threadIdx.x == 0) {
    scratch[0] = t;
    scratch[1] = sqrt(t);
    if (norm_out) *norm_out = sqrt(t);
    *ticket = 0u;
  }
}

// device address of the scratch array on the CURRENT device (taken once per device; the array's name is only a handle here)
static double* scratch_of_current_device() {
  static std::mutex mu;
  static double* ptr[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) throw std::runtime_error(std::string("fk_residual_norm_noalloc: no CUDA device (") + cudaGetErrorString(e) + ")");
  if (dev < 0 || dev >= kMaxDevices) throw std::runtime_error("fk_residual_norm_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lk(mu);
  if (!ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_scratch);
    if (e != cudaSuccess || !p)
      throw std::runtime_error(std::string("fk_residual_norm_noalloc: cudaGetSymbolAddress failed (") + cudaGetErrorString(e) + ")");
    ptr[dev] = static_cast<double*>(p);
  }
  return ptr[dev];
}
}  // namespace noalloc
}  // namespace fk

extern "C" int fk_residual_norm_noalloc(const double* b_dev, const double* Ax_dev, long long n, double* r_dev, double* norm_dev_out,
                                        double* norm_host_out, void* stream) {
  using namespace fk::noalloc;
  try {
    if (n < 0) throw std::runtime_error("fk_residual_norm_noalloc: n < 0");
    if (n > (long long)FK_NOALLOC_MAX_ROWS)
      throw std::runtime_error("fk_residual_norm_noalloc: n = " + std::to_string(n) + " exceeds FK_NOALLOC_MAX_ROWS = " +
                               std::to_string((long long)FK_NOALLOC_MAX_ROWS) + " (the static scratch holds one partial per 4096 rows); "
                               "use fk_residual_dev for larger vectors");
    if (n > 0 && (!b_dev || !Ax_dev)) throw std::runtime_error("fk: NULL argument");
    double* scratch = scratch_of_current_device();
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const unsigned grid = (unsigned)std::max<long long>(1, (n + kRowsPerCta - 1) / kRowsPerCta);
    cudaError_t e = cudaMemsetAsync(scratch + 2, 0, sizeof(unsigned), s);   // the ticket must start at zero whatever the last call left
    if (e == cudaSuccess) {
      residual_norm_kernel<<<grid, kThreads, 0, s>>>(b_dev, Ax_dev, n, r_dev, scratch, norm_dev_out);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess && norm_host_out) {
      e = cudaMemcpyAsync(norm_host_out, scratch + 1, sizeof(double), cudaMemcpyDeviceToHost, s);
      if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    }
    if (e != cudaSuccess) throw std::runtime_error(std::string("fk_residual_norm_noalloc: CUDA error: ") + cudaGetErrorString(e));
    return 0;
  } catch (const std::exception& ex) {
    fk::set_error(ex.what());
    return 1;
  }
}
