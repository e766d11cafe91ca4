// C ABI of libfemeko_cuda.so (see include/femeko_cuda.h).  Host-side orchestration only: every
// floating-point operation of the hot path runs in the kernels of kernels.cuh on the GPU; there is
// no CPU fallback (fk_create fails when no CUDA device is present).
#include <chrono>
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "comm.hpp"
#include "handle.hpp"

#ifdef _OPENMP
#include <omp.h>
#endif

namespace fk {

static thread_local std::string g_err;
void set_error(const std::string& s) { g_err = s; }

template <class F>
static int guard(F&& f) {
  try {
    return f();
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  } catch (...) {
    set_error("fk: unknown error");
    return 1;
  }
}

static void need(bool ok, const char* msg) {
  if (!ok) throw std::runtime_error(msg);
}
static void check_slot(int s) { need(s >= 0 && s < FK_NUM_SLOTS, "fk: matrix slot out of range"); }

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    FK_CUDA(cudaGetDevice(&prev));
    if (prev != dev) FK_CUDA(cudaSetDevice(dev));
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

static inline int cdiv(long long a, int b) { return (int)((a + b - 1) / b); }

static double* slot_vals(fk_handle h, int slot) {
  check_slot(slot);
  if (!h->vals[slot].ptr) {  // padding entries of the SELL layout must be (and stay) zero
    h->vals[slot].ensure((size_t)h->nnz_sell + kSlack);
    FK_CUDA(cudaMemsetAsync(h->vals[slot].ptr, 0, sizeof(double) * ((size_t)h->nnz_sell + kSlack), h->stream));
  }
  return h->vals[slot].ptr;
}

// ------------------------------------------------------------------------------------------------
// assembly launcher
// ------------------------------------------------------------------------------------------------
static AsmArgs base_args(fk_handle h, double* vals) {
  AsmArgs a{};
  a.VE = h->user_ve ? h->VE.ptr : nullptr;
  a.tile_hdr = reinterpret_cast<const int4*>(h->tile_hdr.ptr);
  a.tile_blob = reinterpret_cast<const uint4*>(h->tile_blob.ptr);
  a.ntiles = h->sym.ntiles;
  a.vals = vals;
  a.e_pad = h->e_pad;
  a.nno_max = h->sym.max_tile_nodes;
  a.blob_i_max = h->sym.max_blob_i;
  a.blob_g_max = h->sym.max_blob_g;
  a.blob_b_max = h->sym.max_blob_b;
  a.dbg = h->asm_dbg;
  return a;
}

// The persisting-L2 set-aside that keeps the PCG vectors resident (configure_l2) slows every OTHER kernel down while it
// is configured -- measured on B200 with identical DRAM bytes, L2 hit rates and clocks: the sweep's element kernel 150 ->
// 293 us, its streamed sums 325 -> 550 us (profiles/r2_sweep_regimes_l2_set_aside.csv), whether or not any line is
// actually persisting.  So assembly work takes the set-aside away (cudaLimitPersistingL2CacheSize = 0) and the next solve
// puts it back (pcg_dev): Picard sweep inside the bench step 0.63 -> 0.49 ms at N=119, 1.57 -> 0.94 ms at N=150.
// fk_tune(15): 0 leaves the set-aside alone, bit 0 also resets the persisting lines, bit 1 (default) = the above.
static void l2_release_for_assembly(fk_handle h) {
  if (!h->l2_window_set) return;
  if (h->l2_sweep_mode & 1) cudaCtxResetPersistingL2Cache();
  if ((h->l2_sweep_mode & 2) && !h->l2_limit_zeroed) {
    const auto t0 = std::chrono::steady_clock::now();
    if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, 0) == cudaSuccess) h->l2_limit_zeroed = true;
    else cudaGetLastError();
    h->l2_toggle_ns[0] += std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
    h->l2_toggles++;
  }
}

// launch configurations of the assembly kernel: {threads per CTA, min resident CTAs per SM}
constexpr int kNumAsmCfg = 8;
template <int MODE, int THREADS, int MINB>
static void launch_assemble_cfg(fk_handle h, const AsmArgs& a, bool setup) {
  if (setup) {  // persistent grid: every resident CTA slot of the device, fixed per configuration
    FK_CUDA(cudaFuncSetAttribute(assemble_kernel<MODE, THREADS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->asm_smem));
    int occ = 1;
    FK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, assemble_kernel<MODE, THREADS, MINB>, THREADS, h->asm_smem));
    h->asm_grid[MODE] = std::max(1, std::min(h->sym.ntiles, h->num_sms * std::max(1, occ)));
    return;
  }
  assemble_kernel<MODE, THREADS, MINB><<<h->asm_grid[MODE], THREADS, h->asm_smem, h->stream>>>(a);
}
template <int MODE>
static void launch_assemble(fk_handle h, const AsmArgs& a, bool setup = false) {
  if (!setup) need(h->sym.npe == 4, "fk: this entry point assembles 4-node elements; the handle holds P2 elements (use fk_assemble_*_p2)");
  if (h->sym.ntiles == 0) return;
  if (!setup) l2_release_for_assembly(h);
  // 0 = measured default: {256, 2} for every mode (two CTAs per SM is what shared memory allows; the 128-register
  // budget keeps phase 1 free of spills)
  const int cfg = h->asm_cfg > 0 ? h->asm_cfg : 2;
  switch (cfg) {
    case 2: launch_assemble_cfg<MODE, 256, 2>(h, a, setup); break;
    case 3: launch_assemble_cfg<MODE, 384, 2>(h, a, setup); break;
    case 4: launch_assemble_cfg<MODE, 512, 2>(h, a, setup); break;
    case 5: launch_assemble_cfg<MODE, 128, 6>(h, a, setup); break;
    case 6: launch_assemble_cfg<MODE, 512, 1>(h, a, setup); break;
    case 7: launch_assemble_cfg<MODE, 768, 1>(h, a, setup); break;
    default: launch_assemble_cfg<MODE, 256, 3>(h, a, setup); break;
  }
  if (setup) return;
  FK_CUDA(cudaGetLastError());
  h->launches++;
}

static void ensure_geo(fk_handle h);
static void launch_geo_rescale(fk_handle h, double* vals, int grid_cap = 0, const double* g = nullptr);
static void assemble_stiffness_dev(fk_handle h, int slot, const double* f_dev) {
  h->dir_applied &= ~(1u << slot);
  double* vals = slot_vals(h, slot);
  if (h->geo_mode == 2 || (h->geo_mode == 1 && h->geo_ready)) {   // rescale path: s = VE f, then the streamed sums
    ensure_geo(h);
    geo_scale_kernel<<<cdiv(h->nt, 256), 256, 0, h->stream>>>(h->VE.ptr, f_dev, h->nt, h->geo_s0.ptr);
    h->launches++;
    h->s0_dirty = true;   // the scales now belong to f, not to mu
    launch_geo_rescale(h, vals);
    return;
  }
  AsmArgs a = base_args(h, vals);
  a.coef = f_dev;
  launch_assemble<ASM_STIFF>(h, a);
}
static void assemble_mass_dev(fk_handle h, int slot, const double* rho_dev) {
  h->dir_applied &= ~(1u << slot);
  AsmArgs a = base_args(h, slot_vals(h, slot));
  a.coef = rho_dev;
  launch_assemble<ASM_MASS>(h, a);
}
static void assemble_tangential_dev(fk_handle h, int slot, const double* H_dev, const double* dmu_dev) {
  h->dir_applied &= ~(1u << slot);
  AsmArgs a = base_args(h, slot_vals(h, slot));
  a.coef = dmu_dev;
  a.Hvec = H_dev;
  launch_assemble<ASM_TANGENT>(h, a);
}

// ---- rescale assembly (rescale.cuh) ------------------------------------------------------------------
constexpr int kNumGeoVariants = 4;
template <typename F>
static void geo_variant_dispatch(int variant, F&& f) {   // {contributions per batch, min CTAs per SM}
  switch (variant) {
    case 1: f(std::integral_constant<int, 4>{}, std::integral_constant<int, 6>{}); break;
    case 2: f(std::integral_constant<int, 8>{}, std::integral_constant<int, 3>{}); break;
    case 3: f(std::integral_constant<int, 6>{}, std::integral_constant<int, 5>{}); break;
    default: f(std::integral_constant<int, 8>{}, std::integral_constant<int, 4>{}); break;
  }
}
static void configure_geo(fk_handle h) {
  int occ = 1;
  geo_variant_dispatch(h->geo_variant, [&](auto B, auto MINB) {
    FK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, geo_rescale_kernel<B(), MINB()>, kGeoThreads, 0));
  });
  const int groups = (h->geo_nslices + kGeoThreads / 32 - 1) / (kGeoThreads / 32);
  h->geo_grid = std::max(1, std::min(groups, h->num_sms * std::max(1, occ)));
}
// the one-off part: plan on the host (symbolic.cpp: build_rescale_plan), factors on the device
static void ensure_geo(fk_handle h) {
  if (h->geo_ready) return;
  need(h->sym.npe == 4, "fk: the rescale path of P1 elements was asked for on a P2 handle");
  RescalePlan R;
  build_rescale_plan(h->sym, h->geo_window, R, h->geo_rowsum_diag);
  h->geo_plan_rowsum = h->geo_rowsum_diag;
  cudaStream_t s = h->stream;
  h->geo_nslices = R.nslices;
  h->geo_slots = R.n_slots;
  h->geo_items = R.n_items;
  h->geo_contrib = R.n_contrib;
  h->geo_slice_ptr.upload(R.slice_ptr, s);
  h->geo_dst1.upload(R.dst1, s);
  h->geo_dst2.upload(R.dst2, s);
  h->geo_ref.upload(R.cref, s);
  h->geo_g.ensure((size_t)std::max<long long>(R.n_slots, 1));
  h->geo_s0.ensure((size_t)h->nt);
  if (R.n_slots > 0) {
    geo_build_kernel<<<std::min(cdiv(R.n_slots, 256), h->num_sms * 16), 256, 0, s>>>(h->p4.ptr, h->t4.ptr, h->geo_ref.ptr, R.n_slots,
                                                                                     h->geo_g.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
  }
  FK_CUDA(cudaStreamSynchronize(s));   // the host plan dies with this scope
  configure_geo(h);
  h->geo_ready = true;
}
// vals = sum_k s0[k] G_k on the plan's items (every stored entry is an item or the mirror of one)
static void launch_geo_rescale(fk_handle h, double* vals, int grid_cap, const double* g) {
  if (h->geo_nslices == 0) return;
  l2_release_for_assembly(h);
  if (!g) g = h->geo_g.ptr;
  const int grid = grid_cap > 0 ? std::min(h->geo_grid, grid_cap) : h->geo_grid;
  geo_variant_dispatch(h->geo_variant, [&](auto B, auto MINB) {
    geo_rescale_kernel<B(), MINB()><<<grid, kGeoThreads, 0, h->stream>>>(h->geo_slice_ptr.ptr, h->geo_dst1.ptr, h->geo_dst2.ptr,
                                                                               h->geo_ref.ptr, g, h->geo_s0.ptr,
                                                                               h->geo_nslices, vals);
  });
  FK_CUDA(cudaGetLastError());
  h->launches++;
  if (h->geo_plan_rowsum && h->sym.npe == 4 && !(h->geo_dbg & 2)) {   // the plan carries no diagonal contributions: a_rr = -(sum of the row's other entries)
    diag_from_rowsum_kernel<<<cdiv(h->sym.nslices, 8), 256, 0, h->stream>>>(reinterpret_cast<const int4*>(h->slice_meta.ptr), h->srow.ptr,
                                                                          h->sell_diag.ptr, h->sym.nslices, h->n_rows, vals);
    FK_CUDA(cudaGetLastError());
    h->launches++;
  }
}

static void ensure_state(fk_handle h) {
  if (h->mu.ptr) return;
  h->mu.ensure((size_t)h->nt);
  h->dmu.ensure((size_t)h->nt);
  h->Hn.ensure((size_t)h->nt);
  h->Hfield.ensure((size_t)3 * h->nt);
  h->matid.ensure((size_t)h->nt);
  h->tile_max.ensure((size_t)std::max(1, 8 * h->num_sms));   // per-CTA maxima of material_update_kernel
  h->errflag.ensure(1);
  h->scal.ensure(16);
  const int g = std::min(cdiv(h->nt, 256), 65535);
  fill_kernel<<<g, 256, 0, h->stream>>>(h->mu.ptr, 4e-7 * 3.14159265358979323846, h->nt);  // mu0 (magneticField.jl:38)
  FK_CUDA(cudaMemsetAsync(h->dmu.ptr, 0, sizeof(double) * (size_t)h->nt, h->stream));
  FK_CUDA(cudaMemsetAsync(h->Hn.ptr, 0, sizeof(double) * (size_t)h->nt, h->stream));
  FK_CUDA(cudaMemsetAsync(h->Hfield.ptr, 0, sizeof(double) * (size_t)3 * h->nt, h->stream));
  FK_CUDA(cudaMemsetAsync(h->matid.ptr, 0, (size_t)h->nt, h->stream));
  FK_CUDA(cudaMemsetAsync(h->errflag.ptr, 0, sizeof(int), h->stream));
  h->launches++;
}

// ------------------------------------------------------------------------------------------------
// SpMV / PCG launchers
// ------------------------------------------------------------------------------------------------
// launch with (pdl) or without the programmatic-stream-serialization attribute (kernels.cuh: griddep_wait)
template <typename... KArgs, typename... Args>
static void launch_k(bool pdl, void (*k)(KArgs...), int grid, int block, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  FK_CUDA(cudaLaunchKernelEx(&cfg, k, KArgs(std::forward<Args>(args))...));
}

// list == nullptr: every slice.  Otherwise the `nlist` slices named by `list`, partial p.Ap sums at part[0 .. grid).
static void launch_spmv_list(fk_handle h, cudaStream_t s, const double* vals, const double* x, double* y, bool dot, const PcgCtl* ctl,
                             int j, const RedArgs& ra, const int* list, int nlist, int grid, double* part) {
  if (nlist <= 0) return;
  const int4* meta = reinterpret_cast<const int4*>(h->slice_meta.ptr);
  spmv_variant_dispatch(h->spmv_variant, [&](auto G, auto MINB) {
    if (dot && list != nullptr && list == h->slices_bnd.ptr)   // halo-coupled rows beside the interior kernel: coherent gathers
      spmv_sell_kernel<true, G(), MINB(), true><<<grid, kSpmvThreads, 0, s>>>(meta, h->srow.ptr, h->sell_colz.ptr, vals, x, y, nlist, part,
                                                                              ra, ctl, j, list);
    else if (dot)
      launch_k(h->pdl && !h->comm, spmv_sell_kernel<true, G(), MINB()>, grid, kSpmvThreads, s, meta, h->srow.ptr, h->sell_colz.ptr, vals,
               x, y, nlist, part, ra, ctl, j, list);
    else
      spmv_sell_kernel<false, G(), MINB()><<<grid, kSpmvThreads, 0, s>>>(meta, h->srow.ptr, h->sell_colz.ptr, vals, x, y, nlist,
                                                                         nullptr, RedArgs{}, ctl, j, list);
  });
  FK_CUDA(cudaGetLastError());
}
void launch_spmv(fk_handle h, cudaStream_t s, const double* vals, const double* x, double* y, bool dot, const PcgCtl* ctl,
                 int j, bool publish_only) {
  // partitioned: p.Ap is finished (and all-reduced over peer memory) inside the kernel -> scal[0]
  const RedArgs ra = (dot && h->comm && !publish_only) ? comm_red_args(h, h->tickets.ptr, h->scal.ptr, 0) : RedArgs{};
  launch_spmv_list(h, s, vals, x, y, dot, ctl, j, ra, nullptr, h->sym.nslices, h->grid_spmv, h->part.ptr);
}

// every cached PCG chunk graph bakes in launch shapes, the L2 window and (on partitions) peer pointers: drop them
// whenever any of those changes
void invalidate_graphs(fk_handle h) {
  for (int s = 0; s < FK_NUM_SLOTS; s++)
    for (int r = 0; r < 2; r++)
      if (h->graph[s][r]) { cudaGraphExecDestroy(h->graph[s][r]); h->graph[s][r] = nullptr; }
}

static size_t part_words(fk_handle h) { return (size_t)8 * std::max(std::max(h->grid_spmv, h->grid_vec), h->grid_pcg) + 16; }

// launch geometry of the persistent SpMV: full residency, fixed per handle (deterministic p.Ap)
static void configure_spmv(fk_handle h, int ctas_per_sm) {
  int occ = 1;
  spmv_variant_dispatch(h->spmv_variant, [&](auto G, auto MINB) {
    FK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, spmv_sell_kernel<true, G(), MINB()>, kSpmvThreads, 0));
  });
  occ = std::max(1, occ);
  if (ctas_per_sm > 0) occ = std::min(occ, ctas_per_sm);
  const int groups = (h->sym.nslices + (kSpmvThreads / 32) - 1) / (kSpmvThreads / 32);
  h->grid_spmv = std::max(1, std::min(groups, h->num_sms * occ));
  const int wpb = kSpmvThreads / 32;
  // the interior rows leave `spmv_int_reserve` CTA slots per SM free, so the halo branch (push, wait, halo-coupled
  // rows) is scheduled beside them instead of behind a full persistent grid
  h->grid_spmv_int = std::max(1, std::min((h->n_slices_int + wpb - 1) / wpb, h->num_sms * std::max(1, occ - h->spmv_int_reserve)));
  h->grid_spmv_bnd = std::max(1, std::min((h->n_slices_bnd + wpb - 1) / wpb, h->num_sms * occ));
  invalidate_graphs(h);
  if (h->part.ptr) h->part.ensure(part_words(h));
}

// B200 has 126 MB of L2: the PCG vectors of a 10 M-tet problem (14 MB each) fit, the matrix stream (307 MB per
// product) does not.  Pin the vectors an iteration streams in L2 (persisting access-policy window) and let everything
// else -- the matrix is read with evict-first loads anyway -- stream past them, so the vector kernels and the x gather /
// y write of the SpMV run at L2 speed instead of HBM speed.  Slab order: dinv | x | r | p | Ap | s | z | u.
//   recurrence 0 (Eigen):            dinv x r p Ap; when they exceed what may persist, whole vectors only, always p
//   recurrence 1 (single reduction): (dinv x) r p Ap s z; shrinking windows always end at z (the SpMV's gather source)
// A partially persisting window (hit ratio < 1) turns the rest of the vectors into evict-first lines and was measured
// far slower than no window at 20 M tets (profiles/r03r_*).
static std::atomic<int> g_l2_users{0};
static void configure_l2(fk_handle h, int recurrence) {
  const bool was_set = h->l2_window_set;
  h->l2_window_set = false;
  h->l2_recurrence = recurrence;
  if (!h->pcg_slab.ptr) return;
  int max_persist = 0, max_window = 0;
  cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, h->device);
  cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, h->device);
  if (std::getenv("FK_DEBUG"))
    std::fprintf(stderr, "fk: L2 max persisting %d B, max window %d B, vector %zu B, recurrence %d\n", max_persist, max_window,
                 h->slab_stride * sizeof(double), recurrence);
  cudaStreamAttrValue attr{};
  attr.accessPolicyWindow.num_bytes = 0;  // disables the window
  attr.accessPolicyWindow.hitProp = cudaAccessPropertyNormal;
  attr.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
  if (h->l2_persist && max_persist > 0 && max_window > 0) {
    const size_t vec = h->slab_stride * sizeof(double), cap = std::min((size_t)max_persist, (size_t)max_window);
    // slab order: dinv | x | r | p | Ap | s | z | u
    const size_t lim = h->l2_max_vecs > 0 ? (size_t)h->l2_max_vecs : 7;
    int first = 0, count = 0;
    if (recurrence == 0) {   // dinv x r p Ap, then whole vectors only, always p
      const int opts[5][2] = {{0, 5}, {1, 4}, {2, 3}, {3, 2}, {3, 1}};
      for (const auto& o : opts)
        if ((size_t)o[1] * vec <= cap && (size_t)o[1] <= lim) { first = o[0]; count = o[1]; break; }
    } else {                 // windows that end at z: with five, dinv (one pass) and x (two) are the ones left to stream
      for (int c = 7; c >= 1 && count == 0; c--)
        if ((size_t)c * vec <= cap && (size_t)c <= lim) { first = 7 - c; count = c; }
    }
    if (count > 0) {
      const size_t win = (size_t)count * vec;
      if (!h->l2_limit_changed) {   // remember the application's limit: fk_destroy puts it back
        size_t cur = 0;
        if (cudaDeviceGetLimit(&cur, cudaLimitPersistingL2CacheSize) == cudaSuccess) h->l2_limit_saved = cur;
      }
      size_t cur = 0;
      cudaDeviceGetLimit(&cur, cudaLimitPersistingL2CacheSize);
      if (cur >= win || cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, win) == cudaSuccess) {   // only ever raise it
        if (cur < win) h->l2_limit_changed = true;
        attr.accessPolicyWindow.base_ptr = h->pcg_slab.ptr + (size_t)first * h->slab_stride;
        attr.accessPolicyWindow.num_bytes = win;
        attr.accessPolicyWindow.hitRatio = 1.0f;
        attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        attr.accessPolicyWindow.missProp = h->l2_miss_normal ? cudaAccessPropertyNormal : cudaAccessPropertyStreaming;
        h->l2_window_set = true;
      } else {
        cudaGetLastError();
      }
    }
  }
  if (h->l2_window_set && !was_set) g_l2_users++;
  if (!h->l2_window_set && was_set) g_l2_users--;
  h->l2_window = attr.accessPolicyWindow;
  // The window rides on the kernel nodes of the PCG chunk graphs (build_graph).  It is NOT left on the handle's stream any
  // more: a stream attribute covers every kernel launched there, and under it the assembly sweep's gathers (scales, node
  // coordinates: everything outside the window) were treated as streaming (evict-first) lines.  fk_tune(14, 1) restores
  // the stream attribute (the path without graphs -- NCCL fallback -- always sets it around its launches).
  cudaStreamAttrValue off{};
  off.accessPolicyWindow.num_bytes = 0;
  off.accessPolicyWindow.hitProp = cudaAccessPropertyNormal;
  off.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
  const cudaStreamAttrValue& sa = (h->l2_on_stream || (h->comm && !comm_capturable(h))) ? attr : off;
  if (h->stream && cudaStreamSetAttribute(h->stream, cudaStreamAttributeAccessPolicyWindow, &sa) != cudaSuccess) cudaGetLastError();
  if (h->halo_stream && cudaStreamSetAttribute(h->halo_stream, cudaStreamAttributeAccessPolicyWindow, &sa) != cudaSuccess) cudaGetLastError();
  invalidate_graphs(h);
}

void ensure_pcg(fk_handle h) {
  if (h->x.ptr) return;
  const size_t nvec = (size_t)h->nv;  // vectors carry halo rows too (SpMV input)
  const size_t stride = (nvec + 31) & ~(size_t)31;
  h->slab_stride = stride;
  h->pcg_slab.ensure(kSlabVecs * stride + kCtrlWords);   // + the control block of the peer-memory path (comm.cu)
  h->dinv.ptr = h->pcg_slab.ptr;
  h->x.ptr = h->dinv.ptr + stride;
  h->r.ptr = h->x.ptr + stride;
  h->p.ptr = h->r.ptr + stride;
  h->Ap.ptr = h->p.ptr + stride;
  h->s.ptr = h->Ap.ptr + stride;
  h->z.ptr = h->s.ptr + stride;
  h->u.ptr = h->z.ptr + stride;
  configure_l2(h, 0);
  h->bvec.ensure(nvec);
  h->part.ensure(part_words(h));
  h->ctl.ensure(1);
  h->scal.ensure(16);
  h->tickets.ensure(8);
  FK_CUDA(cudaMemsetAsync(h->tickets.ptr, 0, 8 * sizeof(unsigned int), h->stream));
  FK_CUDA(cudaMemsetAsync(h->pcg_slab.ptr, 0, (kSlabVecs * stride + kCtrlWords) * sizeof(double), h->stream));
  if (!h->ctl_host) FK_CUDA(cudaMallocHost((void**)&h->ctl_host, sizeof(PcgCtl)));
}

// One PCG iteration (index j inside a chunk) issued into stream s.  Three kernels on one GPU; on a partitioned
// problem the halo rows of p are exchanged first (two more kernels over peer memory) and the dot products
// are all-reduced inside the SpMV / update kernels.  NCCL fallback: halo by send/recv, all-reduce by NCCL.
static void issue_iteration(fk_handle h, cudaStream_t s, const double* vals, int j) {
  const int n = h->n_rows;
  PcgCtl* ctl = h->ctl.ptr;
  double* part = h->part.ptr;
  double* scal = h->scal.ptr;
  const int gs = h->grid_spmv, gv = h->grid_pcg;
  if (h->comm && comm_capturable(h)) {
    // peer memory: the producers publish per-CTA partials like on one GPU; each consumer folds them and exchanges
    // this GPU's sums with the peers in its prologue (peer_fold) -- no last-CTA tail, no extra kernel
    const PeerRed pr = comm_peer_red(h);
    int npart = gs;
    if (h->halo_mode == 2) {
      // the exchange rides inside the SpMV (spmv.cuh: spmv_sell_fused_kernel): no push / wait launches, no second stream
      const HaloFused hf = comm_halo_fused(h, h->p.ptr, h->tickets.ptr + 4);
      const int4* meta = reinterpret_cast<const int4*>(h->slice_meta.ptr);
      spmv_variant_dispatch(h->spmv_variant, [&](auto G, auto MINB) {
        spmv_sell_fused_kernel<G(), MINB()><<<gs, kSpmvThreads, 0, s>>>(meta, h->srow.ptr, h->sell_colz.ptr, vals, h->p.ptr, h->Ap.ptr,
                                                                         h->sym.nslices, part, ctl, j, h->slices_all.ptr, hf);
      });
    } else if (h->n_slices_bnd > 0 && h->n_slices_int > 0 && h->overlap_halo && h->halo_stream) {
      // the rows that read no halo column start at once; the halo exchange and the rows that need it run beside
      // them on a second stream (a fork / join inside the captured graph)
      const int gi = h->grid_spmv_int, gb = h->grid_spmv_bnd;
      FK_CUDA(cudaEventRecord(h->ev_fork, s));
      FK_CUDA(cudaStreamWaitEvent(h->halo_stream, h->ev_fork, 0));
      comm_halo_exchange(h, h->halo_stream, h->p.ptr);
      launch_spmv_list(h, h->halo_stream, vals, h->p.ptr, h->Ap.ptr, true, ctl, j, RedArgs{}, h->slices_bnd.ptr, h->n_slices_bnd, gb,
                       part + gi);
      FK_CUDA(cudaEventRecord(h->ev_join, h->halo_stream));
      launch_spmv_list(h, s, vals, h->p.ptr, h->Ap.ptr, true, ctl, j, RedArgs{}, h->slices_int.ptr, h->n_slices_int, gi, part);
      FK_CUDA(cudaStreamWaitEvent(s, h->ev_join, 0));
      npart = gi + gb;
    } else {
      comm_halo_exchange(h, s, h->p.ptr);
      launch_spmv(h, s, vals, h->p.ptr, h->Ap.ptr, true, ctl, j, /*publish_only=*/true);
    }
    double* part2 = part + 2 * gs;   // >= npart: both split grids are capped like the full one
    pcg_update_kernel<<<gv, kVecThreads, 0, s>>>(h->x.ptr, h->r.ptr, h->p.ptr, h->Ap.ptr, h->dinv.ptr, n, part, npart, part2,
                                                 RedArgs{}, ctl, j, pr);
    pcg_direction_kernel<<<gv, kVecThreads, 0, s>>>(h->p.ptr, h->r.ptr, h->dinv.ptr, n, part2, gv, gv, ctl, j, pr);
  } else if (h->comm) {   // NCCL fallback: reduce over this GPU inside the kernels, then ncclAllReduce
    comm_halo_exchange(h, s, h->p.ptr);
    launch_spmv(h, s, vals, h->p.ptr, h->Ap.ptr, true, ctl, j);                    // -> scal[0] = p.Ap (this GPU)
    comm_nccl_allreduce(h, s, scal, 1, 0);
    pcg_update_kernel<<<gv, kVecThreads, 0, s>>>(h->x.ptr, h->r.ptr, h->p.ptr, h->Ap.ptr, h->dinv.ptr, n, scal, 1, part + gs,
                                                 comm_red_args(h, h->tickets.ptr + 1, scal + 1, 0), ctl, j, PeerRed{});   // -> scal[1..2] = r.r, r.z
    comm_nccl_allreduce(h, s, scal + 1, 2, 0);
    pcg_direction_kernel<<<gv, kVecThreads, 0, s>>>(h->p.ptr, h->r.ptr, h->dinv.ptr, n, scal + 1, 1, 1, ctl, j, PeerRed{});
  } else {  // one GPU: the consumers fold the per-CTA partials themselves
    launch_spmv(h, s, vals, h->p.ptr, h->Ap.ptr, true, ctl, j);
    launch_k(h->pdl, pcg_update_kernel, gv, kVecThreads, s, h->x.ptr, h->r.ptr, h->p.ptr, h->Ap.ptr, h->dinv.ptr, n, part, gs, part + gs,
             RedArgs{}, ctl, j, PeerRed{});
    launch_k(h->pdl, pcg_direction_kernel, gv, kVecThreads, s, h->p.ptr, h->r.ptr, h->dinv.ptr, n, part + gs, gv, gv, ctl, j, PeerRed{});
  }
  FK_CUDA(cudaGetLastError());
}

// Partial-sum layout of the single-reduction recurrence inside h->part (M = stride = max grid): delta partials of the
// SpMV at [0, M), buffer A = (gamma | r.r | b.b) at [M, 4M), buffer B = (gamma | r.r) at [4M, 6M).
struct CgParts {
  int M;
  double *ps, *A, *B;
};
static CgParts cg_parts(fk_handle h) {
  const int M = std::max(std::max(h->grid_spmv, h->grid_vec), h->grid_pcg);
  return CgParts{M, h->part.ptr, h->part.ptr + M, h->part.ptr + 4 * (size_t)M};
}

// S kernel of the single-reduction recurrence: w = A z fused with (w, z); on a partition also the halo exchange of z.
// jS: iteration index the kernel's activity test and exchange sequence number are derived from.
static void launch_cg_spmv(fk_handle h, cudaStream_t s, const double* vals, const PcgCtl* ctl, int jS) {
  const CgParts cp = cg_parts(h);
  if (h->comm) {
    const HaloFused hf = comm_halo_fused(h, h->z.ptr, h->tickets.ptr + 4);
    const int4* meta = reinterpret_cast<const int4*>(h->slice_meta.ptr);
    spmv_variant_dispatch(h->spmv_variant, [&](auto G, auto MINB) {
      spmv_sell_fused_kernel<G(), MINB()><<<h->grid_spmv, kSpmvThreads, 0, s>>>(meta, h->srow.ptr, h->sell_colz.ptr, vals, h->z.ptr,
                                                                                h->Ap.ptr, h->sym.nslices, cp.ps, ctl, jS,
                                                                                h->slices_all.ptr, hf);
    });
    FK_CUDA(cudaGetLastError());
  } else {
    launch_spmv(h, s, vals, h->z.ptr, h->Ap.ptr, true, ctl, jS);
  }
}

// One iteration of the single-reduction recurrence (kernels.cuh: cg_update_kernel): V_j then S_j.
static void issue_iteration_cg(fk_handle h, cudaStream_t s, const double* vals, int j) {
  const CgParts cp = cg_parts(h);
  const PeerRed pr = h->comm ? comm_peer_red(h) : PeerRed{0, 1, nullptr, nullptr};
  cg_update_kernel<<<h->grid_pcg, kVecThreads, 0, s>>>(h->x.ptr, h->r.ptr, h->p.ptr, h->s.ptr, h->Ap.ptr, h->dinv.ptr, h->z.ptr, h->n_rows,
                                                       cp.A, cp.B, h->grid_pcg, cp.M, cp.ps, h->grid_spmv, h->ctl.ptr, j, pr);
  launch_cg_spmv(h, s, vals, h->ctl.ptr, j + 1);
  FK_CUDA(cudaGetLastError());
}

static void build_graph(fk_handle h, int slot, int chunk, int rec) {
  if (h->graph[slot][rec] && h->graph_chunk[slot][rec] == chunk) return;
  if (h->graph[slot][rec]) {
    cudaGraphExecDestroy(h->graph[slot][rec]);
    h->graph[slot][rec] = nullptr;
  }
  if (!h->cap_stream) FK_CUDA(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
  cudaGraph_t g = nullptr;
  const long long launches0 = h->launches;   // capturing is not launching
  FK_CUDA(cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
  try {
    if (rec == 0) {
      for (int j = 0; j < chunk; j++) issue_iteration(h, h->cap_stream, h->vals[slot].ptr, j);
      pcg_advance_kernel<<<1, 1, 0, h->cap_stream>>>(h->ctl.ptr, chunk);
    } else {
      for (int j = 0; j < chunk; j++) issue_iteration_cg(h, h->cap_stream, h->vals[slot].ptr, j);
      const CgParts cp = cg_parts(h);
      cg_chunk_end_kernel<<<1, 256, 0, h->cap_stream>>>(cp.A, cp.B, h->grid_pcg, cp.M, h->ctl.ptr, chunk,
                                                        h->comm ? comm_peer_red(h) : PeerRed{0, 1, nullptr, nullptr});
    }
  } catch (...) {
    cudaStreamEndCapture(h->cap_stream, &g);
    if (g) cudaGraphDestroy(g);
    throw;
  }
  FK_CUDA(cudaStreamEndCapture(h->cap_stream, &g));
  h->launches = launches0;
  {  // the L2 policy window is a per-launch attribute: put it on every kernel node of the chunk
    size_t nn = 0;
    FK_CUDA(cudaGraphGetNodes(g, nullptr, &nn));
    std::vector<cudaGraphNode_t> nodes(nn);
    if (nn) FK_CUDA(cudaGraphGetNodes(g, nodes.data(), &nn));
    cudaLaunchAttributeValue val{};
    val.accessPolicyWindow = h->l2_window;
    for (cudaGraphNode_t nd : nodes) {
      cudaGraphNodeType ty;
      if (cudaGraphNodeGetType(nd, &ty) == cudaSuccess && ty == cudaGraphNodeTypeKernel)
        if (cudaGraphKernelNodeSetAttribute(nd, cudaLaunchAttributeAccessPolicyWindow, &val) != cudaSuccess) cudaGetLastError();
    }
  }
  cudaError_t e = cudaGraphInstantiate(&h->graph[slot][rec], g, 0);
  cudaGraphDestroy(g);
  FK_CUDA(e);
  h->graph_chunk[slot][rec] = chunk;
}

// sums (a | a.w, w) -> out[0..1] over every rank (w == nullptr: ones)
static void ns_sums(fk_handle h, cudaStream_t s, const double* a, const double* w, int mode, double* out) {
  const int g = h->grid_vec;
  ns_partials_kernel<<<g, kVecThreads, 0, s>>>(a, w, h->n_rows, mode, h->part.ptr, g);
  FK_CUDA(cudaGetLastError());
  h->launches++;
  comm_allreduce_partials(h, s, h->part.ptr, g, g, 2, 0, out, nullptr, 0);
}

static void pcg_dev(fk_handle h, int slot, const double* b_dev, double* x_dev, const fk_pcg_opts* o, int* iters_out,
                    double* relres_out) {
  check_slot(slot);
  need(h->vals[slot].ptr != nullptr, "fk_pcg: matrix slot has not been assembled");
  ensure_pcg(h);
  const int n = h->n_rows;
  fk_pcg_opts opt{};
  if (o) opt = *o;
  // tol_disable: tol = 0 -> threshold = DBL_MIN, never undercut by a non-zero residual
  const double tol = opt.tol_disable ? 0.0 : (opt.tol > 0 ? opt.tol : DBL_EPSILON);
  // Eigen's default 2n (IterativeSolverBase.h:251) counts the rows of the WHOLE system: on a partition every rank
  // must use the same limit, or the ranks fall out of step in the exchanges
  const long long n_glob = h->comm ? comm_global_rows(h) : (long long)n;
  const int max_iter = opt.max_iter > 0 ? opt.max_iter : (int)std::min<long long>(2 * n_glob, 0x7fffffff);
  int chunk = opt.check_every > 0 ? opt.check_every : 32;
  chunk = std::min(chunk, std::max(2, max_iter));
  if (chunk & 1) chunk++;  // scalar ping-pong parity must line up across chunks
  // single-reduction recurrence: one GPU, or partitions on the peer-memory path (the NCCL fallback keeps Eigen's)
  const int rec = (opt.recurrence == 1 && (!h->comm || comm_capturable(h))) ? 1 : 0;
  need(opt.recurrence == 0 || opt.recurrence == 1, "fk_pcg: unknown recurrence");
  if (rec != h->l2_recurrence) configure_l2(h, rec);
  cudaStream_t s = h->stream;
  const int gv = h->grid_vec;

  // singular Neumann systems: b <- b - (sum b / sum w) w (fk_pcg_opts.project_nullspace)
  const double* w_ns = (h->ns_user && h->ns_w.ptr) ? h->ns_w.ptr : nullptr;
  if (opt.project_nullspace) {
    h->bproj.ensure((size_t)h->nv);
    ns_sums(h, s, b_dev, w_ns, 0, h->scal.ptr + 8);
    ns_apply_kernel<<<gv, kVecThreads, 0, s>>>(b_dev, w_ns, n, 0, h->scal.ptr + 8, h->bproj.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    b_dev = h->bproj.ptr;
  }

  jacobi_kernel<<<cdiv(n, 256), 256, 0, s>>>(h->vals[slot].ptr, h->sell_diag.ptr, n, opt.no_precond, h->dinv.ptr);
  if (h->trace_cap > 0) trace_reset_kernel<<<cdiv((long long)h->trace_cap * kTraceKernels, 256), 256, 0, s>>>(h->trace.ptr, h->trace_cap * kTraceKernels);
  const double* Ax0 = nullptr;
  // constrained slot (fk_apply_dirichlet): rows and columns of the fixed DOFs are identity, so with x0[fixed] = gD the
  // residual, the direction and A p vanish on them EXACTLY in every iteration and x[fixed] == gD bit for bit on return
  // (Fluids/fluids3D.jl:173 writes UP[fixed] = gD after the solve).  x0 = g needs no product: A g == g on such a matrix.
  const bool constrained = ((h->dir_applied >> slot) & 1u) != 0 && h->n_fixed > 0 && h->dir_g.ptr != nullptr;
  if (opt.warm_start) {
    FK_CUDA(cudaMemcpyAsync(h->x.ptr, x_dev, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, s));
    if (constrained) {
      copy_rows_kernel<<<cdiv(h->n_fixed, 256), 256, 0, s>>>(h->x.ptr, h->dir_ids.ptr, h->dir_g.ptr, h->n_fixed);
      h->launches++;
    }
    if (h->comm) comm_halo_exchange(h, s, h->x.ptr);
    launch_spmv(h, s, h->vals[slot].ptr, h->x.ptr, h->Ap.ptr, false, nullptr, 0);
    Ax0 = h->Ap.ptr;
    h->launches++;
  } else if (constrained) {
    FK_CUDA(cudaMemcpyAsync(h->x.ptr, h->dir_g.ptr, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, s));
    Ax0 = h->dir_g.ptr;
  } else {
    FK_CUDA(cudaMemsetAsync(h->x.ptr, 0, sizeof(double) * (size_t)n, s));
  }
  const unsigned long long epoch = ++h->pcg_epoch;
  // LL mailbox tags (kernels.cuh: peer_fold): this solve owns [tag_base, tag_base + 2 max_iter + 8); every rank advances alike
  need(max_iter < (1 << 29), "fk_pcg: max_iter too large");
  const unsigned long long tag_base_next = h->tag_base + 2ULL * (unsigned long long)max_iter + 8ULL;
  // in-iteration projection of the residual's mean: one GPU (either recurrence) or the single-reduction recurrence on
  // partitions (its fused reduction carries the fourth value); the two-reduction partitioned path only projects b and x
  const int proj = (opt.project_nullspace && (!h->comm || rec == 1)) ? 1 : 0;
  const double inv_n = 1.0 / (double)n_glob;
  if (rec == 1) {
    const CgParts cp = cg_parts(h);
    const int g2 = h->grid_pcg;
    cg_init_kernel<<<g2, kVecThreads, 0, s>>>(b_dev, Ax0, h->dinv.ptr, n, h->r.ptr, h->z.ptr, cp.A, cp.M);   // partials of (r.z, r.r, b.b)
    if (h->comm) {
      comm_allreduce_partials(h, s, cp.A, g2, cp.M, 3, 0, h->scal.ptr + 4, nullptr, 0);   // -> scal[4..6] over every rank
      pcg_init_fin_kernel<<<1, 256, 0, s>>>(h->scal.ptr + 6, h->scal.ptr + 5, h->scal.ptr + 4, 1, tol, max_iter, epoch, proj, inv_n, h->ctl.ptr, h->trace.ptr, h->trace_cap, (unsigned)h->tag_base);
    } else {
      pcg_init_fin_kernel<<<1, 256, 0, s>>>(cp.A + 2 * (size_t)cp.M, cp.A + cp.M, cp.A, g2, tol, max_iter, epoch, proj, inv_n, h->ctl.ptr, h->trace.ptr, h->trace_cap, (unsigned)h->tag_base);
    }
    launch_cg_spmv(h, s, h->vals[slot].ptr, h->ctl.ptr, 0);   // w0 = A z0, delta partials (and the first halo exchange)
    h->launches++;
  } else if (h->comm) {
    pcg_init_kernel<<<gv, kVecThreads, 0, s>>>(b_dev, Ax0, h->dinv.ptr, n, h->r.ptr, h->p.ptr, h->part.ptr,
                                               comm_red_args(h, h->tickets.ptr + 2, h->scal.ptr + 4, 0));   // -> scal[4..6] = b.b, r.r, r.p
    comm_nccl_allreduce(h, s, h->scal.ptr + 4, 3, 0);
    pcg_init_fin_kernel<<<1, 256, 0, s>>>(h->scal.ptr + 4, h->scal.ptr + 5, h->scal.ptr + 6, 1, tol, max_iter, epoch, 0, 0.0, h->ctl.ptr, h->trace.ptr, h->trace_cap, (unsigned)h->tag_base);
  } else {
    pcg_init_kernel<<<gv, kVecThreads, 0, s>>>(b_dev, Ax0, h->dinv.ptr, n, h->r.ptr, h->p.ptr, h->part.ptr, RedArgs{});
    pcg_init_fin_kernel<<<1, 256, 0, s>>>(h->part.ptr, h->part.ptr + gv, h->part.ptr + 2 * (size_t)gv, gv, tol, max_iter, epoch, proj, inv_n, h->ctl.ptr, h->trace.ptr, h->trace_cap, (unsigned)h->tag_base);
  }
  FK_CUDA(cudaGetLastError());
  h->launches += 3;

  if (rec == 0 && h->halo_mode != 2 && h->comm && comm_capturable(h) && !h->halo_stream) {   // second stream of the split SpMV (created outside any capture)
    FK_CUDA(cudaStreamCreateWithFlags(&h->halo_stream, cudaStreamNonBlocking));
    FK_CUDA(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
    FK_CUDA(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
    if (h->l2_window_set) {
      cudaStreamAttrValue attr{};
      attr.accessPolicyWindow = h->l2_window;
      if (cudaStreamSetAttribute(h->halo_stream, cudaStreamAttributeAccessPolicyWindow, &attr) != cudaSuccess) cudaGetLastError();
    }
  }
  if (h->l2_limit_zeroed) {   // the sweep took the persisting set-aside away (fk_tune 15): the solve gets it back
    const auto t0 = std::chrono::steady_clock::now();
    if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, h->l2_window.num_bytes) != cudaSuccess) cudaGetLastError();
    h->l2_toggle_ns[1] += std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
    h->l2_limit_zeroed = false;
  }
  const bool graphed = !h->comm || comm_capturable(h);
  if (graphed) build_graph(h, slot, chunk, rec);
  else if (h->l2_window_set) {   // no graph to carry the window (NCCL fallback): the stream does
    cudaStreamAttrValue attr{};
    attr.accessPolicyWindow = h->l2_window;
    if (s && cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &attr) != cudaSuccess) cudaGetLastError();
  }
  int launched = 0;
  while (true) {
    if (!graphed) {
      for (int j = 0; j < chunk; j++) issue_iteration(h, s, h->vals[slot].ptr, j);
      pcg_advance_kernel<<<1, 1, 0, s>>>(h->ctl.ptr, chunk);
      h->launches += (long long)chunk * 3 + 1;   // the halo pack kernel is counted in comm_halo_exchange
    } else {
      FK_CUDA(cudaGraphLaunch(h->graph[slot][rec], s));
      const bool split = h->comm && h->halo_stream && h->overlap_halo && h->n_slices_bnd > 0 && h->n_slices_int > 0;
      const int per_it = rec == 1 ? 2 : (!h->comm || h->halo_mode == 2) ? 3 : (split ? 6 : 5);
      h->launches += (long long)chunk * per_it + 1;
    }
    launched += chunk;
    FK_CUDA(cudaMemcpyAsync(h->ctl_host, h->ctl.ptr, sizeof(PcgCtl), cudaMemcpyDeviceToHost, s));
    FK_CUDA(cudaStreamSynchronize(s));
    if (h->ctl_host->done || launched >= max_iter) break;
    if (!(h->ctl_host->r2 == h->ctl_host->r2)) break;   // NaN residual (e.g. non-finite matrix values): it would never pass the test
  }
  h->tag_base = tag_base_next;
  if (h->comm) comm_check(h);
  const PcgCtl& c = *h->ctl_host;
  const int iters = c.done ? c.iters : max_iter;
  if (c.rhs2 == 0.0) {  // x = 0 (ConjugateGradient.h:43-49)
    FK_CUDA(cudaMemsetAsync(x_dev, 0, sizeof(double) * (size_t)n, s));
  } else if (opt.project_nullspace == 1) {   // x <- x - (w.x / sum w)
    ns_sums(h, s, h->x.ptr, w_ns, 1, h->scal.ptr + 8);
    ns_apply_kernel<<<gv, kVecThreads, 0, s>>>(h->x.ptr, nullptr, n, 1, h->scal.ptr + 8, x_dev);
    FK_CUDA(cudaGetLastError());
    h->launches++;
  } else {
    FK_CUDA(cudaMemcpyAsync(x_dev, h->x.ptr, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, s));
  }
  if (iters_out) *iters_out = iters;
  if (relres_out) *relres_out = c.rhs2 > 0.0 ? std::sqrt(c.r2 / c.rhs2) : 0.0;
}

// ------------------------------------------------------------------------------------------------
// host spline helpers (FITPACK curfit(k=3,s=0) knots + collocation solve; fpbspl evaluation)
// ------------------------------------------------------------------------------------------------
static int find_interval(const double* t, int n, int k, double x) {
  int lo = k, hi = n - k - 2;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (t[mid] <= x) lo = mid; else hi = mid - 1;
  }
  return lo;
}
static void bspline_basis(const double* t, int k, int l, double x, double* h) {
  double hh[6];
  h[0] = 1.0;
  for (int j = 1; j <= k; j++) {
    for (int i = 0; i < j; i++) hh[i] = h[i];
    h[0] = 0.0;
    for (int i = 0; i < j; i++) {
      const double tli = t[l + i + 1], tlj = t[l + i + 1 - j];
      if (tli == tlj) { h[i + 1] = 0.0; continue; }
      const double f = hh[i] / (tli - tlj);
      h[i] = h[i] + f * (tli - x);
      h[i + 1] = f * (x - tlj);
    }
  }
}
static int spline_eval_host(const double* t, const double* c, int n, int k, int bc, double x, double* out) {
  const double tb = t[k], te = t[n - k - 1];
  if (x < tb || x > te || x != x) {
    if (bc == FK_BC_ZERO) { *out = 0.0; return 0; }
    if (bc == FK_BC_ERROR || x != x) { *out = NAN; return 2; }
    if (bc == FK_BC_NEAREST) x = (x < tb) ? tb : te;
  }
  const int l = find_interval(t, n, k, x);
  double h[6];
  bspline_basis(t, k, l, x, h);
  double sp = 0.0;
  for (int j = 0; j <= k; j++) sp += c[l - k + j] * h[j];
  *out = sp;
  return 0;
}

}  // namespace fk

using namespace fk;

// =================================================================================================
struct fk_symbolic_s { fk::Symbolic S; };

static SymbolicOpts to_symbolic_opts(const fk_create_opts& o) {
  SymbolicOpts so;
  if (o.tile_max_elems > 0) so.tile_max_elems = o.tile_max_elems;
  if (o.tile_max_rows > 0) so.tile_max_rows = o.tile_max_rows;
  so.natural_order = o.natural_order != 0;
  if (o.sell_sigma > 0) so.sell_sigma = o.sell_sigma;
  if (o.n_owned > 0) so.n_owned = o.n_owned;
  so.wide_cols = o.sell_wide_cols != 0;
  if (o.nodes_per_element != 0) so.npe = o.nodes_per_element;
  return so;
}

template <class T, class U>
static void copy_vec(const std::vector<T>& v, U* dst) {
  static_assert(sizeof(T) == sizeof(U), "size mismatch");
  if (dst && !v.empty()) std::memcpy(dst, v.data(), sizeof(T) * v.size());
}

extern "C" {

static void ensure_p2(fk_handle h);   // P2 contribution plan (defined with the P2 entry points below)
static void ensure_ned(fk_handle h);  // edge elements: the same, 21 local pairs

const char* fk_last_error(void) { return fk::g_err.c_str(); }
int fk_set_host_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
  return omp_get_max_threads();
#else
  (void)n;
  return 1;
#endif
}
int fk_version(void) { return 100; }
// Page-lock a caller's host array (cudaHostRegister) so that the host-pointer entry points copy at pinned speed: a Julia Array
// is pageable, and a caller without CUDA.jl has no other way to pin it.  The array must be unregistered before it is freed.
int fk_host_register(void* ptr, long long bytes) {
  return guard([&]() {
    need(ptr != nullptr && bytes > 0, "fk_host_register: NULL pointer or empty range");
    FK_CUDA(cudaHostRegister(ptr, (size_t)bytes, cudaHostRegisterPortable));
    return 0;
  });
}
int fk_host_unregister(void* ptr) {
  return guard([&]() {
    need(ptr != nullptr, "fk_host_unregister: NULL pointer");
    FK_CUDA(cudaHostUnregister(ptr));
    return 0;
  });
}

// fk_create and fk_create_edge.  t = the connectivity of the DEGREES OF FREEDOM (npe x nt, ids < nv); t_vertices = the four
// vertices of every element when t does not start with them (edge elements), n_points = points in p
static int create_impl(const double* p, int n_points, const int* t, int nv, int nt, const double* VE, const fk_create_opts* opts,
                       const int* t_vertices, fk_handle* out) {
  {
    need(out != nullptr, "fk_create: out is NULL");
    *out = nullptr;
    need(p && t, "fk_create: p and t must not be NULL");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
      throw std::runtime_error("fk_create: no CUDA device available (this library has no CPU path)");
    fk_create_opts o{};
    if (opts) o = *opts;
    int dev = o.device;
    if (opts == nullptr || dev < 0) FK_CUDA(cudaGetDevice(&dev));
    need(dev < ndev, "fk_create: device ordinal out of range");
    DeviceGuard dg(dev);
    std::unique_ptr<fk_handle_s> h(new fk_handle_s());
    h->device = dev;
    cudaDeviceProp prop;
    FK_CUDA(cudaGetDeviceProperties(&prop, dev));
    h->num_sms = prop.multiProcessorCount;
    h->nv = nv;
    h->nt = nt;
    h->n_points = n_points;
    h->edge_dofs = t_vertices != nullptr;

    SymbolicOpts so = to_symbolic_opts(o);
    if (h->edge_dofs) so.npe = 6;
    else need(so.npe != 6, "fk_create: 6 degrees of freedom per element are edge elements: use fk_create_edge");
    build_symbolic(h->edge_dofs ? nullptr : p, t, nv, nt, so, h->sym);
    const Symbolic& S = h->sym;
    h->n_rows = S.n_rows;
    h->nnz = S.nnz;
    h->nnz_sell = S.nnz_sell;

    cudaStream_t s = h->stream;
    // mesh: nodes padded to 32 B, connectivity as int4
    h->st_a.upload(p, (size_t)3 * n_points, s);
    h->p4.ensure((size_t)n_points);
    pad_nodes_kernel<<<cdiv(n_points, 256), 256, 0, s>>>(h->st_a.ptr, n_points, h->p4.ptr);
    h->t4.ensure((size_t)nt);
    std::vector<int> t4host;   // P2: the vertices of every element (first four of ten) for the geometry kernels
    if (S.npe == 4) {
      FK_CUDA(cudaMemcpyAsync(h->t4.ptr, t, sizeof(int) * 4 * (size_t)nt, cudaMemcpyHostToDevice, s));
    } else {
      t4host.resize(4 * (size_t)nt);
      for (int k = 0; k < nt; k++)
        for (int i = 0; i < 4; i++) {
          const int v = t_vertices ? t_vertices[4 * (size_t)k + i] : t[(size_t)S.npe * k + i];
          need(v >= 0 && v < n_points, "fk_create: vertex id out of range");
          t4host[4 * (size_t)k + i] = v;
        }
      FK_CUDA(cudaMemcpyAsync(h->t4.ptr, t4host.data(), sizeof(int) * 4 * (size_t)nt, cudaMemcpyHostToDevice, s));
      h->tP2.upload(t, (size_t)S.npe * nt, s);
      FK_CUDA(cudaStreamSynchronize(s));
    }
    h->VE.ensure((size_t)nt);
    if (VE) {
      h->user_ve = true;
      FK_CUDA(cudaMemcpyAsync(h->VE.ptr, VE, sizeof(double) * (size_t)nt, cudaMemcpyHostToDevice, s));
    } else {
      volumes_kernel<<<cdiv(nt, 256), 256, 0, s>>>(h->p4.ptr, h->t4.ptr, nt, h->VE.ptr);
    }
    h->rowptr.ensure(S.rowptr.size() + kSlack);
    h->rowptr.upload(S.rowptr, s);
    h->colind.ensure(S.colind.size() + kSlack);
    h->colind.upload(S.colind, s);
    h->slice_ptr.upload(S.slice_ptr, s);
    h->srow.upload(S.srow, s);
    h->slice_meta.upload(S.slice_meta, s);
    if (o.n_owned > 0 && o.n_owned < nv) {   // partition: slices that read halo columns vs slices that do not
      std::vector<int> li, lb;
      for (int sl = 0; sl < S.nslices; sl++) {
        bool halo = false;
        for (int e = S.slice_ptr[sl]; e < S.slice_ptr[sl + 1] && !halo; e++) halo = S.sell_col[(size_t)e] >= o.n_owned;
        (halo ? lb : li).push_back(sl);
      }
      h->n_slices_int = (int)li.size();
      h->n_slices_bnd = (int)lb.size();
      std::vector<int> la(li);
      la.insert(la.end(), lb.begin(), lb.end());
      h->slices_int.upload(li, s);
      h->slices_bnd.upload(lb, s);
      h->slices_all.upload(la, s);
      FK_CUDA(cudaStreamSynchronize(s));
    }
    h->sell_colz.ensure(S.sell_colz.size() + 2 * kSlack);
    h->sell_colz.upload(S.sell_colz, s);
    h->csr2sell.upload(S.csr2sell, s);
    h->sell_diag.upload(S.sell_diag, s);
    h->tile_hdr.upload(S.tile_hdr, s);
    h->tile_blob.upload(S.tile_blob, s);
    FK_CUDA(cudaStreamSynchronize(s));
    // the device only needs the packed blobs; drop the host copies of the unpacked plan
    h->sym.tile_blob = std::vector<uint32_t>();
    h->e_pad = S.e_pad;
    need(o.asm_cfg >= 0 && o.asm_cfg < kNumAsmCfg, "fk_create: asm_cfg out of range");
    h->asm_cfg = o.asm_cfg;
    h->asm_smem = asm_smem_bytes(h->e_pad, S.max_tile_nodes, S.max_blob_i, S.max_blob_g, S.max_blob_b);
    need(h->asm_smem <= (size_t)prop.sharedMemPerBlockOptin, "fk_create: assembly tile does not fit in shared memory");
    {
      AsmArgs dummy{};
      launch_assemble<ASM_STIFF>(h.get(), dummy, true);
      launch_assemble<ASM_MASS>(h.get(), dummy, true);
      launch_assemble<ASM_TANGENT>(h.get(), dummy, true);
      launch_assemble<ASM_PICARD>(h.get(), dummy, true);
      launch_assemble<ASM_LAGRANGE>(h.get(), dummy, true);
      launch_assemble<ASM_STIFF2D>(h.get(), dummy, true);
      launch_assemble<ASM_MASS2D>(h.get(), dummy, true);
    }

    // launch geometry of the persistent SpMV / vector kernels: full residency, fixed for determinism
    h->spmv_ctas = o.spmv_ctas_per_sm;
    configure_spmv(h.get(), o.spmv_ctas_per_sm);
    h->grid_vec = std::max(1, std::min(cdiv(std::max(1, h->n_rows), kVecThreads), h->num_sms * 8));
    h->grid_pcg = std::max(1, std::min(cdiv(std::max(1, h->n_rows), 2 * kVecUnroll * kVecThreads), h->num_sms * kPcgVecCtas));
    FK_CUDA(cudaStreamSynchronize(s));
    h->st_a.release();
    h->launches += 2;
    *out = h.release();
    return 0;
  }
}
int fk_create(const double* p, const int* t, int nv, int nt, const double* VE, const fk_create_opts* opts, fk_handle* out) {
  return guard([&]() { return create_impl(p, nv, t, nv, nt, VE, opts, nullptr, out); });
}
int fk_create_edge(const double* p, int n_points, const int* t4, const int* edges6, int n_edges, int nt, const double* VE,
                   const fk_create_opts* opts, fk_handle* out) {
  return guard([&]() {
    need(t4 != nullptr, "fk_create_edge: t4 is NULL");
    need(n_points > 0 && n_edges > 0, "fk_create_edge: empty mesh");
    return create_impl(p, n_points, edges6, n_edges, nt, VE, opts, t4, out);
  });
}

int fk_destroy(fk_handle h) {
  return guard([&]() {
    if (!h) return 0;
    DeviceGuard dg(h->device);
    cudaDeviceSynchronize();
    if (std::getenv("FK_DEBUG") && h->l2_toggles)
      std::fprintf(stderr, "fk: L2 set-aside toggled %lld times: release %.1f us, restore %.1f us per call (host time, waits included)\n",
                   h->l2_toggles, h->l2_toggle_ns[0] / 1e3 / h->l2_toggles, h->l2_toggle_ns[1] / 1e3 / h->l2_toggles);
    invalidate_graphs(h);
    if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
    if (h->ctl_host) cudaFreeHost(h->ctl_host);
    if (h->pin8) cudaFreeHost(h->pin8);
    if (h->ev_fork) { cudaEventDestroy(h->ev_fork); h->ev_fork = nullptr; }
    if (h->ev_join) { cudaEventDestroy(h->ev_join); h->ev_join = nullptr; }
    if (h->halo_stream) { cudaStreamDestroy(h->halo_stream); h->halo_stream = nullptr; }   // comm_destroy below must not destroy it again
    if (h->asm_fork) cudaEventDestroy(h->asm_fork);
    if (h->asm_join) cudaEventDestroy(h->asm_join);
    if (h->asm_stream) cudaStreamDestroy(h->asm_stream);
    if (h->copy_ev) cudaEventDestroy(h->copy_ev);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    // leave the device as the host application had it: the persisting lines are released when the last handle with a
    // window goes away, and the persisting-L2 limit this handle raised is put back
    if (h->l2_window_set && --g_l2_users <= 0) { cudaCtxResetPersistingL2Cache(); g_l2_users = 0; }
    if (h->l2_limit_changed && g_l2_users <= 0) cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, h->l2_limit_saved);
    comm_destroy(h);
    delete h;
    return 0;
  });
}

int fk_set_stream(fk_handle h, void* stream) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    h->stream = (cudaStream_t)stream;
    configure_l2(h, h->l2_recurrence);
    return 0;
  });
}

int fk_tune(fk_handle h, int what, int value) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    FK_CUDA(cudaStreamSynchronize(h->stream));
    switch (what) {
      case 0: {
        need(value >= 0 && value < kNumAsmCfg, "fk_tune: asm_cfg out of range");
        h->asm_cfg = value;
        AsmArgs dummy{};
        launch_assemble<ASM_STIFF>(h, dummy, true);
        launch_assemble<ASM_MASS>(h, dummy, true);
        launch_assemble<ASM_TANGENT>(h, dummy, true);
        launch_assemble<ASM_PICARD>(h, dummy, true);
        launch_assemble<ASM_LAGRANGE>(h, dummy, true);
        launch_assemble<ASM_STIFF2D>(h, dummy, true);
        launch_assemble<ASM_MASS2D>(h, dummy, true);
        break;
      }
      case 1: h->spmv_ctas = value; configure_spmv(h, value); break;
      case 2: h->l2_persist = value != 0; configure_l2(h, h->l2_recurrence); break;
      case 3: h->asm_dbg = value; break;   // profiling experiments: skip assembly phases (results are garbage)
      case 4:
        need(value >= 0 && value < kNumSpmvVariants, "fk_tune: spmv variant out of range");
        h->spmv_variant = value;
        configure_spmv(h, h->spmv_ctas);
        break;
      case 7:   // single GPU: programmatic dependent launch between the kernels of a PCG iteration (1 = on; default off)
        h->pdl = value != 0;
        invalidate_graphs(h);
        break;
      case 6:
        need(value >= 0 && value < 8, "fk_tune: reserve out of range");
        h->spmv_int_reserve = value;
        configure_spmv(h, h->spmv_ctas);
        break;
      case 5:   // partitions, Eigen recurrence: 2 (default) = exchange fused into the SpMV kernel; 1 = halo exchange + halo-coupled
                // rows on a second stream beside the interior rows; 0 = serial
        need(value >= 0 && value <= 2, "fk_tune: halo mode out of range");
        h->halo_mode = value;
        h->overlap_halo = value != 0;
        invalidate_graphs(h);
        break;
      case 8:   // L2 policy of everything outside the persisting window: 0 = streaming (default), 1 = normal
        h->l2_miss_normal = value != 0;
        configure_l2(h, h->l2_recurrence);
        break;
      case 10:  // rescale assembly: 0 = never (tile kernel only), 1 = from the second Picard sweep on (default), 2 = at once
        need(value >= 0 && value <= 2, "fk_tune: geo mode out of range");
        h->geo_mode = value;
        break;
      case 14:  // 1: keep the L2 window on the handle's stream too (the behaviour before round 2); 0 (default): graph nodes only
        h->l2_on_stream = value != 0;
        configure_l2(h, h->l2_recurrence);
        break;
      case 15:  // the assembly sweep and the PCG's persisting L2 set-aside: 0 leave it; bit 0: reset the persisting lines before a
                // sweep; bit 1: take the set-aside away for the sweep (cudaLimitPersistingL2CacheSize = 0) and put it back
                // before the next solve
        need(value >= 0 && value <= 3, "fk_tune: sweep L2 mode out of range");
        h->l2_sweep_mode = value;
        break;
      case 19:  // profiling experiments only (results are garbage): bit 0 skips the field-recovery kernel of the side-by-side sweep,
                // bit 1 the row-sum diagonal, bit 2 the streamed sums
        h->geo_dbg = value;
        break;
      case 18:  // rescale path: 1 (default) = diagonal from the row sums (6 contributions per tetrahedron); 0 = diagonal summed from its
                // own contributions (10 per tetrahedron, bit-identical to the tile kernel).  Rebuilds the plan.
        h->geo_rowsum_diag = value != 0;
        if (h->geo_ready && h->geo_plan_rowsum != h->geo_rowsum_diag) h->geo_ready = false;
        break;
      case 17:  // rescale sweep: 0 = element kernel then sums, one stream; otherwise 16 a + b = field recovery on b CTAs per SM
                // beside the streamed sums on a CTAs per SM (default 16*4 + 1).  With the PCG's L2 set-aside released (fk_tune 15)
                // 4 + 1 measures 0.40 ms against 0.45 in sequence (profiles/r2_asm_bench_rowsum_diag_side_by_side.jsonl); while the
                // set-aside was still active no split gained anything (profiles/r2_asm_bench_side_by_side.jsonl)
        need(value >= 0 && value < 256, "fk_tune: split out of range");
        h->geo_overlap = value != 0;
        if (value) { h->geo_split_resc = std::max(1, value >> 4); h->geo_split_elem = std::max(1, value & 15); }
        break;
      case 13:  // rescale plan: items are sorted by list length inside windows of this many items (32 = plan order kept); rebuilds
        need(value >= 32 && value <= (1 << 20), "fk_tune: geo window out of range");
        h->geo_window = value;
        h->geo_ready = false;
        break;
      case 12:  // launch shape of geo_rescale_kernel {batch, min CTAs per SM}: 0 {8,4} (default) 1 {4,6} 2 {8,3} 3 {6,5}
        need(value >= 0 && value < kNumGeoVariants, "fk_tune: geo variant out of range");
        h->geo_variant = value;
        if (h->geo_ready) configure_geo(h);
        break;
      case 11:  // device-side timeline of the PCG kernels: record the last `value` iterations of every solve (0 = off)
        need(value >= 0 && value <= 1 << 20, "fk_tune: trace length out of range");
        h->trace_cap = value;
        if (value > 0) h->trace.ensure((size_t)value * kTraceKernels * 4);
        break;
      case 9:   // cap on the number of PCG vectors pinned in L2 (0 = as many as may persist)
        need(value >= 0 && value <= 7, "fk_tune: vector count out of range");
        h->l2_max_vecs = value;
        configure_l2(h, h->l2_recurrence);
        break;
      default: throw std::runtime_error("fk_tune: unknown option");
    }
    return 0;
  });
}

int fk_trace_read(fk_handle h, unsigned long long* out, int max_records, int* n_records) {
  return guard([&]() {
    need(h && out && n_records, "fk: NULL argument");
    DeviceGuard dg(h->device);
    const int n = std::min(max_records, h->trace_cap * kTraceKernels);
    *n_records = n;
    if (n > 0) {
      FK_CUDA(cudaMemcpyAsync(out, h->trace.ptr, sizeof(unsigned long long) * 4 * (size_t)n, cudaMemcpyDeviceToHost, h->stream));
      FK_CUDA(cudaStreamSynchronize(h->stream));
    }
    return 0;
  });
}

int fk_synchronize(fk_handle h) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    FK_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
  });
}

int fk_launch_count(fk_handle h, long long* n) {
  return guard([&]() {
    need(h && n, "fk: NULL argument");
    *n = h->launches;
    return 0;
  });
}

int fk_sizes(fk_handle h, int* nv, int* nt, long long* nnz, int* ntiles) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    if (nv) *nv = h->nv;
    if (nt) *nt = h->nt;
    if (nnz) *nnz = h->nnz;
    if (ntiles) *ntiles = h->sym.ntiles;
    return 0;
  });
}

int fk_csr_pattern(fk_handle h, int* rowptr, int* colind) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    if (rowptr) std::memcpy(rowptr, h->sym.rowptr.data(), sizeof(int) * ((size_t)h->nv + 1));
    if (colind) std::memcpy(colind, h->sym.colind.data(), sizeof(int) * (size_t)h->nnz);
    return 0;
  });
}

int fk_elem2nnz(fk_handle h, int* map) {
  return guard([&]() {
    need(h && map, "fk: NULL argument");
    std::memcpy(map, h->sym.e2nnz.data(), sizeof(int) * h->sym.e2nnz.size());   // 16 nt (P1) or 100 nt (P2) entries
    return 0;
  });
}

int fk_sell_stats(fk_handle h, long long* nnz_sell, long long* n_slices, long long* n_slices16, long long* col_bytes) {
  return guard([&]() {
    need(h, "fk: NULL argument");
    if (nnz_sell) *nnz_sell = h->sym.nnz_sell;
    if (n_slices) *n_slices = h->sym.nslices;
    if (n_slices16) *n_slices16 = h->sym.n_slices16;
    if (col_bytes) *col_bytes = 2 * (long long)h->sym.sell_colz.size();
    return 0;
  });
}
int fk_geo_stats(fk_handle h, long long* items, long long* contributions, long long* slots, long long* bytes) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    if (h->sym.npe == 10) ensure_p2(h); else if (h->sym.npe == 6) ensure_ned(h); else ensure_geo(h);
    if (items) *items = h->geo_items;
    if (contributions) *contributions = h->geo_contrib;
    if (slots) *slots = h->geo_slots;
    if (bytes) *bytes = 12 * h->geo_slots + 8LL * 32 * h->geo_nslices + 4LL * (h->geo_nslices + 1) + 8LL * h->nt;
    return 0;
  });
}
int fk_tile_stats(fk_handle h, long long* staged_elems, int* max_elems, int* max_nnz, int* max_ents) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    if (staged_elems) *staged_elems = (long long)h->sym.tile_elems.size();
    if (max_elems) *max_elems = h->sym.max_tile_elems;
    if (max_nnz) *max_nnz = h->sym.max_tile_nnz;
    if (max_ents) *max_ents = h->sym.max_tile_ents;
    return 0;
  });
}

int fk_volumes(fk_handle h, double* VE) {
  return guard([&]() {
    need(h && VE, "fk: NULL argument");
    DeviceGuard dg(h->device);
    FK_CUDA(cudaMemcpyAsync(VE, h->VE.ptr, sizeof(double) * (size_t)h->nt, cudaMemcpyDeviceToHost, h->stream));
    FK_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
  });
}

int fk_device_ptr(fk_handle h, int which, void** ptr) {
  return guard([&]() {
    need(h && ptr, "fk: NULL argument");
    DeviceGuard dg(h->device);
    if (which >= 16 && which < 16 + FK_NUM_SLOTS) { *ptr = slot_vals(h, which - 16); return 0; }
    if (which >= 3 && which <= 6) ensure_state(h);
    switch (which) {
      case 0: *ptr = h->rowptr.ptr; break;
      case 1: *ptr = h->colind.ptr; break;
      case 2: *ptr = h->VE.ptr; break;
      case 3: *ptr = h->mu.ptr; h->s0_dirty = true; break;   // the caller may write mu through this pointer
      case 4: *ptr = h->Hn.ptr; break;
      case 5: *ptr = h->Hfield.ptr; break;
      case 6: ensure_pcg(h); *ptr = h->u.ptr; break;
      case 7: *ptr = h->dmu.ptr; break;
      default: throw std::runtime_error("fk_device_ptr: unknown array id");
    }
    return 0;
  });
}

// ---- symbolic phase alone (host) ------------------------------------------------------------------
int fk_symbolic_create(const double* p, const int* t, int nv, int nt, const fk_create_opts* opts, fk_symbolic* out) {
  return guard([&]() {
    need(out && p && t, "fk: NULL argument");
    fk_create_opts o{};
    if (opts) o = *opts;
    std::unique_ptr<fk_symbolic_s> s(new fk_symbolic_s());
    build_symbolic(p, t, nv, nt, to_symbolic_opts(o), s->S);
    *out = s.release();
    return 0;
  });
}
int fk_symbolic_sizes(fk_symbolic s, long long* nnz, int* ntiles, long long* n_tile_elems, long long* nnz_tiles,
                      long long* n_ents, int* nslices, long long* nnz_sell) {
  return guard([&]() {
    need(s, "fk: NULL argument");
    if (nnz) *nnz = s->S.nnz;
    if (ntiles) *ntiles = s->S.ntiles;
    if (n_tile_elems) *n_tile_elems = (long long)s->S.tile_elems.size();
    if (nnz_tiles) *nnz_tiles = (long long)s->S.tile_gidx.size();
    if (n_ents) *n_ents = (long long)s->S.tile_ents.size();
    if (nslices) *nslices = s->S.nslices;
    if (nnz_sell) *nnz_sell = s->S.nnz_sell;
    return 0;
  });
}
int fk_symbolic_export(fk_symbolic s, int* rowptr, int* colind, int* diag, int* e2nnz, long long* tile_elem_ptr,
                       unsigned int* tile_elems, long long* tile_nnz_ptr, int* tile_gidx, unsigned short* tile_coff,
                       long long* tile_ent_ptr, unsigned short* tile_ents, int* slice_ptr, int* srow, int* sell_col,
                       int* csr2sell) {
  return guard([&]() {
    need(s, "fk: NULL argument");
    const Symbolic& S = s->S;
    copy_vec(S.rowptr, rowptr);
    copy_vec(S.colind, colind);
    copy_vec(S.diag, diag);
    copy_vec(S.e2nnz, e2nnz);
    copy_vec(S.tile_elem_ptr, tile_elem_ptr);
    copy_vec(S.tile_elems, tile_elems);
    copy_vec(S.tile_nnz_ptr, tile_nnz_ptr);
    copy_vec(S.tile_gidx, tile_gidx);
    copy_vec(S.tile_coff, tile_coff);
    copy_vec(S.tile_ent_ptr, tile_ent_ptr);
    copy_vec(S.tile_ents, tile_ents);
    copy_vec(S.slice_ptr, slice_ptr);
    copy_vec(S.srow, srow);
    copy_vec(S.sell_col, sell_col);
    copy_vec(S.csr2sell, csr2sell);
    return 0;
  });
}
int fk_symbolic_colz(fk_symbolic s, long long* n_colz, long long* n_slices16, int* slice_meta, unsigned short* colz) {
  return guard([&]() {
    need(s, "fk: NULL argument");
    const Symbolic& S = s->S;
    if (n_colz) *n_colz = (long long)S.sell_colz.size();
    if (n_slices16) *n_slices16 = S.n_slices16;
    if (slice_meta) copy_vec(S.slice_meta, slice_meta);
    if (colz) copy_vec(S.sell_colz, colz);
    return 0;
  });
}
// rescale plan of a host-side symbolic object (tests): sizes with every pointer NULL-able, then the arrays
int fk_symbolic_rescale(fk_symbolic s, long long* n_items, int* n_slices, long long* n_slots, unsigned int* slice_ptr, int* dst1,
                        int* dst2, unsigned int* cref) {
  return guard([&]() {
    need(s, "fk: NULL argument");
    RescalePlan R;
    build_rescale_plan(s->S, 4096, R);
    if (n_items) *n_items = R.n_items;
    if (n_slices) *n_slices = R.nslices;
    if (n_slots) *n_slots = R.n_slots;
    copy_vec(R.slice_ptr, slice_ptr);
    copy_vec(R.dst1, dst1);
    copy_vec(R.dst2, dst2);
    copy_vec(R.cref, cref);
    return 0;
  });
}
int fk_symbolic_destroy(fk_symbolic s) {
  delete s;
  return 0;
}

// ---- assembly -----------------------------------------------------------------------------------
// matrix values leave the device in CSR order (the pattern fk_csr_pattern returns)
static void copy_out_matrix(fk_handle h, const double* vals, double* host) {
  if (host) {
    h->st_m.ensure((size_t)h->nnz);
    sell_to_csr_kernel<<<std::min(cdiv(h->nnz, 256), h->num_sms * 16), 256, 0, h->stream>>>(vals, h->csr2sell.ptr, h->nnz, h->st_m.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    FK_CUDA(cudaMemcpyAsync(host, h->st_m.ptr, sizeof(double) * (size_t)h->nnz, cudaMemcpyDeviceToHost, h->stream));
  }
  FK_CUDA(cudaStreamSynchronize(h->stream));
}
static void copy_out(fk_handle h, const double* dev, double* host, size_t n) {
  if (host) FK_CUDA(cudaMemcpyAsync(host, dev, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
  FK_CUDA(cudaStreamSynchronize(h->stream));
}

int fk_assemble_stiffness_dev(fk_handle h, int slot, const double* f_dev) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    assemble_stiffness_dev(h, slot, f_dev);
    return 0;
  });
}
int fk_assemble_mass_dev(fk_handle h, int slot, const double* rho_dev) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    assemble_mass_dev(h, slot, rho_dev);
    return 0;
  });
}
int fk_assemble_tangential_dev(fk_handle h, int slot, const double* H_dev, const double* dmu_dev) {
  return guard([&]() {
    need(h && H_dev && dmu_dev, "fk: NULL argument");
    DeviceGuard dg(h->device);
    assemble_tangential_dev(h, slot, H_dev, dmu_dev);
    return 0;
  });
}
int fk_assemble_stiffness(fk_handle h, int slot, const double* f, double* vals_out) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    if (f) h->st_a.upload(f, (size_t)h->nt, h->stream);
    assemble_stiffness_dev(h, slot, f ? h->st_a.ptr : nullptr);
    copy_out_matrix(h, h->vals[slot].ptr, vals_out);
    return 0;
  });
}
int fk_assemble_mass(fk_handle h, int slot, const double* rho, double* vals_out) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    if (rho) h->st_a.upload(rho, (size_t)h->nt, h->stream);
    assemble_mass_dev(h, slot, rho ? h->st_a.ptr : nullptr);
    copy_out_matrix(h, h->vals[slot].ptr, vals_out);
    return 0;
  });
}
int fk_assemble_tangential(fk_handle h, int slot, const double* Hfield, const double* dmu, double* vals_out) {
  return guard([&]() {
    need(h && Hfield && dmu, "fk: NULL argument");
    DeviceGuard dg(h->device);
    h->st_a.upload(dmu, (size_t)h->nt, h->stream);
    h->st_b.upload(Hfield, (size_t)3 * h->nt, h->stream);
    assemble_tangential_dev(h, slot, h->st_b.ptr, h->st_a.ptr);
    copy_out_matrix(h, h->vals[slot].ptr, vals_out);
    return 0;
  });
}
// ---- P2 tetrahedra (p2.cuh) ---------------------------------------------------------------------------
// sum_q w_q phi_i phi_j on the reference element, 15-point rule of GaussQuadrature3D(4) (src/FEM3D.jl:711-745)
struct P2Quad {
  const double (*pts)[3];
  const double* wts;
};
static P2Quad p2_quadrature() {
  static const double pts[15][3] = {
      {0.2500000000000000, 0.2500000000000000, 0.2500000000000000}, {0.0000000000000000, 0.3333333333333333, 0.3333333333333333},
      {0.3333333333333333, 0.3333333333333333, 0.3333333333333333}, {0.3333333333333333, 0.3333333333333333, 0.0000000000000000},
      {0.3333333333333333, 0.0000000000000000, 0.3333333333333333}, {0.7272727272727273, 0.0909090909090909, 0.0909090909090909},
      {0.0909090909090909, 0.0909090909090909, 0.0909090909090909}, {0.0909090909090909, 0.0909090909090909, 0.7272727272727273},
      {0.0909090909090909, 0.7272727272727273, 0.0909090909090909}, {0.4334498464263357, 0.0665501535736643, 0.0665501535736643},
      {0.0665501535736643, 0.4334498464263357, 0.0665501535736643}, {0.0665501535736643, 0.0665501535736643, 0.4334498464263357},
      {0.0665501535736643, 0.4334498464263357, 0.4334498464263357}, {0.4334498464263357, 0.0665501535736643, 0.4334498464263357},
      {0.4334498464263357, 0.4334498464263357, 0.0665501535736643}};
  static const double wts[15] = {0.1817020685825351, 0.0361607142857143, 0.0361607142857143, 0.0361607142857143, 0.0361607142857143,
                                 0.0698714945161738, 0.0698714945161738, 0.0698714945161738, 0.0698714945161738, 0.0656948493683187,
                                 0.0656948493683187, 0.0656948493683187, 0.0656948493683187, 0.0656948493683187, 0.0656948493683187};
  return P2Quad{pts, wts};
}
static P2Ref p2_reference_mass() {
  const P2Quad Q = p2_quadrature();
  const double (*pts)[3] = Q.pts;
  const double* wts = Q.wts;
  static const int ea[6] = {0, 1, 0, 0, 2, 1}, eb[6] = {1, 2, 2, 3, 3, 3};
  P2Ref R{};
  for (int q = 0; q < 15; q++) {
    const double l[4] = {1.0 - pts[q][0] - pts[q][1] - pts[q][2], pts[q][0], pts[q][1], pts[q][2]};
    double phi[10];
    for (int v = 0; v < 4; v++) phi[v] = l[v] * (2.0 * l[v] - 1.0);
    for (int e = 0; e < 6; e++) phi[4 + e] = 4.0 * l[ea[e]] * l[eb[e]];
    const double w = wts[q] / 6.0;
    int n = 0;
    for (int i = 0; i < 10; i++)
      for (int j = i; j < 10; j++) R.m[n++] += w * phi[i] * phi[j];
  }
  return R;
}
static void ensure_p2(fk_handle h) {
  if (h->geo_ready) return;
  need(h->sym.npe == 10, "fk: the handle does not hold P2 elements (fk_create_opts.nodes_per_element = 10)");
  RescalePlan R;
  build_rescale_plan(h->sym, h->geo_window, R);
  cudaStream_t s = h->stream;
  h->geo_nslices = R.nslices;
  h->geo_slots = R.n_slots;
  h->geo_items = R.n_items;
  h->geo_contrib = R.n_contrib;
  h->geo_slice_ptr.upload(R.slice_ptr, s);
  h->geo_dst1.upload(R.dst1, s);
  h->geo_dst2.upload(R.dst2, s);
  h->geo_ref.upload(R.cref, s);
  h->ned_ref.upload(R.cref, s);   // the copy that keeps its pair bits (convection: per-call slot values)
  h->ned_shift = R.pair_shift;
  h->geo_g.ensure((size_t)std::max<long long>(R.n_slots, 1));
  h->p2_gmass.ensure((size_t)std::max<long long>(R.n_slots, 1));
  h->geo_s0.ensure((size_t)h->nt);
  if (R.n_slots > 0) {
    p2_geo_build_kernel<<<std::min(cdiv(R.n_slots, 128), h->num_sms * 16), 128, 0, s>>>(h->p4.ptr, h->tP2.ptr, h->VE.ptr, h->geo_ref.ptr,
                                                                                        R.n_slots, R.pair_shift, p2_reference_mass(),
                                                                                        h->geo_g.ptr, h->p2_gmass.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
  }
  FK_CUDA(cudaStreamSynchronize(s));
  configure_geo(h);
  h->geo_ready = true;
}
static int assemble_p2(fk_handle h, int slot, const double* coef, double* vals_out, bool mass) {
  need(h, "fk: NULL handle");
  DeviceGuard dg(h->device);
  check_slot(slot);
  ensure_p2(h);
  const double* cd = nullptr;
  if (coef) { h->st_a.upload(coef, (size_t)h->nt, h->stream); cd = h->st_a.ptr; }
  h->dir_applied &= ~(1u << slot);
  double* vals = slot_vals(h, slot);
  p2_scale_kernel<<<cdiv(h->nt, 256), 256, 0, h->stream>>>(cd, h->nt, h->geo_s0.ptr);
  h->launches++;
  launch_geo_rescale(h, vals, 0, mass ? h->p2_gmass.ptr : h->geo_g.ptr);
  copy_out_matrix(h, vals, vals_out);
  return 0;
}
int fk_assemble_stiffness_p2(fk_handle h, int slot, const double* mu, double* vals_out) {
  return guard([&]() { return assemble_p2(h, slot, mu, vals_out, false); });
}
int fk_assemble_mass_p2(fk_handle h, int slot, const double* F, double* vals_out) {
  return guard([&]() { return assemble_p2(h, slot, F, vals_out, true); });
}

// T[i][j][m] = sum_q w_q phi_i (d phi_j / d lambda_m) on the reference element (p2.cuh: P2ConvRef), same 15-point rule
static void p2_reference_convection(P2ConvRef& R) {
  P2Quad Q = p2_quadrature();
  std::memset(&R, 0, sizeof(R));
  static const int ea[6] = {0, 1, 0, 0, 2, 1}, eb[6] = {1, 2, 2, 3, 3, 3};
  for (int q = 0; q < 15; q++) {
    const double l[4] = {1.0 - Q.pts[q][0] - Q.pts[q][1] - Q.pts[q][2], Q.pts[q][0], Q.pts[q][1], Q.pts[q][2]};
    double phi[10], dphi[10][4] = {};
    for (int v = 0; v < 4; v++) { phi[v] = l[v] * (2.0 * l[v] - 1.0); dphi[v][v] = 4.0 * l[v] - 1.0; }
    for (int e = 0; e < 6; e++) {
      phi[4 + e] = 4.0 * l[ea[e]] * l[eb[e]];
      dphi[4 + e][ea[e]] = 4.0 * l[eb[e]];
      dphi[4 + e][eb[e]] = 4.0 * l[ea[e]];
    }
    const double w = Q.wts[q] / 6.0;
    for (int i = 0; i < 10; i++)
      for (int j = 0; j < 10; j++)
        for (int m = 0; m < 4; m++) R.T[i][j][m] += w * phi[i] * dphi[j][m];
  }
}
int fk_assemble_convection_p2(fk_handle h, int slot, const double* vel, const double* F, int compat_bug, double* vals_out) {
  return guard([&]() {
    need(h && vel, "fk: NULL argument");
    DeviceGuard dg(h->device);
    check_slot(slot);
    ensure_p2(h);
    cudaStream_t s = h->stream;
    if (!h->p2_conv_ref.ptr) {
      P2ConvRef R;
      p2_reference_convection(R);
      h->p2_conv_ref.upload(&R.T[0][0][0], 400, s);
      FK_CUDA(cudaStreamSynchronize(s));   // R lives on this stack frame
    }
    const size_t ns = (size_t)std::max<long long>(h->geo_slots, 1);
    h->p2_c1.ensure(ns);
    h->p2_c2.ensure(ns);
    h->st_b.upload(vel, (size_t)3 * h->nv, s);
    const double* cd = nullptr;
    if (F) { h->st_a.upload(F, (size_t)h->nt, s); cd = h->st_a.ptr; }
    h->dir_applied &= ~(1u << slot);
    double* vals = slot_vals(h, slot);
    p2_scale_kernel<<<cdiv(h->nt, 256), 256, 0, s>>>(cd, h->nt, h->geo_s0.ptr);
    if (h->geo_slots > 0) {
      p2_conv_slot_kernel<<<std::min(cdiv(h->geo_slots, 128), h->num_sms * 16), 128, 0, s>>>(
          h->p4.ptr, h->tP2.ptr, h->VE.ptr, h->ned_ref.ptr, h->geo_slots, h->ned_shift, h->st_b.ptr,
          reinterpret_cast<const P2ConvRef*>(h->p2_conv_ref.ptr), compat_bug, h->p2_c1.ptr, h->p2_c2.ptr);
      FK_CUDA(cudaGetLastError());
      l2_release_for_assembly(h);
      geo_rescale2_kernel<<<std::max(1, std::min(cdiv(h->geo_nslices, kGeoThreads / 32), h->num_sms * 3)), kGeoThreads, 0, s>>>(
          h->geo_slice_ptr.ptr, h->geo_dst1.ptr, h->geo_dst2.ptr, h->geo_ref.ptr, h->p2_c1.ptr, h->p2_c2.ptr, h->geo_s0.ptr, h->geo_nslices, vals);
      FK_CUDA(cudaGetLastError());
    }
    h->launches += 3;
    copy_out_matrix(h, vals, vals_out);
    return 0;
  });
}

// ---- Nedelec (edge) elements (nedelec.cuh; src/FEM3D.jl:153-271, Magnetostatics/nedelec3D.jl:104-250) ---------------------
static void ensure_ned(fk_handle h) {
  need(h->edge_dofs && h->sym.npe == 6, "fk: the handle does not hold edge elements (fk_create_edge)");
  if (h->geo_ready) return;
  RescalePlan R;
  build_rescale_plan(h->sym, h->geo_window, R);
  cudaStream_t s = h->stream;
  h->geo_nslices = R.nslices;
  h->geo_slots = R.n_slots;
  h->geo_items = R.n_items;
  h->geo_contrib = R.n_contrib;
  h->ned_shift = R.pair_shift;
  h->geo_slice_ptr.upload(R.slice_ptr, s);
  h->geo_dst1.upload(R.dst1, s);
  h->geo_dst2.upload(R.dst2, s);
  h->ned_ref.upload(R.cref, s);
  const size_t ns = (size_t)std::max<long long>(R.n_slots, 1);
  h->geo_ref.ensure(ns);
  h->geo_g.ensure(ns);
  h->p2_gmass.ensure(ns);
  h->geo_s0.ensure((size_t)h->nt);
  if (R.n_slots > 0) {
    const int grid = std::min(cdiv(R.n_slots, 128), h->num_sms * 16);
    ned_strip_kernel<<<std::min(cdiv(R.n_slots, 256), h->num_sms * 16), 256, 0, s>>>(h->ned_ref.ptr, R.n_slots, R.pair_shift, h->geo_ref.ptr);
    ned_slot_kernel<0><<<grid, 128, 0, s>>>(h->p4.ptr, h->t4.ptr, h->VE.ptr, h->ned_ref.ptr, R.n_slots, R.pair_shift, nullptr, nullptr,
                                           nullptr, 0.0, 0.0, 0.0, h->geo_g.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches += 2;
  }
  FK_CUDA(cudaStreamSynchronize(s));
  configure_geo(h);
  h->geo_ready = true;
}
int fk_nedelec_local_stiffness(fk_handle h, double* Ak) {
  return guard([&]() {
    need(h && Ak, "fk: NULL argument");
    need(h->edge_dofs, "fk: the handle does not hold edge elements (fk_create_edge)");
    DeviceGuard dg(h->device);
    h->st_b.ensure((size_t)36 * h->nt);
    ned_local_kernel<<<cdiv(h->nt, 128), 128, 0, h->stream>>>(h->p4.ptr, h->t4.ptr, h->VE.ptr, h->nt, h->st_b.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    copy_out(h, h->st_b.ptr, Ak, (size_t)36 * h->nt);
    return 0;
  });
}
int fk_assemble_nedelec_stiffness(fk_handle h, int slot, const double* nu, double* vals_out) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    check_slot(slot);
    ensure_ned(h);
    const double* cd = nullptr;
    if (nu) { h->st_a.upload(nu, (size_t)h->nt, h->stream); cd = h->st_a.ptr; }
    h->dir_applied &= ~(1u << slot);
    double* vals = slot_vals(h, slot);
    p2_scale_kernel<<<cdiv(h->nt, 256), 256, 0, h->stream>>>(cd, h->nt, h->geo_s0.ptr);
    h->launches++;
    launch_geo_rescale(h, vals, 0, h->geo_g.ptr);
    copy_out_matrix(h, vals, vals_out);
    return 0;
  });
}
int fk_assemble_nedelec_tangential(fk_handle h, int slot, const double* dnu, const double* Bfield, const double* B, double* vals_out) {
  return guard([&]() {
    need(h && dnu && Bfield && B, "fk: NULL argument");
    DeviceGuard dg(h->device);
    check_slot(slot);
    ensure_ned(h);
    cudaStream_t s = h->stream;
    h->st_a.upload(dnu, (size_t)h->nt, s);
    h->st_b.upload(Bfield, (size_t)3 * h->nt, s);
    h->st_c.upload(B, (size_t)h->nt, s);
    h->dir_applied &= ~(1u << slot);
    double* vals = slot_vals(h, slot);
    if (h->geo_slots > 0) {
      ned_slot_kernel<1><<<std::min(cdiv(h->geo_slots, 128), h->num_sms * 16), 128, 0, s>>>(
          h->p4.ptr, h->t4.ptr, h->VE.ptr, h->ned_ref.ptr, h->geo_slots, h->ned_shift, h->st_a.ptr, h->st_b.ptr, h->st_c.ptr, 0.0, 0.0, 0.0,
          h->p2_gmass.ptr);
      FK_CUDA(cudaGetLastError());
    }
    p2_scale_kernel<<<cdiv(h->nt, 256), 256, 0, s>>>(nullptr, h->nt, h->geo_s0.ptr);
    h->launches += 2;
    launch_geo_rescale(h, vals, 0, h->p2_gmass.ptr);
    copy_out_matrix(h, vals, vals_out);
    return 0;
  });
}
int fk_nedelec_load_vector(fk_handle h, const double* Bext, double* rhs_out) {
  return guard([&]() {
    need(h && Bext && rhs_out, "fk: NULL argument");
    DeviceGuard dg(h->device);
    ensure_ned(h);
    cudaStream_t s = h->stream;
    h->ned_vals.ensure((size_t)h->nnz_sell + kSlack);
    FK_CUDA(cudaMemsetAsync(h->ned_vals.ptr, 0, sizeof(double) * ((size_t)h->nnz_sell + kSlack), s));
    if (h->geo_slots > 0) {
      ned_slot_kernel<2><<<std::min(cdiv(h->geo_slots, 128), h->num_sms * 16), 128, 0, s>>>(
          h->p4.ptr, h->t4.ptr, h->VE.ptr, h->ned_ref.ptr, h->geo_slots, h->ned_shift, nullptr, nullptr, nullptr, Bext[0], Bext[1], Bext[2],
          h->p2_gmass.ptr);
      FK_CUDA(cudaGetLastError());
    }
    p2_scale_kernel<<<cdiv(h->nt, 256), 256, 0, s>>>(nullptr, h->nt, h->geo_s0.ptr);
    h->launches += 2;
    launch_geo_rescale(h, h->ned_vals.ptr, 0, h->p2_gmass.ptr);
    h->st_d.ensure((size_t)h->nv);
    gather_diag_kernel<<<cdiv(h->n_rows, 256), 256, 0, s>>>(h->ned_vals.ptr, h->sell_diag.ptr, h->n_rows, h->st_d.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    copy_out(h, h->st_d.ptr, rhs_out, (size_t)h->n_rows);
    return 0;
  });
}
int fk_nedelec_recover_field(fk_handle h, const double* u, double* Bfield_out, double* B_out) {
  return guard([&]() {
    need(h && u, "fk: NULL argument");
    need(h->edge_dofs, "fk: the handle does not hold edge elements (fk_create_edge)");
    DeviceGuard dg(h->device);
    cudaStream_t s = h->stream;
    h->st_c.upload(u, (size_t)h->nv, s);
    h->st_b.ensure((size_t)3 * h->nt);
    h->st_a.ensure((size_t)h->nt);
    ned_recover_kernel<<<cdiv(h->nt, 128), 128, 0, s>>>(h->p4.ptr, h->t4.ptr, h->tP2.ptr, h->st_c.ptr, h->nt, h->st_b.ptr, h->st_a.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    if (Bfield_out) copy_out(h, h->st_b.ptr, Bfield_out, (size_t)3 * h->nt);
    if (B_out) copy_out(h, h->st_a.ptr, B_out, (size_t)h->nt);
    FK_CUDA(cudaStreamSynchronize(s));
    return 0;
  });
}

// 2-D P1 triangles (src/FEM2D.jl:48-76 stiffnessMatrix2D, :139-161 massMatrix2D) on a handle created from the elements
// (n0, n1, n2, n2): same plan, same kernels, 3 x 3 local matrices
int fk_assemble_stiffness2d(fk_handle h, int slot, const double* f, double* vals_out) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    const double* fd = nullptr;
    if (f) { h->st_a.upload(f, (size_t)h->nt, h->stream); fd = h->st_a.ptr; }
    h->dir_applied &= ~(1u << slot);
    AsmArgs a = base_args(h, slot_vals(h, slot));
    a.coef = fd;
    launch_assemble<ASM_STIFF2D>(h, a);
    if (vals_out) copy_out_matrix(h, h->vals[slot].ptr, vals_out);
    else FK_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
  });
}
int fk_assemble_mass2d(fk_handle h, int slot, const double* rho, double* vals_out) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    const double* rd = nullptr;
    if (rho) { h->st_a.upload(rho, (size_t)h->nt, h->stream); rd = h->st_a.ptr; }
    h->dir_applied &= ~(1u << slot);
    AsmArgs a = base_args(h, slot_vals(h, slot));
    a.coef = rd;
    launch_assemble<ASM_MASS2D>(h, a);
    if (vals_out) copy_out_matrix(h, h->vals[slot].ptr, vals_out);
    else FK_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
  });
}
int fk_local_stiffness(fk_handle h, const double* f, double* Ak) {
  return guard([&]() {
    need(h && Ak, "fk: NULL argument");
    DeviceGuard dg(h->device);
    if (f) h->st_a.upload(f, (size_t)h->nt, h->stream);
    h->st_b.ensure((size_t)16 * h->nt);
    local_stiffness_kernel<<<cdiv(h->nt, 128), 128, 0, h->stream>>>(h->p4.ptr, h->t4.ptr, h->user_ve ? h->VE.ptr : nullptr,
                                                                    f ? h->st_a.ptr : nullptr, h->nt, h->st_b.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    copy_out(h, h->st_b.ptr, Ak, (size_t)16 * h->nt);
    return 0;
  });
}

int fk_matrix_axpby(fk_handle h, int dst, double alpha, int x, double beta, int y) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    check_slot(x);
    DeviceGuard dg(h->device);
    need(h->vals[x].ptr != nullptr, "fk_matrix_axpby: slot x is empty");
    if (beta != 0.0) { check_slot(y); need(h->vals[y].ptr != nullptr, "fk_matrix_axpby: slot y is empty"); }
    double* d = slot_vals(h, dst);
    h->dir_applied &= ~(1u << dst);
    axpby_kernel<<<std::min(cdiv(h->nnz_sell, 256), h->num_sms * 8), 256, 0, h->stream>>>(
        d, alpha, h->vals[x].ptr, beta, beta != 0.0 ? h->vals[y].ptr : nullptr, h->nnz_sell);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    return 0;
  });
}
int fk_matrix_get(fk_handle h, int slot, double* vals_out) {
  return guard([&]() {
    need(h && vals_out, "fk: NULL argument");
    check_slot(slot);
    need(h->vals[slot].ptr != nullptr, "fk_matrix_get: slot is empty");
    DeviceGuard dg(h->device);
    copy_out_matrix(h, h->vals[slot].ptr, vals_out);
    return 0;
  });
}
int fk_matrix_set(fk_handle h, int slot, const double* vals_in) {
  return guard([&]() {
    need(h && vals_in, "fk: NULL argument");
    DeviceGuard dg(h->device);
    double* d = slot_vals(h, slot);
    h->dir_applied &= ~(1u << slot);
    h->st_m.upload(vals_in, (size_t)h->nnz, h->stream);
    csr_to_sell_kernel<<<std::min(cdiv(h->nnz, 256), h->num_sms * 16), 256, 0, h->stream>>>(h->st_m.ptr, h->csr2sell.ptr, h->nnz, d);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    FK_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
  });
}

// ---- right-hand sides ---------------------------------------------------------------------------
int fk_lagrange(fk_handle h, double* C_out) {
  return guard([&]() {
    need(h && C_out, "fk: NULL argument");
    DeviceGuard dg(h->device);
    // diag-only "matrix" VE/4 per local node, assembled through the same gather plan: the diagonal
    // then holds sum_k VE[k]/4 in ascending element order = the reference loop (FEM3D.jl:274-281)
    h->st_b.ensure((size_t)h->nnz_sell);
    AsmArgs a = base_args(h, h->st_b.ptr);
    a.VE = h->VE.ptr;
    launch_assemble<ASM_LAGRANGE>(h, a);
    h->st_c.ensure((size_t)h->nv);
    FK_CUDA(cudaMemsetAsync(h->st_c.ptr, 0, sizeof(double) * (size_t)h->nv, h->stream));
    gather_diag_kernel<<<cdiv(h->n_rows, 256), 256, 0, h->stream>>>(h->st_b.ptr, h->sell_diag.ptr, h->n_rows, h->st_c.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    copy_out(h, h->st_c.ptr, C_out, (size_t)h->nv);
    return 0;
  });
}

int fk_boundary_integral(fk_handle h, const int* surfaceT, const double* normal, const double* AE, int ns, const double* F,
                         const int* shell_ids, int nshell, double* RHS_out) {
  return guard([&]() {
    need(h && surfaceT && normal && F && RHS_out, "fk: NULL argument");
    need(ns >= 0 && (nshell == 0 || shell_ids), "fk_boundary_integral: bad shell id list");
    DeviceGuard dg(h->device);
    const int nv = h->nv;
    // integer bookkeeping on the host: node -> selected surface triangles, ascending triangle order
    std::vector<int> ptr((size_t)nv + 1, 0);
    std::vector<char> sel((size_t)ns, 0);
    for (int s = 0; s < ns; s++) {
      const int id = surfaceT[4 * (size_t)s + 3];
      for (int m = 0; m < nshell; m++) if (shell_ids[m] == id) sel[s] = 1;   // `in shell_id`, FEM3D.jl:289
      if (!sel[s]) continue;
      for (int i = 0; i < 3; i++) {
        const int n = surfaceT[4 * (size_t)s + i];
        need(n >= 0 && n < nv, "fk_boundary_integral: surface node out of range");
        ptr[(size_t)n + 1]++;
      }
    }
    for (int n = 0; n < nv; n++) ptr[n + 1] += ptr[n];
    std::vector<int> n2s((size_t)ptr[nv] + 1), fill(ptr.begin(), ptr.end() - 1);
    for (int s = 0; s < ns; s++)
      if (sel[s])
        for (int i = 0; i < 3; i++) n2s[fill[surfaceT[4 * (size_t)s + i]]++] = s;
    cudaStream_t st = h->stream;
    h->st_i.upload(surfaceT, (size_t)4 * ns, st);
    h->st_j.upload(ptr, st);
    h->st_k.upload(n2s, st);
    h->st_a.upload(normal, (size_t)3 * ns, st);
    if (AE) h->st_b.upload(AE, (size_t)ns, st);
    h->st_c.ensure((size_t)nv);
    boundary_integral_kernel<<<cdiv(nv, 256), 256, 0, st>>>(h->p4.ptr, h->st_i.ptr, h->st_a.ptr, AE ? h->st_b.ptr : nullptr,
                                                           h->st_j.ptr, h->st_k.ptr, nv, F[0], F[1], F[2], h->st_c.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    copy_out(h, h->st_c.ptr, RHS_out, (size_t)nv);
    return 0;
  });
}

// ---- SpMV / PCG ---------------------------------------------------------------------------------
int fk_spmv_dev(fk_handle h, int slot, const double* x_dev, double* y_dev) {
  return guard([&]() {
    need(h && x_dev && y_dev, "fk: NULL argument");
    check_slot(slot);
    need(h->vals[slot].ptr != nullptr, "fk_spmv: matrix slot has not been assembled");
    DeviceGuard dg(h->device);
    launch_spmv(h, h->stream, h->vals[slot].ptr, x_dev, y_dev, false, nullptr, 0);
    h->launches++;
    return 0;
  });
}
int fk_spmv(fk_handle h, int slot, const double* x, double* y) {
  return guard([&]() {
    need(h && x && y, "fk: NULL argument");
    check_slot(slot);
    need(h->vals[slot].ptr != nullptr, "fk_spmv: matrix slot has not been assembled");
    DeviceGuard dg(h->device);
    h->st_c.upload(x, (size_t)h->nv, h->stream);
    h->st_d.ensure((size_t)h->nv);
    FK_CUDA(cudaMemsetAsync(h->st_d.ptr, 0, sizeof(double) * (size_t)h->nv, h->stream));
    launch_spmv(h, h->stream, h->vals[slot].ptr, h->st_c.ptr, h->st_d.ptr, false, nullptr, 0);
    h->launches++;
    copy_out(h, h->st_d.ptr, y, (size_t)h->nv);
    return 0;
  });
}
int fk_pcg_dev(fk_handle h, int slot, const double* b_dev, double* x_dev, const fk_pcg_opts* opts, int* iters_out,
               double* relres_out) {
  return guard([&]() {
    need(h && b_dev && x_dev, "fk: NULL argument");
    DeviceGuard dg(h->device);
    pcg_dev(h, slot, b_dev, x_dev, opts, iters_out, relres_out);
    return 0;
  });
}
int fk_pcg(fk_handle h, int slot, const double* b, double* x, const fk_pcg_opts* opts, int* iters_out, double* relres_out) {
  return guard([&]() {
    need(h && b && x, "fk: NULL argument");
    DeviceGuard dg(h->device);
    ensure_pcg(h);
    const size_t n = (size_t)h->n_rows;
    FK_CUDA(cudaMemcpyAsync(h->bvec.ptr, b, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    h->st_d.ensure((size_t)h->nv);
    if (opts && opts->warm_start) FK_CUDA(cudaMemcpyAsync(h->st_d.ptr, x, sizeof(double) * n, cudaMemcpyHostToDevice, h->stream));
    pcg_dev(h, slot, h->bvec.ptr, h->st_d.ptr, opts, iters_out, relres_out);
    copy_out(h, h->st_d.ptr, x, n);
    return 0;
  });
}

// ||b - A x|| (and the residual itself) over every rank; x_dev must be a vector of nv entries (halo rows are refreshed)
static double residual_dev(fk_handle h, int slot, const double* x_dev, const double* b_dev, double* r_dev) {
  check_slot(slot);
  need(h->vals[slot].ptr != nullptr, "fk_residual: matrix slot has not been assembled");
  ensure_pcg(h);
  cudaStream_t s = h->stream;
  const double* xin = x_dev;
  if (h->comm) {   // the product reads halo rows: stage x in a slab vector and exchange
    FK_CUDA(cudaMemcpyAsync(h->z.ptr, x_dev, sizeof(double) * (size_t)h->n_rows, cudaMemcpyDeviceToDevice, s));
    comm_halo_exchange(h, s, h->z.ptr);
    xin = h->z.ptr;
  }
  launch_spmv(h, s, h->vals[slot].ptr, xin, h->Ap.ptr, false, nullptr, 0);
  const int g = h->grid_vec;
  residual_kernel<<<g, kVecThreads, 0, s>>>(b_dev, h->Ap.ptr, h->n_rows, r_dev, h->part.ptr);
  FK_CUDA(cudaGetLastError());
  h->launches += 2;
  comm_allreduce_partials(h, s, h->part.ptr, g, g, 1, 0, h->scal.ptr + 10, nullptr, 0);
  if (!h->pin8) FK_CUDA(cudaMallocHost((void**)&h->pin8, 16 * sizeof(double)));
  FK_CUDA(cudaMemcpyAsync(h->pin8 + 10, h->scal.ptr + 10, sizeof(double), cudaMemcpyDeviceToHost, s));
  FK_CUDA(cudaStreamSynchronize(s));
  return std::sqrt(h->pin8[10]);
}
int fk_residual_dev(fk_handle h, int slot, const double* x_dev, const double* b_dev, double* r_dev, double* norm_out) {
  return guard([&]() {
    need(h && x_dev && b_dev, "fk: NULL argument");
    DeviceGuard dg(h->device);
    const double nr = residual_dev(h, slot, x_dev, b_dev, r_dev);
    if (norm_out) *norm_out = nr;
    return 0;
  });
}
int fk_residual(fk_handle h, int slot, const double* x, const double* b, double* r, double* norm_out) {
  return guard([&]() {
    need(h && x && b, "fk: NULL argument");
    DeviceGuard dg(h->device);
    h->st_c.upload(x, (size_t)h->nv, h->stream);
    h->st_a.upload(b, (size_t)h->n_rows, h->stream);
    h->st_d.ensure((size_t)h->nv);
    const double nr = residual_dev(h, slot, h->st_c.ptr, h->st_a.ptr, r ? h->st_d.ptr : nullptr);
    if (norm_out) *norm_out = nr;
    if (r) copy_out(h, h->st_d.ptr, r, (size_t)h->n_rows);
    return 0;
  });
}

// lineSearch of src/FEM3D.jl:123-152 (same trial sequence, same acceptance rule); every vector stays on the device
static void line_search_dev(fk_handle h, int slot, double* u_dev, const double* du_dev, const double* rhs_dev, double min_step,
                            double* residue_out, double* alpha_out) {
  if (!(min_step > 0.0)) min_step = 1e-4;
  cudaStream_t s = h->stream;
  const int n = h->n_rows;
  const int g = std::min(cdiv(std::max(n, 1), 256), h->num_sms * 8);
  h->bproj.ensure((size_t)h->nv);   // trial vector
  double* trial = h->bproj.ptr;
  const double residual_old = residual_dev(h, slot, u_dev, rhs_dev, nullptr);
  double alf = 1.0;
  axpby_kernel<<<g, 256, 0, s>>>(trial, 1.0, u_dev, alf, du_dev, n);
  double residue = residual_dev(h, slot, trial, rhs_dev, nullptr);
  while (residue > residual_old && alf > min_step) {
    alf *= 0.9;
    axpby_kernel<<<g, 256, 0, s>>>(trial, 1.0, u_dev, alf, du_dev, n);
    h->launches++;
    residue = residual_dev(h, slot, trial, rhs_dev, nullptr);
  }
  h->launches++;
  if (alf < min_step) {
    alf = 0.0;   // "N-R could not decrease the residue further": u is left as it was
  } else {
    FK_CUDA(cudaMemcpyAsync(u_dev, trial, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, s));
  }
  if (residue_out) *residue_out = residue;
  if (alpha_out) *alpha_out = alf;
}
int fk_line_search_dev(fk_handle h, int slot, double* u_dev, const double* du_dev, const double* rhs_dev, double min_step,
                       double* residue_out, double* alpha_out) {
  return guard([&]() {
    need(h && u_dev && du_dev && rhs_dev, "fk: NULL argument");
    DeviceGuard dg(h->device);
    line_search_dev(h, slot, u_dev, du_dev, rhs_dev, min_step, residue_out, alpha_out);
    return 0;
  });
}
int fk_line_search(fk_handle h, int slot, double* u, const double* du, const double* rhs, double min_step, double* residue_out,
                   double* alpha_out) {
  return guard([&]() {
    need(h && u && du && rhs, "fk: NULL argument");
    DeviceGuard dg(h->device);
    h->st_c.upload(u, (size_t)h->nv, h->stream);
    h->st_a.upload(rhs, (size_t)h->n_rows, h->stream);
    h->st_b.upload(du, (size_t)h->n_rows, h->stream);
    line_search_dev(h, slot, h->st_c.ptr, h->st_b.ptr, h->st_a.ptr, min_step, residue_out, alpha_out);
    copy_out(h, h->st_c.ptr, u, (size_t)h->n_rows);
    return 0;
  });
}

// ---- field recovery / nonlinear step --------------------------------------------------------------
int fk_recover_field_dev(fk_handle h, const double* u_dev, const double* mu_dev, int compat_bug, double* Hfield_dev,
                         double* Hnorm_dev, double* B_dev) {
  return guard([&]() {
    need(h && u_dev, "fk: NULL argument");
    need(!B_dev || mu_dev, "fk_recover_field: B requested without mu");
    DeviceGuard dg(h->device);
    recover_field_kernel<<<cdiv(h->nt, 256), 256, 0, h->stream>>>(h->p4.ptr, h->t4.ptr, u_dev, mu_dev, h->nt, compat_bug,
                                                                  Hfield_dev, Hnorm_dev, B_dev);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    return 0;
  });
}
int fk_recover_field(fk_handle h, const double* u, const double* mu, int compat_bug, double* Hfield_out, double* Hnorm_out,
                     double* B_out) {
  return guard([&]() {
    need(h && u, "fk: NULL argument");
    need(!B_out || mu, "fk_recover_field: B requested without mu");
    DeviceGuard dg(h->device);
    const size_t nt = (size_t)h->nt;
    h->st_c.upload(u, (size_t)h->nv, h->stream);
    if (mu) h->st_a.upload(mu, nt, h->stream);
    h->st_b.ensure(5 * nt);
    double* Hf = h->st_b.ptr;
    double* Hn = Hf + 3 * nt;
    double* B = Hn + nt;
    recover_field_kernel<<<cdiv(h->nt, 256), 256, 0, h->stream>>>(h->p4.ptr, h->t4.ptr, h->st_c.ptr, mu ? h->st_a.ptr : nullptr,
                                                                  h->nt, compat_bug, Hfield_out ? Hf : nullptr,
                                                                  Hnorm_out ? Hn : nullptr, B_out ? B : nullptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    if (Hfield_out) FK_CUDA(cudaMemcpyAsync(Hfield_out, Hf, sizeof(double) * 3 * nt, cudaMemcpyDeviceToHost, h->stream));
    if (Hnorm_out) FK_CUDA(cudaMemcpyAsync(Hnorm_out, Hn, sizeof(double) * nt, cudaMemcpyDeviceToHost, h->stream));
    if (B_out) FK_CUDA(cudaMemcpyAsync(B_out, B, sizeof(double) * nt, cudaMemcpyDeviceToHost, h->stream));
    FK_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
  });
}

int fk_spline_fit(const double* x, const double* y, int m, double* t_out, double* c_out) {
  return guard([&]() {
    need(x && y && t_out && c_out, "fk: NULL argument");
    need(m >= 4, "fk_spline_fit: a cubic interpolating spline needs at least 4 points");
    for (int i = 1; i < m; i++) need(x[i] > x[i - 1], "fk_spline_fit: x must be strictly increasing");
    const int k = 3, n = m + k + 1;
    for (int i = 0; i <= k; i++) { t_out[i] = x[0]; t_out[n - 1 - i] = x[m - 1]; }
    for (int i = 0; i < m - k - 1; i++) t_out[k + 1 + i] = x[i + 2];   // interior knots x[3..m-2] (1-based)
    // collocation N c = y, N[i][j] = B_j(x_i): banded, solved densely with partial pivoting (m is table-sized)
    std::vector<double> N((size_t)m * m, 0.0), rhs(y, y + m);
    for (int i = 0; i < m; i++) {
      const int l = find_interval(t_out, n, k, x[i]);
      double hb[6];
      bspline_basis(t_out, k, l, x[i], hb);
      for (int j = 0; j <= k; j++) N[(size_t)i * m + (l - k + j)] = hb[j];
    }
    for (int c = 0; c < m; c++) {
      int piv = c;
      for (int r = c + 1; r < m; r++) if (std::fabs(N[(size_t)r * m + c]) > std::fabs(N[(size_t)piv * m + c])) piv = r;
      need(N[(size_t)piv * m + c] != 0.0, "fk_spline_fit: singular collocation matrix");
      if (piv != c) {
        for (int q = 0; q < m; q++) std::swap(N[(size_t)c * m + q], N[(size_t)piv * m + q]);
        std::swap(rhs[c], rhs[piv]);
      }
      for (int r = c + 1; r < m; r++) {
        const double f = N[(size_t)r * m + c] / N[(size_t)c * m + c];
        if (f == 0.0) continue;
        for (int q = c; q < m; q++) N[(size_t)r * m + q] -= f * N[(size_t)c * m + q];
        rhs[r] -= f * rhs[c];
      }
    }
    for (int r = m - 1; r >= 0; r--) {
      double s = rhs[r];
      for (int q = r + 1; q < m; q++) s -= N[(size_t)r * m + q] * c_out[q];
      c_out[r] = s / N[(size_t)r * m + r];
    }
    for (int i = m; i < n; i++) c_out[i] = 0.0;
    return 0;
  });
}

int fk_spline_eval(const double* t, const double* c, int n, int k, int bc, const double* x, int nx, double* y_out) {
  int status = 0;
  int rc = guard([&]() {
    need(t && c && x && y_out, "fk: NULL argument");
    need(k >= 1 && k <= 5 && n >= 2 * (k + 1), "fk_spline_eval: bad spline");
    for (int i = 0; i < nx; i++) status |= spline_eval_host(t, c, n, k, bc, x[i], y_out + i);
    return 0;
  });
  if (rc) return rc;
  if (status) set_error("fk_spline_eval: argument outside the knot range with bc = error");
  return status;
}

int fk_set_spline(fk_handle h, int id, const double* t, const double* c, int n, int k, int bc) {
  return guard([&]() {
    need(h && t && c, "fk: NULL argument");
    need(id >= 0 && id < FK_MAX_SPLINES, "fk_set_spline: table id out of range");
    need(k >= 1 && k <= 5 && n >= 2 * (k + 1), "fk_set_spline: bad spline (need 1 <= k <= 5 and n >= 2k+2)");
    need(bc >= 0 && bc <= 3, "fk_set_spline: bad bc");
    DeviceGuard dg(h->device);
    h->spl_t[id].upload(t, (size_t)n, h->stream);
    h->spl_c[id].upload(c, (size_t)n, h->stream);
    FK_CUDA(cudaStreamSynchronize(h->stream));
    h->spl.s[id] = SplineDev{h->spl_t[id].ptr, h->spl_c[id].ptr, n, k, bc};
    return 0;
  });
}

static int spline_apply_dev(fk_handle h, int id, const double* x_dev, int nx, double* y_dev) {
  need(id >= 0 && id < FK_MAX_SPLINES && h->spl.s[id].n > 0, "fk_spline_apply: table has not been set (fk_set_spline)");
  if (nx <= 0) return 0;
  h->errflag.ensure(1);
  FK_CUDA(cudaMemsetAsync(h->errflag.ptr, 0, sizeof(int), h->stream));
  splev_kernel<<<cdiv(nx, 256), 256, 0, h->stream>>>(h->spl.s[id], x_dev, nx, y_dev, h->errflag.ptr);
  FK_CUDA(cudaGetLastError());
  h->launches++;
  int err = 0;
  FK_CUDA(cudaMemcpyAsync(&err, h->errflag.ptr, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  FK_CUDA(cudaStreamSynchronize(h->stream));
  if (err) { set_error("fk_spline_apply: argument outside the knot range with bc = error"); return 2; }
  return 0;
}
int fk_spline_apply_dev(fk_handle h, int id, const double* x_dev, int nx, double* y_dev) {
  return guard([&]() {
    need(h && x_dev && y_dev, "fk: NULL argument");
    DeviceGuard dg(h->device);
    return spline_apply_dev(h, id, x_dev, nx, y_dev);
  });
}
int fk_spline_apply(fk_handle h, int id, const double* x, int nx, double* y_out) {
  return guard([&]() {
    need(h && x && y_out && nx >= 0, "fk: NULL argument");
    DeviceGuard dg(h->device);
    h->st_a.upload(x, (size_t)nx, h->stream);
    h->st_b.ensure((size_t)std::max(nx, 1));
    const int rc = spline_apply_dev(h, id, h->st_a.ptr, nx, h->st_b.ptr);
    copy_out(h, h->st_b.ptr, y_out, (size_t)nx);
    return rc;
  });
}

int fk_set_material(fk_handle h, const unsigned char* matid) {
  return guard([&]() {
    need(h && matid, "fk: NULL argument");
    for (int k = 0; k < h->nt; k++) need(matid[k] <= FK_MAX_SPLINES / 2, "fk_set_material: material id too large");
    DeviceGuard dg(h->device);
    ensure_state(h);
    FK_CUDA(cudaMemcpyAsync(h->matid.ptr, matid, (size_t)h->nt, cudaMemcpyHostToDevice, h->stream));
    std::vector<int> list;
    h->mat_used = 0;
    for (int k = 0; k < h->nt; k++) if (matid[k] > 0) { list.push_back(k); h->mat_used |= 1u << (matid[k] - 1); }
    h->n_mat = (int)list.size();
    if (h->n_mat > 0) h->mat_list.upload(list, h->stream);
    FK_CUDA(cudaStreamSynchronize(h->stream));
    h->have_material = true;
    return 0;
  });
}
int fk_set_inside(fk_handle h, const int* ids, int n) {
  return guard([&]() {
    need(h && (ids || n == 0), "fk: NULL argument");
    for (int i = 0; i < n; i++) need(ids[i] >= 0 && ids[i] < h->nt, "fk_set_inside: element id out of range");
    DeviceGuard dg(h->device);
    ensure_state(h);
    FK_CUDA(cudaMemsetAsync(h->matid.ptr, 0, (size_t)h->nt, h->stream));
    h->n_mat = n;
    h->mat_used = n > 0 ? 1u : 0u;
    if (n > 0) {
      h->mat_list.upload(ids, (size_t)n, h->stream);
      set_material_kernel<<<cdiv(n, 256), 256, 0, h->stream>>>(h->matid.ptr, h->mat_list.ptr, n);
      FK_CUDA(cudaGetLastError());
      h->launches++;
    }
    FK_CUDA(cudaStreamSynchronize(h->stream));
    h->have_material = true;
    return 0;
  });
}
int fk_set_mu(fk_handle h, const double* mu) {
  return guard([&]() {
    need(h && mu, "fk: NULL argument");
    DeviceGuard dg(h->device);
    ensure_state(h);
    FK_CUDA(cudaMemcpyAsync(h->mu.ptr, mu, sizeof(double) * (size_t)h->nt, cudaMemcpyHostToDevice, h->stream));
    h->s0_dirty = true;
    FK_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
  });
}
int fk_get_mu(fk_handle h, double* mu_out, double* dmu_out) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    ensure_state(h);
    if (mu_out) FK_CUDA(cudaMemcpyAsync(mu_out, h->mu.ptr, sizeof(double) * (size_t)h->nt, cudaMemcpyDeviceToHost, h->stream));
    if (dmu_out) FK_CUDA(cudaMemcpyAsync(dmu_out, h->dmu.ptr, sizeof(double) * (size_t)h->nt, cudaMemcpyDeviceToHost, h->stream));
    FK_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
  });
}
int fk_reset_field(fk_handle h) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    if (!h->mu.ptr) { ensure_state(h); return 0; }
    FK_CUDA(cudaMemsetAsync(h->dmu.ptr, 0, sizeof(double) * (size_t)h->nt, h->stream));
    FK_CUDA(cudaMemsetAsync(h->Hn.ptr, 0, sizeof(double) * (size_t)h->nt, h->stream));
    FK_CUDA(cudaMemsetAsync(h->Hfield.ptr, 0, sizeof(double) * (size_t)3 * h->nt, h->stream));
    return 0;
  });
}
int fk_get_field(fk_handle h, double* Hfield_out, double* Hnorm_out) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    ensure_state(h);
    if (Hfield_out) FK_CUDA(cudaMemcpyAsync(Hfield_out, h->Hfield.ptr, sizeof(double) * 3 * (size_t)h->nt, cudaMemcpyDeviceToHost, h->stream));
    if (Hnorm_out) FK_CUDA(cudaMemcpyAsync(Hnorm_out, h->Hn.ptr, sizeof(double) * (size_t)h->nt, cudaMemcpyDeviceToHost, h->stream));
    FK_CUDA(cudaStreamSynchronize(h->stream));
    return 0;
  });
}

// scalars of the last enqueued sweep -> host (one 128-byte copy + one synchronisation); status as fk_picard_step
static int picard_collect(fk_handle h, double* div_out) {
  if (!h->pin8) FK_CUDA(cudaMallocHost((void**)&h->pin8, 16 * sizeof(double)));
  FK_CUDA(cudaMemcpyAsync(h->pin8, h->scal.ptr, 8 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  FK_CUDA(cudaStreamSynchronize(h->stream));
  h->picard_pending = false;
  if (h->comm) comm_check(h);
  const double dmax = h->pin8[0];
  int err = 0;
  std::memcpy(&err, h->pin8 + 7, sizeof(int));
  if (div_out) *div_out = (4e-7 * 3.14159265358979323846) * dmax;   // mu0*max|H - Hold| (NewtonRaphson.jl:200)
  if (err & 1) { set_error("fk_picard_step: |H| outside the spline's knot range with bc = error"); return 2; }
  if (err & 2) { set_error("fk_picard_step: Nans/Infs in mu"); return 3; }   // NewtonRaphson.jl:210-212
  return 0;
}

// sync = false: only enqueue (the scalars stay on the device until picard_collect)
static int picard_step_dev(fk_handle h, int slot, const double* u_dev, int compat_bug, double* div_out, bool sync = true) {
  ensure_state(h);
  AsmArgs a = base_args(h, slot_vals(h, slot));
  h->dir_applied &= ~(1u << slot);
  a.u = u_dev;
  if (h->comm) {  // partitions: the sweep reads the potential on halo nodes too -> refresh them from their owners
    ensure_pcg(h);
    if (u_dev != h->u.ptr)
      FK_CUDA(cudaMemcpyAsync(h->u.ptr, u_dev, sizeof(double) * (size_t)h->n_rows, cudaMemcpyDeviceToDevice, h->stream));
    comm_halo_exchange(h, h->stream, h->u.ptr);
    a.u = h->u.ptr;
  }
  a.coef = h->mu.ptr;
  a.Hn = h->Hn.ptr;
  a.Hfield = h->Hfield.ptr;
  a.compat = compat_bug;
  int* errflag = reinterpret_cast<int*>(h->scal.ptr + 7);   // rides in the scalar block: one copy brings back both
  FK_CUDA(cudaMemsetAsync(errflag, 0, sizeof(double), h->stream));
  const bool have_mat = h->have_material && h->n_mat > 0;
  // A handle that sweeps repeatedly (a Picard loop) switches to the rescale path (rescale.cuh): same bits, streamed
  // instead of recomputed.
  h->geo_sweeps++;
  l2_release_for_assembly(h);
  const bool geo = h->geo_mode == 2 || (h->geo_mode == 1 && (h->geo_ready || h->geo_sweeps >= 2));
  if (geo) ensure_geo(h);
  const bool side = geo && h->geo_overlap;   // field recovery beside the streamed sums (needs the scales kept current)
  if (side && h->s0_dirty) {   // mu may have changed behind the handle's back (fk_set_mu, a tile sweep, the device pointer)
    geo_scale_kernel<<<cdiv(h->nt, 256), 256, 0, h->stream>>>(h->VE.ptr, h->mu.ptr, h->nt, h->geo_s0.ptr);
    h->launches++;
    h->s0_dirty = false;
  }
  if (!side) h->s0_dirty = true;
  // 1. material elements: mu(|H|), dmu/dH from the spline and the convergence metric (per-CTA maxima)
  int npart = 1;
  if (have_mat) {
    // every material id in use needs its table (a missing one would dereference NULL on the device)
    for (int m = 0; m < FK_MAX_SPLINES / 2; m++)
      need(!(h->mat_used & (1u << m)) || h->spl.s[m].n > 0, "fk_picard_step: a material id in use has no spline (fk_set_spline)");
    npart = std::min(cdiv(h->n_mat, 256), 4 * h->num_sms);
    h->tile_max.ensure((size_t)npart);
    material_update_kernel<<<npart, 256, 0, h->stream>>>(h->p4.ptr, h->t4.ptr, h->mat_list.ptr, h->n_mat,
                                                        h->matid.ptr, a.u, h->spl, compat_bug, h->Hn.ptr, h->mu.ptr, h->dmu.ptr,
                                                        h->tile_max.ptr, errflag, h->VE.ptr, side ? h->geo_s0.ptr : nullptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
  } else {
    FK_CUDA(cudaMemsetAsync(h->tile_max.ptr, 0, sizeof(double), h->stream));
  }
  // 2. the sweep: stiffness with the refreshed mu + H, |H| of every element
  bool folded = false;
  if (geo) {
    if (h->elem_grid == 0) {
      int occ = 1;
      FK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, picard_element_kernel<true>, kElemThreads, 0));
      h->elem_grid = h->num_sms * std::max(1, occ);
    }
    if (side) {
      // The two streaming kernels of the sweep are independent once the material pass has set the scales: the sums go first
      // on the main stream (their persistent grid takes its CTA slots evenly), field recovery runs beside them on a side
      // stream on the slots that are left, followed there by the fold of the convergence metric; the diagonal follows the
      // sums.  (Starting field recovery BEFORE the material pass was measured slower, 0.48 vs 0.40 ms: the persistent grid
      // of the sums then lands unevenly on the SMs.)
      if (!h->asm_stream) {
        FK_CUDA(cudaStreamCreateWithFlags(&h->asm_stream, cudaStreamNonBlocking));
        FK_CUDA(cudaEventCreateWithFlags(&h->asm_fork, cudaEventDisableTiming));
        FK_CUDA(cudaEventCreateWithFlags(&h->asm_join, cudaEventDisableTiming));
      }
      FK_CUDA(cudaEventRecord(h->asm_fork, h->stream));
      FK_CUDA(cudaStreamWaitEvent(h->asm_stream, h->asm_fork, 0));
      if (!(h->geo_dbg & 4)) launch_geo_rescale(h, a.vals, h->num_sms * h->geo_split_resc);
      const int ge = std::min(cdiv(h->nt, kElemThreads), h->num_sms * h->geo_split_elem);
      if (!(h->geo_dbg & 1))
        picard_element_kernel<false><<<ge, kElemThreads, 0, h->asm_stream>>>(h->p4.ptr, h->t4.ptr, a.u, nullptr, nullptr, h->nt, compat_bug,
                                                                             h->Hn.ptr, h->Hfield.ptr, nullptr);
      FK_CUDA(cudaGetLastError());
      h->launches++;
      comm_allreduce_partials(h, h->asm_stream, h->tile_max.ptr, npart, npart, 1, 1, h->scal.ptr, nullptr, 0);
      folded = true;
      FK_CUDA(cudaEventRecord(h->asm_join, h->asm_stream));
      FK_CUDA(cudaStreamWaitEvent(h->stream, h->asm_join, 0));
    } else {
      picard_element_kernel<true><<<std::min(cdiv(h->nt, kElemThreads), h->elem_grid), kElemThreads, 0, h->stream>>>(
          h->p4.ptr, h->t4.ptr, a.u, h->mu.ptr, a.VE, h->nt, compat_bug, h->Hn.ptr, h->Hfield.ptr, h->geo_s0.ptr);
      FK_CUDA(cudaGetLastError());
      h->launches++;
      launch_geo_rescale(h, a.vals);
    }
  } else {
    launch_assemble<ASM_PICARD>(h, a);
  }
  // div = mu0 * max over every rank's material elements (NewtonRaphson.jl:200)
  if (!folded) comm_allreduce_partials(h, h->stream, h->tile_max.ptr, npart, npart, 1, 1, h->scal.ptr, nullptr, 0);
  h->picard_pending = true;
  if (!sync) return 0;
  return picard_collect(h, div_out);
}
int fk_picard_step_dev(fk_handle h, int slot, const double* u_dev, int compat_bug, double* div_out) {
  return guard([&]() {
    need(h && u_dev, "fk: NULL argument");
    DeviceGuard dg(h->device);
    return picard_step_dev(h, slot, u_dev, compat_bug, div_out, div_out != nullptr);
  });
}
int fk_picard_result(fk_handle h, double* div_out) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    need(h->scal.ptr != nullptr, "fk_picard_result: no sweep has been run");
    return picard_collect(h, div_out);
  });
}
int fk_picard_step(fk_handle h, int slot, const double* u, int compat_bug, double* div_out) {
  return guard([&]() {
    need(h && u, "fk: NULL argument");
    DeviceGuard dg(h->device);
    ensure_pcg(h);
    FK_CUDA(cudaMemcpyAsync(h->u.ptr, u, sizeof(double) * (size_t)h->nv, cudaMemcpyHostToDevice, h->stream));
    return picard_step_dev(h, slot, h->u.ptr, compat_bug, div_out);
  });
}

// ---- constraints ------------------------------------------------------------------------------------
int fk_set_dirichlet(fk_handle h, const int* fixed_ids, int n_fixed, const double* gD) {
  return guard([&]() {
    need(h && (fixed_ids || n_fixed == 0) && n_fixed >= 0, "fk_set_dirichlet: bad argument");
    for (int i = 0; i < n_fixed; i++) need(fixed_ids[i] >= 0 && fixed_ids[i] < h->nv, "fk_set_dirichlet: DOF id out of range (ids are 0-based)");
    DeviceGuard dg(h->device);
    cudaStream_t s = h->stream;
    h->n_fixed = n_fixed;
    h->dir_mask.ensure((size_t)h->nv);
    h->dir_g.ensure((size_t)h->nv);
    FK_CUDA(cudaMemsetAsync(h->dir_mask.ptr, 0, (size_t)h->nv, s));
    FK_CUDA(cudaMemsetAsync(h->dir_g.ptr, 0, sizeof(double) * (size_t)h->nv, s));
    if (n_fixed > 0) {
      h->dir_ids.upload(fixed_ids, (size_t)n_fixed, s);
      if (gD) h->dir_val.upload(gD, (size_t)n_fixed, s);
      scatter_flag_kernel<<<cdiv(n_fixed, 256), 256, 0, s>>>(h->dir_mask.ptr, h->dir_ids.ptr, n_fixed);
      scatter_kernel<<<cdiv(n_fixed, 256), 256, 0, s>>>(h->dir_g.ptr, h->dir_ids.ptr, gD ? h->dir_val.ptr : nullptr, n_fixed);
      FK_CUDA(cudaGetLastError());
      h->launches += 2;
    }
    FK_CUDA(cudaStreamSynchronize(s));
    return 0;
  });
}
static void apply_dirichlet_dev(fk_handle h, int slot, double* b_dev) {
  check_slot(slot);
  need(h->vals[slot].ptr != nullptr, "fk_apply_dirichlet: matrix slot has not been assembled");
  need(h->dir_mask.ptr != nullptr, "fk_apply_dirichlet: call fk_set_dirichlet first");
  if (h->n_fixed == 0) return;
  cudaStream_t s = h->stream;
  if (b_dev) {   // lift with the unconstrained matrix: b[free] -= A[free,fixed] gD, b[fixed] = gD
    h->st_d.ensure((size_t)h->nv);
    FK_CUDA(cudaMemsetAsync(h->st_d.ptr, 0, sizeof(double) * (size_t)h->nv, s));
    launch_spmv(h, s, h->vals[slot].ptr, h->dir_g.ptr, h->st_d.ptr, false, nullptr, 0);
    dirichlet_lift_kernel<<<cdiv(h->n_rows, 256), 256, 0, s>>>(b_dev, h->st_d.ptr, h->dir_g.ptr, h->dir_mask.ptr, h->n_rows);
    h->launches += 2;
  }
  const int4* meta = reinterpret_cast<const int4*>(h->slice_meta.ptr);
  dirichlet_mask_kernel<<<cdiv(h->sym.nslices, 8), 256, 0, s>>>(meta, h->srow.ptr, h->sell_colz.ptr, h->sell_diag.ptr, h->dir_mask.ptr,
                                                               h->sym.nslices, h->n_rows, h->vals[slot].ptr);
  FK_CUDA(cudaGetLastError());
  h->launches++;
  h->dir_applied |= 1u << slot;   // fk_pcg on this slot starts from x0[fixed] = gD: the fixed rows then stay exact
}
int fk_apply_dirichlet_dev(fk_handle h, int slot, double* b_dev) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    apply_dirichlet_dev(h, slot, b_dev);
    return 0;
  });
}
int fk_apply_dirichlet(fk_handle h, int slot, double* b) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    if (b) h->st_c.upload(b, (size_t)h->nv, h->stream);
    apply_dirichlet_dev(h, slot, b ? h->st_c.ptr : nullptr);
    copy_out(h, h->st_c.ptr, b, (size_t)h->nv);
    return 0;
  });
}
int fk_set_nullspace(fk_handle h, const double* w) {
  return guard([&]() {
    need(h, "fk: NULL handle");
    DeviceGuard dg(h->device);
    h->ns_user = w != nullptr;
    if (w) {
      h->ns_w.upload(w, (size_t)h->nv, h->stream);
      FK_CUDA(cudaStreamSynchronize(h->stream));
    }
    return 0;
  });
}

// One nonlinear (Picard) iteration from host buffers: fk_picard_step(u) followed by fk_pcg(b -> x) -- the body of
// the reference loop (NewtonRaphson.jl:148-214, manyBodyIntersecting.jl:141-210).  Same kernels, same results;
// the right-hand side rides up on a second stream while the sweep runs, so the step pays for one upload, not two.
int fk_picard_pcg(fk_handle h, int slot, const double* u, int compat_bug, const double* b, double* x, const fk_pcg_opts* opts,
                  double* div_out, int* iters_out, double* relres_out) {
  return guard([&]() {
    need(h && u && b && x, "fk: NULL argument");
    DeviceGuard dg(h->device);
    ensure_pcg(h);
    const size_t n = (size_t)h->n_rows;
    if (!h->copy_stream) {
      FK_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
      FK_CUDA(cudaEventCreateWithFlags(&h->copy_ev, cudaEventDisableTiming));
    }
    h->st_d.ensure((size_t)h->nv);
    l2_release_for_assembly(h);   // before anything is in flight: the limit change waits for the device
    FK_CUDA(cudaMemcpyAsync(h->u.ptr, u, sizeof(double) * (size_t)h->nv, cudaMemcpyHostToDevice, h->stream));
    // bvec / st_d were last touched by work on h->stream: order the side copies after it
    FK_CUDA(cudaEventRecord(h->copy_ev, h->stream));
    FK_CUDA(cudaStreamWaitEvent(h->copy_stream, h->copy_ev, 0));
    FK_CUDA(cudaMemcpyAsync(h->bvec.ptr, b, sizeof(double) * n, cudaMemcpyHostToDevice, h->copy_stream));
    if (opts && opts->warm_start) FK_CUDA(cudaMemcpyAsync(h->st_d.ptr, x, sizeof(double) * n, cudaMemcpyHostToDevice, h->copy_stream));
    FK_CUDA(cudaEventRecord(h->copy_ev, h->copy_stream));
    // the sweep is only enqueued: its scalars (div, error flags) are parked in a second block so the PCG can reuse the
    // first, and come back with the solve's own synchronisation -- one host round trip less per nonlinear iteration
    picard_step_dev(h, slot, h->u.ptr, compat_bug, nullptr, false);
    FK_CUDA(cudaMemcpyAsync(h->scal.ptr + 12, h->scal.ptr, sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    FK_CUDA(cudaMemcpyAsync(h->scal.ptr + 13, h->scal.ptr + 7, sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    FK_CUDA(cudaStreamWaitEvent(h->stream, h->copy_ev, 0));
    pcg_dev(h, slot, h->bvec.ptr, h->st_d.ptr, opts, iters_out, relres_out);
    FK_CUDA(cudaMemcpyAsync(h->scal.ptr, h->scal.ptr + 12, sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    FK_CUDA(cudaMemcpyAsync(h->scal.ptr + 7, h->scal.ptr + 13, sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    FK_CUDA(cudaMemcpyAsync(x, h->st_d.ptr, sizeof(double) * n, cudaMemcpyDeviceToHost, h->stream));
    return picard_collect(h, div_out);
  });
}

// ---- implicit heat step (Heat/heat3D.jl:83-86) --------------------------------------------------------
// T <- (M + dt A)^-1 (M T): mass product (with the halo rows of T refreshed on a partition), then Jacobi-PCG on the
// slot that holds M + dt A (fk_matrix_axpby), warm-started from T.  One call per time step, no host round trip but the
// solve's own convergence polls.
static void heat_step_dev(fk_handle h, int slot_mass, int slot_lhs, double* T_dev, const fk_pcg_opts* opts, int* iters_out,
                          double* relres_out) {
  check_slot(slot_mass);
  check_slot(slot_lhs);
  need(h->vals[slot_mass].ptr && h->vals[slot_lhs].ptr, "fk_heat_step: matrix slot has not been assembled");
  ensure_pcg(h);
  cudaStream_t s = h->stream;
  if (T_dev != h->u.ptr) FK_CUDA(cudaMemcpyAsync(h->u.ptr, T_dev, sizeof(double) * (size_t)h->n_rows, cudaMemcpyDeviceToDevice, s));
  if (h->comm) comm_halo_exchange(h, s, h->u.ptr);
  launch_spmv(h, s, h->vals[slot_mass].ptr, h->u.ptr, h->bvec.ptr, false, nullptr, 0);
  h->launches++;
  fk_pcg_opts o{};
  if (opts) o = *opts;
  o.warm_start = 1;
  pcg_dev(h, slot_lhs, h->bvec.ptr, T_dev, &o, iters_out, relres_out);
}
int fk_heat_step_dev(fk_handle h, int slot_mass, int slot_lhs, double* T_dev, const fk_pcg_opts* opts, int* iters_out,
                     double* relres_out) {
  return guard([&]() {
    need(h && T_dev, "fk: NULL argument");
    DeviceGuard dg(h->device);
    heat_step_dev(h, slot_mass, slot_lhs, T_dev, opts, iters_out, relres_out);
    return 0;
  });
}
int fk_heat_step(fk_handle h, int slot_mass, int slot_lhs, double* T, const fk_pcg_opts* opts, int* iters_out, double* relres_out) {
  return guard([&]() {
    need(h && T, "fk: NULL argument");
    DeviceGuard dg(h->device);
    ensure_pcg(h);
    h->st_d.ensure((size_t)h->nv);
    FK_CUDA(cudaMemcpyAsync(h->st_d.ptr, T, sizeof(double) * (size_t)h->n_rows, cudaMemcpyHostToDevice, h->stream));
    heat_step_dev(h, slot_mass, slot_lhs, h->st_d.ptr, opts, iters_out, relres_out);
    copy_out(h, h->st_d.ptr, T, (size_t)h->n_rows);
    return 0;
  });
}

// ---- micromagnetics demagnetising field -----------------------------------------------
// LL::LL + LL::magnetostaticField of Femeko_jl_C++/Micromagnetics/LL.cpp: the constructor's work (node volumes
// :29-36, stiffnessMatrix(p,t,VE) :40) is fk_demag_setup; each fk_demag_field call is one magnetostaticField()
// (:287-348): load vector from M, Jacobi-PCG on the persistent matrix, element field, box method to the nodes.
int fk_demag_setup(fk_handle h, int slot, const int* inside_elements, int n_inside, double* volumes_out) {
  return guard([&]() {
    need(h && (inside_elements || n_inside == 0), "fk: NULL argument");
    need(n_inside >= 0, "fk_demag_setup: negative element count");
    DeviceGuard dg(h->device);
    cudaStream_t s = h->stream;
    const int nv = h->nv, nt = h->nt;
    // node -> (4 * list position + local index), ascending list position = the reference's scatter order
    std::vector<int> t((size_t)4 * nt);
    FK_CUDA(cudaMemcpyAsync(t.data(), h->t4.ptr, sizeof(int) * 4 * (size_t)nt, cudaMemcpyDeviceToHost, s));
    FK_CUDA(cudaStreamSynchronize(s));
    std::vector<int> ptr((size_t)nv + 1, 0);
    for (int ik = 0; ik < n_inside; ik++) {
      const int k = inside_elements[ik];
      need(k >= 0 && k < nt, "fk_demag_setup: element id out of range (ids are 0-based, permalloy.jl:70-72)");
      for (int i = 0; i < 4; i++) ptr[(size_t)t[4 * (size_t)k + i] + 1]++;
    }
    for (int n = 0; n < nv; n++) ptr[n + 1] += ptr[n];
    std::vector<int> lst((size_t)4 * n_inside), fill(ptr.begin(), ptr.end() - 1);
    for (int ik = 0; ik < n_inside; ik++) {
      const int k = inside_elements[ik];
      for (int i = 0; i < 4; i++) lst[fill[t[4 * (size_t)k + i]]++] = 4 * ik + i;
    }
    h->dm_inside.upload(inside_elements, (size_t)n_inside, s);
    h->dm_n2i_ptr.upload(ptr, s);
    h->dm_n2i.upload(lst, s);
    h->dm_vol.ensure((size_t)nv);
    h->dm_w.ensure((size_t)4 * std::max(n_inside, 1));
    h->dm_M.ensure((size_t)3 * nv);
    h->dm_Hd.ensure((size_t)3 * nv);
    h->dm_n_inside = n_inside;
    demag_node_gather_kernel<0><<<cdiv(std::max(nv, 1), 256), 256, 0, s>>>(h->dm_n2i_ptr.ptr, h->dm_n2i.ptr, nv, h->VE.ptr,
                                                                          h->dm_inside.ptr, nullptr, 0, h->dm_vol.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    assemble_stiffness_dev(h, slot, nullptr);   // A = stiffnessMatrix(p, t, VE), LL.cpp:40
    if (volumes_out) copy_out(h, h->dm_vol.ptr, volumes_out, (size_t)nv);
    else FK_CUDA(cudaStreamSynchronize(s));     // ptr / lst are read by the pending uploads
    return 0;
  });
}

static void demag_field_dev(fk_handle h, int slot, const double* M_dev, const fk_pcg_opts* opts, int nan_outside,
                            double* Hd_dev, double* u_dev, int* iters_out, double* relres_out) {
  need(h->dm_n_inside >= 0, "fk_demag_field: call fk_demag_setup first");
  ensure_pcg(h);
  cudaStream_t s = h->stream;
  const int nv = h->nv, ni = h->dm_n_inside;
  // load vector, LL.cpp:292-312
  if (ni) demag_elem_rhs_kernel<<<cdiv(ni, 256), 256, 0, s>>>(h->p4.ptr, h->t4.ptr, h->VE.ptr, h->dm_inside.ptr, ni, M_dev, h->dm_w.ptr);
  demag_node_gather_kernel<1><<<cdiv(nv, 256), 256, 0, s>>>(h->dm_n2i_ptr.ptr, h->dm_n2i.ptr, nv, h->dm_w.ptr, nullptr, nullptr, 0,
                                                            h->bvec.ptr);
  FK_CUDA(cudaGetLastError());
  h->launches += ni ? 2 : 1;
  // u = CG.solve(RHS), LL.cpp:315 (warm_start: from the potential of the previous call, kept in the handle)
  pcg_dev(h, slot, h->bvec.ptr, h->u.ptr, opts, iters_out, relres_out);
  // partitions: the element field of a ghost element reads the potential on halo nodes (every rank is in this call: fk_pcg's rule)
  if (h->comm) comm_halo_exchange(h, s, h->u.ptr);
  // element field + box method, LL.cpp:317-346
  if (ni) demag_elem_field_kernel<<<cdiv(ni, 256), 256, 0, s>>>(h->p4.ptr, h->t4.ptr, h->VE.ptr, h->dm_inside.ptr, ni, h->u.ptr, h->dm_w.ptr);
  demag_node_gather_kernel<2><<<cdiv(nv, 256), 256, 0, s>>>(h->dm_n2i_ptr.ptr, h->dm_n2i.ptr, nv, h->dm_w.ptr, nullptr, h->dm_vol.ptr,
                                                            nan_outside, Hd_dev);
  FK_CUDA(cudaGetLastError());
  h->launches += ni ? 2 : 1;
  if (u_dev) FK_CUDA(cudaMemcpyAsync(u_dev, h->u.ptr, sizeof(double) * (size_t)nv, cudaMemcpyDeviceToDevice, s));
}
int fk_demag_field_dev(fk_handle h, int slot, const double* M_dev, const fk_pcg_opts* opts, int nan_outside, double* Hd_dev,
                       double* u_dev, int* iters_out, double* relres_out) {
  return guard([&]() {
    need(h && M_dev && Hd_dev, "fk: NULL argument");
    DeviceGuard dg(h->device);
    demag_field_dev(h, slot, M_dev, opts, nan_outside, Hd_dev, u_dev, iters_out, relres_out);
    return 0;
  });
}
int fk_demag_field(fk_handle h, int slot, const double* M, const fk_pcg_opts* opts, int nan_outside, double* Hd_out,
                   double* u_out, int* iters_out, double* relres_out) {
  return guard([&]() {
    need(h && M && Hd_out, "fk: NULL argument");
    need(h->dm_n_inside >= 0, "fk_demag_field: call fk_demag_setup first");
    DeviceGuard dg(h->device);
    const size_t nv = (size_t)h->nv;
    FK_CUDA(cudaMemcpyAsync(h->dm_M.ptr, M, sizeof(double) * 3 * nv, cudaMemcpyHostToDevice, h->stream));
    demag_field_dev(h, slot, h->dm_M.ptr, opts, nan_outside, h->dm_Hd.ptr, nullptr, iters_out, relres_out);
    if (u_out) FK_CUDA(cudaMemcpyAsync(u_out, h->u.ptr, sizeof(double) * nv, cudaMemcpyDeviceToHost, h->stream));
    copy_out(h, h->dm_Hd.ptr, Hd_out, 3 * nv);
    return 0;
  });
}

// ---- legacy drop-in -----------------------------------------------------------------------------
// ABI-identical to the reference's scalarPotential (magnetostatics.cpp:49-95): BoundaryIntegral,
// stiffnessMatrix(p,t,VE,mu), Jacobi-PCG with Eigen's defaults on -RHS, result copied into u_in.
// It cannot return a status (void); failures are reported on stderr and u_in is left untouched.
void scalarPotential(double* u_in, double* p_in, int* t_in, int* surfaceT_in, double* normal_in, int nv, int nt, int ns,
                     double* VE, double* mu, double* Hext, int shell_id) {
  fk_handle h = nullptr;
  int rc = fk_create(p_in, t_in, nv, nt, VE, nullptr, &h);
  std::vector<double> rhs((size_t)std::max(nv, 1));
  if (!rc) rc = fk_boundary_integral(h, surfaceT_in, normal_in, nullptr, ns, Hext, &shell_id, 1, rhs.data());
  if (!rc) rc = fk_assemble_stiffness(h, 0, mu, nullptr);
  if (!rc) {
    for (int i = 0; i < nv; i++) rhs[i] = -rhs[i];
    int iters = 0;
    double relres = 0.0;
    // Eigen's defaults (tol = machine epsilon, 2n iterations, x0 = 0) like magnetostatics.cpp:26-38.  The system is
    // the singular pure-Neumann one: the residual's null-space component (rounding of sum(b) and of every A p) is
    // projected out so the recursion can actually reach epsilon; the iterate itself is not shifted.
    fk_pcg_opts po{};
    po.project_nullspace = 2;
    rc = fk_pcg(h, 0, rhs.data(), u_in, &po, &iters, &relres);
    if (!rc && !(relres <= DBL_EPSILON))
      std::fprintf(stderr, "femeko_cuda scalarPotential: PCG stopped after %d iterations at relative residual %.3e "
                           "(Eigen's default tolerance is machine epsilon; the reference throws here)\n", iters, relres);
  }
  if (rc) std::fprintf(stderr, "femeko_cuda scalarPotential failed: %s\n", fk_last_error());
  if (h) fk_destroy(h);
}

}  // extern "C"
