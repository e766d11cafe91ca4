// K1: tile-gather assembly (included inside namespace fk by kernels.cuh).
//
// Persistent CTAs walk the tiles of the symbolic phase (symbolic.hpp).  Everything a tile needs is ONE
// packed blob in HBM -- node coordinates, local connectivity, element ids, and the gather plan -- that
// thread 0 fetches with two bulk (TMA, cp.async.bulk) copies completing on mbarriers, one tile AHEAD of
// the compute (double-buffered), so the warps never wait on a dependent chain of global gathers:
//   copy A  x|y|z of the tile's nodes, global node ids, element ids, local connectivity  -> phase 1
//   copy B  gidx | coff | contributor lists                                              -> phase 2
// Phase 1: one thread per staged element computes the 10 unique entries of the symmetric 4x4 local
// matrix into shared memory (SoA: pair-major) from shared-memory coordinates.
// Phase 2: one thread per CSR entry owned by the tile sums its contributors (slot, pair) in
// ascending element order out of shared memory and writes the value once.  Entries are issued
// longest list first so a warp's 32 lists have near-equal length.  No atomics; bitwise reproducible.
// Reference math: src/FEM3D.jl:59-76 (stiffness), :466-476 (mass), :79-105 (tangential),
// :274-281 (lagrange), NewtonRaphson.jl:178-207 (Picard: H recovery + spline + rescale).

enum AsmMode { ASM_STIFF = 0, ASM_MASS = 1, ASM_TANGENT = 2, ASM_PICARD = 3, ASM_LAGRANGE = 4 };

struct AsmArgs {
  const double* VE;       // nullable: volume from the determinant
  const double* coef;     // STIFF: f (nullable = 1) | MASS: rho (nullable) | TANGENT: dmu | PICARD: mu (already updated)
  const double* Hvec;     // TANGENT: 3 x nt
  // PICARD: the element's owner tile also stores H = -grad(u) and |H| (field recovery fused into the sweep)
  const double* u;
  double* Hn;
  double* Hfield;         // nullable, 3 x nt
  int compat;
  // plan
  const int4* tile_hdr;   // {blob offset / 16, ne | nno << 16, nn, nents}
  const uint4* tile_blob; // 16-byte units
  int ntiles;
  double* vals;
  int e_pad;              // staging stride (>= max tile elements)
  int nno_max;            // max nodes per tile (u staging)
  int blob_a_max;         // bytes reserved in shared memory for section A / B (per buffer)
  int blob_b_max;
  int dbg;                // profiling experiments only (fk_tune 3): bit 0 skips phase 1, bit 1 skips phase 2
};

// ---- mbarrier + bulk-copy primitives (PTX ISA: mbarrier, cp.async.bulk; sm_90+) -----------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!ok);
}
// global -> shared bulk copy; dst/src 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// ---- one element: the 10 unique entries of its symmetric local matrix --------------------------
// xs/ys/zs: coordinates of the tile's nodes, us: potential at the tile's nodes (PICARD), l: local ids
template <int MODE>
__device__ __forceinline__ void element_local(const AsmArgs& a, const int k, const bool owner, const double* __restrict__ xs,
                                              const double* __restrict__ ys, const double* __restrict__ zs,
                                              const double* __restrict__ us, const uint2 l, const double fpre, const double vpre,
                                              double (&m)[10]) {
  const int l0 = l.x & 0xffff, l1 = l.x >> 16, l2 = l.y & 0xffff, l3 = l.y >> 16;
  TetGeom G;
  if (MODE == ASM_MASS || MODE == ASM_LAGRANGE) {
    double v;
    if (a.VE) v = vpre;
    else {
      tet_geometry_xyz(xs[l0], ys[l0], zs[l0], xs[l1], ys[l1], zs[l1], xs[l2], ys[l2], zs[l2], xs[l3], ys[l3], zs[l3], G);
      v = G.vol;
    }
    if (MODE == ASM_MASS) {
      if (a.coef) v = v * fpre;
      const double d = v * ((1.0 / 20.0) * 2.0), o = v * ((1.0 / 20.0) * 1.0);
      m[0] = d; m[1] = o; m[2] = o; m[3] = o; m[4] = d; m[5] = o; m[6] = o; m[7] = d; m[8] = o; m[9] = d;
    } else {
      const double d = v / 4.0;
      m[0] = d; m[1] = 0; m[2] = 0; m[3] = 0; m[4] = d; m[5] = 0; m[6] = 0; m[7] = d; m[8] = 0; m[9] = d;
    }
    return;
  }
  tet_geometry_xyz(xs[l0], ys[l0], zs[l0], xs[l1], ys[l1], zs[l1], xs[l2], ys[l2], zs[l2], xs[l3], ys[l3], zs[l3], G);
  const double V = a.VE ? vpre : G.vol;
  if (MODE == ASM_TANGENT) {
    const double hx = __ldg(a.Hvec + 3 * (size_t)k), hy = __ldg(a.Hvec + 3 * (size_t)k + 1), hz = __ldg(a.Hvec + 3 * (size_t)k + 2);
    const double H = sqrt(hx * hx + hy * hy + hz * hz);
    const double s0 = V * fpre / H;
    double gh[4];
#pragma unroll
    for (int i = 0; i < 4; i++) gh[i] = hx * G.g[i][0] + hy * G.g[i][1] + hz * G.g[i][2];
    int q = 0;
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
      for (int j = i; j < 4; j++) m[q++] = s0 * gh[i] * gh[j];
    return;
  }
  const double f = fpre;  // 1.0 when the coefficient vector is absent (fetch_coef)
  if (MODE == ASM_PICARD) {
    // field recovery rides along: the tile that owns the element stores H and |H| (NewtonRaphson.jl:178-197);
    // the recomputed copies -- about half of the staged elements, grouped after the owned ones by the plan -- skip
    // it.  mu(|H|) of the material elements was refreshed by material_update_kernel just before this sweep.
    if (owner) {
      const double u0 = us[l0], u1 = us[l1], u2 = us[l2], u3 = us[l3];
      double hx = -(u0 * G.g[0][0]) - u1 * G.g[1][0] - u2 * G.g[2][0] - u3 * G.g[3][0];
      double hy = -(u0 * G.g[0][1]) - u1 * G.g[1][1] - u2 * G.g[2][1] - u3 * G.g[3][1];
      double hz = -(u0 * G.g[0][2]) - u1 * G.g[1][2] - u2 * G.g[2][2] - u3 * G.g[3][2];
      if (a.compat) { hy += hz; hz = 0.0; }  // NewtonRaphson.jl:187-189: d-term lands in row 2
      a.Hn[k] = sqrt(hx * hx + hy * hy + hz * hz);
      if (a.Hfield) { a.Hfield[3 * (size_t)k] = hx; a.Hfield[3 * (size_t)k + 1] = hy; a.Hfield[3 * (size_t)k + 2] = hz; }
    }
  }
  const double s0 = V * f;  // VE[k]*f[k]*( b b' + c c' + d d' ), src/FEM3D.jl:71
  int q = 0;
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = i; j < 4; j++)
      m[q++] = s0 * ((G.g[i][0] * G.g[j][0] + G.g[i][1] * G.g[j][1]) + G.g[i][2] * G.g[j][2]);
}

// per-element coefficient gathers (the only global gathers left in phase 1): issued one tile ahead
template <int MODE>
__device__ __forceinline__ void fetch_coef(const AsmArgs& a, const int k, double& f, double& v) {
  f = (MODE != ASM_LAGRANGE && a.coef) ? __ldg(a.coef + k) : 1.0;
  v = a.VE ? __ldg(a.VE + k) : 0.0;
}

// ---- the kernel ----------------------------------------------------------------------------------
// Shared memory: [ stage 10 x e_pad f64 | us 2 x nno_max f64 | (A | B) buffer 0 | (A | B) buffer 1 ].
__host__ __device__ inline size_t asm_smem_bytes(int e_pad, int nno_max, int blob_a_max, int blob_b_max) {
  return (size_t)80 * e_pad + 2 * (size_t)pad16(8 * nno_max) + 2 * ((size_t)blob_a_max + blob_b_max);
}

struct TileView {  // pointers into one shared-memory copy of section A
  int ne, nno, nn;
  const double *xs, *ys, *zs;
  const int* gnode;
  const uint32_t* ek;
  const uint2* loc;
  __device__ __forceinline__ explicit TileView(const unsigned char* sA) {
    const int4 h = *reinterpret_cast<const int4*>(sA);
    ne = h.x; nno = h.y; nn = h.z;
    const unsigned char* q = sA + 16;
    xs = reinterpret_cast<const double*>(q); q += pad16(8 * nno);
    ys = reinterpret_cast<const double*>(q); q += pad16(8 * nno);
    zs = reinterpret_cast<const double*>(q); q += pad16(8 * nno);
    gnode = reinterpret_cast<const int*>(q); q += pad16(4 * nno);
    ek = reinterpret_cast<const uint32_t*>(q); q += pad16(4 * ne);
    loc = reinterpret_cast<const uint2*>(q);
  }
};

constexpr int kAsmPre = 2;  // per-thread register prefetch depth (elements / nodes of the NEXT tile)

// Two CTA-wide barriers per tile: one between the phases, one at the end of the tile (which also publishes the
// next tile's potential values and releases the staging array and the blob buffer).
template <int MODE, int THREADS, int MINB>
static __global__ void __launch_bounds__(THREADS, MINB) assemble_kernel(const __grid_constant__ AsmArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t barA[2], barB[2];
  const int tid = threadIdx.x;
  const int EP = a.e_pad;
  double* stage = reinterpret_cast<double*>(smem_raw);
  double* us0 = stage + (size_t)10 * EP;
  const int us_stride = pad16(8 * a.nno_max) / 8;
  unsigned char* buf0 = reinterpret_cast<unsigned char*>(us0 + 2 * (size_t)us_stride);
  const size_t bufsz = (size_t)a.blob_a_max + a.blob_b_max;

  const int mine = (a.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // tiles of this CTA
  if (mine <= 0) return;
  auto tile_of = [&](int i) { return (int)blockIdx.x + i * (int)gridDim.x; };
  auto issue = [&](int i, const int4 hdr) {  // thread 0: fetch tile i's blob into buffer i & 1
    unsigned char* dst = buf0 + (size_t)(i & 1) * bufsz;
    const uint4* src = a.tile_blob + (size_t)hdr.x;
    mbar_expect_tx(&barA[i & 1], (uint32_t)hdr.y);
    bulk_g2s(dst, src, (uint32_t)hdr.y, &barA[i & 1]);
    mbar_expect_tx(&barB[i & 1], (uint32_t)hdr.z);
    bulk_g2s(dst + a.blob_a_max, src + (hdr.y >> 4), (uint32_t)hdr.z, &barB[i & 1]);
  };

  double fpre[kAsmPre], vpre[kAsmPre], upre[kAsmPre];
  auto prefetch = [&](const TileView& T) {  // coefficient / potential gathers of a tile whose section A has landed
#pragma unroll
    for (int j = 0; j < kAsmPre; j++) {
      const int s = tid + j * THREADS;
      fpre[j] = 1.0; vpre[j] = 0.0; upre[j] = 0.0;
      if (s < T.ne) fetch_coef<MODE>(a, (int)(T.ek[s] & 0x7fffffffu), fpre[j], vpre[j]);
      if (MODE == ASM_PICARD && s < T.nno) upre[j] = __ldg(a.u + T.gnode[s]);
    }
  };
  auto stage_u = [&](const TileView& T, double* us) {  // potential at a tile's nodes: one gather per node
#pragma unroll
    for (int j = 0; j < kAsmPre; j++)
      if (tid + j * THREADS < T.nno) us[tid + j * THREADS] = upre[j];
    for (int n = tid + kAsmPre * THREADS; n < T.nno; n += THREADS) us[n] = __ldg(a.u + T.gnode[n]);
  };

  int4 hnext = make_int4(0, 0, 0, 0);  // thread 0: header of tile i + 1, loaded one tile ahead
  if (tid == 0) {
    mbar_init(&barA[0], 1); mbar_init(&barA[1], 1);
    mbar_init(&barB[0], 1); mbar_init(&barB[1], 1);
    mbar_init_fence();
    issue(0, __ldg(a.tile_hdr + tile_of(0)));
    if (mine > 1) hnext = __ldg(a.tile_hdr + tile_of(1));
  }
  __syncthreads();  // barrier init visible to every waiter
  mbar_wait(&barA[0], 0);
  prefetch(TileView(buf0));
  if (MODE == ASM_PICARD) {
    stage_u(TileView(buf0), us0);
    __syncthreads();
  }

  for (int i = 0; i < mine; i++) {
    if (tid == 0 && i + 1 < mine) {  // buffer (i+1)&1 was released by the barrier that ended tile i-1
      issue(i + 1, hnext);
      if (i + 2 < mine) hnext = __ldg(a.tile_hdr + tile_of(i + 2));
    }
    const unsigned char* sA = buf0 + (size_t)(i & 1) * bufsz;
    const unsigned char* sB = sA + a.blob_a_max;
    const TileView T(sA);  // section A of tile i was awaited before its prefetch
    const int ne = T.ne, nn = T.nn;
    const double* us = us0 + (size_t)(i & 1) * us_stride;

    // ---- phase 1 ------------------------------------------------------------------------------
    if (!(a.dbg & 1)) {
#pragma unroll
    for (int j = 0; j < kAsmPre; j++) {
      const int s = tid + j * THREADS;
      if (s < ne) {
        const uint32_t raw = T.ek[s];
        double m[10];
        element_local<MODE>(a, (int)(raw & 0x7fffffffu), (raw >> 31) != 0, T.xs, T.ys, T.zs, us, T.loc[s], fpre[j], vpre[j], m);
#pragma unroll
        for (int q = 0; q < 10; q++) stage[q * EP + s] = m[q];
      }
    }
    for (int s = tid + kAsmPre * THREADS; s < ne; s += THREADS) {  // tiles larger than the prefetch depth
      const uint32_t raw = T.ek[s];
      double f, v, m[10];
      fetch_coef<MODE>(a, (int)(raw & 0x7fffffffu), f, v);
      element_local<MODE>(a, (int)(raw & 0x7fffffffu), (raw >> 31) != 0, T.xs, T.ys, T.zs, us, T.loc[s], f, v, m);
#pragma unroll
      for (int q = 0; q < 10; q++) stage[q * EP + s] = m[q];
    }
    }
    if (i + 1 < mine) {  // next tile's gathers fly during this tile's phase 2
      mbar_wait(&barA[(i + 1) & 1], (uint32_t)(((i + 1) >> 1) & 1));
      prefetch(TileView(buf0 + (size_t)((i + 1) & 1) * bufsz));
    }
    __syncthreads();
    mbar_wait(&barB[i & 1], (uint32_t)((i >> 1) & 1));

    // ---- phase 2 ------------------------------------------------------------------------------
    const int* gidx = reinterpret_cast<const int*>(sB);
    const uint16_t* coff = reinterpret_cast<const uint16_t*>(sB + pad16(4 * nn));
    const uint16_t* ents = reinterpret_cast<const uint16_t*>(sB + pad16(4 * nn) + pad16(2 * (nn + 1)));
    if (!(a.dbg & 2))
    for (int q = tid; q < nn; q += THREADS) {
      const int beg = coff[q], end = coff[q + 1];
      double acc = 0.0;
      for (int e = beg; e < end; e++) acc += stage[ents[e]];  // ents: pair * e_pad + slot, ascending element order
      a.vals[gidx[q]] = acc;
    }
    if (MODE == ASM_PICARD && i + 1 < mine)
      stage_u(TileView(buf0 + (size_t)((i + 1) & 1) * bufsz), us0 + (size_t)((i + 1) & 1) * us_stride);
    __syncthreads();  // stage, this tile's blob buffer and us[i & 1] may be overwritten from here on
  }
}
