// K1: tile-gather assembly (included inside namespace fk by kernels.cuh).
//
// Persistent CTAs walk the tiles of the symbolic phase (symbolic.hpp).  Everything a tile needs is ONE
// packed blob in HBM, fetched by thread 0 with three bulk (TMA, cp.async.bulk) copies completing on
// mbarriers, ahead of the compute, so the warps never wait on a dependent chain of global gathers:
//   section I  sizes, global node ids, element ids (two tiles ahead: feeds the next tile's gathers)
//   section G  x|y|z of the tile's nodes, local connectivity (one tile ahead)               -> phase 1
//   section B  gidx | coff | contributor lists (one tile ahead)                             -> phase 2
// Phase 1: one thread per staged element computes the 10 unique entries of the symmetric 4x4 local
// matrix into shared memory (SoA: pair-major) from shared-memory coordinates.  The elements a tile owns
// come first in the staging order; in the fused Picard sweep only those also store H and |H|.
// Phase 2: one thread per CSR entry owned by the tile sums its contributors (staging index = pair *
// e_pad + slot) in ascending element order out of shared memory and writes the value once.  Entries are
// issued longest list first so a warp's 32 lists have near-equal length.  No atomics; bitwise reproducible.
// mu(|H|) of the material elements is refreshed by material_update_kernel (kernels.cuh) before the sweep,
// so this kernel carries no spline code.
// Reference math: src/FEM3D.jl:59-76 (stiffness), :466-476 (mass), :79-105 (tangential),
// :274-281 (lagrange), NewtonRaphson.jl:178-197 (Picard: H recovery + rescale by mu).

// ASM_STIFF2D / ASM_MASS2D: 2-D P1 triangles (src/FEM2D.jl:48-76,139-161) through the same plan.  A triangle (n0,n1,n2)
// enters the handle as the element (n0,n1,n2,n2): the pattern is the triangle's node adjacency, every local pair that
// involves the fourth slot carries an exact 0.0 (x + 0.0 == x, so the sums are those of the 3 x 3 local matrices).
enum AsmMode { ASM_STIFF = 0, ASM_MASS = 1, ASM_TANGENT = 2, ASM_PICARD = 3, ASM_LAGRANGE = 4, ASM_STIFF2D = 5, ASM_MASS2D = 6 };

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
  const int4* tile_hdr;   // {blob offset / 16, bytes of I, bytes of G, bytes of B}
  const uint4* tile_blob; // 16-byte units
  int ntiles;
  double* vals;
  int e_pad;              // staging stride (>= max tile elements)
  int nno_max;            // max nodes per tile (u staging)
  int blob_i_max;         // bytes reserved in shared memory for sections I / G / B (per buffer)
  int blob_g_max;
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
  if (MODE == ASM_STIFF2D || MODE == ASM_MASS2D) {
    // abc of src/FEM2D.jl:5-25 in closed form: f_i = a + b x + c y, b_i = (y_j - y_k) / det, c_i = (x_k - x_j) / det (i,j,k cyclic),
    // det = (x1-x0)(y2-y0) - (x2-x0)(y1-y0); area = |det| / 2 (mesh.VE of a 2-D mesh)
    const double x0 = xs[l0], y0 = ys[l0], x1 = xs[l1], y1 = ys[l1], x2 = xs[l2], y2 = ys[l2];
    const double det = (x1 - x0) * (y2 - y0) - (x2 - x0) * (y1 - y0);
    const double V = a.VE ? vpre : fabs(det) / 2.0;
#pragma unroll
    for (int q = 0; q < 10; q++) m[q] = 0.0;
    if (MODE == ASM_MASS2D) {   // VE/12 [2 1 1; 1 2 1; 1 1 2], src/FEM2D.jl:139-161
      const double v = a.coef ? V * fpre : V;
      const double d = v * ((1.0 / 12.0) * 2.0), o = v * ((1.0 / 12.0) * 1.0);
      m[0] = d; m[1] = o; m[2] = o; m[4] = d; m[5] = o; m[7] = d;
      return;
    }
    const double inv = 1.0 / det;
    const double b0 = (y1 - y2) * inv, b1 = (y2 - y0) * inv, b2 = (y0 - y1) * inv;
    const double c0 = (x2 - x1) * inv, c1 = (x0 - x2) * inv, c2 = (x1 - x0) * inv;
    const double s0 = V * fpre;   // VE[k]*mu[k]*(b b' + c c'), src/FEM2D.jl:62
    m[0] = __dmul_rn(s0, __fma_rn(c0, c0, __dmul_rn(b0, b0)));
    m[1] = __dmul_rn(s0, __fma_rn(c0, c1, __dmul_rn(b0, b1)));
    m[2] = __dmul_rn(s0, __fma_rn(c0, c2, __dmul_rn(b0, b2)));
    m[4] = __dmul_rn(s0, __fma_rn(c1, c1, __dmul_rn(b1, b1)));
    m[5] = __dmul_rn(s0, __fma_rn(c1, c2, __dmul_rn(b1, b2)));
    m[7] = __dmul_rn(s0, __fma_rn(c2, c2, __dmul_rn(b2, b2)));
    return;
  }
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
      a.Hn[k] = norm3(hx, hy, hz);
      if (a.Hfield) { a.Hfield[3 * (size_t)k] = hx; a.Hfield[3 * (size_t)k + 1] = hy; a.Hfield[3 * (size_t)k + 2] = hz; }
    }
  }
  const double s0 = V * f;  // VE[k]*f[k]*( b b' + c c' + d d' ), src/FEM3D.jl:71
  int q = 0;
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = i; j < 4; j++)
      m[q++] = __dmul_rn(s0, grad_dot3(G.g[i][0], G.g[i][1], G.g[i][2], G.g[j][0], G.g[j][1], G.g[j][2]));
}

// per-element coefficient gathers (the only global gathers left in phase 1): issued one tile ahead
template <int MODE>
__device__ __forceinline__ void fetch_coef(const AsmArgs& a, const int k, double& f, double& v) {
  f = (MODE != ASM_LAGRANGE && a.coef) ? __ldg(a.coef + k) : 1.0;
  v = a.VE ? __ldg(a.VE + k) : 0.0;
}

// ---- the kernel ----------------------------------------------------------------------------------
// Shared memory: [ stage 10 x e_pad f64 | us 2 x nno_max f64 | I x (depth + 1) | (G | B) x depth ].
constexpr int kAsmDepth = 2;  // ring depth of the geometry / gather-plan sections (ids: one more); deeper rings measured slower
__host__ __device__ inline size_t asm_smem_bytes(int e_pad, int nno_max, int blob_i_max, int blob_g_max, int blob_b_max) {
  return (size_t)80 * e_pad + 2 * (size_t)pad16(8 * nno_max) + (size_t)(kAsmDepth + 1) * blob_i_max +
         (size_t)kAsmDepth * ((size_t)blob_g_max + blob_b_max);
}

struct TileView {  // pointers into the shared-memory copies of sections I and G of one tile
  int ne, nno, nn;
  const int* gnode;
  const uint32_t* ek;
  const double *xs, *ys, *zs;
  const uint2* loc;
  __device__ __forceinline__ TileView(const unsigned char* sI, const unsigned char* sG) {
    const int4 h = *reinterpret_cast<const int4*>(sI);
    ne = h.x; nno = h.y; nn = h.z;
    const unsigned char* q = sI + 32;
    gnode = reinterpret_cast<const int*>(q); q += pad16(4 * nno);
    ek = reinterpret_cast<const uint32_t*>(q);
    q = sG;
    xs = reinterpret_cast<const double*>(q); q += pad16(8 * nno);
    ys = reinterpret_cast<const double*>(q); q += pad16(8 * nno);
    zs = reinterpret_cast<const double*>(q); q += pad16(8 * nno);
    loc = reinterpret_cast<const uint2*>(q);
  }
};

constexpr int kAsmPre = 2;  // per-thread register prefetch depth (elements / nodes of the NEXT tile)

// Pipeline (per persistent CTA, ring depth D = kAsmDepth): while tile i computes, the ids of tiles i+1 .. i+D and the
// geometry + gather plans of tiles i+1 .. i+D-1 are resident or in flight, so HBM always has this CTA's requests
// queued whatever phase the SM is in; the coefficient / potential gathers of tile i+1 are issued (from ids that
// landed at least a tile ago) before phase 2 of tile i.  Two CTA-wide barriers per tile: one between the phases,
// one at the end of the tile (which also publishes the next tile's potential values and releases the staging array
// and the ring slots of tile i).
template <int MODE, int THREADS, int MINB>
static __global__ void __launch_bounds__(THREADS, MINB) assemble_kernel(const __grid_constant__ AsmArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t barI[kAsmDepth + 1], barG[kAsmDepth], barB[kAsmDepth];
  const int tid = threadIdx.x;
  const int EP = a.e_pad;
  constexpr int D = kAsmDepth, DI = kAsmDepth + 1;
  double* stage = reinterpret_cast<double*>(smem_raw);
  double* us0 = stage + (size_t)10 * EP;
  const int us_stride = pad16(8 * a.nno_max) / 8;
  unsigned char* bufI = reinterpret_cast<unsigned char*>(us0 + 2 * (size_t)us_stride);
  unsigned char* bufG = bufI + (size_t)DI * a.blob_i_max;
  const size_t gbsz = (size_t)a.blob_g_max + a.blob_b_max;

  const int mine = (a.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // tiles of this CTA
  if (mine <= 0) return;
  auto tile_of = [&](int i) { return (int)blockIdx.x + i * (int)gridDim.x; };
  auto ids_of = [&](int i) { return bufI + (size_t)(i % DI) * a.blob_i_max; };
  auto gb_of = [&](int i) { return bufG + (size_t)(i % D) * gbsz; };
  auto issue_ids = [&](int i, const int2 hdr) {  // thread 0: ids of tile i -> ring slot i % (D + 1)
    mbar_expect_tx(&barI[i % DI], (uint32_t)hdr.y);
    bulk_g2s(ids_of(i), a.tile_blob + (size_t)hdr.x, (uint32_t)hdr.y, &barI[i % DI]);
  };
  auto issue_gb = [&](int i) {  // thread 0: geometry + gather plan of tile i -> ring slot i % D; sizes come from its ids
    mbar_wait(&barI[i % DI], (uint32_t)((i / DI) & 1));  // requested at least a tile ago
    const int4 h = *reinterpret_cast<const int4*>(ids_of(i) + 16);  // {G offset / 16, bytes I, bytes G, bytes B}
    unsigned char* dst = gb_of(i);
    const uint4* src = a.tile_blob + (size_t)h.x;
    mbar_expect_tx(&barG[i % D], (uint32_t)h.z);
    bulk_g2s(dst, src, (uint32_t)h.z, &barG[i % D]);
    mbar_expect_tx(&barB[i % D], (uint32_t)h.w);
    bulk_g2s(dst + a.blob_g_max, src + (h.z >> 4), (uint32_t)h.w, &barB[i % D]);
  };
  const int2* hdr2 = reinterpret_cast<const int2*>(a.tile_hdr);  // first half of each int4 header: {offset / 16, bytes of I}

  double fpre[kAsmPre], vpre[kAsmPre], upre[kAsmPre];
  auto prefetch = [&](const TileView& T) {  // coefficient / potential gathers of a tile whose ids have landed
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

  int2 hnext = make_int2(0, 0);  // thread 0: header of the next tile whose ids will be requested
  if (tid == 0) {
    for (int b = 0; b < DI; b++) mbar_init(&barI[b], 1);
    for (int b = 0; b < D; b++) { mbar_init(&barG[b], 1); mbar_init(&barB[b], 1); }
    mbar_init_fence();
    for (int j = 0; j < D && j < mine; j++) issue_ids(j, __ldg(hdr2 + 2 * (size_t)tile_of(j)));
    for (int j = 0; j < D - 1 && j < mine; j++) issue_gb(j);
    if (D < mine) hnext = __ldg(hdr2 + 2 * (size_t)tile_of(D));
  }
  __syncthreads();  // barrier init visible to every waiter
  mbar_wait(&barI[0], 0);
  prefetch(TileView(ids_of(0), bufG));
  if (MODE == ASM_PICARD) {
    stage_u(TileView(ids_of(0), bufG), us0);
    __syncthreads();
  }

  for (int i = 0; i < mine; i++) {
    if (tid == 0) {  // the ring slots written here were released by the barrier that ended tile i-1
      if (i + D - 1 < mine) issue_gb(i + D - 1);
      if (i + D < mine) {
        issue_ids(i + D, hnext);
        if (i + D + 1 < mine) hnext = __ldg(hdr2 + 2 * (size_t)tile_of(i + D + 1));
      }
    }
    mbar_wait(&barG[i % D], (uint32_t)((i / D) & 1));
    const TileView T(ids_of(i), gb_of(i));  // the ids of tile i were awaited before its prefetch
    const unsigned char* sB = gb_of(i) + a.blob_g_max;
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
    if (i + 1 < mine) {  // next tile's gathers fly during this tile's phase 2; its ids were requested >= a tile ago
      mbar_wait(&barI[(i + 1) % DI], (uint32_t)(((i + 1) / DI) & 1));
      prefetch(TileView(ids_of(i + 1), bufG));   // ids only: the geometry pointers are not used
    }
    __syncthreads();
    mbar_wait(&barB[i % D], (uint32_t)((i / D) & 1));

    // ---- phase 2 ------------------------------------------------------------------------------
    const int* gidx = reinterpret_cast<const int*>(sB);
    const uint16_t* coff = reinterpret_cast<const uint16_t*>(sB + pad16(4 * nn));
    const uint16_t* ents = reinterpret_cast<const uint16_t*>(sB + pad16(4 * nn) + pad16(2 * (nn + 1)));
    // two entries per thread at a time (independent load -> add chains); entries are sorted longest list first, so
    // the second entry's list is never the longer one.  ents: pair * e_pad + slot, in ascending element order.
    if (!(a.dbg & 2))
    for (int q = tid; q < nn; q += 2 * THREADS) {
      const int q2 = q + THREADS;
      const bool two = q2 < nn;
      const int b1 = coff[q], n1 = coff[q + 1] - b1;
      const int b2 = two ? coff[q2] : 0, n2 = two ? coff[q2 + 1] - b2 : 0;
      double acc1 = 0.0, acc2 = 0.0;
      int e = 0;
      for (; e < n2; e++) {
        acc1 += stage[ents[b1 + e]];
        acc2 += stage[ents[b2 + e]];
      }
      for (; e < n1; e++) acc1 += stage[ents[b1 + e]];
      a.vals[gidx[q]] = acc1;
      if (two) a.vals[gidx[q2]] = acc2;
    }
    if (MODE == ASM_PICARD && i + 1 < mine) stage_u(TileView(ids_of(i + 1), bufG), us0 + (size_t)((i + 1) & 1) * us_stride);
    __syncthreads();  // stage, this tile's ring slots and us[i & 1] may be overwritten from here on
  }
}
