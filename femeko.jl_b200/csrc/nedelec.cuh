// Nedelec (lowest-order edge) elements through the contribution plan (included inside namespace fk by kernels.cuh).
//
// Reference: nedelecLocalStiffness / nedelecTangentialStiffness (src/FEM3D.jl:153-271) and the vector-potential Newton loop
// of Magnetostatics/nedelec3D.jl:104-250 (load vector :108-146, global curl-curl matrix :163-178, B recovery :196-236).
// Degrees of freedom = the edges of the mesh (edge2localMap[t[5:10, :]], src/mesh.jl:170-185); the geometry comes from the
// four vertices.  Local edge ie joins the vertices NodesFromLocalEdge gives (src/mesh.jl:603-637):
//   (0,1) (1,2) (2,0) (3,0) (2,3) (1,3), each pair then ORIENTED from the lower to the higher global vertex label, and
//   curl N_ie = 2 L_ie grad(lambda_i) x grad(lambda_j)     (constant on the element)
// Assembly = the contribution plan of the rescale path (symbolic.hpp: RescalePlan, 21 local pairs): the curl-curl factors
//   g = VE curlN_a . curlN_b
// are computed once, every stiffness assembly is geo_rescale_kernel with s = nu; the tangential matrix
//   VE (curlN_a . B)(curlN_b . B) dnu / |B|      (src/FEM3D.jl:238-247; 0/0 = NaN where |B| = 0, as in the reference)
// changes with B, so its per-slot values are recomputed every call and summed by the same kernel with s = 1; the load
// vector is the diagonal of the same sum over the slots of the diagonal pairs (ascending element order = the reference's
// loop order).  Deterministic, no atomics.

struct NedCurls {
  double c[6][3];
};

__device__ __forceinline__ void ned_edge(const int ie, int& a, int& b) {
  a = (ie == 0) ? 0 : (ie == 1 ? 1 : (ie == 2 ? 2 : (ie == 3 ? 3 : (ie == 4 ? 2 : 1))));
  b = (ie == 0) ? 1 : (ie == 1 ? 2 : (ie == 2 ? 0 : (ie == 3 ? 0 : 3)));
}

// the six curls of one element
__device__ __forceinline__ void ned_curls(const double4* __restrict__ p4, const int4 nd, NedCurls& C) {
  double x[4], y[4], z[4];
  load_node(p4, nd.x, x[0], y[0], z[0]);
  load_node(p4, nd.y, x[1], y[1], z[1]);
  load_node(p4, nd.z, x[2], y[2], z[2]);
  load_node(p4, nd.w, x[3], y[3], z[3]);
  TetGeom G;
  tet_geometry_xyz(x[0], y[0], z[0], x[1], y[1], z[1], x[2], y[2], z[2], x[3], y[3], z[3], G);
  const int id[4] = {nd.x, nd.y, nd.z, nd.w};
#pragma unroll
  for (int ie = 0; ie < 6; ie++) {
    int a, b;
    ned_edge(ie, a, b);
    if (id[a] > id[b]) { const int q = a; a = b; b = q; }   // "the edge nodes must be sorted" (src/mesh.jl:630-634)
    double ex = 0, ey = 0, ez = 0, ga[3] = {0, 0, 0}, gb[3] = {0, 0, 0};
#pragma unroll
    for (int q = 0; q < 4; q++) {   // static indexing keeps everything in registers
      if (q == a) { ex -= x[q]; ey -= y[q]; ez -= z[q]; ga[0] = G.g[q][0]; ga[1] = G.g[q][1]; ga[2] = G.g[q][2]; }
      if (q == b) { ex += x[q]; ey += y[q]; ez += z[q]; gb[0] = G.g[q][0]; gb[1] = G.g[q][1]; gb[2] = G.g[q][2]; }
    }
    const double L2 = 2.0 * sqrt(ex * ex + ey * ey + ez * ez);
    C.c[ie][0] = L2 * (ga[1] * gb[2] - ga[2] * gb[1]);
    C.c[ie][1] = L2 * (ga[2] * gb[0] - ga[0] * gb[2]);
    C.c[ie][2] = L2 * (ga[0] * gb[1] - ga[1] * gb[0]);
  }
}
__device__ __forceinline__ void ned_pick(const NedCurls& C, const int e, double (&v)[3]) {
  v[0] = v[1] = v[2] = 0.0;
#pragma unroll
  for (int q = 0; q < 6; q++)
    if (q == e) { v[0] = C.c[q][0]; v[1] = C.c[q][1]; v[2] = C.c[q][2]; }
}
// packed pair q of a 6 x 6 symmetric local matrix -> (a <= b); row a holds 6 - a entries
__device__ __forceinline__ void ned_pair(const int q, int& a, int& b) {
  a = 0;
  int rem = q;
  while (rem >= 6 - a) { rem -= 6 - a; a++; }
  b = a + rem;
}

// nedelecLocalStiffness: Ak[6 ie + je, k] = VE[k] curlN_ie . curlN_je  (36 x nt, column-major)
static __global__ void __launch_bounds__(128) ned_local_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4,
                                                               const double* __restrict__ VE, int nt, double* __restrict__ Ak) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nt) return;
  NedCurls C;
  ned_curls(p4, __ldg(t4 + k), C);
  const double V = VE[k];
  double* out = Ak + 36 * (size_t)k;
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j < 6; j++) out[6 * i + j] = V * (C.c[i][0] * C.c[j][0] + C.c[i][1] * C.c[j][1] + C.c[i][2] * C.c[j][2]);
}

// per-slot values of one sum over the plan (cref keeps its pair bits: nothing is stripped here):
//   MODE 0  curl-curl factor        VE curlN_a . curlN_b
//   MODE 1  tangential entry        VE (curlN_a . B)(curlN_b . B) dnu / |B|
//   MODE 2  load-vector term        VE Bext . curlN_a on the slots of the diagonal pairs (a == b), 0 elsewhere
template <int MODE>
static __global__ void __launch_bounds__(128) ned_slot_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4,
                                                              const double* __restrict__ VE, const uint32_t* __restrict__ cref,
                                                              long long nslots, int shift, const double* __restrict__ dnu,
                                                              const double* __restrict__ Bfield, const double* __restrict__ Bn,
                                                              double bx, double by, double bz, double* __restrict__ g) {
  const uint32_t emask = (1u << shift) - 1u;
  for (long long s = blockIdx.x * (long long)blockDim.x + threadIdx.x; s < nslots; s += (long long)gridDim.x * blockDim.x) {
    const uint32_t c = cref[s];
    if (c == 0xffffffffu) { g[s] = 0.0; continue; }
    const int k = (int)(c & emask);
    int a, b;
    ned_pair((int)(c >> shift), a, b);
    if (MODE == 2 && a != b) { g[s] = 0.0; continue; }
    NedCurls C;
    ned_curls(p4, __ldg(t4 + k), C);
    double ca[3], cb[3];
    ned_pick(C, a, ca);
    ned_pick(C, b, cb);
    const double V = VE[k];
    if (MODE == 0) {
      g[s] = V * (ca[0] * cb[0] + ca[1] * cb[1] + ca[2] * cb[2]);
    } else if (MODE == 1) {
      const double B0 = Bfield[3 * (size_t)k], B1 = Bfield[3 * (size_t)k + 1], B2 = Bfield[3 * (size_t)k + 2];
      const double da = ca[0] * B0 + ca[1] * B1 + ca[2] * B2, db = cb[0] * B0 + cb[1] * B1 + cb[2] * B2;
      g[s] = V * da * db * dnu[k] / Bn[k];
    } else {
      g[s] = V * (bx * ca[0] + by * ca[1] + bz * ca[2]);
    }
  }
}

// strips the pair bits of the element references (what geo_rescale_kernel reads)
static __global__ void __launch_bounds__(256) ned_strip_kernel(const uint32_t* __restrict__ cref, long long nslots, int shift,
                                                               uint32_t* __restrict__ out) {
  const uint32_t emask = (1u << shift) - 1u;
  for (long long s = blockIdx.x * (long long)blockDim.x + threadIdx.x; s < nslots; s += (long long)gridDim.x * blockDim.x) {
    const uint32_t c = cref[s];
    out[s] = (c == 0xffffffffu) ? c : (c & emask);
  }
}

// B = curl of the vector potential: Bfield[:, k] = sum_i u[e_i] curlN_i (i ascending), |B|  (nedelec3D.jl:196-236)
static __global__ void __launch_bounds__(128) ned_recover_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4,
                                                                 const int* __restrict__ edges6, const double* __restrict__ u, int nt,
                                                                 double* __restrict__ Bfield, double* __restrict__ Bn) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nt) return;
  NedCurls C;
  ned_curls(p4, __ldg(t4 + k), C);
  double b0 = 0.0, b1 = 0.0, b2 = 0.0;
#pragma unroll
  for (int i = 0; i < 6; i++) {
    const double ue = __ldg(u + edges6[6 * (size_t)k + i]);
    b0 += ue * C.c[i][0];
    b1 += ue * C.c[i][1];
    b2 += ue * C.c[i][2];
  }
  if (Bfield) { Bfield[3 * (size_t)k] = b0; Bfield[3 * (size_t)k + 1] = b1; Bfield[3 * (size_t)k + 2] = b2; }
  if (Bn) Bn[k] = norm3(b0, b1, b2);
}
