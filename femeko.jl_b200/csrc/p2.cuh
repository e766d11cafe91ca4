// P2 (10-node) tetrahedra through the contribution plan (included inside namespace fk by kernels.cuh).
//
// Reference: quadraticStiffnessMatrix / quadraticLocalStiffnessMatrix (src/FEM3D.jl:324-395) and quadraticMassMatrix
// (src/FEM3D.jl:494-561, 15-point rule of GaussQuadrature3D(4), :711-745).  The reference fits every basis function by a
// 10 x 10 solve per node and element (quadraticBasis, :21-37); on straight-sided elements whose mid nodes sit on the edge
// midpoints those fits ARE the Lagrange P2 functions, so the kernels use them in closed form over the P1 gradients:
//   vertex v:       phi = l_v (2 l_v - 1),   grad phi = (4 l_v - 1) grad l_v
//   edge (a, b):    phi = 4 l_a l_b,         grad phi = 4 (l_a grad l_b + l_b grad l_a)
// Local node order = Gmsh's 10-node tetrahedron: vertices 0..3, then the mid nodes of the edges
// (0,1) (1,2) (0,2) (0,3) (2,3) (1,3).
// Assembly = the contribution plan of the rescale path (symbolic.hpp: RescalePlan, 55 local pairs): the per-slot factors
//   g_stiff = VE/10 sum_n grad phi_i(x_n) . grad phi_j(x_n)     (the reference's "10 node quadrature", :344-361)
//   g_mass  = VE sum_q w_q phi_i(x_q) phi_j(x_q)                (weights sum to 1/6 and multiply VE, as the reference's do)
// are computed once; every assembly is then geo_rescale_kernel with s = mu (or F): deterministic, no atomics, same
// summation order as the reference's triplets (ascending element).

constexpr int kP2Pairs = 55;
struct P2Ref {
  double m[kP2Pairs];   // sum_q w_q phi_i phi_j on the reference element, packed pair order (row a starts at 10 a - a (a - 1) / 2)
};

__device__ __forceinline__ void p2_edge(const int e, int& a, int& b) {   // mid node 4 + e sits on edge (a, b)
  a = (e == 1) ? 1 : (e == 4 ? 2 : (e == 5 ? 1 : 0));
  b = (e == 0) ? 1 : (e == 1 ? 2 : (e == 2 ? 2 : 3));
}
// barycentric coordinates of local node n
__device__ __forceinline__ void p2_node_lambda(const int n, double (&l)[4]) {
  l[0] = l[1] = l[2] = l[3] = 0.0;
  if (n < 4) { l[n] = 1.0; return; }
  int a, b;
  p2_edge(n - 4, a, b);
  l[a] = 0.5;
  l[b] = 0.5;
}
// gradient of basis function i at barycentric point l, from the P1 gradients gl[4][3]
__device__ __forceinline__ void p2_grad(const int i, const double (&l)[4], const double (&gl)[4][3], double (&g)[3]) {
  if (i < 4) {
    const double c = 4.0 * l[i] - 1.0;
    g[0] = c * gl[i][0]; g[1] = c * gl[i][1]; g[2] = c * gl[i][2];
  } else {
    int a, b;
    p2_edge(i - 4, a, b);
    g[0] = 4.0 * (l[a] * gl[b][0] + l[b] * gl[a][0]);
    g[1] = 4.0 * (l[a] * gl[b][1] + l[b] * gl[a][1]);
    g[2] = 4.0 * (l[a] * gl[b][2] + l[b] * gl[a][2]);
  }
}

// factors of every contribution slot (one-off); strips the pair bits of the slot's element reference
static __global__ void __launch_bounds__(128) p2_geo_build_kernel(const double4* __restrict__ p4, const int* __restrict__ t10,
                                                                  const double* __restrict__ VE, uint32_t* __restrict__ cref,
                                                                  long long nslots, int shift, const P2Ref ref,
                                                                  double* __restrict__ g_stiff, double* __restrict__ g_mass) {
  const uint32_t emask = (1u << shift) - 1u;
  for (long long s = blockIdx.x * (long long)blockDim.x + threadIdx.x; s < nslots; s += (long long)gridDim.x * blockDim.x) {
    const uint32_t c = cref[s];
    if (c == 0xffffffffu) { g_stiff[s] = 0.0; g_mass[s] = 0.0; continue; }
    const int k = (int)(c & emask), q = (int)(c >> shift);
    int pi = 0, rem = q;   // unpack the pair: row a holds 10 - a entries
    while (rem >= 10 - pi) { rem -= 10 - pi; pi++; }
    const int pj = pi + rem;
    const int* e = t10 + 10 * (size_t)k;
    TetGeom G;
    tet_geometry(p4, make_int4(__ldg(e), __ldg(e + 1), __ldg(e + 2), __ldg(e + 3)), G);
    const double V = VE[k];
    double aux = 0.0;
    for (int n = 0; n < 10; n++) {
      double l[4], gi[3], gj[3];
      p2_node_lambda(n, l);
      p2_grad(pi, l, G.g, gi);
      p2_grad(pj, l, G.g, gj);
      aux += gi[0] * gj[0] + gi[1] * gj[1] + gi[2] * gj[2];
    }
    g_stiff[s] = (aux / 10.0) * V;
    g_mass[s] = ref.m[q] * V;
    cref[s] = (uint32_t)k;
  }
}

// s = f (or 1 where f is absent)
static __global__ void __launch_bounds__(256) p2_scale_kernel(const double* __restrict__ f, int nt, double* __restrict__ s0) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nt) s0[k] = f ? f[k] : 1.0;
}

// quadraticConvectionMatrix (src/FEM3D.jl:563-648): C_k[i, j] = F VE sum_q w_q phi_i(x_q) (u_avg . grad phi_j(x_q)), u_avg = the mean
// of the velocity over the four vertices (:594).  With grad phi_j = sum_m (d phi_j / d lambda_m) grad lambda_m the quadrature
// collapses onto a constant tensor of the reference element,
//   C_k[i, j] = F VE sum_m (u_avg . grad lambda_m) T[i][j][m],     T[i][j][m] = sum_q w_q phi_i d phi_j / d lambda_m   (15-point rule)
// The matrix is NOT symmetric: a slot of the pair (a <= b) yields both C_k[a, b] and C_k[b, a]; g1 goes to the item's canonical
// entry -- the one with the lower row, i.e. the lower global node label --, g2 to its mirror (geo_rescale2_kernel).
// NOTE the reference assembles the TRANSPOSE of its local matrices (Clocal[:, k] = Ck[:] flattens column-major, the assembly
// loop n = 10 (i - 1) + j reads row-major, src/FEM3D.jl:632-645): global entry (t_i, t_j) += C_k[j, i].  `transposed` = 1
// reproduces that (fk_assemble_convection_p2: compat_bug), 0 gives entry (t_i, t_j) += C_k[i, j].
struct P2ConvRef {
  double T[10][10][4];
};
static __global__ void __launch_bounds__(128) p2_conv_slot_kernel(const double4* __restrict__ p4, const int* __restrict__ t10,
                                                                  const double* __restrict__ VE, const uint32_t* __restrict__ cref,
                                                                  long long nslots, int shift, const double* __restrict__ vel,
                                                                  const P2ConvRef* __restrict__ ref, int transposed,
                                                                  double* __restrict__ g1, double* __restrict__ g2) {
  const uint32_t emask = (1u << shift) - 1u;
  for (long long s = blockIdx.x * (long long)blockDim.x + threadIdx.x; s < nslots; s += (long long)gridDim.x * blockDim.x) {
    const uint32_t c = cref[s];
    if (c == 0xffffffffu) { g1[s] = 0.0; g2[s] = 0.0; continue; }
    const int k = (int)(c & emask), q = (int)(c >> shift);
    int a = 0, rem = q;
    while (rem >= 10 - a) { rem -= 10 - a; a++; }
    const int b = a + rem;
    const int* e = t10 + 10 * (size_t)k;
    const int n0 = __ldg(e), n1 = __ldg(e + 1), n2 = __ldg(e + 2), n3 = __ldg(e + 3);
    TetGeom G;
    tet_geometry(p4, make_int4(n0, n1, n2, n3), G);
    double ua[3];
#pragma unroll
    for (int d = 0; d < 3; d++)
      ua[d] = (vel[3 * (size_t)n0 + d] + vel[3 * (size_t)n1 + d] + vel[3 * (size_t)n2 + d] + vel[3 * (size_t)n3 + d]) / 4.0;
    const double V = VE[k];
    double cab = 0.0, cba = 0.0;
#pragma unroll
    for (int m = 0; m < 4; m++) {
      const double dm = ua[0] * G.g[m][0] + ua[1] * G.g[m][1] + ua[2] * G.g[m][2];
      cab += dm * ref->T[a][b][m];
      cba += dm * ref->T[b][a][m];
    }
    cab *= V;
    cba *= V;
    // row of the canonical entry = the lower global label.  transposed (the reference's assembly, src/FEM3D.jl:632-645:
    // Clocal[:, k] = Ck[:] is column-major while the loop over (i, j) reads it row-major): entry (t_i, t_j) receives Ck[j, i]
    const bool ab_is_canonical = (__ldg(e + a) <= __ldg(e + b)) != (transposed != 0);
    g1[s] = ab_is_canonical ? cab : cba;
    g2[s] = ab_is_canonical ? cba : cab;
  }
}
