// Micromagnetics demagnetising field (included inside namespace fk by kernels.cuh).
// Reference: Femeko_jl_C++/Micromagnetics/LL.cpp -- box-method node volumes :29-36, load vector :292-312,
// element field :317-326, box-method mapping to the nodes :328-346; Micromagnetics/LL.jl:162-211.
// The reference scatters element contributions into node arrays in list order of the magnetic elements; here each
// node GATHERS its incident magnetic elements in that same order (node -> (list position, local index) lists built
// once by fk_demag_setup), so there are no atomics and the sums carry the reference's order.
#pragma once

// per magnetic element e = inside[ik]: w[4 ik + i] = VE * (grad(lambda_i) . <M>),  <M> = sum_i M(:, t_i) / 4
static __global__ void __launch_bounds__(256) demag_elem_rhs_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4,
                                                             const double* __restrict__ VE, const int* __restrict__ inside,
                                                             int n_inside, const double* __restrict__ M,
                                                             double* __restrict__ w) {
  const int ik = blockIdx.x * blockDim.x + threadIdx.x;
  if (ik >= n_inside) return;
  const int k = __ldg(inside + ik);
  const int4 nd = __ldg(t4 + k);
  TetGeom G;
  tet_geometry(p4, nd, G);
  const int n4[4] = {nd.x, nd.y, nd.z, nd.w};
  double mx = 0.0, my = 0.0, mz = 0.0;
#pragma unroll
  for (int i = 0; i < 4; i++) {  // LL.cpp:299-304
    mx += __ldg(M + 3 * (size_t)n4[i]) / 4.0;
    my += __ldg(M + 3 * (size_t)n4[i] + 1) / 4.0;
    mz += __ldg(M + 3 * (size_t)n4[i] + 2) / 4.0;
  }
  const double v = __ldg(VE + k);
#pragma unroll
  for (int i = 0; i < 4; i++) w[4 * (size_t)ik + i] = v * (G.g[i][0] * mx + G.g[i][1] * my + G.g[i][2] * mz);  // LL.cpp:309
}

// per magnetic element: w3[3 ik + c] = VE * Hd_k(c),  Hd_k = -sum_i grad(lambda_i) u(t_i)   (LL.cpp:317-337)
static __global__ void __launch_bounds__(256) demag_elem_field_kernel(const double4* __restrict__ p4, const int4* __restrict__ t4,
                                                               const double* __restrict__ VE, const int* __restrict__ inside,
                                                               int n_inside, const double* __restrict__ u,
                                                               double* __restrict__ w3) {
  const int ik = blockIdx.x * blockDim.x + threadIdx.x;
  if (ik >= n_inside) return;
  const int k = __ldg(inside + ik);
  const int4 nd = __ldg(t4 + k);
  TetGeom G;
  tet_geometry(p4, nd, G);
  const double u0 = __ldg(u + nd.x), u1 = __ldg(u + nd.y), u2 = __ldg(u + nd.z), u3 = __ldg(u + nd.w);
  const double hx = -(u0 * G.g[0][0]) - u1 * G.g[1][0] - u2 * G.g[2][0] - u3 * G.g[3][0];
  const double hy = -(u0 * G.g[0][1]) - u1 * G.g[1][1] - u2 * G.g[2][1] - u3 * G.g[3][1];
  const double hz = -(u0 * G.g[0][2]) - u1 * G.g[1][2] - u2 * G.g[2][2] - u3 * G.g[3][2];
  const double v = __ldg(VE + k);
  w3[3 * (size_t)ik] = v * hx;
  w3[3 * (size_t)ik + 1] = v * hy;
  w3[3 * (size_t)ik + 2] = v * hz;
}

// node gathers over n2i (entries = 4 * list position + local index, ascending list position)
// MODE 0: vol[n] = sum VE[inside[ik]]                    (LL.cpp:29-36)
// MODE 1: rhs[n] = sum w[4 ik + i]                       (LL.cpp:306-310)
// MODE 2: Hd[:, n] = (sum w3[3 ik + :]) / vol[n]         (LL.cpp:328-346; vol == 0: NaN like the reference if
//         nan_outside, else 0)
template <int MODE>
static __global__ void __launch_bounds__(256) demag_node_gather_kernel(const int* __restrict__ n2i_ptr, const int* __restrict__ n2i,
                                                                int nv, const double* __restrict__ src,
                                                                const int* __restrict__ inside, const double* __restrict__ vol,
                                                                int nan_outside, double* __restrict__ out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= nv) return;
  const int b = __ldg(n2i_ptr + n), e = __ldg(n2i_ptr + n + 1);
  if (MODE == 2) {
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (int q = b; q < e; q++) {
      const size_t ik = (size_t)(__ldg(n2i + q) >> 2);
      ax += __ldg(src + 3 * ik); ay += __ldg(src + 3 * ik + 1); az += __ldg(src + 3 * ik + 2);
    }
    const double v = __ldg(vol + n);
    if (v == 0.0 && !nan_outside) { ax = ay = az = 0.0; }
    else { ax /= v; ay /= v; az /= v; }
    out[3 * (size_t)n] = ax; out[3 * (size_t)n + 1] = ay; out[3 * (size_t)n + 2] = az;
  } else {
    double acc = 0.0;
    for (int q = b; q < e; q++) {
      const int idx = __ldg(n2i + q);
      acc += (MODE == 0) ? __ldg(src + __ldg(inside + (idx >> 2))) : __ldg(src + idx);
    }
    out[n] = acc;
  }
}
