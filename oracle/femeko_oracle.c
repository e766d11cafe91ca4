/* TEST INFRASTRUCTURE ONLY -- the CPU oracle.  Never imported, linked or executed by the product
 * path (femeko.jl_b200/); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may use it, and only as the checker.
 *
 * Plain-C restatement of the reference algorithm for the hot path (P1 tetrahedra, fp64).
 * Citations are relative to /root/reference.  Parity status: the reference ships no golden vectors
 * or tests, so this restatement is pinned against OUTPUTS OF THE REFERENCE ITSELF, i.e.
 * the unmodified src/FEMc.cpp + vendored Eigen compiled into oracle/_ref/ (tests/test_oracle.py,
 * fixtures under tests/golden/ made by tests/golden/make_golden.py).  The Julia-only pieces
 * (summation order of `A += sparse(...)`, Dierckx splines) cannot be executed here (no Julia):
 * for those, parity is pinned only through scipy's FITPACK wrapper and the C++ twin.
 *
 * Conventions: 0-based indices (run.jl:50-52), p is 3 x nv column-major (xyz contiguous per node),
 * t is 4 x nt column-major, all matrices CSR with ascending columns over the structural
 * node-adjacency pattern (== Eigen's setFromTriplets pattern, == CSC by symmetry).
 * Compiled with -ffp-contract=off so that the stated evaluation orders are what runs.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------- basis: abcd ------------- */
/* src/FEM3D.jl:5-17 / src/FEMc.cpp:18-42: solve [1 x y z] r = e1, rows ordered (nd, others in
 * element order).  Julia's `\` on a square dense matrix is LU with partial pivoting. */
void orc_abcd(const double* p, const int* nodes, int nd, double* r) {
  int ord[4], n = 1, i, j, k;
  double M[4][5];
  ord[0] = nd;
  for (i = 0; i < 4; i++) if (nodes[i] != nd && n < 4) ord[n++] = nodes[i];
  for (i = 0; i < 4; i++) {
    M[i][0] = 1.0;
    M[i][1] = p[3 * ord[i] + 0];
    M[i][2] = p[3 * ord[i] + 1];
    M[i][3] = p[3 * ord[i] + 2];
    M[i][4] = (i == 0) ? 1.0 : 0.0;
  }
  for (k = 0; k < 4; k++) {
    int piv = k;
    for (i = k + 1; i < 4; i++) if (fabs(M[i][k]) > fabs(M[piv][k])) piv = i;
    if (piv != k) for (j = 0; j < 5; j++) { double tmp = M[k][j]; M[k][j] = M[piv][j]; M[piv][j] = tmp; }
    for (i = k + 1; i < 4; i++) {
      double l = M[i][k] / M[k][k];
      for (j = k; j < 5; j++) M[i][j] -= l * M[k][j];
    }
  }
  for (i = 3; i >= 0; i--) {
    double s = M[i][4];
    for (j = i + 1; j < 4; j++) s -= M[i][j] * r[j];
    r[i] = s / M[i][i];
  }
}

/* ---------------------------------------------------------------- local matrices ---------- */
/* localStiffnessMatrix(mesh, f) src/FEM3D.jl:59-76: Ak[:,k] = vec(VE[k]*f[k]*(b b' + c c' + d d')).
 * Julia folds VE*f first, and the three outer products left to right.  f may be NULL (ones).
 * The C++ twin (FEMc.cpp:76-99) is the f == NULL case; it multiplies by mu[k] at triplet time
 * (FEMc.cpp:122) -- see orc_assemble_eigen. */
void orc_local_stiffness(const double* p, const int* t, const double* VE, const double* f,
                         int nt, double* Ak) {
  int k, i, j;
  for (k = 0; k < nt; k++) {
    double b[4], c[4], d[4], r[4];
    double s = f ? VE[k] * f[k] : VE[k];
    for (i = 0; i < 4; i++) {
      orc_abcd(p, t + 4 * k, t[4 * k + i], r);
      b[i] = r[1]; c[i] = r[2]; d[i] = r[3];
    }
    for (j = 0; j < 4; j++)
      for (i = 0; i < 4; i++)
        Ak[16 * (size_t)k + 4 * j + i] = s * ((b[i] * b[j] + c[i] * c[j]) + d[i] * d[j]);
  }
}

/* massMatrix src/FEM3D.jl:466-476: Mk[:,k] = VE[k] * (1/20*[2 1 1 1;...])[:]  (rho optional) */
void orc_local_mass(const double* VE, const double* rho, int nt, double* Mk) {
  int k, i, j;
  for (k = 0; k < nt; k++) {
    double v = rho ? VE[k] * rho[k] : VE[k];
    for (j = 0; j < 4; j++)
      for (i = 0; i < 4; i++)
        Mk[16 * (size_t)k + 4 * j + i] = v * ((1.0 / 20.0) * (i == j ? 2.0 : 1.0));
  }
}

/* tangentialStiffnessMatrix src/FEM3D.jl:79-105:
 * aux[i,j] = VE*dmu/|H| * (H.grad_i) * (H.grad_j), Hvec is 3 x nt. */
void orc_local_tangential(const double* p, const int* t, const double* VE, const double* Hvec,
                          const double* dmu, int nt, double* Ak) {
  int k, i, j;
  for (k = 0; k < nt; k++) {
    double b[4], c[4], d[4], r[4], g[4];
    const double* h = Hvec + 3 * (size_t)k;
    double H = sqrt(h[0] * h[0] + h[1] * h[1] + h[2] * h[2]);
    for (i = 0; i < 4; i++) {
      orc_abcd(p, t + 4 * k, t[4 * k + i], r);
      b[i] = r[1]; c[i] = r[2]; d[i] = r[3];
      g[i] = h[0] * b[i] + h[1] * c[i] + h[2] * d[i];
    }
    for (i = 0; i < 4; i++)
      for (j = i; j < 4; j++) {
        double v = VE[k] * dmu[k] / H * g[i] * g[j];
        Ak[16 * (size_t)k + 4 * j + i] = v;
        Ak[16 * (size_t)k + 4 * i + j] = v;
      }
  }
}

/* ---------------------------------------------------------------- symbolic ---------------- */
static int cmp_int(const void* a, const void* b) {
  int x = *(const int*)a, y = *(const int*)b;
  return (x > y) - (x < y);
}

/* Canonical pattern: row r holds every node sharing an element with r (incl. r),
 * ascending.  Returns nnz; colind may be NULL to size it (rowptr is always written). */
long orc_pattern(const int* t, int nv, int nt, int* rowptr, int* colind) {
  int* cnt = (int*)calloc((size_t)nv + 1, sizeof(int));
  int* ptr = (int*)malloc(((size_t)nv + 1) * sizeof(int));
  int* inc;
  int k, i, r;
  long nnz = 0;
  for (k = 0; k < nt; k++) for (i = 0; i < 4; i++) cnt[t[4 * k + i]]++;
  ptr[0] = 0;
  for (r = 0; r < nv; r++) ptr[r + 1] = ptr[r] + cnt[r];
  inc = (int*)malloc((size_t)ptr[nv] * sizeof(int) + 4);
  memset(cnt, 0, ((size_t)nv + 1) * sizeof(int));
  for (k = 0; k < nt; k++) for (i = 0; i < 4; i++) { r = t[4 * k + i]; inc[ptr[r] + cnt[r]++] = k; }
  rowptr[0] = 0;
  for (r = 0; r < nv; r++) {
    int m = 0, u = 0, q, deg = ptr[r + 1] - ptr[r];
    int* buf = (int*)malloc((size_t)(4 * deg + 1) * sizeof(int));
    for (q = ptr[r]; q < ptr[r + 1]; q++) for (i = 0; i < 4; i++) buf[m++] = t[4 * inc[q] + i];
    qsort(buf, (size_t)m, sizeof(int), cmp_int);
    for (q = 0; q < m; q++) if (q == 0 || buf[q] != buf[q - 1]) { if (colind) colind[nnz + u] = buf[q]; u++; }
    nnz += u;
    rowptr[r + 1] = (int)nnz;
    free(buf);
  }
  free(cnt); free(ptr); free(inc);
  return nnz;
}

static int find_col(const int* rowptr, const int* colind, int r, int c) {
  int lo = rowptr[r], hi = rowptr[r + 1] - 1;
  while (lo <= hi) {
    int mid = (lo + hi) >> 1;
    if (colind[mid] == c) return mid;
    if (colind[mid] < c) lo = mid + 1; else hi = mid - 1;
  }
  return -1;
}

/* element -> nnz map: e2nnz[16k + 4i + j] = CSR index of (t[i,k], t[j,k]) (the
 * reference's nearest analogue is the n-major triplet layout of NewtonRaphson.jl:115-127). */
void orc_elem2nnz(const int* t, int nt, const int* rowptr, const int* colind, int* e2nnz) {
  int k, i, j;
  for (k = 0; k < nt; k++)
    for (i = 0; i < 4; i++)
      for (j = 0; j < 4; j++)
        e2nnz[16 * (size_t)k + 4 * i + j] = find_col(rowptr, colind, t[4 * k + i], t[4 * k + j]);
}

/* ---------------------------------------------------------------- global assembly ---------- */
/* Julia order, src/FEM3D.jl:47-53: for n = (i outer, j inner): A += sparse(t[i,:], t[j,:], Ak[n,:]).
 * sparse() combines duplicates in input (k ascending) order; the 16 partial matrices are folded
 * left to right.  `scale` (optional, per element) reproduces NewtonRaphson.jl:153-161
 * (Acsc = Ak[n,k]*mu[k], one sparse() over the n-major triplet list) when n_major_single != 0:
 * then every nnz is one left-to-right sum over (n ascending, k ascending). */
void orc_assemble_julia(const int* e2nnz, int nt, long nnz, const double* Ak, const double* scale,
                        int n_major_single, double* vals) {
  double* part = (double*)malloc((size_t)nnz * sizeof(double));
  char* touched = (char*)malloc((size_t)nnz);
  int n, k;
  long q;
  for (q = 0; q < nnz; q++) vals[q] = 0.0;
  if (n_major_single) {
    for (n = 0; n < 16; n++)
      for (k = 0; k < nt; k++) {
        double v = Ak[16 * (size_t)k + n];
        if (scale) v = v * scale[k];
        vals[e2nnz[16 * (size_t)k + n]] += v;
      }
  } else {
    for (n = 0; n < 16; n++) {
      memset(touched, 0, (size_t)nnz);
      for (q = 0; q < nnz; q++) part[q] = 0.0;
      for (k = 0; k < nt; k++) {
        double v = Ak[16 * (size_t)k + n];
        long id = e2nnz[16 * (size_t)k + n];
        if (scale) v = v * scale[k];
        part[id] += v;
        touched[id] = 1;
      }
      for (q = 0; q < nnz; q++) if (touched[q]) vals[q] += part[q];
    }
  }
  free(part); free(touched);
}

/* Eigen order, src/FEMc.cpp:111-129 + SparseMatrix.h:1160-1168,1486-1500: triplets
 * (t(i,k), t(j,k), Ak(n,k)*mu[k]) in (k, i, j) order; duplicates summed left to right. */
void orc_assemble_eigen(const int* e2nnz, int nt, long nnz, const double* Ak, const double* mu,
                        double* vals) {
  int k, n;
  long q;
  for (q = 0; q < nnz; q++) vals[q] = 0.0;
  for (k = 0; k < nt; k++)
    for (n = 0; n < 16; n++) {
      double v = Ak[16 * (size_t)k + n];
      if (mu) v = v * mu[k];
      vals[e2nnz[16 * (size_t)k + n]] += v;
    }
}

/* ---------------------------------------------------------------- vectors ------------------ */
/* lagrange src/FEM3D.jl:274-281 / FEMc.cpp:57-73 */
void orc_lagrange(const int* t, const double* VE, int nv, int nt, double* C) {
  int k, i;
  for (i = 0; i < nv; i++) C[i] = 0.0;
  for (k = 0; k < nt; k++) for (i = 0; i < 4; i++) C[t[4 * k + i]] += VE[k] / 4.0;
}

/* BoundaryIntegral, Julia form src/FEM3D.jl:284-297: id list + stored AE */
void orc_boundary_integral_julia(const int* surfaceT, const double* normal, const double* AE, int ns,
                                 const double* F, const int* shell_ids, int nshell, int nv, double* RHS) {
  int s, i, m;
  for (i = 0; i < nv; i++) RHS[i] = 0.0;
  for (s = 0; s < ns; s++) {
    int in = 0;
    double dotv;
    for (m = 0; m < nshell; m++) if (surfaceT[4 * s + 3] == shell_ids[m]) in = 1;
    if (!in) continue;
    dotv = normal[3 * s] * F[0] + normal[3 * s + 1] * F[1] + normal[3 * s + 2] * F[2];
    for (i = 0; i < 3; i++) RHS[surfaceT[4 * s + i]] += dotv * AE[s] / 3;
  }
}

/* BoundaryIntegral, C++ form src/FEMc.cpp:170-200: single id, area recomputed by cross product */
void orc_boundary_integral_cpp(const double* p, const int* surfaceT, const double* normal, int ns,
                               const double* F, int shell_id, int nv, double* RHS) {
  int s, i;
  for (i = 0; i < nv; i++) RHS[i] = 0.0;
  for (s = 0; s < ns; s++) {
    const double *a, *b, *c;
    double ab[3], ac[3], cr[3], area;
    if (surfaceT[4 * s + 3] != shell_id) continue;
    a = p + 3 * surfaceT[4 * s]; b = p + 3 * surfaceT[4 * s + 1]; c = p + 3 * surfaceT[4 * s + 2];
    for (i = 0; i < 3; i++) { ab[i] = b[i] - a[i]; ac[i] = c[i] - a[i]; }
    cr[0] = ab[1] * ac[2] - ab[2] * ac[1];
    cr[1] = ab[2] * ac[0] - ab[0] * ac[2];
    cr[2] = ab[0] * ac[1] - ab[1] * ac[0];
    area = 0.5 * sqrt(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]);
    for (i = 0; i < 3; i++)
      RHS[surfaceT[4 * s + i]] += (normal[3 * s] * F[0] + normal[3 * s + 1] * F[1] + normal[3 * s + 2] * F[2]) * area / 3;
  }
}

/* ---------------------------------------------------------------- field recovery ----------- */
/* H_k = -sum_i u[t_i] grad(lambda_i); |H_k|.  Correct form: magneticField.jl:111-132,
 * manyBodyIntersecting.jl:156-174.  compat_bug != 0 reproduces NewtonRaphson.jl:187-189 /
 * run.jl:94-96 where the d-term is subtracted from row 2 and row 3 stays 0. */
void orc_recover_field(const double* p, const int* t, const double* u, int nt, int compat_bug,
                       double* Hfield, double* H) {
  int k, i;
  for (k = 0; k < nt; k++) {
    double h[3] = {0.0, 0.0, 0.0}, r[4];
    for (i = 0; i < 4; i++) {
      int nd = t[4 * k + i];
      orc_abcd(p, t + 4 * k, nd, r);
      h[0] -= u[nd] * r[1];
      h[1] -= u[nd] * r[2];
      if (compat_bug) h[1] -= u[nd] * r[3]; else h[2] -= u[nd] * r[3];
    }
    Hfield[3 * (size_t)k] = h[0]; Hfield[3 * (size_t)k + 1] = h[1]; Hfield[3 * (size_t)k + 2] = h[2];
    H[k] = sqrt(h[0] * h[0] + h[1] * h[1] + h[2] * h[2]);
  }
}

/* ---------------------------------------------------------------- material tables ---------- */
/* gradient(x, y) src/FEM.jl:66-92 */
void orc_gradient(const double* x, const double* y, int n, double* z) {
  int i;
  z[0] = (y[1] - y[0]) / (x[1] - x[0]);
  z[n - 1] = (y[n - 1] - y[n - 2]) / (x[n - 1] - x[n - 2]);
  for (i = 1; i < n - 1; i++) z[i] = (y[i + 1] - y[i - 1]) / (x[i + 1] - x[i - 1]);
}

/* materialPermeability src/magneticProperties.jl:157-188 */
void orc_material_permeability(const double* H, const double* B, int n, double* mu, double* dmu,
                               double* nu) {
  const double mu0 = 3.14159265358979323846 * 4e-7;
  int i;
  double m;
  char* bad = (char*)malloc((size_t)n);
  for (i = 0; i < n; i++) { mu[i] = B[i] / H[i]; bad[i] = !isfinite(mu[i]); }
  for (i = 0; i < n; i++) if (bad[i]) mu[i] = 0.0;                 /* :166-169 */
  m = -INFINITY;
  for (i = 0; i < n; i++) if (mu[i] > m) m = mu[i];
  for (i = 0; i < n; i++) if (bad[i]) mu[i] = m;                   /* :170 */
  orc_gradient(H, mu, n, dmu);                                     /* :173-175 */
  for (i = 0; i < n; i++) bad[i] = !isfinite(dmu[i]);
  for (i = 0; i < n; i++) if (bad[i]) dmu[i] = 0.0;                /* :178-180 */
  m = INFINITY;
  for (i = 0; i < n; i++) if (dmu[i] < m) m = dmu[i];
  for (i = 0; i < n; i++) if (bad[i]) dmu[i] = m;                  /* :181 */
  free(bad);
  for (i = 0; i < n; i++) { nu[i] = mu0 * (H[i] / B[i]); if (!isfinite(nu[i])) nu[i] = 0.0; }
}

/* ---------------------------------------------------------------- B-spline evaluation ------ */
/* FITPACK splev/fpbspl (the routine behind Dierckx.jl's Spline1D call, NewtonRaphson.jl:203-207).
 * t[0..n) knots, c[0..n-k-1) coefficients, degree k<=5.  e: 0 extrapolate, 1 zero, 2 error,
 * 3 nearest (Dierckx.jl bc strings).  Returns 0, or 1 when e == 2 and x is out of range. */
int orc_splev(const double* t, int n, const double* c, int k, double x, int e, double* out) {
  double tb = t[k], te = t[n - k - 1], h[6], hh[5];
  int l = k, i, j;
  if (x < tb || x > te) {
    if (e == 1) { *out = 0.0; return 0; }
    if (e == 2) { *out = NAN; return 1; }
    if (e == 3) x = (x < tb) ? tb : te;
  }
  while (!(x < t[l + 1]) && l != n - k - 2) l++;   /* t[l] <= x < t[l+1] */
  while (x < t[l] && l != k) l--;
  h[0] = 1.0;
  for (j = 1; j <= k; j++) {
    for (i = 0; i < j; i++) hh[i] = h[i];
    h[0] = 0.0;
    for (i = 0; i < j; i++) {
      int li = l + i + 1, lj = li - j;
      if (t[li] == t[lj]) { h[i + 1] = 0.0; continue; }
      {
        double f = hh[i] / (t[li] - t[lj]);
        h[i] = h[i] + f * (t[li] - x);
        h[i + 1] = f * (x - t[lj]);
      }
    }
  }
  {
    double sp = 0.0;
    for (j = 0; j <= k; j++) sp += c[l - k + j] * h[j];
    *out = sp;
  }
  return 0;
}

/* ---------------------------------------------------------------- SpMV + Jacobi-PCG -------- */
void orc_spmv(int n, const int* rowptr, const int* colind, const double* vals, const double* x, double* y) {
  int r, q;
  for (r = 0; r < n; r++) {
    double s = 0.0;
    for (q = rowptr[r]; q < rowptr[r + 1]; q++) s += vals[q] * x[colind[q]];
    y[r] = s;
  }
}

/* Eigen's conjugate_gradient with DiagonalPreconditioner, x0 = 0
 * (extern/eigen/Eigen/src/IterativeLinearSolvers/ConjugateGradient.h:29-85,
 *  BasicPreconditioners.h:62-75).  Returns 0 on convergence, 1 if max_iter reached. */
int orc_pcg(int n, const int* rowptr, const int* colind, const double* vals, const double* b,
            double* x, double tol, int max_iter, int* iters_out, double* err_out) {
  double *r = (double*)malloc(sizeof(double) * n), *p = (double*)malloc(sizeof(double) * n),
         *z = (double*)malloc(sizeof(double) * n), *tmp = (double*)malloc(sizeof(double) * n),
         *dinv = (double*)malloc(sizeof(double) * n);
  double rhs2 = 0.0, thr, r2 = 0.0, absNew = 0.0;
  int i, it = 0, q;
  for (i = 0; i < n; i++) {
    double d = 0.0; int found = 0;
    for (q = rowptr[i]; q < rowptr[i + 1]; q++) if (colind[q] == i) { d = vals[q]; found = 1; }
    dinv[i] = (found && d != 0.0) ? 1.0 / d : 1.0;
    x[i] = 0.0; r[i] = b[i];
    rhs2 += b[i] * b[i];
  }
  if (rhs2 == 0.0) { *iters_out = 0; *err_out = 0.0; goto done; }
  thr = tol * tol * rhs2;
  if (thr < 2.2250738585072014e-308) thr = 2.2250738585072014e-308;
  r2 = rhs2;
  if (r2 < thr) { *iters_out = 0; *err_out = sqrt(r2 / rhs2); goto done; }
  for (i = 0; i < n; i++) { p[i] = dinv[i] * r[i]; absNew += r[i] * p[i]; }
  while (it < max_iter) {
    double pAp = 0.0, alpha, absOld, beta;
    orc_spmv(n, rowptr, colind, vals, p, tmp);
    for (i = 0; i < n; i++) pAp += p[i] * tmp[i];
    alpha = absNew / pAp;
    r2 = 0.0;
    for (i = 0; i < n; i++) { x[i] += alpha * p[i]; r[i] -= alpha * tmp[i]; r2 += r[i] * r[i]; }
    if (r2 < thr) break;
    absOld = absNew; absNew = 0.0;
    for (i = 0; i < n; i++) { z[i] = dinv[i] * r[i]; absNew += r[i] * z[i]; }
    beta = absNew / absOld;
    for (i = 0; i < n; i++) p[i] = z[i] + beta * p[i];
    it++;
  }
  *iters_out = it; *err_out = sqrt(r2 / rhs2);
done:
  free(r); free(p); free(z); free(tmp); free(dinv);
  return (it >= max_iter) ? 1 : 0;
}

/* ---------------------------------------------------------------- micromagnetics demag ------ */
/* Femeko_jl_C++/Micromagnetics/LL.cpp.  Box-method node volumes, constructor :29-36:
   volumes[t(i,k)] += VE[k] over the magnetic elements, in list order. */
void orc_node_volumes(const int* t, const double* VE, const int* inside, int n_inside, int nv, double* vol) {
  for (int n = 0; n < nv; n++) vol[n] = 0.0;
  for (int ik = 0; ik < n_inside; ik++) {
    int k = inside[ik];
    for (int i = 0; i < 4; i++) vol[t[4 * k + i]] += VE[k];
  }
}

/* Load vector of magnetostaticField, LL.cpp:292-312: m_avg += M(:,nd)/4 over the element's nodes, then
   RHS(nd) += VE[k]*( b*mx + c*my + d*mz ) with (b,c,d) from abcd (linearBasis :139-160). M is 3 x nv. */
void orc_demag_rhs(const double* p, const int* t, const double* VE, const int* inside, int n_inside, int nv,
                   const double* M, double* rhs) {
  for (int n = 0; n < nv; n++) rhs[n] = 0.0;
  for (int ik = 0; ik < n_inside; ik++) {
    int k = inside[ik];
    double m[3] = {0.0, 0.0, 0.0};
    for (int i = 0; i < 4; i++) {
      int nd = t[4 * k + i];
      m[0] += M[3 * nd] / 4.0; m[1] += M[3 * nd + 1] / 4.0; m[2] += M[3 * nd + 2] / 4.0;
    }
    for (int i = 0; i < 4; i++) {
      int nd = t[4 * k + i];
      double r[4];
      orc_abcd(p, t + 4 * k, nd, r);
      rhs[nd] += VE[k] * (r[1] * m[0] + r[2] * m[1] + r[3] * m[2]);
    }
  }
}

/* Field on the nodes, LL.cpp:317-346: Hdk = -sum_i (b,c,d)_i u(t_i) per element, Hd(nd) += VE[k]*Hdk over the
   magnetic elements, Hd(nd) /= volumes(nd) for EVERY node (0/0 = NaN outside the magnetic region). Hd is 3 x nv. */
void orc_demag_box(const double* p, const int* t, const double* VE, const int* inside, int n_inside, int nv,
                   const double* u, const double* vol, double* Hd) {
  for (int n = 0; n < 3 * nv; n++) Hd[n] = 0.0;
  for (int ik = 0; ik < n_inside; ik++) {
    int k = inside[ik];
    double h[3] = {0.0, 0.0, 0.0};
    for (int i = 0; i < 4; i++) {
      int nd = t[4 * k + i];
      double r[4];
      orc_abcd(p, t + 4 * k, nd, r);
      h[0] -= r[1] * u[nd]; h[1] -= r[2] * u[nd]; h[2] -= r[3] * u[nd];
    }
    for (int i = 0; i < 4; i++) {
      int nd = t[4 * k + i];
      Hd[3 * nd] += VE[k] * h[0]; Hd[3 * nd + 1] += VE[k] * h[1]; Hd[3 * nd + 2] += VE[k] * h[2];
    }
  }
  for (int n = 0; n < nv; n++) { Hd[3 * n] /= vol[n]; Hd[3 * n + 1] /= vol[n]; Hd[3 * n + 2] /= vol[n]; }
}
