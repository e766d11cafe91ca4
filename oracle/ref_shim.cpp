// TEST INFRASTRUCTURE ONLY (oracle). Not part of the product path.
//
// Thin extern "C" shim around the UNMODIFIED reference sources.  It is compiled
// *in place* against /root/reference (see oracle/Makefile, -I$(REF)), never copied:
//   src/FEMc.cpp            : abcd :18, lagrange :57, localStiffnessMatrix :76,
//                             stiffnessMatrix(p,t,VE,mu) :102, stiffnessMatrix(p,t,VE) :136,
//                             BoundaryIntegral :170
//   extern/eigen ConjugateGradient.h:29-85 (Jacobi/DiagonalPreconditioner default)
// The reference functions are not extern "C", so this TU includes FEMc.cpp (exactly like
// Femeko_jl_C++/Magnetostatics/magnetostatics.cpp:2 does) and re-exports the results as plain
// arrays.  Output: oracle/_ref/libfemeko_ref.so (git-ignored).
#include "src/FEMc.cpp"
#include <cstring>
#include <chrono>

static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

extern "C" {

// abcd for one (element, node) -> r[4]                         (FEMc.cpp:18-42)
void ref_abcd(double* p_in, int nv, const int* nodes4, int nd, double* r_out) {
  Eigen::Map<Eigen::MatrixXd> p(p_in, 3, nv);
  Eigen::Vector4i nodes(nodes4[0], nodes4[1], nodes4[2], nodes4[3]);
  Eigen::Vector4d r = abcd(p, nodes, nd);
  for (int i = 0; i < 4; i++) r_out[i] = r[i];
}

// dense 16 x nt local stiffness (col-major)                    (FEMc.cpp:76-99)
void ref_local_stiffness(double* p_in, int* t_in, int nv, int nt, double* VE, double* Ak_out) {
  Eigen::Map<Eigen::MatrixXd> p(p_in, 3, nv);
  Eigen::Map<Eigen::MatrixXi> t(t_in, 4, nt);
  Eigen::MatrixXd Ak = localStiffnessMatrix(p, t, VE);
  std::memcpy(Ak_out, Ak.data(), sizeof(double) * 16 * (size_t)nt);
}

// global stiffness; returns nnz. Call with outer==NULL to query nnz only.  (FEMc.cpp:102-167)
// mu may be NULL -> the no-permeability overload.  seconds_out (optional) = assembly wall time.
long ref_stiffness_csc(double* p_in, int* t_in, int nv, int nt, double* VE, double* mu,
                       int* outer, int* inner, double* values, double* seconds_out) {
  Eigen::Map<Eigen::MatrixXd> p(p_in, 3, nv);
  Eigen::Map<Eigen::MatrixXi> t(t_in, 4, nt);
  double t0 = now_s();
  Eigen::SparseMatrix<double> A = mu ? stiffnessMatrix(p, t, VE, mu) : stiffnessMatrix(p, t, VE);
  double t1 = now_s();
  if (seconds_out) *seconds_out = t1 - t0;
  A.makeCompressed();
  long nnz = A.nonZeros();
  if (outer) {
    std::memcpy(outer, A.outerIndexPtr(), sizeof(int) * (size_t)(nv + 1));
    std::memcpy(inner, A.innerIndexPtr(), sizeof(int) * (size_t)nnz);
    std::memcpy(values, A.valuePtr(), sizeof(double) * (size_t)nnz);
  }
  return nnz;
}

void ref_lagrange(int* t_in, double* VE, int nv, int nt, double* C_out) {        // FEMc.cpp:57-73
  Eigen::Map<Eigen::MatrixXi> t(t_in, 4, nt);
  Eigen::VectorXd C = lagrange(t, VE, nv, nt);
  std::memcpy(C_out, C.data(), sizeof(double) * (size_t)nv);
}

void ref_boundary_integral(double* p_in, int* surfaceT_in, double* normal_in, int nv, int ns,
                           double* F, int shell_id, double* RHS_out) {           // FEMc.cpp:170-200
  Eigen::Map<Eigen::MatrixXd> p(p_in, 3, nv);
  Eigen::Map<Eigen::MatrixXi> surfaceT(surfaceT_in, 4, ns);
  Eigen::Map<Eigen::MatrixXd> normal(normal_in, 3, ns);
  Eigen::VectorXd RHS = BoundaryIntegral(p, surfaceT, normal, F, shell_id);
  std::memcpy(RHS_out, RHS.data(), sizeof(double) * (size_t)nv);
}

// Eigen's Jacobi-PCG exactly as magnetostatics.cpp:26-38 instantiates it
// (ConjugateGradient<SparseMatrix<double>>: UpLo=Lower, DiagonalPreconditioner), on a CSC matrix
// handed in as arrays, with explicit tolerance / max iterations (<=0 keeps Eigen's defaults:
// tol = eps, maxit = 2n; IterativeSolverBase.h:251,368).  x0 = 0.  Returns Eigen's info().
int ref_cg(int n, const int* outer, const int* inner, const double* values, const double* b,
           double* x, double tol, int max_iter, int* iters_out, double* err_out,
           double* seconds_out) {
  Eigen::Map<const Eigen::SparseMatrix<double>> Am(n, n, outer[n], outer, inner, values);
  Eigen::SparseMatrix<double> A = Am;
  Eigen::Map<const Eigen::VectorXd> bv(b, n);
  Eigen::ConjugateGradient<Eigen::SparseMatrix<double>> solver;
  if (tol > 0) solver.setTolerance(tol);
  if (max_iter > 0) solver.setMaxIterations(max_iter);
  double t0 = now_s();
  solver.compute(A);
  Eigen::VectorXd xv = solver.solve(bv);
  double t1 = now_s();
  if (seconds_out) *seconds_out = t1 - t0;
  std::memcpy(x, xv.data(), sizeof(double) * (size_t)n);
  if (iters_out) *iters_out = (int)solver.iterations();
  if (err_out) *err_out = solver.error();
  return (int)solver.info();
}

}  // extern "C"
