// TEST INFRASTRUCTURE ONLY (oracle). Not part of the product path.
//
// extern "C" access to the UNMODIFIED reference micromagnetics class:
//   Femeko_jl_C++/Micromagnetics/LL.cpp  constructor :4-66 (box-method node volumes :29-36, stiffness :40,
//                                        CG.compute :44), magnetostaticField :287-348
// The TU includes LL.cpp where it lies (exactly like micromagnetics.cpp:1 does); nothing is copied.
// Output: oracle/_ref/libfemeko_ref_ll.so (git-ignored).
#include "Femeko_jl_C++/Micromagnetics/LL.cpp"
#include <cstring>

extern "C" {

// One demagnetising-field evaluation for the magnetisation M (3 x nv): Hd (3 x nv, box method) and the
// box-method node volumes.  Nodes outside the magnetic region get 0/0 = NaN in the reference (LL.cpp:341-345).
void ref_demag_field(double* p_ptr, int* t_ptr, double* VE, int* inside_ptr, int* inside_nodes_ptr, int nv, int nt,
                     int nInside, int nInsideNodes, double* M_ptr, double* Hd_out, double* volumes_out) {
  Eigen::Map<Eigen::MatrixXd> p(p_ptr, 3, nv);
  Eigen::Map<Eigen::MatrixXi> t(t_ptr, 4, nt);
  Eigen::Map<Eigen::VectorXi> InsideElements(inside_ptr, nInside);
  Eigen::Map<Eigen::VectorXi> InsideNodes(inside_nodes_ptr, nInsideNodes);
  Eigen::Map<Eigen::MatrixXd> M(M_ptr, 3, nv);
  LL solver(p, t, InsideElements, InsideNodes, VE, M);
  solver.magnetostaticField();
  std::memcpy(Hd_out, solver.Hd.data(), sizeof(double) * 3 * (size_t)nv);
  if (volumes_out) std::memcpy(volumes_out, solver.volumes.data(), sizeof(double) * (size_t)nv);
}

}  // extern "C"
