// Multi-GPU plumbing: one process per GPU, NCCL (loaded at run time with dlopen so single-GPU users
// need no NCCL) over NVLink.  Only two exchanges exist on the hot path: the halo rows of
// the search direction before each SpMV, and the dot-product all-reduce of the PCG.
#pragma once
#include <cuda_runtime.h>

struct fk_handle_s;

namespace fk {
struct PcgCtl;
void launch_spmv(fk_handle_s* h, cudaStream_t s, const double* vals, const double* x, double* y, bool dot, const PcgCtl* ctl,
                 int j);
void comm_destroy(fk_handle_s* h);
void comm_halo_exchange(fk_handle_s* h, cudaStream_t s, double* v);
void comm_allreduce_sum(fk_handle_s* h, cudaStream_t s, double* v, int n);
void comm_allreduce_max(fk_handle_s* h, cudaStream_t s, double* v, int n);
void comm_issue_iteration(fk_handle_s* h, cudaStream_t s, const double* vals, int j);
}  // namespace fk
