// Multi-GPU plumbing: one process per GPU, NCCL (loaded at run time with dlopen so single-GPU users
// need no NCCL) over NVLink.  Only two exchanges exist on the hot path: the halo rows of
// the search direction before each SpMV, and the dot-product all-reduce of the PCG.
#pragma once
#include <cuda_runtime.h>

struct fk_handle_s;

namespace fk {
struct PcgCtl;
struct RedArgs;
struct PeerRed;
PeerRed comm_peer_red(fk_handle_s* h);   // peer-memory descriptor of the consumer-side fold (kernels.cuh: peer_fold)
RedArgs comm_red_args(fk_handle_s* h, unsigned int* ticket, double* out, int op);
void comm_nccl_allreduce(fk_handle_s* h, cudaStream_t s, double* v, int n, int op);   // no-op unless on the NCCL path
int comm_kind(fk_handle_s* h);   // 0 single GPU, 1 NCCL, 2 peer memory
void launch_spmv(fk_handle_s* h, cudaStream_t s, const double* vals, const double* x, double* y, bool dot, const PcgCtl* ctl,
                 int j, bool publish_only = false);
void comm_destroy(fk_handle_s* h);
void comm_halo_exchange(fk_handle_s* h, cudaStream_t s, double* v);
// out[0..nvals) = reduction over ranks (op 0 sum / 1 max) of the folded per-CTA partials part[v*stride + 0..npart);
// also the single-GPU fold (no communicator).  One kernel on the peer-memory path, fold + ncclAllReduce otherwise.
void comm_allreduce_partials(fk_handle_s* h, cudaStream_t s, const double* part, int npart, int stride, int nvals, int op, double* out,
                             const PcgCtl* ctl, int j);
void comm_check(fk_handle_s* h);   // throws if a peer timed out
bool comm_capturable(fk_handle_s* h);   // peer-memory path: the whole iteration is kernels -> CUDA-graph capturable
void ensure_pcg(fk_handle_s* h);   // api.cu: allocates the PCG vector slab
void invalidate_graphs(fk_handle_s* h);   // api.cu: cached chunk graphs bake in peer pointers and launch shapes
struct HaloFused;
HaloFused comm_halo_fused(fk_handle_s* h, double* v, unsigned int* ticket);   // descriptor of the in-kernel exchange (spmv.cuh)
long long comm_global_rows(fk_handle_s* h);   // rows owned by all ranks together
}  // namespace fk
