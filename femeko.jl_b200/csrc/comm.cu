// Multi-GPU PCG plumbing: one process per GPU.  NCCL is bound at run time (dlopen) so
// the single-GPU path has no NCCL dependency, and a process that already loaded torch's bundled
// libnccl.so.2 reuses it.  The reference has no counterpart (single process, single thread).
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>

#include "comm.hpp"
#include "handle.hpp"

namespace fk {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi& nccl() {
  static NcclApi api;
  if (api.lib) return api;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* nm : names) {
    api.lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  if (!api.lib) throw std::runtime_error("fk: cannot load libnccl.so.2 (needed only for multi-GPU runs)");
#define FK_SYM(field, name)                                                        \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.lib, name));          \
  if (!api.field) throw std::runtime_error(std::string("fk: NCCL symbol missing: ") + name)
  FK_SYM(GetUniqueId, "ncclGetUniqueId");
  FK_SYM(CommInitRank, "ncclCommInitRank");
  FK_SYM(CommDestroy, "ncclCommDestroy");
  FK_SYM(Send, "ncclSend");
  FK_SYM(Recv, "ncclRecv");
  FK_SYM(AllReduce, "ncclAllReduce");
  FK_SYM(GroupStart, "ncclGroupStart");
  FK_SYM(GroupEnd, "ncclGroupEnd");
  FK_SYM(GetErrorString, "ncclGetErrorString");
#undef FK_SYM
  return api;
}

#define FK_NCCL(call)                                                                                   \
  do {                                                                                                  \
    ncclResult_t r_ = (call);                                                                           \
    if (r_ != ncclSuccess)                                                                              \
      throw std::runtime_error(std::string("NCCL error: ") + nccl().GetErrorString(r_) + " (" #call ")"); \
  } while (0)

struct Comm {
  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1;
  // halo plan
  int nneigh = 0;
  std::vector<int> neigh_rank, send_ptr, recv_start, recv_count;
  DevBuf<int> send_idx;
  DevBuf<double> sendbuf;
  bool have_halo = false;
};

void comm_destroy(fk_handle_s* h) {
  if (!h->comm) return;
  if (h->comm->comm) nccl().CommDestroy(h->comm->comm);
  delete h->comm;
  h->comm = nullptr;
}

void comm_halo_exchange(fk_handle_s* h, cudaStream_t s, double* v) {
  Comm* c = h->comm;
  if (!c || !c->have_halo || c->nneigh == 0) return;
  const int nsend = c->send_ptr[c->nneigh];
  if (nsend > 0) {
    pack_kernel<<<(nsend + 255) / 256, 256, 0, s>>>(v, c->send_idx.ptr, nsend, c->sendbuf.ptr);
    FK_CUDA(cudaGetLastError());
  }
  FK_NCCL(nccl().GroupStart());
  for (int q = 0; q < c->nneigh; q++) {
    const int ns = c->send_ptr[q + 1] - c->send_ptr[q];
    if (ns > 0) FK_NCCL(nccl().Send(c->sendbuf.ptr + c->send_ptr[q], (size_t)ns, ncclDouble, c->neigh_rank[q], c->comm, s));
    if (c->recv_count[q] > 0)
      FK_NCCL(nccl().Recv(v + c->recv_start[q], (size_t)c->recv_count[q], ncclDouble, c->neigh_rank[q], c->comm, s));
  }
  FK_NCCL(nccl().GroupEnd());
}

void comm_allreduce_sum(fk_handle_s* h, cudaStream_t s, double* v, int n) {
  Comm* c = h->comm;
  if (!c || c->nranks == 1) return;
  FK_NCCL(nccl().AllReduce(v, v, (size_t)n, ncclDouble, ncclSum, c->comm, s));
}
void comm_allreduce_max(fk_handle_s* h, cudaStream_t s, double* v, int n) {
  Comm* c = h->comm;
  if (!c || c->nranks == 1) return;
  FK_NCCL(nccl().AllReduce(v, v, (size_t)n, ncclDouble, ncclMax, c->comm, s));
}

// One PCG iteration with the two exchanges: halo rows of p before the SpMV, and the all-reduce of
// p.Ap, then of (r.r, r.z).  The per-CTA partial sums are folded on the device first, so each
// all-reduce carries 1-2 doubles; the update/direction kernels then read the reduced scalars
// (npart == 1) instead of the partial arrays.
void comm_issue_iteration(fk_handle_s* h, cudaStream_t s, const double* vals, int j) {
  const int n = h->n_rows;
  PcgCtl* ctl = h->ctl.ptr;
  double* part = h->part.ptr;
  double* scal = h->scal.ptr;
  const int gs = h->grid_spmv, gv = h->grid_vec;
  comm_halo_exchange(h, s, h->p.ptr);
  launch_spmv(h, s, vals, h->p.ptr, h->Ap.ptr, true, ctl, j);
  fold_partials_kernel<<<1, 256, 0, s>>>(part, gs, gs, 1, scal);
  comm_allreduce_sum(h, s, scal, 1);
  pcg_update_kernel<<<gv, kVecThreads, 0, s>>>(h->x.ptr, h->r.ptr, h->p.ptr, h->Ap.ptr, h->dinv.ptr, n, scal, 1, part + gs, ctl, j);
  fold_partials_kernel<<<1, 256, 0, s>>>(part + gs, gv, gv, 2, scal + 1);
  comm_allreduce_sum(h, s, scal + 1, 2);
  pcg_direction_kernel<<<gv, kVecThreads, 0, s>>>(h->p.ptr, h->r.ptr, h->dinv.ptr, n, scal + 1, 1, 1, ctl, j);
  FK_CUDA(cudaGetLastError());
}

}  // namespace fk

using namespace fk;

extern "C" {

int fk_nccl_unique_id(unsigned char* id128) {
  try {
    if (!id128) throw std::runtime_error("fk: NULL argument");
    ncclUniqueId id;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is expected to be 128 bytes");
    FK_NCCL(nccl().GetUniqueId(&id));
    std::memcpy(id128, &id, 128);
    return 0;
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  }
}

int fk_comm_init(fk_handle h, const unsigned char* id128, int rank, int nranks) {
  try {
    if (!h || !id128) throw std::runtime_error("fk: NULL argument");
    if (nranks < 1 || rank < 0 || rank >= nranks) throw std::runtime_error("fk_comm_init: bad rank / nranks");
    FK_CUDA(cudaSetDevice(h->device));
    comm_destroy(h);
    h->comm = new Comm();
    h->comm->rank = rank;
    h->comm->nranks = nranks;
    ncclUniqueId id;
    std::memcpy(&id, id128, 128);
    FK_NCCL(nccl().CommInitRank(&h->comm->comm, nranks, id, rank));
    h->scal.ensure(8);
    return 0;
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  }
}

int fk_set_halo(fk_handle h, int n_owned, int nneigh, const int* neigh_rank, const int* send_ptr, const int* send_idx,
                const int* recv_start, const int* recv_count) {
  try {
    if (!h) throw std::runtime_error("fk: NULL handle");
    if (!h->comm) throw std::runtime_error("fk_set_halo: call fk_comm_init first");
    if (n_owned != h->n_rows) throw std::runtime_error("fk_set_halo: n_owned differs from the value given to fk_create");
    if (nneigh < 0 || (nneigh > 0 && !(neigh_rank && send_ptr && recv_start && recv_count)))
      throw std::runtime_error("fk_set_halo: NULL argument");
    FK_CUDA(cudaSetDevice(h->device));
    Comm* c = h->comm;
    c->nneigh = nneigh;
    c->neigh_rank.assign(neigh_rank, neigh_rank + nneigh);
    c->send_ptr.assign(1, 0);
    if (nneigh > 0) c->send_ptr.assign(send_ptr, send_ptr + nneigh + 1);
    c->recv_start.assign(recv_start, recv_start + nneigh);
    c->recv_count.assign(recv_count, recv_count + nneigh);
    const int nsend = c->send_ptr[nneigh];
    for (int q = 0; q < nneigh; q++) {
      if (neigh_rank[q] < 0 || neigh_rank[q] >= c->nranks || neigh_rank[q] == c->rank)
        throw std::runtime_error("fk_set_halo: bad neighbour rank");
      if (recv_count[q] < 0 || recv_start[q] < h->n_rows || recv_start[q] + recv_count[q] > h->nv)
        throw std::runtime_error("fk_set_halo: receive range outside the halo rows");
    }
    for (int i = 0; i < nsend; i++)
      if (send_idx[i] < 0 || send_idx[i] >= h->n_rows) throw std::runtime_error("fk_set_halo: send index is not an owned row");
    if (nsend > 0) c->send_idx.upload(send_idx, (size_t)nsend, h->stream);
    c->sendbuf.ensure((size_t)std::max(nsend, 1));
    FK_CUDA(cudaStreamSynchronize(h->stream));
    c->have_halo = true;
    return 0;
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  }
}

}  // extern "C"
