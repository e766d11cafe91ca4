// Multi-GPU PCG plumbing: one process per GPU.  NCCL is bound at run time (dlopen) so
// the single-GPU path has no NCCL dependency, and a process that already loaded torch's bundled
// libnccl.so.2 reuses it.  The reference has no counterpart (single process, single thread).
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
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

constexpr int kPushCtas = 16;     // CTAs per neighbour of the halo push

// Peer-memory state of the fused path.  Every rank exports (CUDA IPC) its PCG vector slab and a small
// control block; peers write halo rows straight into the owner's vector and reduction operands straight
// into the owner's mailbox over NVLink, each followed by a system-scope release of a sequence flag.
struct P2P {
  bool on = false;
  void* peer_slab[kMaxRanks] = {};                 // peer r's pcg_slab (own rank: local pointer)
  unsigned long long* peer_ctrl[kMaxRanks] = {};   // peer r's control block
  size_t peer_stride[kMaxRanks] = {};              // doubles between consecutive vectors of peer r's slab
  bool opened_slab[kMaxRanks] = {};
  int remote_start[kMaxNeigh] = {};                // where my rows land in neighbour q's vectors
  int remote_slot[kMaxNeigh] = {};                 // my index in neighbour q's neighbour list (its flag slot for me)
  unsigned long long* ctrl = nullptr;              // my control block: the kCtrlWords words after the kSlabVecs slab vectors
  DevBuf<unsigned long long*> peer_ctrl_dev;       // device copy of peer_ctrl
};

struct Comm {
  ncclComm_t comm = nullptr;
  int rank = 0, nranks = 1;
  // halo plan
  int nneigh = 0;
  std::vector<int> neigh_rank, send_ptr, recv_start, recv_count;
  DevBuf<int> send_idx;
  DevBuf<double> sendbuf;
  bool have_halo = false;
  long long n_global = 0;   // owned rows of all ranks (Eigen's default iteration limit is 2 n of the WHOLE system)
  P2P p2p;
};

// ---- peer-memory kernels --------------------------------------------------------------------------
struct HaloPushArgs {
  const double* v;                       // local vector (owned rows are read)
  const int* send_idx;
  int send_ptr[kMaxNeigh + 1];
  double* dst[kMaxNeigh];                // neighbour q's copy of the vector, at the rows it receives from me
  unsigned long long* flag[kMaxNeigh];   // neighbour q's halo flag for me
  unsigned long long* my_ctrl;           // sequence number = my count of executed exchanges + 1 (same on every rank)
  int nneigh;
};
// grid (kPushCtas, nneigh): gather my boundary rows and store them into the neighbour's halo rows over NVLink
static __global__ void __launch_bounds__(256) halo_push_kernel(const HaloPushArgs a) {
  const int q = blockIdx.y;
  const int b = a.send_ptr[q], n = a.send_ptr[q + 1] - b;
  double* dst = a.dst[q];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] = a.v[__ldg(a.send_idx + b + i)];
}
// one warp, stream-ordered after the push: release my rows to every neighbour (the push kernel has completed, so
// its stores precede this system-scope release), wait for every neighbour's rows, advance the exchange counter
static __global__ void halo_sync_kernel(const HaloPushArgs a) {
  const unsigned long long seq = a.my_ctrl[kOffHaloCnt] + 1ULL;
  __syncwarp();
  if ((int)threadIdx.x < a.nneigh) {
    __threadfence_system();
    st_release_sys(a.flag[threadIdx.x], seq);
    spin_until(a.my_ctrl + threadIdx.x, seq, a.my_ctrl + kOffErr);
  }
  __syncwarp();
  if (threadIdx.x == 0) a.my_ctrl[kOffHaloCnt] = seq;
}

// stand-alone all-reduce of folded per-CTA partials (the Picard sweep's max): same protocol as the in-kernel one
static __global__ void __launch_bounds__(256) allreduce_p2p_kernel(const double* __restrict__ part, int npart, int stride, int nvals,
                                                             const RedArgs ra) {
  __shared__ double red[33];
  __shared__ double val[3];
  for (int v = 0; v < nvals; v++) {
    double s = (ra.op == 0) ? 0.0 : -DBL_MAX;
    for (int i = threadIdx.x; i < npart; i += blockDim.x) {
      const double x = part[(size_t)v * stride + i];
      s = (ra.op == 0) ? s + x : fmax(s, x);
    }
    s = (ra.op == 0) ? block_sum(s, red) : block_max(s, red);
    if (threadIdx.x == 0) val[v] = s;
  }
  __syncthreads();
  peer_allreduce(val, nvals, ra);
  if ((int)threadIdx.x < nvals) ra.out[threadIdx.x] = val[threadIdx.x];
}

void comm_destroy(fk_handle_s* h) {
  if (!h->comm) return;
  cudaDeviceSynchronize();
  invalidate_graphs(h);   // they hold peer pointers that die with the mappings below
  if (h->halo_stream) {   // carries the L2 window of the old slab: recreated on demand
    cudaStreamDestroy(h->halo_stream);
    h->halo_stream = nullptr;
  }
  P2P& P = h->comm->p2p;
  for (int r = 0; r < kMaxRanks; r++) {
    if (P.opened_slab[r]) cudaIpcCloseMemHandle(P.peer_slab[r]);
  }
  if (h->comm->comm) nccl().CommDestroy(h->comm->comm);
  delete h->comm;
  h->comm = nullptr;
}

static void check_comm_error(fk_handle_s* h) {  // called at host sync points
  Comm* c = h->comm;
  if (!c || !c->p2p.on) return;
  unsigned long long e = 0;
  FK_CUDA(cudaMemcpyAsync(&e, c->p2p.ctrl + kOffErr, sizeof(e), cudaMemcpyDeviceToHost, h->stream));
  FK_CUDA(cudaStreamSynchronize(h->stream));
  if (e) throw std::runtime_error("fk: a peer GPU did not answer within the timeout (halo / all-reduce over NVLink)");
}
void comm_check(fk_handle_s* h) { check_comm_error(h); }
bool comm_capturable(fk_handle_s* h) { return h->comm && h->comm->p2p.on; }

// Fill the halo rows of `v` (a vector inside the PCG slab) from their owners.
void comm_halo_exchange(fk_handle_s* h, cudaStream_t s, double* v) {
  Comm* c = h->comm;
  if (!c || !c->have_halo || c->nneigh == 0) return;
  if (c->p2p.on) {
    P2P& P = c->p2p;
    const size_t off = (size_t)(v - h->pcg_slab.ptr);
    const size_t vec = off / h->slab_stride;           // which slab vector (the stride differs from rank to rank)
    if (v < h->pcg_slab.ptr || vec >= (size_t)kSlabVecs || off != vec * h->slab_stride)
      throw std::runtime_error("fk: halo exchange of a vector outside the PCG slab");
    HaloPushArgs a{};
    a.v = v;
    a.send_idx = c->send_idx.ptr;
    a.my_ctrl = P.ctrl;
    for (int q = 0; q <= c->nneigh; q++) a.send_ptr[q] = c->send_ptr[q];
    for (int q = 0; q < c->nneigh; q++) {
      const int nb = c->neigh_rank[q];
      a.dst[q] = static_cast<double*>(P.peer_slab[nb]) + vec * P.peer_stride[nb] + P.remote_start[q];
      a.flag[q] = P.peer_ctrl[nb] + P.remote_slot[q];
    }
    a.nneigh = c->nneigh;
    halo_push_kernel<<<dim3(kPushCtas, c->nneigh), 256, 0, s>>>(a);
    halo_sync_kernel<<<1, 32, 0, s>>>(a);
    FK_CUDA(cudaGetLastError());
    h->launches += 2;
    return;
  }
  const int nsend = c->send_ptr[c->nneigh];
  if (nsend > 0) {
    pack_kernel<<<(nsend + 255) / 256, 256, 0, s>>>(v, c->send_idx.ptr, nsend, c->sendbuf.ptr);
    FK_CUDA(cudaGetLastError());
    h->launches++;
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

// out[0..nvals) = reduce over ranks of (fold of this rank's per-CTA partials part[v*stride + 0..npart)).
void comm_allreduce_partials(fk_handle_s* h, cudaStream_t s, const double* part, int npart, int stride, int nvals, int op, double* out,
                             const PcgCtl* ctl, int j) {
  Comm* c = h->comm;
  if (c && c->nranks > 1 && c->p2p.on) {
    P2P& P = c->p2p;
    RedArgs ra{nullptr, out, op, c->rank, c->nranks, P.peer_ctrl_dev.ptr, P.ctrl};
    allreduce_p2p_kernel<<<1, 256, 0, s>>>(part, npart, stride, nvals, ra);
    FK_CUDA(cudaGetLastError());
    h->launches++;
    return;
  }
  if (op == 0) fold_partials_kernel<<<1, 256, 0, s>>>(part, npart, stride, nvals, out);
  else max_array_kernel<<<1, 256, 0, s>>>(part, npart, out);
  FK_CUDA(cudaGetLastError());
  h->launches++;
  if (!c || c->nranks == 1) return;
  FK_NCCL(nccl().AllReduce(out, out, (size_t)nvals, ncclDouble, op == 0 ? ncclSum : ncclMax, c->comm, s));
}

// reduction descriptor of the in-kernel grid reductions (kernels.cuh): peer memory when attached, else local
RedArgs comm_red_args(fk_handle_s* h, unsigned int* ticket, double* out, int op) {
  RedArgs ra{ticket, out, op, 0, 1, nullptr, nullptr};
  Comm* c = h->comm;
  if (c && c->nranks > 1 && c->p2p.on) {
    ra.rank = c->rank;
    ra.nranks = c->nranks;
    ra.peer_ctrl = c->p2p.peer_ctrl_dev.ptr;
    ra.my_ctrl = c->p2p.ctrl;
  }
  return ra;
}
PeerRed comm_peer_red(fk_handle_s* h) {
  PeerRed pr{0, 1, nullptr, nullptr};
  Comm* c = h->comm;
  if (c && c->nranks > 1 && c->p2p.on) pr = PeerRed{c->rank, c->nranks, c->p2p.peer_ctrl_dev.ptr, c->p2p.ctrl};
  return pr;
}
// NCCL fallback: the kernels reduced over this GPU only; finish over the ranks with ncclAllReduce
void comm_nccl_allreduce(fk_handle_s* h, cudaStream_t s, double* v, int n, int op) {
  Comm* c = h->comm;
  if (!c || c->nranks == 1 || c->p2p.on) return;
  FK_NCCL(nccl().AllReduce(v, v, (size_t)n, ncclDouble, op == 0 ? ncclSum : ncclMax, c->comm, s));
}
int comm_kind(fk_handle_s* h) { return !h->comm ? 0 : (h->comm->p2p.on ? 2 : 1); }
long long comm_global_rows(fk_handle_s* h) { return (h->comm && h->comm->n_global > 0) ? h->comm->n_global : (long long)h->n_rows; }

// descriptor of the exchange fused into the SpMV of the single-reduction recurrence (spmv.cuh: spmv_sell_fused_kernel)
HaloFused comm_halo_fused(fk_handle_s* h, double* v, unsigned int* ticket) {
  Comm* c = h->comm;
  if (!c || !c->p2p.on) throw std::runtime_error("fk: fused halo exchange needs the peer-memory path");
  P2P& P = c->p2p;
  const size_t off = (size_t)(v - h->pcg_slab.ptr);
  const size_t vec = off / h->slab_stride;
  if (v < h->pcg_slab.ptr || vec >= (size_t)kSlabVecs || off != vec * h->slab_stride)
    throw std::runtime_error("fk: halo exchange of a vector outside the PCG slab");
  HaloFused a{};
  a.send_idx = c->send_idx.ptr;
  a.my_ctrl = P.ctrl;
  a.ticket = ticket;
  for (int q = 0; q <= c->nneigh; q++) a.send_ptr[q] = c->send_ptr[q];
  for (int q = 0; q < c->nneigh; q++) {
    const int nb = c->neigh_rank[q];
    a.dst[q] = static_cast<double*>(P.peer_slab[nb]) + vec * P.peer_stride[nb] + P.remote_start[q];
    a.flag[q] = P.peer_ctrl[nb] + kOffHalo2 + P.remote_slot[q];
  }
  a.nneigh = c->nneigh;
  const int nsend = c->send_ptr[c->nneigh];
  a.npush = std::max(1, std::min(std::min(32, h->grid_spmv), (nsend + 2 * kSpmvThreads - 1) / (2 * kSpmvThreads)));
  a.n_int = h->n_slices_int;
  return a;
}

// ---- one process, several GPUs (multi.cu): the same peer-memory path without NCCL and without CUDA IPC ----------
void comm_init_local(fk_handle_s* h, int rank, int nranks, long long n_global) {
  FK_CUDA(cudaSetDevice(h->device));
  comm_destroy(h);
  h->comm = new Comm();
  h->comm->rank = rank;
  h->comm->nranks = nranks;
  h->comm->n_global = n_global;
  h->scal.ensure(16);
}
// all[r] = handle of rank r (same process); remote_start / remote_slot as in fk_comm_p2p_attach
void comm_attach_local(fk_handle_s* h, fk_handle_s* const* all, const int* remote_start, const int* remote_slot) {
  Comm* c = h->comm;
  if (!c || !c->have_halo) throw std::runtime_error("fk: comm_attach_local before the halo plan");
  if (c->nranks > kMaxRanks || c->nneigh > kMaxNeigh) throw std::runtime_error("fk: too many ranks / neighbours");
  P2P& P = c->p2p;
  for (int r = 0; r < c->nranks; r++) {   // every slab exists before any pointer to it is taken
    FK_CUDA(cudaSetDevice(all[r]->device));
    ensure_pcg(all[r]);
    FK_CUDA(cudaStreamSynchronize(all[r]->stream));
  }
  FK_CUDA(cudaSetDevice(h->device));
  P.ctrl = reinterpret_cast<unsigned long long*>(h->pcg_slab.ptr + (size_t)kSlabVecs * h->slab_stride);
  FK_CUDA(cudaMemset(P.ctrl, 0, kCtrlWords * sizeof(unsigned long long)));
  for (int r = 0; r < c->nranks; r++) {
    if (r != c->rank && all[r]->device != h->device) {
      int can = 0;
      FK_CUDA(cudaDeviceCanAccessPeer(&can, h->device, all[r]->device));
      if (!can) throw std::runtime_error("fk: the GPUs of one fk_multi handle need peer access to each other (NVLink / PCIe P2P)");
      cudaError_t e = cudaDeviceEnablePeerAccess(all[r]->device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) FK_CUDA(e);
      cudaGetLastError();
    }
    P.peer_slab[r] = all[r]->pcg_slab.ptr;
    P.peer_stride[r] = all[r]->slab_stride;
    P.opened_slab[r] = false;
    P.peer_ctrl[r] = reinterpret_cast<unsigned long long*>(static_cast<double*>(P.peer_slab[r]) + (size_t)kSlabVecs * P.peer_stride[r]);
  }
  for (int q = 0; q < c->nneigh; q++) {
    P.remote_start[q] = remote_start[q];
    P.remote_slot[q] = remote_slot[q];
  }
  std::vector<unsigned long long*> ptrs(P.peer_ctrl, P.peer_ctrl + kMaxRanks);
  P.peer_ctrl_dev.upload(ptrs, h->stream);
  FK_CUDA(cudaStreamSynchronize(h->stream));
  P.on = true;
  invalidate_graphs(h);
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
    h->scal.ensure(16);
    {   // owned rows of the whole system (default max_iter = 2 n must be the same on every rank)
      long long* d = nullptr;
      FK_CUDA(cudaMalloc((void**)&d, sizeof(long long)));
      long long mine = h->n_rows;
      FK_CUDA(cudaMemcpy(d, &mine, sizeof(mine), cudaMemcpyHostToDevice));
      FK_NCCL(nccl().AllReduce(d, d, 1, ncclInt64, ncclSum, h->comm->comm, h->stream));
      FK_CUDA(cudaStreamSynchronize(h->stream));
      FK_CUDA(cudaMemcpy(&mine, d, sizeof(mine), cudaMemcpyDeviceToHost));
      cudaFree(d);
      h->comm->n_global = mine;
    }
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
    invalidate_graphs(h);
    return 0;
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  }
}

// ---- peer-memory (CUDA IPC over NVLink) fused path ---------------------------------------------------
// export (128 bytes): the 64-byte IPC handle of the PCG vector slab -- the control block lives in the same
// allocation, right after the six vectors, so one mapping per peer suffices -- then the slab stride (u64)
int fk_comm_p2p_export(fk_handle h, unsigned char* handles128) {
  try {
    if (!h || !handles128) throw std::runtime_error("fk: NULL argument");
    if (!h->comm) throw std::runtime_error("fk_comm_p2p_export: call fk_comm_init first");
    FK_CUDA(cudaSetDevice(h->device));
    ensure_pcg(h);
    P2P& P = h->comm->p2p;
    P.ctrl = reinterpret_cast<unsigned long long*>(h->pcg_slab.ptr + (size_t)kSlabVecs * h->slab_stride);
    FK_CUDA(cudaMemset(P.ctrl, 0, kCtrlWords * sizeof(unsigned long long)));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is expected to be 64 bytes");
    cudaIpcMemHandle_t a;
    FK_CUDA(cudaIpcGetMemHandle(&a, h->pcg_slab.ptr));
    std::memset(handles128, 0, 128);
    std::memcpy(handles128, &a, 64);
    const unsigned long long stride = h->slab_stride;
    std::memcpy(handles128 + 64, &stride, 8);
    return 0;
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  }
}

// attach: exports of ALL ranks (nranks x 128 bytes, rank order); for each neighbour q of this rank,
// remote_start[q] = first row of neighbour q's vectors that receives my rows, remote_slot[q] = my index in
// neighbour q's neighbour list.  From here on the halo exchange and the all-reduces run over peer memory.
int fk_comm_p2p_attach(fk_handle h, const unsigned char* all_handles, const int* remote_start, const int* remote_slot) {
  try {
    if (!h || !all_handles) throw std::runtime_error("fk: NULL argument");
    Comm* c = h->comm;
    if (!c || !c->have_halo) throw std::runtime_error("fk_comm_p2p_attach: call fk_comm_init and fk_set_halo first");
    if (c->nranks > kMaxRanks || c->nneigh > kMaxNeigh) throw std::runtime_error("fk_comm_p2p_attach: too many ranks / neighbours");
    if (c->nneigh > 0 && !(remote_start && remote_slot)) throw std::runtime_error("fk: NULL argument");
    FK_CUDA(cudaSetDevice(h->device));
    P2P& P = c->p2p;
    if (!P.ctrl) throw std::runtime_error("fk_comm_p2p_attach: call fk_comm_p2p_export first");
    for (int r = 0; r < c->nranks; r++) {
      unsigned long long stride = 0;
      std::memcpy(&stride, all_handles + (size_t)r * 128 + 64, 8);
      P.peer_stride[r] = (size_t)stride;
      if (r == c->rank) {
        P.peer_slab[r] = h->pcg_slab.ptr;
      } else {
        cudaIpcMemHandle_t a;
        std::memcpy(&a, all_handles + (size_t)r * 128, 64);
        void* pa = nullptr;
        FK_CUDA(cudaIpcOpenMemHandle(&pa, a, cudaIpcMemLazyEnablePeerAccess));
        P.peer_slab[r] = pa;
        P.opened_slab[r] = true;
      }
      P.peer_ctrl[r] = reinterpret_cast<unsigned long long*>(static_cast<double*>(P.peer_slab[r]) + (size_t)kSlabVecs * P.peer_stride[r]);
    }
    for (int q = 0; q < c->nneigh; q++) {
      if (remote_start[q] < 0 || remote_slot[q] < 0 || remote_slot[q] >= kMaxNeigh) throw std::runtime_error("fk_comm_p2p_attach: bad remote plan");
      P.remote_start[q] = remote_start[q];
      P.remote_slot[q] = remote_slot[q];
    }
    std::vector<unsigned long long*> ptrs(P.peer_ctrl, P.peer_ctrl + kMaxRanks);
    P.peer_ctrl_dev.upload(ptrs, h->stream);
    FK_CUDA(cudaStreamSynchronize(h->stream));
    P.on = true;
    invalidate_graphs(h);
    return 0;
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  }
}

// After a peer timeout ("a peer GPU did not answer"): clear the error word, the exchange counters, the halo flags and the
// mailboxes of THIS rank's control block.  Collective in spirit: every rank calls it between two host barriers of the
// caller (no kernel of the handle may be in flight anywhere), then the next solve starts from a clean protocol state.
int fk_comm_reset(fk_handle h) {
  try {
    if (!h) throw std::runtime_error("fk: NULL handle");
    Comm* c = h->comm;
    if (!c || !c->p2p.on) return 0;
    FK_CUDA(cudaSetDevice(h->device));
    FK_CUDA(cudaDeviceSynchronize());
    unsigned long long lim = 0;
    FK_CUDA(cudaMemcpy(&lim, c->p2p.ctrl + kOffLimit, sizeof(lim), cudaMemcpyDeviceToHost));
    FK_CUDA(cudaMemset(c->p2p.ctrl, 0, kCtrlWords * sizeof(unsigned long long)));
    FK_CUDA(cudaMemcpy(c->p2p.ctrl + kOffLimit, &lim, sizeof(lim), cudaMemcpyHostToDevice));
    FK_CUDA(cudaMemset(h->tickets.ptr, 0, 8 * sizeof(unsigned int)));
    invalidate_graphs(h);
    return 0;
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  }
}
// peer wait limit in milliseconds (0 = default, about 2 s): how long a kernel spins on a peer's flag before it gives up
int fk_comm_set_timeout(fk_handle h, int ms) {
  try {
    if (!h || ms < 0) throw std::runtime_error("fk_comm_set_timeout: bad argument");
    Comm* c = h->comm;
    if (!c || !c->p2p.on) throw std::runtime_error("fk_comm_set_timeout: the handle is not on the peer-memory path");
    FK_CUDA(cudaSetDevice(h->device));
    int khz = 1900000;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, h->device);
    const unsigned long long ticks = (unsigned long long)ms * (unsigned long long)khz;
    FK_CUDA(cudaMemcpy(c->p2p.ctrl + kOffLimit, &ticks, sizeof(ticks), cudaMemcpyHostToDevice));
    return 0;
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  }
}

int fk_comm_mode(fk_handle h, int* mode) {  // 0 = single GPU, 1 = NCCL, 2 = peer memory
  if (!h || !mode) return 1;
  *mode = !h->comm ? 0 : (h->comm->p2p.on ? 2 : 1);
  return 0;
}

}  // extern "C"
