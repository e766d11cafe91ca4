// Internal state behind fk_handle.  One handle = one mesh (or mesh partition) resident on one GPU.
#pragma once
#include <cuda_runtime.h>

#include <stdexcept>
#include <string>
#include <vector>

#include "kernels.cuh"
#include "symbolic.hpp"

namespace fk {

void set_error(const std::string& s);

#define FK_CUDA(call)                                                                                    \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess) {                                                                             \
      (void)cudaGetLastError(); /* clear the (non-sticky) error: a later check must not report it again */ \
      throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(e_) + " at " + __FILE__ +  \
                               ":" + std::to_string(__LINE__) + " (" #call ")");                         \
    }                                                                                                    \
  } while (0)

template <class T>
struct DevBuf {
  T* ptr = nullptr;
  size_t count = 0;
  void ensure(size_t n) {
    if (n <= count && ptr) return;
    release();
    FK_CUDA(cudaMalloc((void**)&ptr, std::max<size_t>(n, 1) * sizeof(T)));
    count = n;
  }
  void upload(const T* src, size_t n, cudaStream_t s) {
    ensure(n);
    if (n) FK_CUDA(cudaMemcpyAsync(ptr, src, n * sizeof(T), cudaMemcpyHostToDevice, s));
  }
  void upload(const std::vector<T>& v, cudaStream_t s) { upload(v.data(), v.size(), s); }
  void release() {
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    count = 0;
  }
  ~DevBuf() { release(); }
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
};

struct Comm;  // multi-GPU state (comm.cu)
constexpr size_t kSlack = 8;  // elements of slack after vals / colind / rowptr (aligned bulk copies may over-read)

}  // namespace fk

struct fk_handle_s {
  int device = 0;
  int num_sms = 148;
  cudaStream_t stream = nullptr;       // all work is issued here (nullptr = legacy default stream)
  cudaStream_t cap_stream = nullptr;   // private stream used only to capture the PCG chunk graph
  int nv = 0, nt = 0, n_rows = 0;
  long long nnz = 0;
  bool user_ve = false;
  fk::Symbolic sym;                    // host copy of the symbolic artefacts (pattern, e2nnz)

  // mesh
  fk::DevBuf<double4> p4;
  fk::DevBuf<int4> t4;
  fk::DevBuf<double> VE;
  // pattern + plans
  fk::DevBuf<int> rowptr, colind;                                  // canonical CSR pattern (exports, diagnostics)
  fk::DevBuf<int> slice_ptr, slice_meta, srow, csr2sell, sell_diag;
  fk::DevBuf<unsigned short> sell_colz;  // packed column stream  // SELL-32-sigma device layout (symbolic.hpp)
  long long nnz_sell = 0;
  fk::DevBuf<int> tile_hdr;            // int4 per tile
  fk::DevBuf<uint32_t> tile_blob;      // packed per-tile plan blobs (symbolic.hpp)
  int e_pad = 0, asm_cfg = 0, asm_dbg = 0;
  // partitions: SELL slices that read no halo column / at least one (spmv split, api.cu: issue_iteration)
  fk::DevBuf<int> slices_int, slices_bnd;
  int n_slices_int = 0, n_slices_bnd = 0, grid_spmv_int = 1, grid_spmv_bnd = 1;
  bool pdl = false;           // programmatic dependent launch inside the single-GPU PCG iteration (measured: 63.8 vs 62.6 us -> off)
  bool overlap_halo = true;
  int halo_mode = 2;          // partitions, Eigen recurrence: 2 = exchange fused into the SpMV kernel, 1 = second stream, 0 = serial
  bool l2_miss_normal = false; // fk_tune 8
  bool l2_on_stream = false;   // fk_tune 14
  int l2_sweep_mode = 2;       // fk_tune 15: what the assembly sweep does about the PCG's persisting L2 set-aside
  bool l2_limit_zeroed = false;
  long long l2_toggles = 0, l2_toggle_ns[2] = {0, 0};   // FK_DEBUG: host time spent in the limit changes
  int l2_max_vecs = 0;         // fk_tune 9
  int spmv_int_reserve = 0;   // CTA slots per SM the interior SpMV leaves to the halo branch (measured: 0 best)
  cudaStream_t halo_stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  cudaStream_t copy_stream = nullptr;  // fk_picard_pcg: uploads that overlap the sweep
  cudaEvent_t copy_ev = nullptr;
  double* pin8 = nullptr;              // pinned host landing zone of the 8-double scalar block
  int asm_grid[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // persistent grid per assembly mode
  size_t asm_smem = 0;
  // matrices
  fk::DevBuf<double> vals[FK_NUM_SLOTS];
  // nonlinear state
  fk::DevBuf<double> mu, dmu, Hn, Hfield, tile_max, scal;
  fk::DevBuf<unsigned char> matid;
  // micromagnetics demag field (fk_demag_*): magnetic elements, node -> (list position, local index) lists, node volumes
  fk::DevBuf<int> dm_inside, dm_n2i_ptr, dm_n2i;
  fk::DevBuf<double> dm_vol, dm_w, dm_M, dm_Hd;
  int dm_n_inside = -1;
  fk::DevBuf<int> mat_list;            // ids of the elements with a B-H curve (matid > 0), ascending
  int n_mat = 0;
  fk::DevBuf<int> errflag;
  fk::DevBuf<double> spl_t[FK_MAX_SPLINES], spl_c[FK_MAX_SPLINES];
  fk::SplineSet spl{};
  bool have_material = false;
  unsigned mat_used = 0;               // bit m-1: material id m occurs in matid
  // PCG workspace
  fk::DevBuf<double> pcg_slab;         // dinv | x | r | p | Ap | s | z | u in ONE allocation (kSlabVecs vectors + control block): one L2
                                       // access-policy window covers the vectors a recurrence streams every iteration, and one
                                       // CUDA-IPC handle exports every vector peers write halo rows into
  struct VecView { double* ptr = nullptr; } x, r, p, Ap, dinv, s, z, u;   // s, z: single-reduction recurrence; u: potential staged for the Picard sweep
  size_t slab_stride = 0;
  bool l2_persist = true, l2_window_set = false;
  cudaAccessPolicyWindow l2_window{};
  int l2_recurrence = 0;               // which recurrence the current window was laid out for
  size_t l2_limit_saved = 0;           // the device's persisting-L2 limit before this handle changed it (restored by fk_destroy)
  bool l2_limit_changed = false;
  fk::DevBuf<double> part, bvec;
  fk::DevBuf<unsigned int> tickets;    // last-CTA tickets of the in-kernel grid reductions (zero between kernels)
  fk::DevBuf<fk::PcgCtl> ctl;
  fk::PcgCtl* ctl_host = nullptr;      // pinned
  unsigned long long tag_base = 16;    // LL mailbox tags handed out so far (kernels.cuh: peer_fold)
  unsigned long long pcg_epoch = 0;   // solves started on this handle (identical on every rank of a partitioned problem)
  int grid_spmv = 0, grid_vec = 0, grid_pcg = 0;   // grid_pcg: the update / direction kernels of an iteration
  int spmv_variant = 3, spmv_ctas = 0;   // SpMV launch shape (spmv.cuh: spmv_variant_dispatch), cap on CTAs per SM
  cudaGraphExec_t graph[FK_NUM_SLOTS][2] = {};   // [slot][recurrence]
  int graph_chunk[FK_NUM_SLOTS][2] = {};
  fk::DevBuf<int> slices_all;          // partitions: slices_int followed by slices_bnd (fused SpMV + halo exchange)
  // constraints
  fk::DevBuf<int> dir_ids;             // fixed DOFs (fk_set_dirichlet)
  fk::DevBuf<double> dir_val, dir_g;   // their values; g = gD scattered over nv (0 on free rows)
  fk::DevBuf<unsigned char> dir_mask;  // nv: 1 = fixed
  unsigned dir_applied = 0;            // bit s: slot s carries the identity rows / columns of fk_apply_dirichlet
  int n_fixed = 0;
  fk::DevBuf<double> ns_w, bproj;      // null-space weights (fk_set_nullspace), projected right-hand side
  bool ns_user = false;
  // deferred result of an enqueued Picard sweep (fk_picard_step_dev with div_out == NULL)
  bool picard_pending = false;
  // generic staging for the host-pointer API
  fk::DevBuf<double> st_a, st_b, st_c, st_d, st_m;
  fk::DevBuf<int> st_i, st_j, st_k;
  // multi-GPU
  fk::Comm* comm = nullptr;
  // rescale assembly (rescale.cuh): geometric factors of every contribution, built on demand
  //   geo_mode 0: never; 1 (default): from the second Picard sweep of a handle on; 2: from the first sweep / assembly on
  int geo_mode = 1, geo_sweeps = 0, geo_nslices = 0, geo_grid = 0, geo_variant = 0, geo_window = 65536, elem_grid = 0;
  bool geo_ready = false, geo_overlap = true, s0_dirty = true;   // s0_dirty: geo_s0 != VE * mu for some element
  int geo_dbg = 0;   // fk_tune 19
  bool geo_rowsum_diag = true, geo_plan_rowsum = false;          // fk_tune 18; what the resident plan was built for
  int geo_split_resc = 4, geo_split_elem = 1;                    // CTAs per SM of the two kernels when they run side by side
  cudaStream_t asm_stream = nullptr;
  cudaEvent_t asm_fork = nullptr, asm_join = nullptr;
  long long geo_slots = 0, geo_items = 0, geo_contrib = 0;
  fk::DevBuf<double> geo_g, geo_s0;
  fk::DevBuf<uint32_t> geo_ref, geo_slice_ptr;
  fk::DevBuf<int> geo_dst1, geo_dst2;
  fk::DevBuf<int> tP2;                 // P2 handles: the 10-node connectivity (t4 holds the vertices)
  fk::DevBuf<double> p2_gmass;         // P2: mass factors per contribution slot (geo_g holds the stiffness factors)
  // Nedelec (edge) handles: nv = number of EDGES (the degrees of freedom), n_points = points of the geometry, t4 = vertices,
  // tP2 = the 6 edge labels of every element; ned_ref keeps the pair bits geo_ref has lost, p2_gmass is the per-call slot scratch
  int n_points = 0;
  bool edge_dofs = false;
  int ned_shift = 26;
  fk::DevBuf<uint32_t> ned_ref;
  fk::DevBuf<double> ned_vals;         // scratch matrix of the load-vector sum
  fk::DevBuf<double> p2_c1, p2_c2, p2_conv_ref;   // P2 convection: per-call slot values (canonical entry / mirror), reference tensor
  // fk_tune(h, 11, n): device-side timeline of the PCG kernels (kernels.cuh: trace_rec)
  fk::DevBuf<unsigned long long> trace;
  int trace_cap = 0;
  // launch counter (kernels of this library launched through this handle)
  long long launches = 0;
};
