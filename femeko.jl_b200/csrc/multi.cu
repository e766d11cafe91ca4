// ONE process driving several GPUs (SURVEY 8b "Threading": the reference's caller is a single Julia process, so the
// 2/4/8-GPU path must be reachable through one @ccall): fk_multi_* partitions the caller's GLOBAL mesh by recursive
// coordinate bisection of the nodes, renumbers every part owned-first / halo-after, builds one fk_handle per GPU, wires
// the handles over peer memory (cudaDeviceEnablePeerAccess: no IPC, no NCCL, no second process) and drives them with one
// host thread per GPU.  Every call takes and returns GLOBAL host arrays in the caller's numbering.  The per-GPU work --
// kernels, halo exchange fused into the SpMV, dot products over the LL mailbox -- is exactly the one-process-per-GPU
// path's (api.cu, comm.cu); this file only adds the host-side decomposition (the C++ twin of partition.py) and the fan-out.
#include <algorithm>
#include <cstring>
#include <functional>
#include <numeric>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "comm.hpp"
#include "handle.hpp"

namespace fk {
void comm_init_local(fk_handle_s* h, int rank, int nranks, long long n_global);
void comm_attach_local(fk_handle_s* h, fk_handle_s* const* all, const int* remote_start, const int* remote_slot);

// ---- recursive coordinate bisection of the nodes (partition.py: rcb) ----------------------------------
// split the longest axis of the current bounding box at the weighted median, ties broken by node id
static void rcb_split(const double* p, std::vector<int>& idx, int lo, int hi, int first, int count, std::vector<int>& part) {
  if (count == 1) {
    for (int i = lo; i < hi; i++) part[idx[i]] = first;
    return;
  }
  double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
  for (int i = lo; i < hi; i++)
    for (int d = 0; d < 3; d++) {
      const double v = p[3 * (size_t)idx[i] + d];
      mn[d] = std::min(mn[d], v);
      mx[d] = std::max(mx[d], v);
    }
  int ax = 0;
  for (int d = 1; d < 3; d++)
    if (mx[d] - mn[d] > mx[ax] - mn[ax]) ax = d;
  const int left_parts = count / 2;
  const int nleft = (int)(((long long)(hi - lo) * left_parts) / count);
  auto less = [&](int a, int b) {
    const double va = p[3 * (size_t)a + ax], vb = p[3 * (size_t)b + ax];
    return va != vb ? va < vb : a < b;
  };
  std::nth_element(idx.begin() + lo, idx.begin() + lo + nleft, idx.begin() + hi, less);
  rcb_split(p, idx, lo, lo + nleft, first, left_parts, part);
  rcb_split(p, idx, lo + nleft, hi, first + left_parts, count - left_parts, part);
}

struct LocalPart {
  int n_owned = 0, nv = 0, nt = 0;
  std::vector<int> l2g;        // local node -> global node (owned ascending, then halo by (owner, global id))
  std::vector<int> elem_l2g;   // local element -> global element (ascending)
  std::vector<double> p;       // 3 x nv
  std::vector<int> t;          // 4 x nt, local ids
  std::vector<int> neigh, send_ptr, send_idx, recv_start, recv_count;
};

// partition.py: partition_mesh + _halo_plan for one rank
static void build_local(const double* p, const int* t, int nv, int nt, const std::vector<int>& part, int rank, LocalPart& L) {
  std::vector<int> g2l((size_t)nv, -1);
  std::vector<char> seen((size_t)nv, 0);
  std::vector<int> halo;
  for (int k = 0; k < nt; k++) {
    const int* e = t + 4 * (size_t)k;
    if (part[e[0]] == rank || part[e[1]] == rank || part[e[2]] == rank || part[e[3]] == rank) {
      L.elem_l2g.push_back(k);
      for (int i = 0; i < 4; i++)
        if (part[e[i]] != rank && !seen[e[i]]) { seen[e[i]] = 1; halo.push_back(e[i]); }
    }
  }
  for (int n = 0; n < nv; n++)
    if (part[n] == rank) { g2l[n] = (int)L.l2g.size(); L.l2g.push_back(n); }
  L.n_owned = (int)L.l2g.size();
  std::sort(halo.begin(), halo.end(), [&](int a, int b) { return part[a] != part[b] ? part[a] < part[b] : a < b; });
  for (int n : halo) { g2l[n] = (int)L.l2g.size(); L.l2g.push_back(n); }
  L.nv = (int)L.l2g.size();
  L.nt = (int)L.elem_l2g.size();
  L.p.resize(3 * (size_t)L.nv);
  for (int i = 0; i < L.nv; i++)
    for (int d = 0; d < 3; d++) L.p[3 * (size_t)i + d] = p[3 * (size_t)L.l2g[i] + d];
  L.t.resize(4 * (size_t)L.nt);
  for (int q = 0; q < L.nt; q++)
    for (int i = 0; i < 4; i++) L.t[4 * (size_t)q + i] = g2l[t[4 * (size_t)L.elem_l2g[q] + i]];
  // neighbours: owners of my halo nodes, and every rank that owns a node of an element holding one of mine
  std::vector<int> nb;
  for (int q = 0; q < L.nt; q++)
    for (int i = 0; i < 4; i++) {
      const int o = part[L.l2g[L.t[4 * (size_t)q + i]]];
      if (o != rank) nb.push_back(o);
    }
  std::sort(nb.begin(), nb.end());
  nb.erase(std::unique(nb.begin(), nb.end()), nb.end());
  L.neigh = nb;
  L.send_ptr.assign(1, 0);
  std::vector<char> mark((size_t)L.n_owned, 0);
  for (int qn : nb) {
    std::fill(mark.begin(), mark.end(), 0);
    for (int q = 0; q < L.nt; q++) {
      const int* e = L.t.data() + 4 * (size_t)q;
      bool has = false;
      for (int i = 0; i < 4; i++) has |= part[L.l2g[e[i]]] == qn;
      if (!has) continue;
      for (int i = 0; i < 4; i++)
        if (e[i] < L.n_owned) mark[e[i]] = 1;
    }
    for (int i = 0; i < L.n_owned; i++)   // owned rows are in ascending global id already
      if (mark[i]) L.send_idx.push_back(i);
    L.send_ptr.push_back((int)L.send_idx.size());
    int first = -1, cnt = 0;
    for (int i = L.n_owned; i < L.nv; i++)
      if (part[L.l2g[i]] == qn) { if (first < 0) first = i; cnt++; }
    L.recv_start.push_back(first < 0 ? L.n_owned : first);
    L.recv_count.push_back(cnt);
  }
}

static void parallel_ranks(int n, const std::function<void(int)>& f) {
  std::vector<std::thread> th;
  std::vector<std::string> err((size_t)n);
  for (int r = 0; r < n; r++)
    th.emplace_back([&, r]() {
      try { f(r); } catch (const std::exception& e) { err[r] = e.what(); if (err[r].empty()) err[r] = "error"; }
    });
  for (auto& x : th) x.join();
  for (int r = 0; r < n; r++)
    if (!err[r].empty()) throw std::runtime_error("GPU rank " + std::to_string(r) + ": " + err[r]);
}
static void ck(int rc) {
  if (rc != 0) throw std::runtime_error(fk_last_error());
}
}  // namespace fk

using namespace fk;

struct fk_multi_s {
  int ngpu = 0, nv = 0, nt = 0;
  std::vector<fk_handle> h;
  std::vector<LocalPart> part;
  std::vector<std::vector<double>> bufn, bufx;   // per-rank node / element staging
};

static thread_local std::string g_merr;
template <class F>
static int mguard(F&& f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  }
}

extern "C" {

int fk_multi_create(const double* p, const int* t, int nv, int nt, const double* VE, int ngpu, const int* devices,
                    const fk_create_opts* opts, fk_multi* out) {
  return mguard([&]() {
    if (!p || !t || !out) throw std::runtime_error("fk: NULL argument");
    if (ngpu < 1 || ngpu > kMaxRanks) throw std::runtime_error("fk_multi_create: ngpu must be in [1, 8]");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) throw std::runtime_error("fk_multi_create: no CUDA device available");
    if (!devices && ngpu > ndev) throw std::runtime_error("fk_multi_create: fewer CUDA devices than ngpu");
    fk_multi m = new fk_multi_s();
    try {
      m->ngpu = ngpu;
      m->nv = nv;
      m->nt = nt;
      std::vector<int> partv((size_t)nv, 0);
      if (ngpu > 1) {
        std::vector<int> idx((size_t)nv);
        std::iota(idx.begin(), idx.end(), 0);
        rcb_split(p, idx, 0, nv, 0, ngpu, partv);
      }
      m->part.resize(ngpu);
      m->h.assign(ngpu, nullptr);
      m->bufn.resize(ngpu);
      m->bufx.resize(ngpu);
      parallel_ranks(ngpu, [&](int r) { build_local(p, t, nv, nt, partv, r, m->part[r]); });
      const int host_threads = fk_set_host_threads(0);
      parallel_ranks(ngpu, [&](int r) {
        LocalPart& L = m->part[r];
        fk_set_host_threads(std::max(1, host_threads / ngpu));   // per calling thread (OpenMP ICV)
        fk_create_opts o{};
        if (opts) o = *opts;
        o.device = devices ? devices[r] : r;
        o.n_owned = ngpu > 1 ? L.n_owned : 0;
        std::vector<double> ve;
        if (VE) {
          ve.resize((size_t)L.nt);
          for (int q = 0; q < L.nt; q++) ve[q] = VE[L.elem_l2g[q]];
        }
        ck(fk_create(L.p.data(), L.t.data(), L.nv, L.nt, VE ? ve.data() : nullptr, &o, &m->h[r]));
      });
      if (ngpu > 1) {
        for (int r = 0; r < ngpu; r++) {
          LocalPart& L = m->part[r];
          comm_init_local(m->h[r], r, ngpu, (long long)nv);
          ck(fk_set_halo(m->h[r], L.n_owned, (int)L.neigh.size(), L.neigh.data(), L.send_ptr.data(), L.send_idx.data(),
                         L.recv_start.data(), L.recv_count.data()));
        }
        for (int r = 0; r < ngpu; r++) {
          LocalPart& L = m->part[r];
          std::vector<int> rstart, rslot;
          for (size_t iq = 0; iq < L.neigh.size(); iq++) {
            const int qn = L.neigh[iq];
            const LocalPart& N = m->part[qn];
            const int me = (int)(std::find(N.neigh.begin(), N.neigh.end(), r) - N.neigh.begin());
            if (me >= (int)N.neigh.size()) throw std::runtime_error("fk_multi_create: asymmetric neighbour lists");
            if (N.recv_count[me] != L.send_ptr[iq + 1] - L.send_ptr[iq])
              throw std::runtime_error("fk_multi_create: halo plans of two ranks disagree");
            rslot.push_back(me);
            rstart.push_back(N.recv_start[me]);
          }
          comm_attach_local(m->h[r], m->h.data(), rstart.data(), rslot.data());
        }
      }
    } catch (...) {
      for (fk_handle hh : m->h)
        if (hh) fk_destroy(hh);
      delete m;
      throw;
    }
    *out = m;
  });
}

// host only (tests): the node partition and one rank's local numbering + halo plan, exactly as fk_multi_create builds them
int fk_rcb(const double* p, int nv, int nparts, int* part_out) {
  return mguard([&]() {
    if (!p || !part_out || nv < 1 || nparts < 1) throw std::runtime_error("fk_rcb: bad argument");
    std::vector<int> part((size_t)nv, 0), idx((size_t)nv);
    std::iota(idx.begin(), idx.end(), 0);
    rcb_split(p, idx, 0, nv, 0, nparts, part);
    std::memcpy(part_out, part.data(), sizeof(int) * (size_t)nv);
  });
}
int fk_partition_local(const double* p, const int* t, int nv, int nt, const int* part, int rank, int* sizes /* n_owned, nv, nt, nneigh, nsend */,
                       int* l2g, int* elem_l2g, int* t_local, int* neigh, int* send_ptr, int* send_idx, int* recv_start, int* recv_count) {
  return mguard([&]() {
    if (!p || !t || !part || !sizes) throw std::runtime_error("fk_partition_local: NULL argument");
    LocalPart L;
    build_local(p, t, nv, nt, std::vector<int>(part, part + nv), rank, L);
    sizes[0] = L.n_owned; sizes[1] = L.nv; sizes[2] = L.nt; sizes[3] = (int)L.neigh.size(); sizes[4] = (int)L.send_idx.size();
    auto cp = [](const std::vector<int>& v, int* d) { if (d && !v.empty()) std::memcpy(d, v.data(), sizeof(int) * v.size()); };
    cp(L.l2g, l2g); cp(L.elem_l2g, elem_l2g); cp(L.t, t_local); cp(L.neigh, neigh); cp(L.send_ptr, send_ptr); cp(L.send_idx, send_idx);
    cp(L.recv_start, recv_start); cp(L.recv_count, recv_count);
  });
}

int fk_multi_destroy(fk_multi m) {
  return mguard([&]() {
    if (!m) return;
    for (fk_handle hh : m->h)
      if (hh) { cudaSetDevice(hh->device); cudaDeviceSynchronize(); }
    for (fk_handle hh : m->h)
      if (hh) fk_destroy(hh);
    delete m;
  });
}

int fk_multi_info(fk_multi m, int* ngpu, int* n_owned /* ngpu */, int* nv_local, int* nt_local, int* nneigh) {
  return mguard([&]() {
    if (!m) throw std::runtime_error("fk: NULL argument");
    if (ngpu) *ngpu = m->ngpu;
    for (int r = 0; r < m->ngpu; r++) {
      if (n_owned) n_owned[r] = m->part[r].n_owned;
      if (nv_local) nv_local[r] = m->part[r].nv;
      if (nt_local) nt_local[r] = m->part[r].nt;
      if (nneigh) nneigh[r] = (int)m->part[r].neigh.size();
    }
  });
}
int fk_multi_handle(fk_multi m, int rank, fk_handle* h) {
  return mguard([&]() {
    if (!m || !h || rank < 0 || rank >= m->ngpu) throw std::runtime_error("fk_multi_handle: bad argument");
    *h = m->h[rank];
  });
}
int fk_multi_partition(fk_multi m, int rank, int* l2g /* nv_local */, int* elem_l2g /* nt_local */) {
  return mguard([&]() {
    if (!m || rank < 0 || rank >= m->ngpu) throw std::runtime_error("fk_multi_partition: bad argument");
    if (l2g) std::memcpy(l2g, m->part[rank].l2g.data(), sizeof(int) * m->part[rank].l2g.size());
    if (elem_l2g) std::memcpy(elem_l2g, m->part[rank].elem_l2g.data(), sizeof(int) * m->part[rank].elem_l2g.size());
  });
}

// ---- fan-out helpers ------------------------------------------------------------------------------------
static const double* elem_local(fk_multi m, int r, const double* g, int width = 1) {   // global per-element array -> rank r's
  if (!g) return nullptr;
  const LocalPart& L = m->part[r];
  std::vector<double>& b = m->bufx[r];
  b.resize((size_t)width * L.nt);
  for (int q = 0; q < L.nt; q++)
    for (int c = 0; c < width; c++) b[(size_t)width * q + c] = g[(size_t)width * L.elem_l2g[q] + c];
  return b.data();
}
static double* node_local(fk_multi m, int r, const double* g) {   // global per-node array -> rank r's (owned + halo)
  const LocalPart& L = m->part[r];
  std::vector<double>& b = m->bufn[r];
  b.resize((size_t)L.nv);
  if (g)
    for (int i = 0; i < L.nv; i++) b[i] = g[L.l2g[i]];
  else
    std::fill(b.begin(), b.end(), 0.0);
  return b.data();
}
static void node_global(fk_multi m, int r, const double* loc, double* g) {   // owned rows back to the caller's numbering
  const LocalPart& L = m->part[r];
  for (int i = 0; i < L.n_owned; i++) g[L.l2g[i]] = loc[i];
}
static void elem_global(fk_multi m, int r, const double* loc, double* g, int width = 1) {
  if (!g) return;
  const LocalPart& L = m->part[r];
  for (int q = 0; q < L.nt; q++) {
    const int k = L.elem_l2g[q];
    // an element held by several ranks (ghost copies) has the same value on each; the copy that owns node t[0] writes
    if (m->ngpu > 1 && L.t[4 * (size_t)q] >= L.n_owned) continue;
    for (int c = 0; c < width; c++) g[(size_t)width * k + c] = loc[(size_t)width * q + c];
  }
}

int fk_multi_assemble_stiffness(fk_multi m, int slot, const double* f /* nt global or NULL */) {
  return mguard([&]() {
    if (!m) throw std::runtime_error("fk: NULL argument");
    parallel_ranks(m->ngpu, [&](int r) { ck(fk_assemble_stiffness(m->h[r], slot, elem_local(m, r, f), nullptr)); });
  });
}
int fk_multi_assemble_mass(fk_multi m, int slot, const double* rho) {
  return mguard([&]() {
    if (!m) throw std::runtime_error("fk: NULL argument");
    parallel_ranks(m->ngpu, [&](int r) { ck(fk_assemble_mass(m->h[r], slot, elem_local(m, r, rho), nullptr)); });
  });
}
int fk_multi_matrix_axpby(fk_multi m, int dst, double alpha, int x, double beta, int y) {
  return mguard([&]() {
    if (!m) throw std::runtime_error("fk: NULL argument");
    parallel_ranks(m->ngpu, [&](int r) { ck(fk_matrix_axpby(m->h[r], dst, alpha, x, beta, y)); });
  });
}
int fk_multi_set_spline(fk_multi m, int id, const double* t, const double* c, int n, int k, int bc) {
  return mguard([&]() {
    if (!m) throw std::runtime_error("fk: NULL argument");
    for (int r = 0; r < m->ngpu; r++) ck(fk_set_spline(m->h[r], id, t, c, n, k, bc));
  });
}
int fk_multi_set_inside(fk_multi m, const int* ids /* global element ids */, int n) {
  return mguard([&]() {
    if (!m || (!ids && n > 0)) throw std::runtime_error("fk: NULL argument");
    std::vector<char> in((size_t)m->nt, 0);
    for (int i = 0; i < n; i++) {
      if (ids[i] < 0 || ids[i] >= m->nt) throw std::runtime_error("fk_multi_set_inside: element id out of range");
      in[ids[i]] = 1;
    }
    parallel_ranks(m->ngpu, [&](int r) {
      const LocalPart& L = m->part[r];
      std::vector<int> loc;
      for (int q = 0; q < L.nt; q++)
        if (in[L.elem_l2g[q]]) loc.push_back(q);
      ck(fk_set_inside(m->h[r], loc.data(), (int)loc.size()));
    });
  });
}
int fk_multi_set_mu(fk_multi m, const double* mu /* nt global */) {
  return mguard([&]() {
    if (!m || !mu) throw std::runtime_error("fk: NULL argument");
    parallel_ranks(m->ngpu, [&](int r) { ck(fk_set_mu(m->h[r], elem_local(m, r, mu))); });
  });
}
int fk_multi_get_mu(fk_multi m, double* mu_out /* nt global */) {
  return mguard([&]() {
    if (!m || !mu_out) throw std::runtime_error("fk: NULL argument");
    parallel_ranks(m->ngpu, [&](int r) {
      std::vector<double> loc((size_t)m->part[r].nt);
      ck(fk_get_mu(m->h[r], loc.data(), nullptr));
      elem_global(m, r, loc.data(), mu_out);
    });
  });
}
int fk_multi_get_field(fk_multi m, double* Hfield_out /* 3 x nt or NULL */, double* Hnorm_out /* nt or NULL */) {
  return mguard([&]() {
    if (!m) throw std::runtime_error("fk: NULL argument");
    parallel_ranks(m->ngpu, [&](int r) {
      const int ntl = m->part[r].nt;
      std::vector<double> hf(Hfield_out ? 3 * (size_t)ntl : 0), hn(Hnorm_out ? (size_t)ntl : 0);
      ck(fk_get_field(m->h[r], Hfield_out ? hf.data() : nullptr, Hnorm_out ? hn.data() : nullptr));
      elem_global(m, r, hf.data(), Hfield_out, 3);
      elem_global(m, r, hn.data(), Hnorm_out, 1);
    });
  });
}
int fk_multi_reset_field(fk_multi m) {
  return mguard([&]() {
    if (!m) throw std::runtime_error("fk: NULL argument");
    for (int r = 0; r < m->ngpu; r++) ck(fk_reset_field(m->h[r]));
  });
}

// Jacobi-PCG on the partitioned matrix: b, x in the caller's global numbering (x: initial guess when opts->warm_start)
int fk_multi_pcg(fk_multi m, int slot, const double* b, double* x, const fk_pcg_opts* opts, int* iters_out, double* relres_out) {
  return mguard([&]() {
    if (!m || !b || !x) throw std::runtime_error("fk: NULL argument");
    std::vector<int> it((size_t)m->ngpu, 0);
    std::vector<double> rr((size_t)m->ngpu, 0.0);
    std::vector<std::vector<double>> xl((size_t)m->ngpu);
    parallel_ranks(m->ngpu, [&](int r) {
      const LocalPart& L = m->part[r];
      xl[r].assign((size_t)L.nv, 0.0);
      if (opts && opts->warm_start)
        for (int i = 0; i < L.nv; i++) xl[r][i] = x[L.l2g[i]];
      ck(fk_pcg(m->h[r], slot, node_local(m, r, b), xl[r].data(), opts, &it[r], &rr[r]));
    });
    for (int r = 0; r < m->ngpu; r++) node_global(m, r, xl[r].data(), x);
    if (iters_out) *iters_out = it[0];
    if (relres_out) *relres_out = rr[0];
  });
}
// One Picard sweep on every part (fk_picard_step): u global; div = mu0 max |dH| over the whole mesh
int fk_multi_picard_step(fk_multi m, int slot, const double* u, int compat_bug, double* div_out) {
  return mguard([&]() {
    if (!m || !u) throw std::runtime_error("fk: NULL argument");
    std::vector<double> div((size_t)m->ngpu, 0.0);
    parallel_ranks(m->ngpu, [&](int r) { ck(fk_picard_step(m->h[r], slot, node_local(m, r, u), compat_bug, &div[r])); });
    if (div_out) *div_out = div[0];   // all-reduced over the parts by the sweep itself
  });
}
// Demagnetising field of the LL loop (fk_demag_setup / fk_demag_field) on every part: the magnetic elements and the nodal
// magnetisation in the caller's global numbering; a part works on the magnetic elements it holds (ghost copies included, in
// ascending global order = the reference's list order when InsideElements is ascending), every owned node gathers all of its
// incident magnetic elements locally, and the potential's halo rows are refreshed before the element field is taken.
int fk_multi_demag_setup(fk_multi m, int slot, const int* inside_elements, int n_inside, double* volumes_out) {
  return mguard([&]() {
    if (!m || (!inside_elements && n_inside > 0)) throw std::runtime_error("fk: NULL argument");
    std::vector<char> in((size_t)m->nt, 0);
    for (int i = 0; i < n_inside; i++) {
      if (inside_elements[i] < 0 || inside_elements[i] >= m->nt) throw std::runtime_error("fk_multi_demag_setup: element id out of range");
      if (i > 0 && inside_elements[i] <= inside_elements[i - 1])
        throw std::runtime_error("fk_multi_demag_setup: the element list must be ascending (the parts keep the reference's summation order that way)");
      in[inside_elements[i]] = 1;
    }
    parallel_ranks(m->ngpu, [&](int r) {
      const LocalPart& L = m->part[r];
      std::vector<int> loc;
      for (int q = 0; q < L.nt; q++)
        if (in[L.elem_l2g[q]]) loc.push_back(q);
      std::vector<double> vol((size_t)L.nv, 0.0);
      ck(fk_demag_setup(m->h[r], slot, loc.data(), (int)loc.size(), vol.data()));
      if (volumes_out) node_global(m, r, vol.data(), volumes_out);
    });
  });
}
int fk_multi_demag_field(fk_multi m, int slot, const double* M, const fk_pcg_opts* opts, int nan_outside, double* Hd_out, double* u_out,
                         int* iters_out, double* relres_out) {
  return mguard([&]() {
    if (!m || !M || !Hd_out) throw std::runtime_error("fk: NULL argument");
    std::vector<int> it((size_t)m->ngpu, 0);
    std::vector<double> rr((size_t)m->ngpu, 0.0);
    parallel_ranks(m->ngpu, [&](int r) {
      const LocalPart& L = m->part[r];
      std::vector<double> Ml(3 * (size_t)L.nv), Hl(3 * (size_t)L.nv), ul(u_out ? (size_t)L.nv : 0);
      for (int i = 0; i < L.nv; i++)
        for (int c = 0; c < 3; c++) Ml[3 * (size_t)i + c] = M[3 * (size_t)L.l2g[i] + c];
      ck(fk_demag_field(m->h[r], slot, Ml.data(), opts, nan_outside, Hl.data(), u_out ? ul.data() : nullptr, &it[r], &rr[r]));
      for (int i = 0; i < L.n_owned; i++)
        for (int c = 0; c < 3; c++) Hd_out[3 * (size_t)L.l2g[i] + c] = Hl[3 * (size_t)i + c];
      if (u_out) node_global(m, r, ul.data(), u_out);
    });
    if (iters_out) *iters_out = it[0];
    if (relres_out) *relres_out = rr[0];
  });
}
int fk_multi_spmv(fk_multi m, int slot, const double* x, double* y) {
  return mguard([&]() {
    if (!m || !x || !y) throw std::runtime_error("fk: NULL argument");
    parallel_ranks(m->ngpu, [&](int r) {
      const LocalPart& L = m->part[r];
      std::vector<double> yl((size_t)L.nv, 0.0);
      ck(fk_spmv(m->h[r], slot, node_local(m, r, x), yl.data()));   // the caller's x carries the halo values already
      node_global(m, r, yl.data(), y);
    });
  });
}

}  // extern "C"
