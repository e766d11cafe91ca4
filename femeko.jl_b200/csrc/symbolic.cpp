// Symbolic phase (host): canonical CSR pattern, element->nnz map, assembly tiles, SpMV row blocks.
// See symbolic.hpp.  Reference context: the pattern equals what Eigen's setFromTriplets produces for
// the 16*nt triplets of src/FEMc.cpp:111-129 (every (t[i,k], t[j,k]) pair, columns ascending), and
// e2nnz is the per-element form of the triplet index lists of NewtonRaphson.jl:114-127.
#include "symbolic.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <climits>
#include <cmath>
#include <cstring>
#include <numeric>
#include <stdexcept>
#include <string>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace fk {
namespace {
// FK_DEBUG: wall-clock split of the symbolic phase on stderr
struct PhaseTimer {
  bool on = std::getenv("FK_DEBUG") != nullptr;
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  void lap(const char* what) {
    if (!on) return;
    const auto t1 = std::chrono::steady_clock::now();
    std::fprintf(stderr, "fk symbolic: %-28s %7.3f s\n", what, std::chrono::duration<double>(t1 - t0).count());
    t0 = t1;
  }
};
}  // namespace


static inline uint64_t spread21(uint64_t v) {
  v &= 0x1fffffULL;
  v = (v | (v << 32)) & 0x1f00000000ffffULL;
  v = (v | (v << 16)) & 0x1f0000ff0000ffULL;
  v = (v | (v << 8)) & 0x100f00f00f00f00fULL;
  v = (v | (v << 4)) & 0x10c30c30c30c30c3ULL;
  v = (v | (v << 2)) & 0x1249249249249249ULL;
  return v;
}

void build_symbolic(const double* p, const int* t, int nv, int nt, const SymbolicOpts& o, Symbolic& S) {
  if (nv <= 0 || nt <= 0) throw std::runtime_error("fk: empty mesh (nv and nt must be positive)");
  if (o.tile_max_elems < 1 || o.tile_max_elems > 4096) throw std::runtime_error("fk: tile_max_elems must be in [1, 4096]");
  const int npe = o.npe;
  if (npe != 4 && npe != 10 && npe != 6)
    throw std::runtime_error("fk: nodes per element must be 4 (P1), 10 (P2) or 6 (the edges of a Nedelec element)");
  if ((int64_t)nt * npe >= (int64_t)1 << 32) throw std::runtime_error("fk: nt too large for 32-bit incidence ids");
  S.npe = npe;
  S.nv = nv;
  S.nt = nt;
  S.n_rows = (o.n_owned < 0 || o.n_owned > nv) ? nv : o.n_owned;
  const int nr = S.n_rows;
  PhaseTimer tm;

  // ---- node -> element incidences, ascending element order (counting sort) ------------------
  std::vector<int64_t> n2e_ptr((size_t)nv + 1, 0);
  for (int64_t q = 0; q < (int64_t)npe * nt; q++) {
    int n = t[q];
    if (n < 0 || n >= nv) throw std::runtime_error("fk: connectivity entry out of range (expected 0-based ids < nv)");
    n2e_ptr[(size_t)n + 1]++;
  }
  for (int r = 0; r < nv; r++) n2e_ptr[r + 1] += n2e_ptr[r];
  std::vector<uint32_t> n2e((size_t)npe * nt);
  {
    std::vector<int64_t> fill(n2e_ptr.begin(), n2e_ptr.end() - 1);
    for (int64_t q = 0; q < (int64_t)npe * nt; q++) n2e[(size_t)fill[t[q]]++] = (uint32_t)q;  // q = npe k + i
  }

  tm.lap("node -> element incidences");
  // ---- row pattern: sorted unique nodes of all incident elements ----------------------------
  S.rowptr.assign((size_t)nv + 1, 0);
  std::vector<int> rowlen((size_t)nv, 0);
#pragma omp parallel
  {
    std::vector<int> buf;
#pragma omp for schedule(dynamic, 1024)
    for (int r = 0; r < nr; r++) {
      buf.clear();
      for (int64_t q = n2e_ptr[r]; q < n2e_ptr[r + 1]; q++) {
        const int* e = t + (size_t)npe * (n2e[q] / (uint32_t)npe);
        for (int i = 0; i < npe; i++) buf.push_back(e[i]);
      }
      std::sort(buf.begin(), buf.end());
      rowlen[r] = (int)(std::unique(buf.begin(), buf.end()) - buf.begin());
    }
  }
  int64_t nnz = 0;
  for (int r = 0; r < nv; r++) {
    S.rowptr[r] = (int)nnz;
    nnz += rowlen[r];
    if (nnz > 0x7fffffffLL) throw std::runtime_error("fk: more than 2^31-1 matrix entries on one GPU; partition the mesh");
  }
  S.rowptr[nv] = (int)nnz;
  S.nnz = nnz;
  S.colind.resize((size_t)nnz);
  S.diag.assign((size_t)nr, -1);
#pragma omp parallel
  {
    std::vector<int> buf;
#pragma omp for schedule(dynamic, 1024)
    for (int r = 0; r < nr; r++) {
      buf.clear();
      for (int64_t q = n2e_ptr[r]; q < n2e_ptr[r + 1]; q++) {
        const int* e = t + (size_t)npe * (n2e[q] / (uint32_t)npe);
        for (int i = 0; i < npe; i++) buf.push_back(e[i]);
      }
      std::sort(buf.begin(), buf.end());
      int m = (int)(std::unique(buf.begin(), buf.end()) - buf.begin());
      int* dst = S.colind.data() + S.rowptr[r];
      for (int c = 0; c < m; c++) {
        dst[c] = buf[c];
        if (buf[c] == r) S.diag[r] = S.rowptr[r] + c;
      }
    }
  }

  tm.lap("row pattern");
  // ---- element -> nnz map ---------------------------------------------------------------------
  S.e2nnz.resize((size_t)npe * npe * nt);
#pragma omp parallel for schedule(static)
  for (int k = 0; k < nt; k++) {
    const int* e = t + (size_t)npe * k;
    for (int i = 0; i < npe; i++) {
      int r = e[i];
      for (int j = 0; j < npe; j++) {
        int idx = -1;
        if (r < nr) {
          const int* b = S.colind.data() + S.rowptr[r];
          const int* en = S.colind.data() + S.rowptr[r + 1];
          idx = (int)(std::lower_bound(b, en, e[j]) - S.colind.data());
        }
        S.e2nnz[(size_t)npe * npe * k + npe * i + j] = idx;
      }
    }
  }

  tm.lap("element -> nnz map");
  // ---- SELL-32-sigma layout (see symbolic.hpp): independent of the assembly tiles, shared by every element type ------
  auto build_sell = [&]() {
    const int C = 32;
    const int sigma = std::max(1, o.sell_sigma);
    std::vector<int> ord((size_t)nr);
    std::iota(ord.begin(), ord.end(), 0);
    if (sigma > 1)
      for (int w0 = 0; w0 < nr; w0 += sigma)
        std::stable_sort(ord.begin() + w0, ord.begin() + std::min(nr, w0 + sigma), [&](int a, int b) { return rowlen[a] > rowlen[b]; });
    S.nslices = (nr + C - 1) / C;
    S.srow.assign((size_t)S.nslices * C, -1);
    std::copy(ord.begin(), ord.end(), S.srow.begin());
    S.slice_ptr.assign((size_t)S.nslices + 1, 0);
    int64_t tot = 0;
    for (int sl = 0; sl < S.nslices; sl++) {
      int w = 0;
      for (int l = 0; l < C; l++) {
        const int r = S.srow[(size_t)sl * C + l];
        if (r >= 0) w = std::max(w, rowlen[r]);
      }
      S.slice_ptr[sl] = (int)tot;
      tot += (int64_t)w * C;
      if (tot > 0x7fffffffLL) throw std::runtime_error("fk: more than 2^31-1 padded matrix entries on one GPU; partition the mesh");
    }
    S.slice_ptr[S.nslices] = (int)tot;
    S.nnz_sell = tot;
    S.sell_col.assign((size_t)tot, 0);
    S.csr2sell.assign((size_t)S.nnz, -1);
    S.sell_diag.assign((size_t)nr, -1);
#pragma omp parallel for schedule(static)
    for (int sl = 0; sl < S.nslices; sl++) {
      const int base = S.slice_ptr[sl], w = (S.slice_ptr[sl + 1] - base) / C;
      for (int l = 0; l < C; l++) {
        const int r = S.srow[(size_t)sl * C + l];
        const int len = r >= 0 ? rowlen[r] : 0;
        for (int k = 0; k < w; k++) {
          const int idx = base + k * C + l;
          if (k < len) {
            const int q = S.rowptr[r] + k;
            S.sell_col[idx] = S.colind[q];
            S.csr2sell[q] = idx;
            if (S.colind[q] == r) S.sell_diag[r] = idx;
          } else {
            S.sell_col[idx] = r >= 0 ? r : 0;  // padding: value stays 0, column is a valid index
          }
        }
      }
    }
    // packed column stream (symbolic.hpp): 16-bit offsets where a slice's columns span < 65536
    S.slice_meta.assign((size_t)4 * S.nslices, 0);
    std::vector<int> cmin((size_t)S.nslices, 0);
    int64_t units = 0, n16 = 0;
    for (int sl = 0; sl < S.nslices; sl++) {
      const int base = S.slice_ptr[sl], w = (S.slice_ptr[sl + 1] - base) / C;
      int lo = INT_MAX, hi = -1;
      for (int l = 0; l < C; l++) {
        const int r = S.srow[(size_t)sl * C + l];
        if (r < 0) continue;
        lo = std::min(lo, std::min(r, S.colind[S.rowptr[r]]));   // columns ascend in a row; padding uses r itself
        hi = std::max(hi, std::max(r, S.colind[S.rowptr[r] + std::max(rowlen[r], 1) - 1]));
      }
      if (hi < 0) lo = hi = 0;
      const bool narrow = !o.wide_cols && (int64_t)hi - lo < 65536;
      if (units > 0x7fffffffLL - 2 * w - 4) throw std::runtime_error("fk: column stream too large for one GPU; partition the mesh");
      S.slice_meta[4 * (size_t)sl + 0] = base;
      S.slice_meta[4 * (size_t)sl + 1] = w;
      S.slice_meta[4 * (size_t)sl + 2] = (int)units;
      S.slice_meta[4 * (size_t)sl + 3] = narrow ? lo : -1;
      cmin[sl] = lo;
      units += narrow ? 4 * ((w + 3) / 4) : 2 * w;
      n16 += narrow;
    }
    S.n_slices16 = n16;
    S.sell_colz.assign((size_t)units * C, 0);
#pragma omp parallel for schedule(static)
    for (int sl = 0; sl < S.nslices; sl++) {
      const int* m = &S.slice_meta[4 * (size_t)sl];
      const int* src = &S.sell_col[(size_t)m[0]];
      uint16_t* dst = &S.sell_colz[(size_t)m[2] * C];
      const int n = m[1] * C;
      if (m[3] >= 0) {   // step k of lane l -> 16-bit word ((k / 4) * 32 + l) * 4 + k % 4 (steps k >= w of the last group: 0)
        for (int e = 0; e < n; e++) {
          const int k = e / C, l = e % C;
          int c = src[e];
          if (c < m[3] || c - m[3] > 65535) c = m[3];   // padding of an unused lane (value 0): any valid column
          dst[((size_t)(k / 4) * C + l) * 4 + (k % 4)] = (uint16_t)(c - m[3]);
        }
      } else {
        std::memcpy(dst, src, sizeof(int) * (size_t)n);
      }
    }
  };

  if (npe != 4) {   // 10-node (P2) elements: pattern, map and SELL layout only; assembly goes through the contribution plan
    build_sell();
    S.ntiles = 0;
    S.tile_elem_ptr.assign(1, 0);
    S.tile_nnz_ptr.assign(1, 0);
    S.tile_ent_ptr.assign(1, 0);
    return;
  }

  // ---- tiles: rows grouped along a space-filling curve so each tile is a compact blob ---------
  std::vector<int> order((size_t)nr);
  std::iota(order.begin(), order.end(), 0);
  if (!o.natural_order && p != nullptr) {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int n = 0; n < nv; n++)
      for (int d = 0; d < 3; d++) {
        double v = p[3 * (size_t)n + d];
        if (!std::isfinite(v)) throw std::runtime_error("fk: non-finite node coordinate");
        lo[d] = std::min(lo[d], v);
        hi[d] = std::max(hi[d], v);
      }
    double ext = std::max({hi[0] - lo[0], hi[1] - lo[1], hi[2] - lo[2], 1e-300});
    std::vector<uint64_t> key((size_t)nr);
#pragma omp parallel for schedule(static)
    for (int r = 0; r < nr; r++) {
      uint64_t c[3];
      for (int d = 0; d < 3; d++) {
        double f = (p[3 * (size_t)r + d] - lo[d]) / ext;
        c[d] = (uint64_t)std::min(2097151.0, std::max(0.0, f * 2097151.0));
      }
      key[r] = (spread21(c[0]) << 2) | (spread21(c[1]) << 1) | spread21(c[2]);
    }
    std::sort(order.begin(), order.end(), [&](int a, int b) { return key[a] != key[b] ? key[a] < key[b] : a < b; });
  }

  tm.lap("Morton order");
  std::vector<int> tile_of_row((size_t)nv, -1);
  std::vector<int64_t> tile_row_ptr;  // -> order positions
  std::vector<int> tile_nelem;
  {
    std::vector<int> marker((size_t)nt, -1);
    int tile = 0, cur_rows = 0, cur_elems = 0;
    int64_t cur_ents = 0, cur_nnz = 0;
    tile_row_ptr.push_back(0);
    for (int pos = 0; pos < nr; pos++) {
      int r = order[pos];
      int64_t deg = n2e_ptr[r + 1] - n2e_ptr[r];
      if (deg > o.tile_max_elems || 4 * deg > 65535)
        throw std::runtime_error("fk: node " + std::to_string(r) + " has " + std::to_string(deg) +
                                 " incident elements, more than one assembly tile can stage");
      // count-and-mark (a degenerate element listing the node twice is counted once)
      auto claim = [&](int id) {
        int c = 0;
        for (int64_t q = n2e_ptr[r]; q < n2e_ptr[r + 1]; q++) {
          int& m = marker[n2e[q] >> 2];
          if (m != id) { m = id; c++; }
        }
        return c;
      };
      int newel = 0;
      for (int64_t q = n2e_ptr[r]; q < n2e_ptr[r + 1]; q++) newel += (marker[n2e[q] >> 2] != tile);
      if (cur_rows > 0 && (cur_elems + newel > o.tile_max_elems || cur_rows + 1 > o.tile_max_rows ||
                           cur_ents + 4 * deg > 65535 || cur_nnz + rowlen[r] > 65535)) {
        tile_nelem.push_back(cur_elems);
        tile_row_ptr.push_back(pos);
        tile++;
        cur_rows = cur_elems = 0;
        cur_ents = cur_nnz = 0;
      }
      newel = claim(tile);
      cur_elems += newel;
      cur_rows++;
      cur_ents += 4 * deg;
      cur_nnz += rowlen[r];
      tile_of_row[r] = tile;
    }
    if (cur_rows > 0) {
      tile_nelem.push_back(cur_elems);
      tile_row_ptr.push_back(nr);
    }
  }
  tm.lap("tile packing");
  const int T = (int)tile_nelem.size();
  S.ntiles = T;
  S.tile_elem_ptr.assign((size_t)T + 1, 0);
  S.tile_nnz_ptr.assign((size_t)T + 1, 0);
  S.tile_ent_ptr.assign((size_t)T + 1, 0);
  for (int tl = 0; tl < T; tl++) {
    int64_t n_nnz = 0, n_ent = 0;
    for (int64_t pos = tile_row_ptr[tl]; pos < tile_row_ptr[tl + 1]; pos++) {
      int r = order[pos];
      n_nnz += rowlen[r];
      n_ent += 4 * (n2e_ptr[r + 1] - n2e_ptr[r]);
    }
    S.tile_elem_ptr[tl + 1] = S.tile_elem_ptr[tl] + tile_nelem[tl];
    S.tile_nnz_ptr[tl + 1] = S.tile_nnz_ptr[tl] + n_nnz;
    S.tile_ent_ptr[tl + 1] = S.tile_ent_ptr[tl] + n_ent;
    S.max_tile_elems = std::max(S.max_tile_elems, tile_nelem[tl]);
    S.max_tile_nnz = std::max<int>(S.max_tile_nnz, (int)n_nnz);
    S.max_tile_ents = std::max<int>(S.max_tile_ents, (int)n_ent);
  }
  S.tile_elems.resize((size_t)S.tile_elem_ptr[T]);
  S.tile_gidx.resize((size_t)S.tile_nnz_ptr[T]);
  S.tile_coff.resize((size_t)S.tile_nnz_ptr[T] + T);
  S.tile_ents.resize((size_t)S.tile_ent_ptr[T]);

  int bad = 0;
#pragma omp parallel
  {
    std::vector<int> rows;
    std::vector<uint32_t> elems;
    std::vector<int> base, cnt, perm, newpos, fillp, slot_of;
#pragma omp for schedule(dynamic, 16)
    for (int tl = 0; tl < T; tl++) {
      rows.assign(order.begin() + tile_row_ptr[tl], order.begin() + tile_row_ptr[tl + 1]);
      std::sort(rows.begin(), rows.end());
      elems.clear();
      for (int r : rows)
        for (int64_t q = n2e_ptr[r]; q < n2e_ptr[r + 1]; q++) elems.push_back(n2e[q] >> 2);
      std::sort(elems.begin(), elems.end());
      elems.erase(std::unique(elems.begin(), elems.end()), elems.end());
      if ((int)elems.size() != tile_nelem[tl]) {
#pragma omp atomic write
        bad = 1;
        continue;
      }
      // staging slots: the elements this tile OWNS first (ascending id), then the ones it only recomputes
      // (ascending id) -- whole warps of the fused sweep can then skip the per-owner work
      uint32_t* te = S.tile_elems.data() + S.tile_elem_ptr[tl];
      slot_of.resize(elems.size());
      {
        int n_own = 0;
        for (size_t s = 0; s < elems.size(); s++) {
          const int* e = t + 4 * (size_t)elems[s];
          int owner_row = -1;
          for (int i = 0; i < 4 && owner_row < 0; i++) if (e[i] < nr) owner_row = e[i];
          const bool own = owner_row >= 0 && tile_of_row[owner_row] == tl;
          slot_of[s] = own ? 1 : 0;
          n_own += own;
        }
        int a = 0, b = n_own;
        for (size_t s = 0; s < elems.size(); s++) {
          const bool own = slot_of[s] != 0;
          const int slot = own ? a++ : b++;
          slot_of[s] = slot;
          te[slot] = elems[s] | (own ? 0x80000000u : 0u);
        }
      }
      // tile-local numbering of the entries of the tile's rows
      int nloc = 0;
      base.resize(rows.size());
      for (size_t m = 0; m < rows.size(); m++) { base[m] = nloc; nloc += rowlen[rows[m]]; }
      cnt.assign((size_t)nloc + 1, 0);
      for (size_t m = 0; m < rows.size(); m++) {
        int r = rows[m];
        for (int64_t q = n2e_ptr[r]; q < n2e_ptr[r + 1]; q++) {
          uint32_t k = n2e[q] >> 2, i = n2e[q] & 3;
          for (int j = 0; j < 4; j++) cnt[base[m] + (S.e2nnz[(size_t)16 * k + 4 * i + j] - S.rowptr[r]) + 1]++;
        }
      }
      // entries are issued longest contributor list first (stable), so the 32 entries a warp sums
      // together have near-equal trip counts (diagonals ~ node degree, off-diagonals ~ edge degree)
      perm.resize((size_t)nloc);
      std::iota(perm.begin(), perm.end(), 0);
      std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return cnt[a + 1] > cnt[b + 1]; });
      newpos.resize((size_t)nloc);
      for (int q = 0; q < nloc; q++) newpos[perm[q]] = q;
      int* gidx = S.tile_gidx.data() + S.tile_nnz_ptr[tl];
      for (size_t m = 0; m < rows.size(); m++)
        for (int c = 0; c < rowlen[rows[m]]; c++) gidx[newpos[base[m] + c]] = S.rowptr[rows[m]] + c;
      uint16_t* coff = S.tile_coff.data() + S.tile_nnz_ptr[tl] + tl;
      fillp.assign((size_t)nloc + 1, 0);
      for (int q = 0; q < nloc; q++) fillp[q + 1] = fillp[q] + cnt[perm[q] + 1];
      for (int q = 0; q <= nloc; q++) coff[q] = (uint16_t)fillp[q];
      uint16_t* ents = S.tile_ents.data() + S.tile_ent_ptr[tl];
      for (size_t m = 0; m < rows.size(); m++) {
        int r = rows[m];
        for (int64_t q = n2e_ptr[r]; q < n2e_ptr[r + 1]; q++) {  // ascending element order
          uint32_t k = n2e[q] >> 2, i = n2e[q] & 3;
          uint32_t slot = (uint32_t)slot_of[std::lower_bound(elems.begin(), elems.end(), k) - elems.begin()];
          for (int j = 0; j < 4; j++) {
            int loc = newpos[base[m] + (S.e2nnz[(size_t)16 * k + 4 * i + j] - S.rowptr[r])];
            ents[fillp[loc]++] = (uint16_t)((slot << 4) | (uint32_t)sym_index((int)i, j));
          }
        }
      }
    }
  }
  if (bad) throw std::runtime_error("fk: internal error building assembly tiles");
  tm.lap("tile plans");

  build_sell();
  tm.lap("SELL layout");

  // ---- packed device blobs (see symbolic.hpp) ---------------------------------------------------
  {
    // per tile: sorted unique nodes of the staged elements (local numbering of the tile)
    std::vector<int64_t> node_ptr((size_t)T + 1, 0);
    std::vector<int> tnodes;
    {
      std::vector<std::vector<int>> per((size_t)T);
#pragma omp parallel for schedule(dynamic, 16)
      for (int tl = 0; tl < T; tl++) {
        std::vector<int>& v = per[tl];
        const uint32_t* te = S.tile_elems.data() + S.tile_elem_ptr[tl];
        v.reserve((size_t)4 * tile_nelem[tl]);
        for (int q = 0; q < tile_nelem[tl]; q++) {
          const int* e = t + 4 * (size_t)(te[q] & 0x7fffffffu);
          v.push_back(e[0]); v.push_back(e[1]); v.push_back(e[2]); v.push_back(e[3]);
        }
        std::sort(v.begin(), v.end());
        v.erase(std::unique(v.begin(), v.end()), v.end());
      }
      for (int tl = 0; tl < T; tl++) {
        if (per[tl].size() > 65535) throw std::runtime_error("fk: an assembly tile touches more than 65535 nodes");
        node_ptr[tl + 1] = node_ptr[tl] + (int64_t)per[tl].size();
      }
      tnodes.resize((size_t)node_ptr[T]);
#pragma omp parallel for schedule(static)
      for (int tl = 0; tl < T; tl++) std::copy(per[tl].begin(), per[tl].end(), tnodes.begin() + node_ptr[tl]);
    }
    tm.lap("  blobs: tile node lists");
    S.e_pad = (S.max_tile_elems + 15) & ~15;
    if (10 * (int64_t)S.e_pad > 65535) throw std::runtime_error("fk: tile_max_elems too large for 16-bit staging indices");
    S.tile_hdr.assign((size_t)4 * T, 0);
    std::vector<int64_t> off((size_t)T + 1, 0);  // bytes
    for (int tl = 0; tl < T; tl++) {
      const int ne = tile_nelem[tl], nn = (int)(S.tile_nnz_ptr[tl + 1] - S.tile_nnz_ptr[tl]),
                nents = (int)(S.tile_ent_ptr[tl + 1] - S.tile_ent_ptr[tl]), nno = (int)(node_ptr[tl + 1] - node_ptr[tl]);
      const int bi = blob_i_bytes(ne, nno), bg = blob_g_bytes(ne, nno), b = blob_b_bytes(nn, nents);
      S.max_blob_i = std::max(S.max_blob_i, bi);
      S.max_blob_g = std::max(S.max_blob_g, bg);
      S.max_blob_b = std::max(S.max_blob_b, b);
      S.max_tile_nodes = std::max(S.max_tile_nodes, nno);
      off[tl + 1] = off[tl] + bi + bg + b;
      if (off[tl + 1] / 16 > 0x7fffffffLL) throw std::runtime_error("fk: assembly plan too large for one GPU; partition the mesh");
      S.tile_hdr[4 * (size_t)tl] = (int)(off[tl] / 16);
      S.tile_hdr[4 * (size_t)tl + 1] = bi;
      S.tile_hdr[4 * (size_t)tl + 2] = bg;
      S.tile_hdr[4 * (size_t)tl + 3] = b;
    }
    S.tile_blob.assign((size_t)(off[T] / 4), 0u);
    tm.lap("  blobs: allocate + zero");
#pragma omp parallel for schedule(static)
    for (int tl = 0; tl < T; tl++) {
      const int ne = tile_nelem[tl], nn = (int)(S.tile_nnz_ptr[tl + 1] - S.tile_nnz_ptr[tl]),
                nents = (int)(S.tile_ent_ptr[tl + 1] - S.tile_ent_ptr[tl]), nno = (int)(node_ptr[tl + 1] - node_ptr[tl]);
      unsigned char* dst = reinterpret_cast<unsigned char*>(S.tile_blob.data()) + off[tl];
      const int* nodes = tnodes.data() + node_ptr[tl];
      // section I: sizes, global node ids, element ids
      {
        int* hd = reinterpret_cast<int*>(dst);
        hd[0] = ne; hd[1] = nno; hd[2] = nn; hd[3] = nents;
        hd[4] = S.tile_hdr[4 * (size_t)tl] + S.tile_hdr[4 * (size_t)tl + 1] / 16;
        hd[5] = S.tile_hdr[4 * (size_t)tl + 1]; hd[6] = S.tile_hdr[4 * (size_t)tl + 2]; hd[7] = S.tile_hdr[4 * (size_t)tl + 3];
        dst += 32;
      }
      std::memcpy(dst, nodes, sizeof(int) * (size_t)nno);
      dst += pad16(4 * nno);
      const uint32_t* te = S.tile_elems.data() + S.tile_elem_ptr[tl];
      std::memcpy(dst, te, sizeof(uint32_t) * (size_t)ne);
      dst += pad16(4 * ne);
      // section G: node coordinates (SoA), local connectivity
      for (int d = 0; d < 3; d++) {
        double* c = reinterpret_cast<double*>(dst);
        for (int i = 0; i < nno; i++) c[i] = p ? p[3 * (size_t)nodes[i] + d] : 0.0;
        dst += pad16(8 * nno);
      }
      uint16_t* loc = reinterpret_cast<uint16_t*>(dst);
      for (int q = 0; q < ne; q++) {
        const int* e = t + 4 * (size_t)(te[q] & 0x7fffffffu);
        for (int i = 0; i < 4; i++) loc[4 * q + i] = (uint16_t)(std::lower_bound(nodes, nodes + nno, e[i]) - nodes);
      }
      dst += pad16(8 * ne);
      // section B: gather plan
      {  // value indices in the device (SELL) layout; the host-side plan (tile_gidx) keeps CSR indices
        int* gd = reinterpret_cast<int*>(dst);
        const int* gs = S.tile_gidx.data() + S.tile_nnz_ptr[tl];
        for (int q = 0; q < nn; q++) gd[q] = S.csr2sell[gs[q]];
      }
      dst += pad16(4 * nn);
      std::memcpy(dst, S.tile_coff.data() + S.tile_nnz_ptr[tl] + tl, sizeof(uint16_t) * ((size_t)nn + 1));
      dst += pad16(2 * (nn + 1));
      {  // contributor lists as direct indices into the staging array: pair * e_pad + slot (< 10 * e_pad <= 65535)
        const uint16_t* es = S.tile_ents.data() + S.tile_ent_ptr[tl];
        uint16_t* ed = reinterpret_cast<uint16_t*>(dst);
        for (int q = 0; q < nents; q++) ed[q] = (uint16_t)((es[q] & 15u) * (uint32_t)S.e_pad + (es[q] >> 4));
      }
    }
  }

  tm.lap("  blobs: fill");
  if (std::getenv("FK_DEBUG_CHECKSUM")) {   // checksums of what the device receives: a regression guard for changes of this file
    auto fnv = [](const void* d, size_t n) {
      uint64_t h = 1469598103934665603ULL;
      const unsigned char* b = static_cast<const unsigned char*>(d);
      for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ULL; }
      return h;
    };
    std::fprintf(stderr, "fk symbolic: checksums blob %016llx hdr %016llx colz %016llx meta %016llx csr2sell %016llx e2nnz %016llx\n",
                 (unsigned long long)fnv(S.tile_blob.data(), 4 * S.tile_blob.size()), (unsigned long long)fnv(S.tile_hdr.data(), 4 * S.tile_hdr.size()),
                 (unsigned long long)fnv(S.sell_colz.data(), 2 * S.sell_colz.size()), (unsigned long long)fnv(S.slice_meta.data(), 4 * S.slice_meta.size()),
                 (unsigned long long)fnv(S.csr2sell.data(), 4 * S.csr2sell.size()), (unsigned long long)fnv(S.e2nnz.data(), 4 * S.e2nnz.size()));
  }
}

// ------------------------------------------------------------------------------------------------
// rescale plan (symbolic.hpp: RescalePlan)
// ------------------------------------------------------------------------------------------------
void build_rescale_plan(const Symbolic& S, int window, RescalePlan& R, bool skip_diag) {
  const int nt = S.nt, npe = S.npe, npe2 = S.npe * S.npe;
  const int64_t nnz = S.nnz;
  const int shift = npe == 4 ? 28 : 26;   // element id below, local pair (10 or 55 of them) above
  R.pair_shift = shift;
  if ((int64_t)nt >= ((int64_t)1 << shift)) throw std::runtime_error("fk: too many elements on one GPU for the contribution plan");
  // index of the unordered local pair (a <= b) in the packed symmetric layout: row a starts at a npe - a (a - 1) / 2
  auto pair_index = [npe](int a, int b) { return a * npe - a * (a - 1) / 2 + (b - a); };
  if (window < 32) window = 32;
  window = (window + 31) & ~31;
  const int* e2 = S.e2nnz.data();
  // canonical entry of local pair (a, b), a <= b, of element k; other = its mirror (or -1)
  auto canon = [&](int64_t k, int a, int b, int& other) -> int {
    const int e1 = e2[(int64_t)npe2 * k + npe * a + b], eb = e2[(int64_t)npe2 * k + npe * b + a];
    if (a == b || e1 == eb) { other = -1; return e1; }
    if (e1 >= 0 && eb >= 0) { other = std::max(e1, eb); return std::min(e1, eb); }
    other = -1;
    return e1 >= 0 ? e1 : eb;
  };
  // ---- contributions per canonical entry ----------------------------------------------------------
  std::vector<int> cnt((size_t)nnz, 0), mirror((size_t)nnz, -1);
#pragma omp parallel for schedule(static)
  for (int64_t k = 0; k < nt; k++)
    for (int a = 0; a < npe; a++)
      for (int b = a + (skip_diag ? 1 : 0); b < npe; b++) {
        int other;
        const int c = canon(k, a, b, other);
        if (c < 0) continue;
#pragma omp atomic
        cnt[c]++;
        if (other >= 0) {
#pragma omp atomic write
          mirror[c] = other;
        }
      }
  // ---- items in SELL order of their canonical entry ------------------------------------------------
  std::vector<int> at((size_t)S.nnz_sell, -1);   // SELL index -> canonical CSR index
#pragma omp parallel for schedule(static)
  for (int64_t e = 0; e < nnz; e++)
    if (cnt[e] > 0) at[S.csr2sell[e]] = (int)e;
  std::vector<int> items;
  items.reserve((size_t)(nnz / 2 + S.n_rows + 64));
  for (int64_t q = 0; q < S.nnz_sell; q++)
    if (at[q] >= 0) items.push_back(at[q]);
  std::vector<int>().swap(at);
  const int64_t ni = (int64_t)items.size();
  R.n_items = ni;
  R.nslices = (int)((ni + 31) / 32);
  // ---- windows sorted by list length, longest first (stable) ---------------------------------------
  const int64_t nwin = (ni + window - 1) / window;
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t w = 0; w < nwin; w++) {
    const int64_t b = w * window, e = std::min(ni, b + window);
    std::stable_sort(items.begin() + b, items.begin() + e, [&](int x, int y) { return cnt[x] > cnt[y]; });
  }
  // ---- slices ----------------------------------------------------------------------------------------
  R.slice_ptr.assign((size_t)R.nslices + 1, 0u);
  R.max_width = 0;
  uint64_t run = 0;
  for (int s = 0; s < R.nslices; s++) {
    int w = 0;
    for (int64_t i = 32 * (int64_t)s; i < std::min(ni, 32 * (int64_t)s + 32); i++) w = std::max(w, cnt[items[i]]);
    R.max_width = std::max(R.max_width, w);
    R.slice_ptr[s] = (uint32_t)run;
    run += (uint64_t)w;
    if (run > 0xffffffffULL) throw std::runtime_error("fk: rescale plan too large for one GPU; partition the mesh");
  }
  R.slice_ptr[R.nslices] = (uint32_t)run;
  R.n_slots = 32 * (int64_t)run;
  R.dst1.assign((size_t)32 * R.nslices, -1);
  R.dst2.assign((size_t)32 * R.nslices, -1);
  std::vector<int64_t> slot_of((size_t)nnz, -1);   // canonical CSR index -> first slot (32 slice_ptr + lane)
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < ni; i++) {
    const int c = items[i];
    R.dst1[i] = S.csr2sell[c];
    R.dst2[i] = mirror[c] >= 0 ? S.csr2sell[mirror[c]] : -1;
    slot_of[c] = 32 * (int64_t)R.slice_ptr[i >> 5] + (i & 31);
  }
  // ---- contribution lists ------------------------------------------------------------------------------
  R.cref.assign((size_t)R.n_slots, 0xffffffffu);
  std::vector<int> fill((size_t)nnz, 0);
  int64_t total = 0;
#pragma omp parallel for schedule(static) reduction(+ : total)
  for (int64_t k = 0; k < nt; k++)
    for (int a = 0; a < npe; a++)
      for (int b = a + (skip_diag ? 1 : 0); b < npe; b++) {
        int other;
        const int c = canon(k, a, b, other);
        if (c < 0) continue;
        int f;
#pragma omp atomic capture
        f = fill[c]++;
        R.cref[(size_t)(slot_of[c] + 32 * (int64_t)f)] = (uint32_t)k | ((uint32_t)pair_index(a, b) << shift);
        total++;
      }
  R.n_contrib = total;
  // ascending element order inside every list (the fill order above depends on the thread schedule)
#pragma omp parallel
  {
    std::vector<uint32_t> buf;
#pragma omp for schedule(static)
    for (int64_t i = 0; i < ni; i++) {
      const int c = items[i], n = cnt[c];
      if (n < 2) continue;
      const int64_t s0 = slot_of[c];
      buf.resize((size_t)n);
      for (int e = 0; e < n; e++) buf[e] = R.cref[(size_t)(s0 + 32 * (int64_t)e)];
      const uint32_t emask = (1u << shift) - 1u;
      std::sort(buf.begin(), buf.end(), [emask](uint32_t x, uint32_t y) {
        const uint32_t kx = x & emask, ky = y & emask;
        return kx != ky ? kx < ky : x < y;
      });
      for (int e = 0; e < n; e++) R.cref[(size_t)(s0 + 32 * (int64_t)e)] = buf[e];
    }
  }
}

}  // namespace fk
