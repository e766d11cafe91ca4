// Symbolic phase of the assembly / SpMV engine (host side, runs once per mesh).
//
// The reference rebuilds its sparsity on every assembly (16 x sparse()+`+=`, src/FEM3D.jl:47-53;
// setFromTriplets, src/FEMc.cpp:111-129) or keeps a triplet index list (NewtonRaphson.jl:114-127).
// Here the pattern, the element->nnz map and the gather plan of the atomic-free assembly kernel are
// computed once and reused by every numeric assembly.
#pragma once
#include <cstdint>
#include <vector>

namespace fk {

struct SymbolicOpts {
  int tile_max_elems = 512;   // elements staged per tile (<= 4096: 12-bit slot in a gather entry)
  int tile_max_rows = 256;
  bool natural_order = false;
  int sell_sigma = 1024;      // SELL-32-sigma: rows are sorted by length inside windows of sigma rows (1 = input order)
  int n_owned = -1;  // rows [0, n_owned) get a pattern (multi-GPU: owned rows first); -1 = all
  bool wide_cols = false;  // true: every slice keeps 32-bit columns in the packed column stream (A/B measurements)
  int npe = 4;             // nodes per element: 4 = P1 tetrahedra (and 2-D P1 as (n0,n1,n2,n2)); 10 = P2 tetrahedra (pattern, map,
                           // SELL layout and the contribution plan only: no assembly tiles)
};

// Gather plan of one assembly tile = a set of rows whose every incident element is staged once in
// shared memory; every CSR entry of those rows is then summed by ONE thread from the staged local
// matrices in ascending element order (deterministic, no atomics).
struct Symbolic {
  int nv = 0, nt = 0, n_rows = 0, npe = 4;
  int64_t nnz = 0;
  std::vector<int> rowptr;   // nv+1
  std::vector<int> colind;   // nnz, ascending per row
  std::vector<int> diag;     // n_rows: CSR index of (r,r)
  std::vector<int> e2nnz;    // npe^2 * nt (16 nt for tetrahedra): index of (t[i,k], t[j,k]); -1 if the row is not assembled here

  int ntiles = 0;
  std::vector<int64_t> tile_elem_ptr;  // ntiles+1 -> tile_elems
  std::vector<uint32_t> tile_elems;    // element id | (owner bit 31): owner = tile holding row t[0,k]
  std::vector<int64_t> tile_nnz_ptr;   // ntiles+1 -> tile_gidx ; coff index = tile_nnz_ptr[T] + T
  std::vector<int> tile_gidx;          // global CSR index of each tile-local entry
  std::vector<uint16_t> tile_coff;     // per tile: (tile nnz + 1) offsets into the tile's entries
  std::vector<int64_t> tile_ent_ptr;   // ntiles+1 -> tile_ents
  std::vector<uint16_t> tile_ents;     // (slot << 4) | sym(i,j): staged element slot + local pair
  int max_tile_elems = 0, max_tile_nnz = 0, max_tile_ents = 0;

  // Device form of the plan: one packed, 16-byte aligned blob per tile in three sections, so the kernel fetches ALL
  // of a tile's inputs with three bulk (TMA) copies and phase 1 issues no dependent global gathers.
  // Section I (ids, fetched two tiles ahead: it feeds the coefficient / potential gathers of the next tile) =
  //   {ne, nno, nn, nents, offset of G in 16-byte units, bytes of I, bytes of G, bytes of B} (8 x i32) | gnode (i32 x nno: global node ids) | ek (u32 x ne: element id, bit 31 = owner).
  // Section G (geometry, one tile ahead) = x | y | z (f64 x nno each: coordinates of the tile's nodes, local
  //   numbering) | loc (4 x u16 x ne: local node ids of each staged element).
  // Section B (gather plan, one tile ahead) = gidx (i32 x nn) | coff (u16 x (nn+1)) | ents (u16 x nents, as
  //   pair * e_pad + slot).
  // Every sub-section is padded to 16 bytes.
  // tile_hdr[T] = {offset of I in 16-byte units, bytes of I, bytes of G, bytes of B}; G follows I, B follows G.
  std::vector<int> tile_hdr;        // 4 * ntiles
  std::vector<uint32_t> tile_blob;  // packed words
  int max_blob_i = 0, max_blob_g = 0, max_blob_b = 0;  // bytes
  int max_tile_nodes = 0;
  int e_pad = 16;                      // staging stride of the assembly kernel: max_tile_elems rounded up to 16

  // Device storage of every matrix on this pattern: SELL-32-sigma (sliced ELLPACK).  Slice s holds the
  // 32 rows srow[32s .. 32s+31] (-1 = padding lane); its entries sit column-major from slice_ptr[s]:
  // entry k of lane l at slice_ptr[s] + 32 k + l, k < width(s) = (slice_ptr[s+1] - slice_ptr[s]) / 32, in
  // ascending column order per row (the CSR order), short rows padded with (value 0, column = own row).
  // One thread owns one row: value / column loads are perfectly coalesced and, on meshes whose numbering
  // has locality, so are the x gathers (neighbouring rows have neighbouring columns).
  int nslices = 0;
  int64_t nnz_sell = 0;              // padded entry count = slice_ptr[nslices]
  std::vector<int> slice_ptr;        // nslices + 1
  std::vector<int> srow;             // 32 * nslices
  std::vector<int> sell_col;         // nnz_sell (host copy, 32-bit; the device reads the packed stream below)
  // Packed column stream of the device SpMV.  A slice whose columns span < 65536 (every slice of a mesh numbered
  // with locality) stores them as 16-bit offsets from the slice's smallest column: 10 instead of 12 bytes per
  // entry on the matrix stream that bounds a PCG iteration.  Any other slice keeps 32-bit columns; the choice is
  // per slice = per warp, so it costs no divergence.  The stream is addressed in 64-byte units.  A 16-bit slice of
  // width w is stored in groups of four steps: lane l reads the offsets of steps 4 g .. 4 g + 3 of its row as ONE
  // 8-byte word at 16-bit index ((g * 32 + l) * 4) (a warp reads a group with one coalesced 256-byte load and has
  // four gathers to issue at once); it takes 4 ceil(w / 4) units, steps >= w of the last group are 0.  A 32-bit slice
  // keeps the plain SELL order (step k of lane l at 32-bit index 32 k + l) and takes 2 w units.
  // slice_meta[4 sl ..] = {slice_ptr[sl], w, offset in 64-byte units, column base (-1: 32-bit columns)}.
  std::vector<int> slice_meta;       // 4 * nslices
  std::vector<uint16_t> sell_colz;   // 32 * (units of all slices)
  int64_t n_slices16 = 0;            // slices stored with 16-bit columns
  std::vector<int> csr2sell;         // nnz: CSR index -> SELL index
  std::vector<int> sell_diag;        // n_rows: SELL index of (r,r), -1 if absent
};

// Plan of the RESCALE assembly (api.cu: ensure_geo): the stiffness matrix of a Picard step is A = sum_k s_k G_k with
// s_k = VE[k] f[k] the only thing that changes between steps (NewtonRaphson.jl:114-127,153-161 keeps the local
// matrices Ak and rescales them by mu every iteration).  The plan lists, for every stored entry pair {(r,c), (c,r)},
// its contributions (element, local pair) in ascending element order -- the summation order of the tile kernel and of
// Eigen's setFromTriplets.  An item = one canonical entry (the smaller CSR index of the pair, or the only one stored on
// this partition); items follow the SELL order of their canonical entry, are sorted by list length inside windows
// (longest first) and are cut into slices of 32: lane l of slice s owns item 32 s + l, its e-th contribution sits at
// 32 (slice_ptr[s] + e) + l, so a warp reads 32 consecutive geometric factors / element ids per step.
struct RescalePlan {
  int64_t n_items = 0;             // canonical entries
  int nslices = 0;
  int64_t n_contrib = 0;           // contributions listed (10 per element on an unpartitioned mesh)
  int64_t n_slots = 0;             // padded contribution slots = 32 * slice_ptr[nslices]
  int max_width = 0;
  int pair_shift = 28;             // cref = element id | local pair << pair_shift (28: 10 pairs of a tetrahedron; 26: 55 pairs of a P2 element)
  std::vector<uint32_t> slice_ptr; // nslices + 1, in units of 32 slots
  std::vector<int> dst1, dst2;     // 32 * nslices: SELL index of the item's entry / of its mirror (-1: none; dst1 -1: padding lane)
  std::vector<uint32_t> cref;      // n_slots: element id | local pair (sym_index) << 28; 0xffffffff = padding
};
// skip_diag: leave the pairs (a, a) out -- the caller obtains the diagonal from the row sums (local matrices with zero row
// sums: stiffness), 6 instead of 10 contributions per tetrahedron
void build_rescale_plan(const Symbolic& S, int window, RescalePlan& R, bool skip_diag = false);

// p: 3 x nv (may be null when natural_order), t: 4 x nt 0-based.  Throws std::runtime_error on bad input.
void build_symbolic(const double* p, const int* t, int nv, int nt, const SymbolicOpts& o, Symbolic& S);

// byte sizes of the packed blob sections (shared by the host packer and the kernel)
#ifdef __CUDACC__
#define FK_HD __host__ __device__
#else
#define FK_HD
#endif
FK_HD inline int pad16(int bytes) { return (bytes + 15) & ~15; }
FK_HD inline int blob_i_bytes(int ne, int nno) { return 32 + pad16(4 * nno) + pad16(4 * ne); }
FK_HD inline int blob_g_bytes(int ne, int nno) { return 3 * pad16(8 * nno) + pad16(8 * ne); }
FK_HD inline int blob_b_bytes(int nn, int nents) { return pad16(4 * nn) + pad16(2 * (nn + 1)) + pad16(2 * nents); }

// index of the unordered local pair (i,j) in the 10-entry symmetric staging layout
inline int sym_index(int i, int j) {
  if (i > j) { int s = i; i = j; j = s; }
  static const int base[4] = {0, 4, 7, 9};
  return base[i] + (j - i);
}

}  // namespace fk
