/* femeko_cuda.h -- C ABI of libfemeko_cuda.so, the B200 (sm_100a, fp64) hot path of Femeko.jl.
 *
 * Boundary convention = the reference's own ccall pattern (Femeko_jl_C++/Magnetostatics/run.jl:50-79,
 * magnetostatics.cpp:49-95): plain pointers + sizes, caller owns every buffer, matrices are
 * column-major (Julia), node/element indices are 0-based `int` (the caller shifts, run.jl:50-52),
 * pointers are only read/written during the call.  Persistent device state lives behind an opaque
 * handle (the way the reference's C++ LL object copies p,t into owned storage, LL.cpp:10-13).
 *
 * Every new entry point returns int (0 = ok, != 0 = error, text via fk_last_error()); nothing throws
 * across the boundary (the reference throws std::runtime_error through extern "C",
 * magnetostatics.cpp:29-38, which aborts the Julia process).
 *
 * Pointer kinds:  *_h / unsuffixed = HOST memory (pageable is fine);  names ending in _dev and every
 * pointer argument of an fk_*_dev function = DEVICE memory on the handle's GPU (e.g. a torch
 * tensor's data_ptr()).  Matrices assembled by the handle live in numbered value slots that share
 * one CSR pattern (slot 0..FK_NUM_SLOTS-1), so M + dt*A (heat3D.jl:85) or A + At
 * (NewtonRaphson.jl:171) are value-only operations.
 */
#ifndef FEMEKO_CUDA_H
#define FEMEKO_CUDA_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fk_handle_s* fk_handle;

#define FK_NUM_SLOTS 4
#define FK_MAX_SPLINES 8

/* spline boundary behaviour = Dierckx.jl `bc` strings (NewtonRaphson.jl:58, manyBodyIntersecting.jl:191) */
enum { FK_BC_EXTRAPOLATE = 0, FK_BC_ZERO = 1, FK_BC_ERROR = 2, FK_BC_NEAREST = 3 };

/* Options of the symbolic phase (all 0 = defaults). */
typedef struct fk_create_opts {
  int tile_max_elems; /* elements staged per assembly tile (<= 4096); default 512             */
  int tile_max_rows;  /* rows per assembly tile; default 256                                  */
  int natural_order;  /* 1: tile rows in input order instead of Morton order of coordinates   */
  int sell_sigma;     /* SELL-32-sigma: rows sorted by length inside windows of sigma rows; default 1024 */
  int device;         /* CUDA device ordinal (0 = the first device, also what an all-zero struct selects); -1 = the
                         calling thread's current device (what opts == NULL selects)            */
  int n_owned;        /* multi-GPU partitions: rows [0, n_owned) are owned (assembled, solved),
                         rows [n_owned, nv) are halo copies; 0 = all rows owned                 */
  int asm_cfg;        /* assembly launch shape {threads, min CTAs/SM}: 0 = default ({256,2}); 1 {256,3} 2 {256,2} 3 {384,2} 4 {512,2} 5 {128,6} 6 {512,1} 7 {768,1} */
  int spmv_ctas_per_sm; /* cap on resident SpMV CTAs per SM (0 = as many as fit)               */
  int sell_wide_cols; /* 1: keep 32-bit columns in every SELL slice (default 0: slices whose columns span
                         < 65536 store 16-bit offsets, ~10 instead of 12 bytes per matrix entry)  */
  int nodes_per_element; /* 0 or 4: P1 tetrahedra, t is 4 x nt; 10: P2 tetrahedra, t is 10 x nt in Gmsh's order (4 vertices,
                         then the mid nodes of the edges (0,1) (1,2) (0,2) (0,3) (2,3) (1,3)): pattern, SpMV, PCG and
                         constraints as for P1, matrices through fk_assemble_stiffness_p2 / fk_assemble_mass_p2 */
} fk_create_opts;

/* PCG options.  Defaults (all 0) reproduce Eigen::ConjugateGradient's: tol = DBL_EPSILON relative
 * residual, max_iter = 2n (IterativeSolverBase.h:251,368), x0 = 0, Jacobi preconditioner. */
typedef struct fk_pcg_opts {
  double tol;           /* relative residual ||r|| / ||b||; <= 0 -> DBL_EPSILON; use tol_disable to run a fixed count */
  int max_iter;         /* <= 0 -> 2n                                                          */
  int warm_start;       /* 1: x holds the initial guess (solveWithGuess); 0: x0 = 0            */
  int check_every;      /* iterations per host convergence poll / CUDA-graph chunk; default 32 */
  int tol_disable;      /* 1: never stop on the residual, run exactly max_iter iterations      */
  int no_precond;       /* 1: identity preconditioner (IterativeSolvers.cg call sites, run.jl:141) */
  int recurrence;       /* 0: Eigen's two-reduction recurrence (ConjugateGradient.h:59-82), three kernels per iteration;
                           1: single-reduction (Chronopoulos-Gear) recurrence -- same Krylov iterates in exact arithmetic,
                           ONE fused reduction and two kernels per iteration (w = A M^-1 r with (w,u) fused; then p, s, x, r,
                           u in one pass); on partitions the halo push and the wait ride inside the same two kernels      */
  int project_nullspace;/* 1: singular Neumann systems (constants in the null space): b <- b - (sum b / sum w) w before the
                           solve and x <- x - (w.x / sum w) after it, w = fk_set_nullspace weights (default ones) -- with
                           w = lagrange(mesh) this is the solution of the reference's bordered system [A C; C' 0]
                           (magneticField.jl:99-107, NewtonRaphson.jl:166).  During the iteration the mean of r (rounding
                           drift out of range(A)) is removed one iteration late, so the recursive residual can reach tol.
                           2: the same without the final shift of x (the iterate Eigen's loop would return)               */
} fk_pcg_opts;

const char* fk_last_error(void);
/* Host threads (OpenMP) of the symbolic phase (fk_create, the rescale plan); n <= 0 only queries.  Returns the count in
 * effect.  Launchers that pin OMP_NUM_THREADS = 1 per process (torch.distributed.run) make fk_create of a 50 M-tet
 * partition take minutes; a driver with one process per GPU gives each process its share of the cores here.        */
int fk_set_host_threads(int n);
int fk_version(void);
/* Host buffers: every host-pointer entry point accepts pageable memory (a Julia Array); its copies then run at about half the
 * speed of pinned memory (bench.py: e2e.pageable).  fk_host_register page-locks a caller's array in place (cudaHostRegister,
 * portable across devices) so that later calls with pointers inside it copy at pinned speed; unregister before the array is
 * freed (in Julia: keep it alive with GC.@preserve / a global until fk_host_unregister).  No counterpart in the reference. */
int fk_host_register(void* ptr, long long bytes);
int fk_host_unregister(void* ptr);

/* ---- lifecycle -------------------------------------------------------------------------------
 * fk_create copies p (3 x nv) and t (4 x nt, 0-based) to the GPU, runs the symbolic phase
 * (canonical CSR pattern = structural node adjacency, ascending columns == Eigen setFromTriplets
 * pattern of FEMc.cpp:111-129; element->nnz map; assembly tiles) and allocates the value slots.
 * VE may be NULL: volumes are then computed as |AB.(ACxAD)|/6 (src/mesh.jl:484-505).            */
int fk_create(const double* p, const int* t, int nv, int nt, const double* VE,
              const fk_create_opts* opts, fk_handle* out);
int fk_destroy(fk_handle h);
/* stream: a cudaStream_t passed as void* (NULL = legacy default stream).  All work of the handle is
 * issued on it.                                                                                  */
int fk_set_stream(fk_handle h, void* stream);
int fk_synchronize(fk_handle h);
/* change a launch option after creation (tuning sweeps): what 0 = asm_cfg, 1 = spmv_ctas_per_sm (values
 * as in fk_create_opts), 2 = pin the PCG vectors in L2 with a persisting access-policy window (1 = on,
 * the default; 0 = off).  Matrix values do not depend on these.  3 = profiling experiments only: bit 0 / 1
 * skip phase 1 / 2 of the assembly kernel (matrix values are then garbage).  4 = SpMV launch shape
 * {entries per batch, min CTAs per SM}: 0 {4,8} 1 {4,6} 2 {8,5} 3 {8,4} (default) 4 {16,3} 5 {16,2}; results do
 * not depend on it.  5 = partitioned handles: 1 (default) overlaps the halo exchange with the SpMV of the rows that
 * read no halo column, 0 serialises them; 6 = CTA slots per SM that SpMV leaves free for the halo branch (default 0); 7 = programmatic dependent launch between the kernels of a single-GPU
 * PCG iteration (default 0: measured 2 % slower on B200).  8 = L2 policy of lines outside the persisting window (0 streaming,
 * 1 normal); 9 = cap on the PCG vectors pinned in L2 (0 = as many as may persist); 11 = device-side timeline of the
 * PCG kernels: keep the records of the last `value` iterations of every solve (0 = off), read with fk_trace_read.   */
int fk_tune(fk_handle h, int what, int value);
/* Timeline records of the last solve (fk_tune 11): record (it % cap) * 4 + k of kernel k (0 SpMV, 1 update, 2 direction)
 * is four 64-bit words {earliest CTA start, latest CTA end, longest cross-GPU wait, wait of CTA 0}, ns of the GPU's
 * %globaltimer; an unused record has start = 2^64 - 1.  Copies min(max_records, 4 cap) records.                    */
/* Rescale assembly (fk_tune 10: 0 never, 1 = from the second Picard sweep of the handle on (default), 2 = at once; fk_tune 12: its
 * launch shape): the geometric factors grad(lambda_i).grad(lambda_j) of every contribution are computed once and every
 * later stiffness assembly / Picard sweep streams sum_k VE[k] f[k] g the way NewtonRaphson.jl:114-127,153-161 keeps Ak and
 * rescales it by mu.  Every off-diagonal entry is bit-identical to the tile kernel's; the diagonal is taken from the row
 * sums (a_rr = -sum_{c != r} a_rc: the local stiffness matrices have zero row sums), which saves the 4 diagonal
 * contributions of every element and differs from the summed diagonal by rounding only (fk_tune 18 = 0: sum it instead,
 * all entries bit-identical).  fk_tune 13: sort window of the plan; 15: what assembly work does about the PCG's
 * persisting-L2 set-aside (default 2: release it, the next solve restores it); 17: field recovery beside the streamed
 * sums (16 a + b CTAs per SM, default 4 + 1; 0 = one after the other).  fk_geo_stats builds the plan if needed and reports
 * its size: entry pairs, contributions, padded slots, device bytes (12 B per slot + 8 B per pair + 8 B per element).   */
int fk_geo_stats(fk_handle h, long long* items, long long* contributions, long long* slots, long long* bytes);
int fk_trace_read(fk_handle h, unsigned long long* out, int max_records, int* n_records);
/* number of kernels of this library launched through the handle so far (graph nodes included) */
int fk_launch_count(fk_handle h, long long* n);

/* ---- symbolic artefacts (host copies) ------------------------------------------------------- */
int fk_sizes(fk_handle h, int* nv, int* nt, long long* nnz, int* ntiles);
int fk_csr_pattern(fk_handle h, int* rowptr /* nv+1 */, int* colind /* nnz */);
/* map[16k + 4i + j] = CSR value index of entry (t[i,k], t[j,k]); the deterministic scatter map
 * (reference analogue: rowIDs/colIDs of NewtonRaphson.jl:115-127).                              */
int fk_elem2nnz(fk_handle h, int* map /* 16 nt */);
int fk_volumes(fk_handle h, double* VE /* nt */);
/* assembly-plan statistics: elements staged over all tiles (>= nt; the excess is halo recompute) */
int fk_tile_stats(fk_handle h, long long* staged_elems, int* max_elems, int* max_nnz, int* max_ents);
/* device matrix layout (SELL-32-sigma): padded entries, slices, slices stored with 16-bit column offsets, and the
 * bytes of the packed column stream one SpMV reads                                                          */
int fk_sell_stats(fk_handle h, long long* nnz_sell, long long* n_slices, long long* n_slices16, long long* col_bytes);
/* device pointers of handle-owned arrays, for zero-copy interop (torch.from_blob-style):
 * which: 0 rowptr(int) 1 colind(int) 2 VE(double) 3 mu(double) 4 Hnorm(double) 5 Hfield(double 3 x nt)
 *        6 u(double nv) 7 dmu(double) 8 csr2sell(int nnz) 16+s values of slot s.  Matrix values live on
 *        the device in the SELL-32-sigma layout of the symbolic phase: the CSR entry q of fk_csr_pattern
 *        is values[csr2sell[q]] (fk_matrix_get / fk_assemble_* return plain CSR order on the host).   */
int fk_device_ptr(fk_handle h, int which, void** ptr);

/* ---- symbolic phase alone (host only, needs no GPU): lets the plan be inspected / validated ---
 * Arrays may be NULL to skip them.  Sizes: tile_*_ptr ntiles+1; tile_elems n_tile_elems (bit 31 =
 * this tile owns the element's per-element outputs); tile_gidx nnz_tiles (CSR indices); tile_coff
 * nnz_tiles+ntiles (tile T's offsets start at tile_nnz_ptr[T]+T); tile_ents 16*nt entries (slot<<4 |
 * pair); SELL-32-sigma device layout: slice_ptr nslices+1, srow 32*nslices (-1 = padding lane),
 * sell_col nnz_sell, csr2sell nnz.                                                               */
typedef struct fk_symbolic_s* fk_symbolic;
int fk_symbolic_create(const double* p, const int* t, int nv, int nt, const fk_create_opts* opts, fk_symbolic* out);
int fk_symbolic_sizes(fk_symbolic s, long long* nnz, int* ntiles, long long* n_tile_elems, long long* nnz_tiles,
                      long long* n_ents, int* nslices, long long* nnz_sell);
int fk_symbolic_export(fk_symbolic s, int* rowptr, int* colind, int* diag, int* e2nnz, long long* tile_elem_ptr,
                       unsigned int* tile_elems, long long* tile_nnz_ptr, int* tile_gidx, unsigned short* tile_coff,
                       long long* tile_ent_ptr, unsigned short* tile_ents, int* slice_ptr, int* srow, int* sell_col,
                       int* csr2sell);
/* packed column stream read by the device SpMV (two calls: sizes with NULL arrays, then the arrays):
 * slice_meta[4 sl ..] = {value base, width, offset in 64-byte units, column base (-1: 32-bit columns)},
 * colz = 16-bit words; a slice with column base >= 0 stores offsets from it in groups of four steps
 * (step k of lane l at ((k / 4) * 32 + l) * 4 + k % 4, 4 ceil(width / 4) units), any other slice 32 x width
 * 32-bit columns (2 width units).  n_slices16 = slices stored with 16-bit columns.                          */
int fk_symbolic_colz(fk_symbolic s, long long* n_colz, long long* n_slices16, int* slice_meta /* 4 x nslices */,
                     unsigned short* colz /* n_colz */);
/* plan of the rescale assembly (symbolic.hpp: RescalePlan): item 32 s + l = lane l of slice s, its e-th contribution at
 * cref[32 (slice_ptr[s] + e) + l] = element | local pair << 28 (0xffffffff: padding); dst1 / dst2 = SELL index of the
 * item's entry and of its mirror (-1: none).  Call with NULL arrays for the sizes first.                            */
int fk_symbolic_rescale(fk_symbolic s, long long* n_items, int* n_slices, long long* n_slots, unsigned int* slice_ptr /* n_slices+1 */,
                        int* dst1 /* 32 n_slices */, int* dst2, unsigned int* cref /* n_slots */);
int fk_symbolic_destroy(fk_symbolic s);

/* ---- assembly: stiffnessMatrix / massMatrix / tangentialStiffnessMatrix ----------------------
 * Host form = the Julia call shape: coefficient vector in, CSR values out (vals_out may be NULL to
 * keep the matrix on the GPU only).  f == NULL means ones(nt) (src/FEM3D.jl:40).                */
int fk_assemble_stiffness(fk_handle h, int slot, const double* f, double* vals_out);     /* FEM3D.jl:40-76, FEMc.cpp:102-167 */
int fk_assemble_mass(fk_handle h, int slot, const double* rho, double* vals_out);       /* FEM3D.jl:466-492 */
int fk_assemble_tangential(fk_handle h, int slot, const double* Hfield /* 3 x nt */,
                           const double* dmu, double* vals_out);                        /* FEM3D.jl:79-119 */
/* 2-D P1 triangles through the same engine (src/FEM2D.jl:48-76 stiffnessMatrix2D, :139-161 massMatrix2D): create the
 * handle from the elements (n0, n1, n2, n2) -- fk_create with the triangle's third node repeated, p still 3 x nv with
 * the z row ignored -- and VE = triangle areas or NULL (|det| / 2).  Local matrices VE mu (b b' + c c') and
 * VE/12 [2 1 1; 1 2 1; 1 1 2]; pattern, plan, SpMV and PCG are the 3-D ones.                                        */
int fk_assemble_stiffness2d(fk_handle h, int slot, const double* f /* nt or NULL */, double* vals_out /* nnz or NULL */);
int fk_assemble_mass2d(fk_handle h, int slot, const double* rho /* nt or NULL */, double* vals_out);
/* P2 (10-node) tetrahedra: quadraticStiffnessMatrix(mesh, S, mu) (src/FEM3D.jl:324-395, the reference's 10-node
 * quadrature) and quadraticMassMatrix(mesh, S, F) (src/FEM3D.jl:494-561, 15-point rule; like the reference's, its weights
 * sum to 1/6 and multiply VE) on a handle created with nodes_per_element = 10.  The fitted basis functions S of the
 * reference are the Lagrange P2 functions on straight-sided elements and are used in closed form.  coef: nt or NULL. */
int fk_assemble_stiffness_p2(fk_handle h, int slot, const double* mu, double* vals_out /* nnz or NULL */);
int fk_assemble_mass_p2(fk_handle h, int slot, const double* F, double* vals_out);
/* quadraticConvectionMatrix(mesh, S, u, F) (src/FEM3D.jl:563-648): local matrices C_k[i, j] = F VE sum_q w_q phi_i (u_avg . grad phi_j),
 * u_avg = mean of the velocity u (3 x nv, column-major) over the element's four vertices, 15-point rule.  compat_bug = 1 assembles
 * them as the reference does -- TRANSPOSED: Clocal[:, k] = Ck[:] flattens column-major while the assembly loop reads row-major
 * (:632-645), so global entry (t_i, t_j) += C_k[j, i]; compat_bug = 0: entry (t_i, t_j) += C_k[i, j].  NON-symmetric values on the
 * (symmetric) pattern of the handle; fk_spmv / fk_matrix_axpby / fk_matrix_get work on it, fk_pcg does not apply.                */
int fk_assemble_convection_p2(fk_handle h, int slot, const double* u /* 3 x nv */, const double* F /* nt or NULL */, int compat_bug,
                              double* vals_out /* nnz or NULL */);
/* Nedelec (lowest-order edge) elements: the magnetic vector potential of Magnetostatics/nedelec3D.jl.  Degrees of freedom
 * = the edges of an order-2 mesh, edge2localMap[t[5:10, :]] (src/mesh.jl:170-185), local edge order and orientation of
 * NodesFromLocalEdge (src/mesh.jl:603-637: (0,1) (1,2) (2,0) (3,0) (2,3) (1,3), each directed from the lower to the higher
 * global vertex label), curl N_e = 2 L_e grad(lambda_i) x grad(lambda_j).
 * fk_create_edge: p = 3 x n_points, t4 = 4 x nt vertex ids, edges6 = 6 x nt edge labels in [0, n_edges); the handle's matrix
 * is n_edges x n_edges (fk_sizes reports nv = n_edges), pattern = sparsity of the reference's sum of sparse(edge1, edge2, ...).
 * fk_spmv / fk_pcg / fk_residual / fk_line_search / fk_matrix_axpby / fk_set_dirichlet work on it as on any handle.
 *   fk_nedelec_local_stiffness      nedelecLocalStiffness (src/FEM3D.jl:153-194): Ak[6 ie + je, k] = VE curlN_ie . curlN_je
 *   fk_assemble_nedelec_stiffness   the global matrix of nedelec3D.jl:163-178: sum of sparse(edge1, edge2, Ak[n, :] .* nu)
 *   fk_assemble_nedelec_tangential  nedelecTangentialStiffness (src/FEM3D.jl:197-271); |B| = 0 gives NaN as in the reference
 *   fk_nedelec_load_vector          RHS[e] += VE Bext . curlN (nedelec3D.jl:108-146)
 *   fk_nedelec_recover_field        Bfield[:, k] = sum_i u[e_i] curlN_i, B = |Bfield| (nedelec3D.jl:196-236)              */
int fk_create_edge(const double* p, int n_points, const int* t4, const int* edges6, int n_edges, int nt, const double* VE /* nt or NULL */,
                   const fk_create_opts* opts, fk_handle* out);
int fk_nedelec_local_stiffness(fk_handle h, double* Ak /* 36 x nt */);
int fk_assemble_nedelec_stiffness(fk_handle h, int slot, const double* nu /* nt or NULL */, double* vals_out /* nnz or NULL */);
int fk_assemble_nedelec_tangential(fk_handle h, int slot, const double* dnu /* nt */, const double* Bfield /* 3 x nt */,
                                   const double* B /* nt */, double* vals_out /* nnz or NULL */);
int fk_nedelec_load_vector(fk_handle h, const double* Bext /* 3 */, double* rhs_out /* n_edges */);
int fk_nedelec_recover_field(fk_handle h, const double* u /* n_edges */, double* Bfield_out /* 3 x nt or NULL */, double* B_out /* nt or NULL */);
int fk_local_stiffness(fk_handle h, const double* f, double* Ak /* 16 x nt */);          /* FEM3D.jl:59-76 */
int fk_assemble_stiffness_dev(fk_handle h, int slot, const double* f_dev);
int fk_assemble_mass_dev(fk_handle h, int slot, const double* rho_dev);
int fk_assemble_tangential_dev(fk_handle h, int slot, const double* Hfield_dev, const double* dmu_dev);

/* slot algebra on the shared pattern: dst = alpha*X + beta*Y (Y ignored when beta == 0) */
int fk_matrix_axpby(fk_handle h, int dst, double alpha, int x, double beta, int y);
int fk_matrix_get(fk_handle h, int slot, double* vals_out);
int fk_matrix_set(fk_handle h, int slot, const double* vals_in);

/* ---- right-hand sides ----------------------------------------------------------------------- */
int fk_lagrange(fk_handle h, double* C_out /* nv */);                                    /* FEM3D.jl:274-281 */
/* BoundaryIntegral (FEM3D.jl:284-297 / FEMc.cpp:170-200).  AE may be NULL: the area is then
 * recomputed by cross product like the C++ reference does.  shell_ids: list of boundary ids.   */
int fk_boundary_integral(fk_handle h, const int* surfaceT /* 4 x ns */, const double* normal /* 3 x ns */,
                         const double* AE, int ns, const double* F /* 3 */, const int* shell_ids,
                         int nshell, double* RHS_out /* nv */);

/* ---- SpMV and Jacobi-PCG (Eigen recurrence, ConjugateGradient.h:29-85) ---------------------- */
int fk_spmv(fk_handle h, int slot, const double* x, double* y);
int fk_spmv_dev(fk_handle h, int slot, const double* x_dev, double* y_dev);
int fk_pcg(fk_handle h, int slot, const double* b, double* x, const fk_pcg_opts* opts,
           int* iters_out, double* relres_out);
int fk_pcg_dev(fk_handle h, int slot, const double* b_dev, double* x_dev, const fk_pcg_opts* opts,
               int* iters_out, double* relres_out);

/* r = b - A x (NewtonRaphson.jl:166 `RHS-A*u`); r may be NULL; norm_out (may be NULL) = ||b - A x||_2 over the owned rows */
int fk_residual(fk_handle h, int slot, const double* x, const double* b, double* r, double* norm_out);
int fk_residual_dev(fk_handle h, int slot, const double* x_dev, const double* b_dev, double* r_dev, double* norm_out);
/* lineSearch(u, du, A, RHS, minStep) of src/FEM3D.jl:123-152, entirely on the device: alpha = 1, then alpha *= 0.9
 * while ||RHS - A (u + alpha du)|| > ||RHS - A u|| and alpha > min_step (<= 0: the reference's 1e-4); u is updated in
 * place unless alpha fell below min_step (the reference's "could not decrease the residue" branch).  Returns the residue
 * of the last trial and the accepted step (0 = u left unchanged).                                                     */
int fk_line_search(fk_handle h, int slot, double* u, const double* du, const double* rhs, double min_step,
                   double* residue_out, double* alpha_out);
int fk_line_search_dev(fk_handle h, int slot, double* u_dev, const double* du_dev, const double* rhs_dev, double min_step,
                       double* residue_out, double* alpha_out);

/* ---- constraints ------------------------------------------------------------------------------
 * Dirichlet elimination as the reference drivers do it on the assembled system (Fluids/fluids3D.jl:142-174:
 * `RHS = -LHS[free,fixed]*gD[fixed]; UP[fixed] = gD[fixed]; UP[free] = LHS[free,free] \ RHS`), kept on the full
 * pattern: fk_set_dirichlet stores the fixed DOFs and their values (gD == NULL: zeros); fk_apply_dirichlet lifts the
 * right-hand side (b[free] -= A[free,fixed] gD, b[fixed] = gD) and turns the fixed rows and columns of `slot` into
 * identity rows / columns, so every later fk_pcg / fk_spmv on that slot solves LHS[free,free] exactly and returns
 * x[fixed] = gD bit for bit (fk_pcg starts a constrained slot from x0[fixed] = gD, so residual, direction and A p are
 * exactly zero on those rows in every iteration).  A re-assembly of the slot needs a new fk_apply_dirichlet.
 * n_fixed = 0 clears the set.
 * fk_set_nullspace: weights w (nv, NULL = ones) of fk_pcg_opts.project_nullspace; lagrange(mesh) gives the
 * bordered-system solution of magneticField.jl:107.                                                             */
int fk_set_dirichlet(fk_handle h, const int* fixed_ids, int n_fixed, const double* gD /* n_fixed or NULL */);
int fk_apply_dirichlet(fk_handle h, int slot, double* b /* nv, in/out; may be NULL: matrix only */);
int fk_apply_dirichlet_dev(fk_handle h, int slot, double* b_dev);
int fk_set_nullspace(fk_handle h, const double* w /* nv or NULL */);

/* ---- field recovery and the nonlinear step --------------------------------------------------
 * H_k = -sum_i u[t_i] grad(lambda_i), |H_k|, B_k = mu_k |H_k| (magneticField.jl:111-132).
 * compat_bug != 0 reproduces NewtonRaphson.jl:187-189 / run.jl:94-96 (d-term written into row 2).
 * Output pointers may be NULL.                                                                  */
int fk_recover_field(fk_handle h, const double* u, const double* mu, int compat_bug,
                     double* Hfield_out /* 3 x nt */, double* Hnorm_out, double* B_out);
int fk_recover_field_dev(fk_handle h, const double* u_dev, const double* mu_dev, int compat_bug,
                         double* Hfield_dev, double* Hnorm_dev, double* B_dev);

/* B-spline table `id` in FITPACK (t, c, k) form = the fields of a Dierckx.jl Spline1D
 * (NewtonRaphson.jl:57-63).  fk_spline_fit builds the interpolating spline curfit(k=3, s=0) returns:
 * n = m + 4 knots [x1 x4 | x3..x_{m-2} | xm x4].                                                */
int fk_spline_fit(const double* x, const double* y, int m, double* t_out /* m+4 */, double* c_out /* m+4 */);
int fk_spline_eval(const double* t, const double* c, int n, int k, int bc, const double* x, int nx,
                   double* y_out);   /* host splev (FITPACK fpbspl), returns 2 if bc=error and out of range */
int fk_set_spline(fk_handle h, int id, const double* t, const double* c, int n, int k, int bc);
/* y = spl_id(x) evaluated by the DEVICE splev (the call `spl(H[mesh.InsideElements])` of NewtonRaphson.jl:203); returns 2
 * if bc = error and an argument lies outside the knots.  x, y: host (fk_spline_apply) or device (fk_spline_apply_dev).  */
int fk_spline_apply(fk_handle h, int id, const double* x, int nx, double* y_out);
int fk_spline_apply_dev(fk_handle h, int id, const double* x_dev, int nx, double* y_dev);
/* material of every element: 0 = linear (mu untouched), m >= 1 = mu from spline table m-1 and
 * dmu from table FK_MAX_SPLINES/2 + m-1 when set.  fk_set_inside marks ids[] as material 1
 * (mesh.InsideElements, NewtonRaphson.jl:204) and everything else 0.                            */
int fk_set_material(fk_handle h, const unsigned char* matid /* nt */);
int fk_set_inside(fk_handle h, const int* ids, int n);
int fk_set_mu(fk_handle h, const double* mu /* nt */);
int fk_get_mu(fk_handle h, double* mu_out, double* dmu_out /* may be NULL */);
int fk_get_field(fk_handle h, double* Hfield_out, double* Hnorm_out);
/* forget the stored field (|H|_old = 0, Hfield = 0, dmu = 0): the state a fresh nonlinear solve starts
 * from (`Hold = zeros(nt)`, NewtonRaphson.jl:139-140); mu is left as set by fk_set_mu.             */
int fk_reset_field(fk_handle h);
/* One fused nonlinear sweep over the elements (manyBodyIntersecting.jl:156-207 + :147, or
 * NewtonRaphson.jl:178-207 + :153-161): from the potential u recover H and |H|, div =
 * mu0*max_{material}| |H| - |H|_old |, mu = spl(|H|) (and dmu) on material elements, and assemble
 * the stiffness matrix with the new mu into `slot` -- one kernel, the mesh is read once.
 * Returns status 2 if a spline with bc = error was evaluated out of range.                      */
int fk_picard_step(fk_handle h, int slot, const double* u, int compat_bug, double* div_out);
/* div_out == NULL: the sweep is only enqueued (no host synchronisation); div and the error status are then
 * delivered by the next fk_picard_result / fk_picard_step* / fk_picard_pcg call on the handle.               */
int fk_picard_step_dev(fk_handle h, int slot, const double* u_dev, int compat_bug, double* div_out);
int fk_picard_result(fk_handle h, double* div_out);
/* One nonlinear iteration from host buffers = fk_picard_step(u) then fk_pcg(b -> x): the body of the reference
 * loop (NewtonRaphson.jl:148-214, manyBodyIntersecting.jl:141-210).  Identical results; the upload of b (and of
 * the warm-start x) overlaps the sweep on a second stream.                                                  */
int fk_picard_pcg(fk_handle h, int slot, const double* u, int compat_bug, const double* b, double* x,
                  const fk_pcg_opts* opts, double* div_out, int* iters_out, double* relres_out);

/* One implicit time step of the heat equation, heat3D.jl:83-86: T <- (M + dt A)^-1 (M T).  slot_mass holds M
 * (fk_assemble_mass), slot_lhs holds M + dt A (fk_matrix_axpby); Jacobi-PCG warm-started from T (the reference
 * re-factorises M + dt A with a sparse Cholesky every step).  T: n_rows values in/out (host / device).          */
int fk_heat_step(fk_handle h, int slot_mass, int slot_lhs, double* T, const fk_pcg_opts* opts, int* iters_out, double* relres_out);
int fk_heat_step_dev(fk_handle h, int slot_mass, int slot_lhs, double* T_dev, const fk_pcg_opts* opts, int* iters_out,
                     double* relres_out);

/* ---- micromagnetics demagnetising field: the repeated solve of the LL time loop ---------------
 * Replaces LL::LL (constructor work) + LL::magnetostaticField of Femeko_jl_C++/Micromagnetics/LL.cpp
 * (:4-66, :287-348; Julia twin Micromagnetics/LL.jl:162-211), reached in the reference through
 * LandauLifshitz(...) (micromagnetics.cpp:6-29, permalloy.jl:75-86).
 * fk_demag_setup: inside_elements = mesh.InsideElements (0-based, permalloy.jl:70-72); computes the
 *   box-method node volumes (LL.cpp:29-36; volumes_out nv, may be NULL) and assembles
 *   stiffnessMatrix(p, t, VE) (LL.cpp:40) into `slot`, which stays resident for every later call.
 * fk_demag_field: one magnetostaticField(): RHS[nd] += VE_k (grad lambda_i . <M>_k) over the magnetic
 *   elements (:292-312), u = Jacobi-PCG(A, RHS) (:315; opts->warm_start = 1 starts from the previous
 *   call's potential), H_k = -sum u grad lambda (:317-326), box method to the nodes (:328-346).
 *   M, Hd are 3 x nv column-major.  Nodes outside the magnetic region have zero node volume: the
 *   reference divides 0/0 there (NaN); nan_outside != 0 reproduces that, 0 writes 0.
 *   u_out (nv) may be NULL.
 *   Partitioned handles (fk_set_halo; every rank makes the same calls): inside_elements = the magnetic elements the part
 *   holds, ghost copies included, in ascending global order; M carries the halo rows too (the load vector of a ghost
 *   element reads them); the potential's halo rows are exchanged before the element field is taken; volumes, Hd and u are
 *   valid on the owned rows.  fk_multi_demag_* does all of that from global arrays.                                      */
int fk_demag_setup(fk_handle h, int slot, const int* inside_elements, int n_inside, double* volumes_out);
int fk_demag_field(fk_handle h, int slot, const double* M, const fk_pcg_opts* opts, int nan_outside,
                   double* Hd_out, double* u_out, int* iters_out, double* relres_out);
int fk_demag_field_dev(fk_handle h, int slot, const double* M_dev, const fk_pcg_opts* opts, int nan_outside,
                       double* Hd_dev, double* u_dev, int* iters_out, double* relres_out);

/* ---- multi-GPU (one process per GPU; NCCL over NVLink) --------------------------------------
 * The handle is built on this rank's partition (owned rows first, then halo rows; ghost elements
 * duplicated so assembly needs no communication).  fk_comm_* wire the halo exchange of the SpMV and
 * the dot-product all-reduce of the PCG into the same kernels.                                   */
int fk_nccl_unique_id(unsigned char* id128);
int fk_comm_init(fk_handle h, const unsigned char* id128, int rank, int nranks);
/* n_owned rows are owned; for each neighbour q: send_idx lists local rows to send, recv rows land in
 * local rows [recv_start[q], recv_start[q]+recv_count[q]).                                       */
int fk_set_halo(fk_handle h, int n_owned, int nneigh, const int* neigh_rank, const int* send_ptr /* nneigh+1 */,
                const int* send_idx, const int* recv_start, const int* recv_count);

/* Peer-memory path (one NVSwitch node, <= 8 ranks): every rank exports CUDA-IPC handles of its PCG vector
 * slab, which also holds a small control block (fk_comm_p2p_export: 128 bytes); after an all-gather of those
 * (any transport) fk_comm_p2p_attach maps the peers.  From then on the halo rows are stored straight into
 * the owner's vectors over NVLink and the all-reduces are single kernels over peer mailboxes (operands
 * combined in rank order: identical bits on every rank); NCCL stays as the fallback when attach fails.
 * remote_start[q] / remote_slot[q]: neighbour q's recv_start for this rank / this rank's index in q's
 * neighbour list.  fk_comm_mode: 0 = single GPU, 1 = NCCL, 2 = peer memory.                        */
int fk_comm_p2p_export(fk_handle h, unsigned char* handles128);
int fk_comm_p2p_attach(fk_handle h, const unsigned char* all_handles /* nranks x 128 */, const int* remote_start,
                       const int* remote_slot);
/* peer-memory path: how long a kernel waits for a peer's flag before it gives up and the next host check reports "a peer
 * GPU did not answer" (milliseconds; 0 = default, about 2 s -- raise it when a rank may stall on the host, e.g. under a
 * debugger), and the way back from such an error: every rank calls fk_comm_reset between two barriers of the caller.  */
int fk_comm_set_timeout(fk_handle h, int ms);
int fk_comm_reset(fk_handle h);
int fk_comm_mode(fk_handle h, int* mode);

/* ---- legacy drop-in: ABI-identical to the reference (magnetostatics.cpp:49-60) --------------- */
void scalarPotential(double* u_in, double* p_in, int* t_in, int* surfaceT_in, double* normal_in,
                     int nv, int nt, int ns, double* VE, double* mu, double* Hext, int shell_id);

/* ---- ONE process, several GPUs --------------------------------------------------------------------------
 * The reference's caller is a single Julia process (Femeko_jl_C++/Magnetostatics/run.jl:66-79): fk_multi_* lets that one
 * process use 2 / 4 / 8 GPUs.  fk_multi_create takes the GLOBAL mesh (as fk_create), partitions its nodes by recursive
 * coordinate bisection, builds one handle per GPU (owned rows first, halo rows after, ghost elements duplicated) and
 * wires them over peer memory (cudaDeviceEnablePeerAccess; no NCCL, no IPC).  Every call below takes and returns GLOBAL
 * host arrays in the caller's numbering; the work runs on one host thread per GPU, and the per-GPU kernels, the halo
 * exchange fused into the SpMV and the dot products over NVLink are those of the one-process-per-GPU path.
 * devices: ngpu CUDA ordinals, or NULL for 0 .. ngpu-1.                                                          */
typedef struct fk_multi_s* fk_multi;
int fk_multi_create(const double* p, const int* t, int nv, int nt, const double* VE, int ngpu, const int* devices,
                    const fk_create_opts* opts, fk_multi* out);
int fk_multi_destroy(fk_multi m);
/* the host-side decomposition on its own (no GPU needed): recursive coordinate bisection of the nodes, and one rank's
 * local numbering + halo plan.  Call fk_partition_local with NULL arrays for the sizes first.                     */
int fk_rcb(const double* p, int nv, int nparts, int* part_out /* nv */);
int fk_partition_local(const double* p, const int* t, int nv, int nt, const int* part, int rank,
                       int* sizes /* n_owned, nv_local, nt_local, nneigh, nsend */, int* l2g, int* elem_l2g, int* t_local /* 4 x nt_local */,
                       int* neigh, int* send_ptr /* nneigh+1 */, int* send_idx, int* recv_start, int* recv_count);
int fk_multi_info(fk_multi m, int* ngpu, int* n_owned /* ngpu */, int* nv_local, int* nt_local, int* nneigh);
int fk_multi_handle(fk_multi m, int rank, fk_handle* h);   /* the part's own handle (per-part diagnostics) */
int fk_multi_partition(fk_multi m, int rank, int* l2g /* nv_local: local node -> global */, int* elem_l2g /* nt_local */);
int fk_multi_assemble_stiffness(fk_multi m, int slot, const double* f /* nt or NULL */);
int fk_multi_assemble_mass(fk_multi m, int slot, const double* rho /* nt or NULL */);
int fk_multi_matrix_axpby(fk_multi m, int dst, double alpha, int x, double beta, int y);
int fk_multi_set_spline(fk_multi m, int id, const double* t, const double* c, int n, int k, int bc);
int fk_multi_set_inside(fk_multi m, const int* ids /* global element ids */, int n);
int fk_multi_set_mu(fk_multi m, const double* mu /* nt */);
int fk_multi_get_mu(fk_multi m, double* mu_out /* nt */);
int fk_multi_get_field(fk_multi m, double* Hfield_out /* 3 x nt or NULL */, double* Hnorm_out /* nt or NULL */);
int fk_multi_reset_field(fk_multi m);
int fk_multi_spmv(fk_multi m, int slot, const double* x /* nv */, double* y /* nv */);
int fk_multi_pcg(fk_multi m, int slot, const double* b /* nv */, double* x /* nv in/out */, const fk_pcg_opts* opts, int* iters_out,
                 double* relres_out);
int fk_multi_picard_step(fk_multi m, int slot, const double* u /* nv */, int compat_bug, double* div_out);
/* fk_demag_setup / fk_demag_field on every part (LL.cpp:4-66,287-348): inside_elements = GLOBAL element ids, ascending; M, Hd, u and
 * the volumes in the caller's global node numbering.                                                                           */
int fk_multi_demag_setup(fk_multi m, int slot, const int* inside_elements, int n_inside, double* volumes_out /* nv or NULL */);
int fk_multi_demag_field(fk_multi m, int slot, const double* M /* 3 x nv */, const fk_pcg_opts* opts, int nan_outside,
                         double* Hd_out /* 3 x nv */, double* u_out /* nv or NULL */, int* iters_out, double* relres_out);

/* ---- allocation-free residual norm -----------------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  r = b - Ax (r_dev may be NULL) and
 * ||r||_2 -- the second half of fk_residual_dev, whose per-CTA partials live in handle-owned buffers -- with the scratch
 * in a file-scope __device__ array instead (csrc/noalloc.cu, 131 328 bytes held by every GPU that loads the library until
 * the process exits).  The call allocates nothing: no cudaMalloc / cudaFree / cudaMallocHost, no handle, no library routine
 * that allocates.  Ax_dev is the caller's product (e.g. from fk_spmv_dev); all four arrays are device pointers of the
 * CURRENT device, `stream` a cudaStream_t of it (NULL = default stream).  norm_dev_out: device double or NULL;
 * norm_host_out: host double or NULL -- when given, the call copies the norm back and synchronises the stream (leave it NULL
 * under stream capture, and make one call outside the capture first: the first call on a device takes the array's address).
 * n > FK_NOALLOC_MAX_ROWS is rejected (return 1, fk_last_error says so).  Every call on a GPU shares the one scratch array:
 * calls that overlap on the same GPU (different streams or host threads) race -- serialise them.  Deterministic: the
 * summation order depends on n alone.                                                                                  */
#define FK_NOALLOC_MAX_ROWS (1LL << 26)   /* 67 108 864 rows: one partial per 4096 rows in the static scratch */
int fk_residual_norm_noalloc(const double* b_dev, const double* Ax_dev, long long n, double* r_dev /* n or NULL */,
                             double* norm_dev_out /* device, or NULL */, double* norm_host_out /* host, or NULL */, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FEMEKO_CUDA_H */
