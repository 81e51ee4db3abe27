/*
 * fempar_b200.h -- C-ABI of libfempar_b200.so, the B200 (sm_100a) implementation of FEMPAR's
 * data-parallel hot path: FE-space numbering, CSR graph, batched Galerkin integration + assembly,
 * SpMV, BLAS-1 vector kernels, Jacobi and the (preconditioned) CG iteration, serial and in the
 * par_* sub-domain layout (NCCL halo exchange + all-reduce).
 *
 * Every entry point is what a Fortran `bind(C)` interface (ISO_C_BINDING) of the FEMPAR type
 * extension named beside it would bind; INTEGRATION.md shows those interfaces.  Plain pointers
 * and sizes only; no torch / C++ types.  Reference file:line citations are relative to
 * /root/reference/Sources/Lib.
 *
 * Conventions:
 *   - all functions return int: 0 = ok, !=0 = error (text in fb2_last_error_string()); the
 *     Fortran shim does `check(ierr==0)` like debug.i90's check macro;
 *   - indices are Fortran 1-based where the reference's are (ja, DOF ids, part ids, irp on export);
 *   - DOF ids are signed int32: >0 free DOF, <0 strong-Dirichlet DOF (-id indexes fixed_dof_values);
 *   - reals are IEEE double; logicals are int;
 *   - host arrays passed in are borrowed for the call only; handles own device memory;
 *   - one host thread per handle; one CUDA device per process (rank).
 *   - there is NO CPU fallback: without a CUDA device every call fails with an error.
 */
#ifndef FEMPAR_B200_H
#define FEMPAR_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct fb2_space_s* fb2_space_t;   /* serial_fe_space_t / par_fe_space_t on a uniform_hex_mesh_t  */
typedef struct fb2_matrix_s* fb2_matrix_t; /* csr_sparse_matrix_t / par_sparse_matrix_t                   */
typedef struct fb2_vector_s* fb2_vector_t; /* serial_scalar_array_t / par_scalar_array_t                  */

/* operator kinds = the user discrete_integration_t plugins on this path (discrete_integration.f90:44-75) */
enum {
  FB2_KIND_POISSON = 0,          /* tutorial_01 ..._modules.f90:71-137, par_test_poisson_discrete_integration.f90:62-142 */
  FB2_KIND_VECTOR_LAPLACIAN = 1, /* Tests/Serial/test_poisson/modules/vector_poisson_discrete_integration.f90:62-140     */
  FB2_KIND_ELASTICITY = 2        /* par_test_pb_bddc_linear_elasticity/.../sbm_irreducible_discrete_integration.i90:93-125 */
};
/* analytic data (source term + boundary function) evaluated on the device */
enum {
  FB2_FUNC_ZERO = 0,
  FB2_FUNC_LINEAR = 1,     /* u = x+y+z, f = 0: par_test_poisson_analytical_functions.f90:71-124 */
  FB2_FUNC_SHARP_WAVE = 2  /* tutorial_01 ..._modules.f90:395-527; params = alpha, r, cx, cy, cz */
};
/* scatter variants of the assembly (north_star: atomic vs conflict-free, chosen by ncu) */
enum { FB2_SCATTER_ATOMIC = 0, FB2_SCATTER_COLORED = 1 };
/* iterative_solver_parameters.f90:37-46 stopping criteria; the CG path supports all ten (cg.f90:240-255).
 * Order: residual-based, preconditioned-norm-based, increment-based, then the AND combinations. */
enum {
  FB2_STOP_RES_RES = 0,                    /* ||r(i)||  <= rtol*||r(0)|| + atol (CG default) */
  FB2_STOP_RES_RHS = 1,                    /* ||r(i)||  <= rtol*||b|| + atol */
  FB2_STOP_RES_NRMGIVEN_RHS_NRMGIVEN = 2,  /* ||r(i)||g <= rtol*||b||g + atol, g = inv(M)-norm */
  FB2_STOP_RES_NRMGIVEN_RES_NRMGIVEN = 3,  /* ||r(i)||g <= rtol*||r(0)||g + atol */
  FB2_STOP_DELTA_RHS = 4,                  /* ||dx(i)|| <= rtol*||b|| + atol */
  FB2_STOP_DELTA_DELTA = 5,                /* ||dx(i)|| <= rtol*||dx(1)|| + atol */
  FB2_STOP_DELTA_RHS_AND_RES_RES = 6,
  FB2_STOP_DELTA_RHS_AND_RES_RHS = 7,
  FB2_STOP_DELTA_DELTA_AND_RES_RES = 8,
  FB2_STOP_DELTA_DELTA_AND_RES_RHS = 9
};

/* ILS_* parameter-list keys (iterative_solver_parameters.f90:167-183) */
typedef struct {
  double rtol;              /* ILS_RTOL, default 1e-6 */
  double atol;              /* ILS_ATOL, default 0    */
  int max_num_iterations;   /* ILS_MAX_NUM_ITERATIONS, default 1000 */
  int stopping_criterium;   /* ILS_STOPPING_CRITERIUM, default RES_RES for CG */
  int track_convergence_history; /* ILS_TRACK_CONVERGENCE_HISTORY */
  int check_every;          /* device iterations between host polls of the convergence flag (>=1) */
} fb2_ils_params_t;

/* ---- runtime ---------------------------------------------------------------------------- */
int fb2_init(int device);                       /* cudaSetDevice + constant tables; FEMPAR_INIT analogue (fempar.f90:126-133) */
const char* fb2_last_error_string(void);
int fb2_device_synchronize(void);
int fb2_version(void);
/* elapsed device time (ms) of the last timed region inside the named phase: "pattern", "scatter_map", "integrate", "spmv", "pcg" */
int fb2_last_phase_ms(const char* phase, double* ms);

/* ---- distributed context: execution_context_t on NCCL (mpi_context.f90:384-474,522-639) -- */
int fb2_comm_unique_id(void* id128);            /* fills a 128-byte ncclUniqueId on the root */
int fb2_comm_init(int rank, int nranks, const void* id128);
int fb2_comm_finalize(void);

/* ---- FE space: uniform_hex_mesh_generator.f90:145-647 + sbm_serial_fe_space.i90:525-715,2623-2875 ---- */
/* part_id: 0 = serial triangulation; 1..prod(num_parts) = this rank's sub-domain (par_triangulation_t).  */
int fb2_fe_space_create(fb2_space_t* out, const int num_cells[3], const int num_parts[3], int part_id, int order,
                        int num_components, uint32_t dirichlet_boundary_mask);
int fb2_fe_space_destroy(fb2_space_t s);
/* sizes[0..7] = num_cells(local+ghost), num_local_cells, dofs_per_cell, num_free_dofs, num_dirichlet_dofs,
 *               num_free_ghost_dofs, num_interior_dofs (par, after I-then-G renumbering), num_interface_dofs */
int fb2_fe_space_get_sizes(fb2_space_t s, int64_t sizes[8]);
/* lst_dofs_gids (fe_space.f90:636-637): [num_cells * dofs_per_cell] signed 1-based ids, cell-major */
int fb2_fe_space_get_lst_dofs_gids(fb2_space_t s, int32_t* host_out);
int fb2_fe_space_get_cells_gids(fb2_space_t s, int64_t* host_out, int32_t* host_mypart_out);
/* fixed_dof_values(1..num_dirichlet) = g(x_node): interpolate_dirichlet_values, sbm_serial_fe_space.i90:997-1176 */
int fb2_fe_space_interpolate_dirichlet(fb2_space_t s, int func, const double* func_params);
/* free DOF values of an fe_function by nodal interpolation of the analytic function:
 * serial_fe_space_interpolate_scalar_function, sbm_serial_fe_space.i90:734-819 (x has num_free_dofs entries) */
int fb2_fe_space_interpolate(fb2_space_t s, int func, const double* func_params, fb2_vector_t x);
/* user source term of the scalar Poisson integrand (the `source_term%get_value_space(quad_coords(qpoint), ...)` calls of
 * integrate_galerkin, tutorial_01_..._modules.f90:100-118): get the physical quadrature points of the local cells
 * (host_out[num_local_cells][nq][3], cells in lst_dofs_gids order, q index x fastest), evaluate any f there on the host and
 * set the samples (host_in[num_local_cells][nq]); fb2_integrate_assemble then uses them instead of the analytic selector.
 * host_in = NULL switches back. */
int fb2_fe_space_get_quadrature_points(fb2_space_t s, double* host_out);
int fb2_fe_space_set_source_values(fb2_space_t s, const double* host_in);
int fb2_fe_space_get_fixed_values(fb2_space_t s, double* host_out);
int fb2_fe_space_set_fixed_values(fb2_space_t s, const double* host_in);

/* dof_import_t of a par space (dof_import.f90:49-90), Fortran conventions (ptrs(1)=1, 1-based local DOF ids in the
 * interior-then-interface numbering).  sizes = num_rcv, num_snd, size(unpack_idx), size(pack_idx); weight may be NULL. */
int fb2_dof_import_get_sizes(fb2_space_t s, int32_t sizes[4]);
int fb2_dof_import_get(fb2_space_t s, int32_t* list_rcv, int32_t* rcv_ptrs, int32_t* unpack_idx, int32_t* list_snd,
                       int32_t* snd_ptrs, int32_t* pack_idx, double* weight);

/* ---- host-built FEMPAR objects handed to the device path (matrix-format, assembler and comm seams) ------ */
/* a space that is only a number of local DOFs (unstructured meshes, other reference FEs: the reference numbers the DOFs) */
int fb2_fe_space_create_bare(fb2_space_t* out, int64_t num_dofs);
/* a space on an UNSTRUCTURED conforming hex mesh that the reference numbered itself: serial_fe_space_t's
 * cell-wise DOF lists (ptr_dofs_x_fe / lst_dofs_gids, sbm_serial_fe_space.i90:2623-2875; here nd = ncomp*(order+1)^3 ids per
 * cell, cell-local order of SURVEY A.3, >0 free, <0 strong Dirichlet, >num_free_dofs ghost-only) and the vertex coordinates
 * of every cell, cell_coords[cell][vertex 0..7, x fastest][xyz].  Cells 0..num_local_cells-1 are integrated.  The device then
 * builds the CSR graph from the lists (fb2_csr_create: all cell pairs, sort, compress -- base_sparse_matrix.f90:4479-4628),
 * integrates with the per-cell Jacobians of the trilinear map (sbm_cell_map.i90:272-296, fb2_integrate_assemble), interpolates
 * and computes error norms; fb2_dof_import_set attaches the par layer. */
int fb2_fe_space_create_from_host(fb2_space_t* out, int32_t num_cells, int32_t num_local_cells, int order, int num_components,
                                  const int32_t* lst_dofs_gids, const double* cell_coords, int64_t num_free_dofs,
                                  int64_t num_dirichlet_dofs);
/* attach the dof_import_t the reference computed (dof_import.f90:49-90; Fortran conventions: part ids, ptrs and local
 * DOF ids 1-based; weight may be NULL = 0 on every packed, i.e. non-owned, DOF) */
int fb2_dof_import_set(fb2_space_t s, int32_t num_rcv, const int32_t* list_rcv, const int32_t* rcv_ptrs,
                       const int32_t* unpack_idx, int32_t num_snd, const int32_t* list_snd, const int32_t* snd_ptrs,
                       const int32_t* pack_idx, const double* weight);
/* base_sparse_matrix_t%move_from_coo_body / convert target (sparse_matrix.f90:783-833): the reference's csr_sparse_matrix_t
 * arrays irp(1)=1 (int32), ja, val copied to the device.  space NULL = serial; val NULL = zero values (graph only). */
int fb2_csr_create_from_host(fb2_matrix_t* out, fb2_space_t space, int64_t num_rows, const int32_t* irp, const int32_t* ja,
                             const double* val, int symmetric_storage);
int fb2_csr_initialize_values(fb2_matrix_t A, double alpha);   /* csr initialize_values */
/* sparse_assembler_t%assembly_matrix / assembly_array for a batch of cells (sparse_assembler.f90:58-126,154-206):
 * fe_dofs[num_cells][nd] signed 1-based ids, elmat[num_cells][nd*nd] column-major (elmat(idof,jdof), test = row), elvec
 * [num_cells][nd]; entries with ids < 1 or > num_rows are skipped (base_sparse_matrix.f90:3469-3513).  elmat / elvec / rhs
 * may be NULL.  With fixed_dof_values != NULL the strong-Dirichlet elimination of fe_cell_iterator%assembly
 * (sbm_fe_cell_iterator.i90:1610-1637) is applied to negative column ids first. */
int fb2_assemble_batch(fb2_matrix_t A, fb2_vector_t rhs, int32_t num_cells, int32_t num_dofs_per_cell, const int32_t* fe_dofs,
                       const double* elmat, const double* elvec, const double* fixed_dof_values, int64_t num_fixed);

/* ---- matrix: csr_sparse_matrix.f90:44-50 ; graph = what coo sort_and_compress would produce ---------- */
int fb2_csr_create(fb2_matrix_t* out, fb2_space_t s, int symmetric_storage);
int fb2_csr_destroy(fb2_matrix_t A);
int fb2_csr_get_sizes(fb2_matrix_t A, int64_t* num_rows, int64_t* nnz);
/* irp is exported 1-based (irp(1)=1) as int64; ja 1-based int32, ascending per row */
int fb2_csr_get_pattern(fb2_matrix_t A, int64_t* irp_host, int32_t* ja_host);
int fb2_csr_get_values(fb2_matrix_t A, double* val_host);
int fb2_csr_set_values(fb2_matrix_t A, const double* val_host);
/* fe_affine_operator%compute: integrate + assemble matrix and RHS (sbm_fe_affine_operator.i90:15-34).
 * func selects source term / boundary function; mat_params = {lambda, mu} for elasticity.           */
int fb2_integrate_assemble(fb2_matrix_t A, fb2_vector_t rhs, int kind, int func, const double* func_params,
                           const double* mat_params, int scatter_variant);
/* assembly kernel variant: 0 = register-blocked contraction + precomputed cell-local -> CSR-slot scatter map where it
 * applies (Poisson / vector Laplacian Q1-Q3, elasticity Q1-Q2; default), 1 = generic warp-per-cell kernel that searches ja (all kinds/orders) */
int fb2_set_assembly_variant(int variant);
/* y = A x : csr apply_body (csr_sparse_matrix.f90:626-672); par: + comm() (par_sparse_matrix.f90:739-762) */
int fb2_spmv(fb2_matrix_t A, fb2_vector_t x, fb2_vector_t y);
/* SpMV kernel variant: 0 = TMA-streamed row tiles (default), 1 = warp-per-row (kept for ncu comparison), 2 = streamed tiles
 * with per-tile dictionary + uint16 column indices (10.6 instead of 12 B per non-zero; bit-identical results; measured slower,
 * experimental; applies to matrices whose first SpMV comes after the call) */
int fb2_set_spmv_variant(int variant);
int fb2_extract_diagonal(fb2_matrix_t A, fb2_vector_t d);   /* csr_sparse_matrix.f90:4231-4256 */

/* ---- vectors: vector_t TBPs (vector.f90:36-52) on serial_scalar_array.f90:213-419,556-588 ---- */
int fb2_vec_create(fb2_vector_t* out, fb2_space_t s);     /* length = num_free_dofs of the space */
int fb2_vec_create_n(fb2_vector_t* out, int64_t n);
int fb2_vec_destroy(fb2_vector_t v);
int fb2_vec_size(fb2_vector_t v, int64_t* n);
int fb2_vec_set(fb2_vector_t v, const double* host_in);    /* insert_subvector / H2D */
int fb2_vec_get(fb2_vector_t v, double* host_out);         /* extract_subvector / D2H */
void* fb2_vec_device_ptr(fb2_vector_t v);
int fb2_vec_init(fb2_vector_t v, double alpha);
int fb2_vec_copy(fb2_vector_t dst, fb2_vector_t src);
int fb2_vec_scal(fb2_vector_t dst, double alpha, fb2_vector_t src);      /* dst = alpha*src */
int fb2_vec_axpby(fb2_vector_t op1, double alpha, fb2_vector_t op2, double beta); /* op1 = alpha*op2 + beta*op1 */
int fb2_vec_dot(fb2_vector_t x, fb2_vector_t y, double* result);        /* par: weighted + all-reduce */
int fb2_vec_nrm2(fb2_vector_t x, double* result);
int fb2_vec_local_dot(fb2_vector_t x, fb2_vector_t y, double* result);  /* vector_t%local_dot: weighted, no all-reduce (vector.f90:41) */
int fb2_vec_nullify_non_owned_dofs(fb2_vector_t x);                     /* vector_t%nullify_non_owned_dofs (vector.f90:52) */
/* vector_t%extract_subvector / insert_subvector on a subset of 1-based indices (vector.f90:161-187) */
int fb2_vec_extract_subvector(fb2_vector_t v, int32_t size_indices, const int32_t* indices, double* values);
int fb2_vec_insert_subvector(fb2_vector_t v, int32_t size_indices, const int32_t* indices, const double* values);
int fb2_vec_entrywise_product(fb2_vector_t z, fb2_vector_t x, fb2_vector_t y); /* z = x .* y */
int fb2_vec_entrywise_invert(fb2_vector_t v);
int fb2_vec_comm(fb2_vector_t v);                          /* par_scalar_array%comm, par_scalar_array.f90:498-533 */

/* ---- Jacobi (jacobi_preconditioner.f90:356-412) and CG (cg.f90:101-238) ---------------------- */
/* dinv = 1 ./ comm(diag(A)) */
int fb2_jacobi_setup(fb2_matrix_t A, fb2_vector_t dinv);
/* Solve A x = b from the initial solution (zero by default, see fb2_csr_set_initial_solution).  dinv == NULL -> identity preconditioner.  history may be NULL; else
 * must hold max_num_iterations doubles.  num_iterations counts loop SpMVs exactly like cg.f90.      */
int fb2_pcg_solve(fb2_matrix_t A, fb2_vector_t dinv, fb2_vector_t b, fb2_vector_t x, const fb2_ils_params_t* params,
                  int* num_iterations, int* did_converge, double* error_estimate, double* history);
/* base_iterative_linear_solver_t%set_initial_solution (base_iterative_linear_solver.f90:340-360): the solves ON THIS OPERATOR
 * that follow (fb2_pcg_solve, fb2_pcg_solve_host, fb2_ils_solve) start from a copy of x0 instead of zero; NULL = back to zero.
 * The copy, like the Krylov work vectors, the device-resident convergence state and the staging vectors of the host-buffer
 * entry point, belongs to the matrix handle: solves on different operators never share state (re-entrant across handles).
 * A vector of another size is ignored like the reference ignores one outside range(A).  The *_multi entry points start from
 * zero and reject parts with an initial solution set. */
int fb2_csr_set_initial_solution(fb2_matrix_t A, fb2_vector_t x0);
/* what the reference's convergence log needs beyond the history (print_convergence_history_header / _new_line / _footer,
 * base_iterative_linear_solver.f90:584-676): out[4] = rhs_convergence_test, rhs_extra_convergence_test,
 * error_estimate_convergence_test, error_estimate_extra_convergence_test of the LAST solve on this operator; history_extra
 * (NULL, or num_iterations doubles; needs a solve with history != NULL) = the estimate of the second criterion per iteration
 * for the four DELTA_*_AND_RES_* criteria. */
int fb2_csr_get_convergence_data(fb2_matrix_t A, double out[4], double* history_extra);
/* Self-test of the NVLink peer-memory scalar all-reduce protocol (fb2_comm.cu) on ONE device: every rank is played by one block
 * of a cooperative launch over `epochs` consecutive all-reduces of `count` (<= 7) doubles; returns the largest deviation from
 * the rank-ordered sum (0 when every rank holds the bit-identical result). */
int fb2_peer_allreduce_selftest(int nranks, int count, int epochs, double* max_abs_error);
/* drops the SpMV tile table / pattern classes of the matrix; the next apply prepares them again (tuning sweeps) */
int fb2_csr_spmv_reset(fb2_matrix_t A);
/* releases the solver workspace of the operator early (fb2_csr_destroy does it too) */
int fb2_csr_free_solver_workspace(fb2_matrix_t A);
/* the same through HOST buffers (what iterative_linear_solver%apply sees from Fortran arrays):
 * H2D of b, solve, D2H of x inside the call. */
int fb2_pcg_solve_host(fb2_matrix_t A, fb2_vector_t dinv, const double* b_host, double* x_host,
                       const fb2_ils_params_t* params, int* num_iterations, int* did_converge,
                       double* error_estimate, double* history);
/* The other Krylov solvers of the reference that share these primitives; keys of the creational-methods
 * dictionary (iterative_linear_solver_creational_methods_dictionary.f90:97-106). */
enum {
  FB2_ILS_CG = 0,          /* cg.f90 */
  FB2_ILS_ICG = 1,         /* icg.f90:106-240, all ten stopping criteria */
  FB2_ILS_RICHARDSON = 2,  /* richardson.f90:101-205, RES_RES / RES_RHS */
  FB2_ILS_RGMRES = 3,      /* rgmres.f90:120-356, RES_NRMGIVEN_* */
  FB2_ILS_FGMRES = 4,      /* fgmres.f90:120-356, RES_NRMGIVEN_* */
  FB2_ILS_LGMRES = 5,      /* lgmres.f90:152-420, RES_NRMGIVEN_*, RES_RES, RES_RHS */
  FB2_ILS_LFOM = 6,        /* lfom.f90:124-330, RES_RES / RES_RHS */
  FB2_ILS_MINRES = 7       /* minres.f90:361-740, RES_RES; istop / Anorm / Acond via fb2_last_phase_ms("minres_istop" ...) */
};
enum { FB2_ORTHO_MGSRO = 0, FB2_ORTHO_ICGSRO = 1 };   /* ILS_ORTHONORM_STRAT: iterative_solver_utils.f90:46-106 / :128-196 */
typedef struct {
  int dkrymax;            /* ILS_MAX_DIM_KRYLOV_BASIS, default 1000 (clamped to max_num_iterations) */
  int orthonorm_strat;    /* ILS_ORTHONORM_STRAT, default ICGSRO */
  double relaxation;      /* ILS_RELAXATION (Richardson), default 1.0 */
} fb2_ils_ext_params_t;
/* iterative_linear_solver_t%apply for any of the types above; ext may be NULL (defaults); x0 = 0 */
int fb2_ils_solve(fb2_matrix_t A, fb2_vector_t dinv, fb2_vector_t b, fb2_vector_t x, int ils_type, const fb2_ils_params_t* params,
                  const fb2_ils_ext_params_t* ext, int* num_iterations, int* did_converge, double* error_estimate,
                  double* history);
/* fixed-iteration variant used by benchmarks: runs exactly n PCG iterations (no convergence exit) and
 * reports device time per phase */
int fb2_pcg_iterate(fb2_matrix_t A, fb2_vector_t dinv, fb2_vector_t b, fb2_vector_t x, int n_iterations,
                    double* ms_total, double* ms_spmv);

/* ---- several sub-domains per process (BASELINE config 3: the 8-part layout on 2 or 4 GPUs) ----------------------------
 * Part p (1-based) lives on rank (p-1)/parts_per_rank.  With parts_per_rank > 1 the single-part entry points above do NOT
 * communicate (fb2_integrate_assemble leaves the local RHS); the *_multi entry points take the arrays of the np co-resident
 * handles and do comm() / weighted dots / the PCG iteration over all of them collectively.  Same semantics as
 * par_scalar_array%comm / %dot, par_sparse_matrix%apply, jacobi numerical_setup and cg_solve_body. */
int fb2_comm_set_parts_per_rank(int parts_per_rank);
int fb2_vec_comm_multi(int np, fb2_vector_t* v);
int fb2_vec_dot_multi(int np, fb2_vector_t* x, fb2_vector_t* y, double* result);
int fb2_spmv_multi(int np, fb2_matrix_t* A, fb2_vector_t* x, fb2_vector_t* y);
int fb2_jacobi_setup_multi(int np, fb2_matrix_t* A, fb2_vector_t* dinv);
int fb2_pcg_solve_multi(int np, fb2_matrix_t* A, fb2_vector_t* dinv /*NULL = identity*/, fb2_vector_t* b, fb2_vector_t* x,
                        const fb2_ils_params_t* params, int* num_iterations, int* did_converge, double* error_estimate,
                        double* history);
int fb2_pcg_iterate_multi(int np, fb2_matrix_t* A, fb2_vector_t* dinv, fb2_vector_t* b, fb2_vector_t* x, int n_iterations,
                          double* ms_total, double* ms_spmv);

/* ---- fe_operator_t residual / tangent at an evaluation point -----------------------------------
 * compute_internal_residual -> discrete_integration%integrate_residual and compute_tangent -> integrate_tangent
 * (Integration/FEOperator/sbm_fe_operator.i90:568-639) for the quasi-linear diffusion operator -div(k(u) grad u) = f,
 * k(u) = 1 + beta u^2 (beta = 0: integrate_residual of Tests/Serial/test_transient_poisson/modules/
 * test_transient_poisson_discrete_integration.f90:248-319).  u = free DOF values of the evaluation point (strong Dirichlet
 * values come from the space's fixed_dof_values); f = the analytic selector, or the user samples of
 * fb2_fe_space_set_source_values.  residual: free rows, comm()'ed in par.  tangent: written into A's values (full storage;
 * derivatives with respect to the free DOFs). */
int fb2_integrate_residual(fb2_matrix_t A, fb2_vector_t u, int func, const double* func_params, double beta,
                           fb2_vector_t residual);
int fb2_integrate_tangent(fb2_matrix_t A, fb2_vector_t u, int func, const double* func_params, double beta);

/* ---- error norms: sqrt(sum_K int |grad(u-u_h)|^2), sqrt(int (u-u_h)^2) ----- */
int fb2_error_norms(fb2_space_t s, fb2_vector_t x, int func, const double* func_params, double* h1_seminorm,
                    double* l2_norm);
/* error_norms_scalar_t%compute for the eleven norm types par_test_poisson checks (par_test_poisson_driver.f90:557-589;
 * Integration/ErrorNorms/error_norms.f90:41-54,160-221, sbm_error_norms_scalar.i90:114-232):
 * norms[] = mean, l1, l2, lp, linfty, h1_seminorm, h1, w1p_seminorm, w1p, w1infty_seminorm, w1infty.
 * exponent <= 0 selects the reference's default 2.0; in par the sums / maxima are all-reduced (l1_sum / l1_max). */
#define FB2_NORM_MEAN 0
#define FB2_NORM_L1 1
#define FB2_NORM_L2 2
#define FB2_NORM_LP 3
#define FB2_NORM_LINFTY 4
#define FB2_NORM_H1_SEMI 5
#define FB2_NORM_H1 6
#define FB2_NORM_W1P_SEMI 7
#define FB2_NORM_W1P 8
#define FB2_NORM_W1INFTY_SEMI 9
#define FB2_NORM_W1INFTY 10
int fb2_error_norms_all(fb2_space_t s, fb2_vector_t x, int func, const double* func_params, double exponent, double norms[11]);

/* ---- allocation-free dot product ------------------------------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  result_host <- sum_i x_dev[i] * y_dev[i], i < n
 * (the plain local dot: no sub-domain weights, no all-reduce).  x_dev / y_dev are device pointers (fb2_vec_device_ptr),
 * result_host a host pointer; the call blocks until the result is there.  The block partials live in a file-scope __device__
 * array of 65 568 doubles (512 KiB + 256 B per GPU that loads the library, held until exit) instead of the library's
 * cudaMalloc'ed scratch; one partial per 4 096 elements bounds the length: n > FB2_NO_ALLOC_MAX_N returns 2 with
 * fb2_last_error_string() set.  No cudaMalloc, cudaFree or allocating routine is called.  Every call shares the array
 * (nothing in it has to start at zero: each slot read was written earlier in the same call): calls that overlap on the GPU race. */
#define FB2_NO_ALLOC_MAX_N 268435456LL
int fb2_vec_dot_no_alloc(const double* x_dev, const double* y_dev, int64_t n, double* result_host);

#ifdef __cplusplus
}
#endif
#endif
