// Internal declarations of libfempar_b200 (not part of the C-ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/fempar_b200.h"

typedef long long i64;

void fb2_set_error(const char* fmt, ...);
#define FB2_CUDA(call)                                                                           \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess) {                                                                     \
      fb2_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return 1;                                                                                  \
    }                                                                                            \
  } while (0)
#define FB2_CHECK(cond, msg)                                       \
  do {                                                             \
    if (!(cond)) {                                                 \
      fb2_set_error("%s:%d: %s", __FILE__, __LINE__, msg);         \
      return 2;                                                    \
    }                                                              \
  } while (0)
#define FB2_TRY(expr)          \
  do {                         \
    int rc_ = (expr);          \
    if (rc_ != 0) return rc_;  \
  } while (0)

cudaStream_t fb2_stream();
cudaStream_t fb2_comm_stream();   // high-priority stream of the overlapped halo exchange
int fb2_num_sms();
void fb2_record_phase(const char* name, double ms);

// Blocking copy ORDERED ON THE LIBRARY STREAM.  The stream is non-blocking, so a plain cudaMemcpy (legacy default stream)
// would neither wait for kernels / NCCL collectives queued on it nor be waited for by them.
static inline cudaError_t fb2_memcpy(void* dst, const void* src, size_t bytes, cudaMemcpyKind kind) {
  cudaError_t e = cudaMemcpyAsync(dst, src, bytes, kind, fb2_stream());
  if (e != cudaSuccess) return e;
  return cudaStreamSynchronize(fb2_stream());
}

template <class T>
int fb2_dalloc(T** p, size_t n) {
  *p = nullptr;
  if (n == 0) n = 1;
  cudaError_t e = cudaMalloc((void**)p, n * sizeof(T));
  if (e != cudaSuccess) {
    fb2_set_error("cudaMalloc of %zu bytes failed: %s", n * sizeof(T), cudaGetErrorString(e));
    return 1;
  }
  return 0;
}

// --- structured local mesh of one part (uniform_hex_mesh_generator.f90:145-647) ---
struct MeshDesc {
  int N[3], P[3], pijk[3];
  int nloc[3], ntot[3], L[3], R[3], first[3];
  int n_local, n_cells, n_itfc;
  int par;  // 1: cell permutation interior -> interface -> ghost is active
};

// --- reference-FE tables shared by host and device (filled on the host once per space) ---
struct RefTables {
  int p, n1, nnode, ncomp, ndofs, nq;
  int m_face[27];          // own nodes (per component) of each n-face, hex polytope order, [26] = cell
  int face_of_code[64];    // polytope index (0-based) from code = free<<3|anchor
  int code_of_face[27];
};

struct fb2_space_s {
  MeshDesc m;
  RefTables rt;
  uint32_t dir_mask;
  int part_id;
  // device arrays
  int* d_perm = nullptr;     // lex (total box) -> local cell id; nullptr when identity
  int* d_lex_of = nullptr;   // local cell id -> lex
  int* d_gids = nullptr;     // lst_dofs_gids [n_cells * ndofs]
  int* d_node_dof = nullptr; // [ncomp][lattice of the LOCAL cell box] -> gid
  int* d_dof_node = nullptr; // [n_free] -> comp * lattice_size + lattice node
  double* d_fixed = nullptr; // fixed_dof_values [n_dirichlet]
  double* d_src_qp = nullptr;// optional user source term sampled at the quadrature points [n_local][nq] (scalar Poisson)
  double* d_coords = nullptr;// host-built (unstructured) spaces only: vertex coordinates [n_cells][8][3], hex vertex order
  // reference tables on device
  double* d_qw = nullptr;    // [nq]
  double* d_phi = nullptr;   // [nq][nnode]
  double* d_dphi = nullptr;  // [nq][nnode][3]
  double* d_gphi = nullptr;  // geometry (Q1) basis at q-points [nq][8]
  double* d_gdphi = nullptr; // [nq][8][3]
  std::vector<double> h_qw, h_phi, h_dphi, h_gphi, h_gdphi;
  std::vector<double> h_b1d, h_c1d;   // 1-D Lagrange values / derivatives at the 1-D Gauss points [q][i]
  i64 n_free = 0, n_dirichlet = 0, n_free_ghost = 0;
  i64 n_interior = 0, n_interface = 0;
  int Mlat[3];               // lattice nodes per direction of the local cell box
  i64 lat_size = 0;
  // par layer (dof_import_t, dof_import.f90:49-90)
  struct fb2_import_s* imp = nullptr;
};

struct fb2_solver_ws;
struct fb2_matrix_s {
  fb2_space_s* s = nullptr;
  fb2_solver_ws* ws = nullptr;   // solver workspace of this operator (fb2_solver.cu)
  int sym = 0;
  i64 n = 0, nnz = 0;
  i64* d_irp = nullptr;   // 0-based offsets, n+1
  int* d_ja = nullptr;    // 1-based columns
  double* d_val = nullptr;
  // cell-local -> CSR-slot scatter map (fb2_assemble_fast.cu): slot = irp[row] + tabs[cell_tab[c]][i*nd + j]
  int* d_cell_tab = nullptr;          // [n_local] table id of each local cell
  unsigned short* d_tabs = nullptr;   // [n_tabs][nd*nd] relative slots, 0xFFFF = entry not stored
  int common_tab = -1;                // table id of the interior position class
  int n_tabs = 0, n_class_tabs = 0, map_state = 0;   // map_state: 0 not built, 1 built, -1 unavailable (search fallback)
  int map_partial = 0;                // 1: tables verified only for the cells the scatter kernel visits (row-owner hybrid)
  // row-owner gather assembly (fb2_assemble_rows.cu): owner cells whose chunk of rows verified against the plan are written by
  // plain stores; the remaining ("slow") rows are zeroed and accumulated by the scatter kernel over the cells touching them
  int rows_state = 0;                     // 0 not built, 1 usable, -1 unavailable
  unsigned short* d_rows_plan = nullptr;  // [units][columns] chunk-relative slots
  unsigned char rows_row_of_node[32] = {0};
  i64* d_chunk_base = nullptr;            // [n_local] irp of the owner cell's first row, -1: not a verified owner cell
  unsigned* d_row_fast = nullptr;         // bit per row: written by the gather kernel
  i64* d_slow_rows = nullptr;             // the other rows, ascending
  int* d_slow_cells = nullptr;            // total-box lexicographic ids of the cells touching a slow row, ascending
  i64 n_fast_cells = 0, n_slow_rows = 0;
  int n_slow_cells = 0;
  // SpMV tile geometry (fb2_spmv.cu)
  int spmv_R = 0, spmv_cap = 0, spmv_grid = 0, spmv_stages = 0, spmv_nc = 0;   // nc: consumer threads per CTA
  int spmv_tr = 0;                          // 1: rows dealt transposed to the lane groups, chunks staggered in the stage
  i64 spmv_ntiles = 0;
  i64* d_tile_row = nullptr;              // first row of every tile
  unsigned char* d_tile_nrows = nullptr;  // rows in the tile (<= spmv_R)
  size_t spmv_smem = 0;
  // compressed column indices of the streaming SpMV (fb2_spmv.cu): per tile a dictionary of its distinct columns + one
  // uint16 dictionary position per stored entry (10 B instead of 12 B per non-zero streamed from HBM)
  int spmv_comp = 0;                        // 0 not built, 1 in use, -1 unavailable / switched off
  int spmv_dcap = 0;                        // largest dictionary (entries, multiple of 4)
  unsigned short* d_idx16 = nullptr;        // [nnz] position of ja[k] in the dictionary of its tile
  int* d_dict = nullptr;                    // concatenated dictionaries, 0-based columns, each padded to 4 entries
  i64* d_dict_ptr = nullptr;                // [ntiles + 1] offsets into d_dict
  size_t spmv_smem_comp = 0;
  // pattern classes of the streaming SpMV (fb2_spmv.cu): tiles whose columns are r0 + class deltas read neither ja nor gather x
  int spmv_cls = 0;                         // 0 not built, 1 in use, -1 unavailable / switched off
  int spmv_ncls = 0;
  i64 spmv_tiles_in_class = 0;
  unsigned char* d_tile_class = nullptr;    // [ntiles] class id, 0xFF: none, 0xFE: private descriptor (d_tile_priv)
  int* d_tile_priv = nullptr;               // [ntiles] descriptor index of a private tile (behind the classes in d_cls_desc / d_cls_idx)
  i64 spmv_tiles_private = 0;
  void* d_cls_desc = nullptr;               // SpmvClassDesc[ncls]
  unsigned short* d_cls_idx = nullptr;      // [ncls][spmv_cap] x-window position of every staged entry
  int owns_space = 0;                       // matrix created from host arrays without a space: it owns a bare one
  i64 spmv_t_itf = -1, spmv_row_itf = -1;   // interface split of the tile table (par overlap)
};

struct fb2_vector_s {
  i64 n = 0;
  double* d = nullptr;
  fb2_space_s* s = nullptr;
};

// reductions
int fb2_dot_device(const double* x, const double* y, const double* w, i64 n, double* d_result /*device*/);
double* fb2_scratch_partials();   // device buffer of FB2_MAX_PARTIALS doubles x 4
#define FB2_MAX_PARTIALS 65536

int fb2_build_scatter_map(fb2_matrix_s* A, const int* only_cells, int n_only);
int fb2_rows_prepare(fb2_matrix_s* A, int kind);     // fb2_assemble_rows.cu: plan + per-cell verification (set-up)
int fb2_rows_zero_slow(fb2_matrix_s* A);
int fb2_assembly_variant();
int fb2_space_setup_reference_fe(fb2_space_s* s, int order, int ncomp);   // reference-FE tables, host + device (fb2_space.cu)
int fb2_csr_graph_from_gids(fb2_matrix_s* A);                             // COO sort/unique graph (fb2_unstructured.cu)
// primitives of fb2_solver.cu used by the host-driven Krylov solvers (fb2_krylov.cu)
bool fb2_matrix_is_par(fb2_matrix_s* A);
const double* fb2_matrix_weight(fb2_matrix_s* A);
int fb2_op_apply(fb2_matrix_s* A, const double* x, double* y);
int fb2_dev_apply_M(const double* dinv, const double* r, double* z, i64 n);
int fb2_dev_axpby(double* op1, double a, const double* op2, double b, i64 n);
int fb2_dev_scal(double* dst, double a, const double* src, i64 n);
int fb2_dev_init(double* x, double a, i64 n);
int fb2_dots_host(fb2_matrix_s* A, int k, const double* const* xs, const double* const* ys, i64 n, double* out);
void fb2_cg_set_icg(int on);
bool fb2_ils_has_initial_solution(fb2_matrix_s* A);
int fb2_ils_load_initial_solution(fb2_matrix_s* A, double* x);   // x <- the initial solution set on this operator, or 0
void fb2_solver_ws_free(fb2_matrix_s* A);
void fb2_solver_record_convergence(fb2_matrix_s* A, double tol, double err, int iters);   // for the host-driven Krylov solvers
int fb2_spmv_stream(fb2_matrix_s* A, const double* x, double* y, bool fuse_dot, const int* done_flag, int* grid_out);
int fb2_spmv_prepare(fb2_matrix_s* A);
void fb2_spmv_set_compress(bool on);
void fb2_spmv_set_classes(bool on);      // pattern-class tiles for matrices prepared from now on (default on)   // dictionary-compressed column indices for matrices prepared from now on
// tiles [t0, t1) only; chunk = 0: persistent CTAs striding over the range, chunk > 0: one CTA per `chunk` consecutive tiles
// (short-lived CTAs, so that a higher-priority stream gets SM resources while the launch is running)
int fb2_spmv_stream_range(fb2_matrix_s* A, const double* x, double* y, bool fuse_dot, const int* done_flag, i64 t0, i64 t1,
                          int chunk, double* partials, int* grid_out);
// first tile holding an interface row (rows >= n_interior) and its first row
int fb2_spmv_interface_split(fb2_matrix_s* A, i64 n_interior, i64* t_itf, i64* row_itf);

// comm layer: dof_import_t on the device (dof_import.f90:49-90)
struct fb2_import_s {
  int num_rcv = 0, num_snd = 0;
  std::vector<int> list_rcv, rcv_ptrs, list_snd, snd_ptrs;   // part ids 1-based, ptrs 0-based offsets
  std::vector<int> h_pack_idx, h_unpack_idx;                 // 1-based local dof ids (host copy for export)
  int *d_pack_idx = nullptr, *d_unpack_idx = nullptr;        // 0-based local dof ids
  double *d_snd_buf = nullptr, *d_rcv_buf = nullptr;
  double* d_weight = nullptr;
  i64 n_pack = 0, n_unpack = 0;
  // stage-1 gather-add plan: unique destination DOFs, their receive-buffer positions in neighbour order
  int n_add_dst = 0;
  int *d_add_dst = nullptr, *d_add_ptr = nullptr, *d_add_pos = nullptr;
};
int fb2_comm_active();
int fb2_comm_nranks();
int fb2_comm_rank();
int fb2_allreduce_sum(double* d_buf, int count);           // in place, device buffer
int fb2_import_exchange(fb2_import_s* imp, double* d_vec); // par_scalar_array%comm()
int fb2_import_exchange_on(fb2_import_s* imp, double* d_vec, cudaStream_t stream);
int fb2_import_exchange_multi(int np, fb2_import_s* const* imps, const int* part_ids, double* const* xs, cudaStream_t s);
int fb2_comm_parts_per_rank();
const double* fb2_import_weight(fb2_import_s* imp);
void fb2_import_destroy(fb2_import_s* imp);
int fb2_import_build_add_plan(fb2_import_s* imp);   // from imp->h_unpack_idx (0-based)
int fb2_build_par_layer(fb2_space_s* s);                   // fb2_par.cu
int fb2_build_node_maps(fb2_space_s* s);                   // fb2_space.cu                   // interface detection + dof_import + I-then-G renumbering
