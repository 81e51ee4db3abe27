// Drop-in seams for host-built FEMPAR objects: a matrix that FEMPAR assembled itself
// (base_sparse_matrix_t%move_from_coo_body -> device CSR), a dof_import_t that FEMPAR computed itself, and the
// element-wise assembler entry points (assembler_t%assembly_matrix / assembly_array) fed in batches.
// With these the device SpMV / BLAS-1 / Jacobi / PCG path serves any conforming FE space of the reference
// (unstructured meshes, other reference FEs), not only the structured spaces fb2_fe_space_create numbers itself.
#include <algorithm>
#include "fb2_internal.h"

// ------------------------------------------------------------------------------------------
// bare space: only a number of local DOFs (+ optionally a dof_import attached afterwards)
// ------------------------------------------------------------------------------------------
extern "C" int fb2_fe_space_create_bare(fb2_space_t* out, int64_t num_dofs) {
  FB2_CHECK(out && num_dofs >= 0, "fb2_fe_space_create_bare: bad argument");
  fb2_space_s* s = new fb2_space_s();
  memset(&s->m, 0, sizeof(s->m));
  memset(&s->rt, 0, sizeof(s->rt));
  s->dir_mask = 0; s->part_id = 0;
  s->n_free = num_dofs; s->n_interior = num_dofs; s->n_interface = 0;
  *out = s;
  return 0;
}

// dof_import_t computed by the reference (dof_import.f90:49-90), Fortran conventions on input:
// list_* hold 1-based part ids, *_ptrs(1) = 1, pack_idx / unpack_idx are 1-based local DOF ids.
extern "C" int fb2_dof_import_set(fb2_space_t s, int32_t num_rcv, const int32_t* list_rcv, const int32_t* rcv_ptrs,
                                  const int32_t* unpack_idx, int32_t num_snd, const int32_t* list_snd, const int32_t* snd_ptrs,
                                  const int32_t* pack_idx, const double* weight) {
  FB2_CHECK(s && num_rcv >= 0 && num_snd >= 0, "fb2_dof_import_set: bad argument");
  if (s->imp) { fb2_import_destroy(s->imp); s->imp = nullptr; }
  fb2_import_s* imp = new fb2_import_s();
  imp->num_rcv = num_rcv; imp->num_snd = num_snd;
  imp->list_rcv.assign(list_rcv, list_rcv + num_rcv);
  imp->list_snd.assign(list_snd, list_snd + num_snd);
  imp->rcv_ptrs.resize(num_rcv + 1); imp->snd_ptrs.resize(num_snd + 1);
  for (int i = 0; i <= num_rcv; ++i) imp->rcv_ptrs[i] = rcv_ptrs[i] - 1;
  for (int i = 0; i <= num_snd; ++i) imp->snd_ptrs[i] = snd_ptrs[i] - 1;
  imp->n_unpack = imp->rcv_ptrs[num_rcv]; imp->n_pack = imp->snd_ptrs[num_snd];
  const i64 n = s->n_free;
  imp->h_unpack_idx.resize((size_t)imp->n_unpack); imp->h_pack_idx.resize((size_t)imp->n_pack);
  i64 min_itf = n;
  for (i64 i = 0; i < imp->n_unpack; ++i) {
    FB2_CHECK(unpack_idx[i] >= 1 && unpack_idx[i] <= n, "fb2_dof_import_set: unpack_idx out of range");
    imp->h_unpack_idx[i] = unpack_idx[i] - 1; min_itf = std::min<i64>(min_itf, unpack_idx[i] - 1);
  }
  for (i64 i = 0; i < imp->n_pack; ++i) {
    FB2_CHECK(pack_idx[i] >= 1 && pack_idx[i] <= n, "fb2_dof_import_set: pack_idx out of range");
    imp->h_pack_idx[i] = pack_idx[i] - 1; min_itf = std::min<i64>(min_itf, pack_idx[i] - 1);
  }
  FB2_TRY(fb2_dalloc(&imp->d_pack_idx, (size_t)imp->n_pack)); FB2_TRY(fb2_dalloc(&imp->d_unpack_idx, (size_t)imp->n_unpack));
  if (imp->n_pack) FB2_CUDA(fb2_memcpy(imp->d_pack_idx, imp->h_pack_idx.data(), sizeof(int) * imp->n_pack, cudaMemcpyHostToDevice));
  if (imp->n_unpack) FB2_CUDA(fb2_memcpy(imp->d_unpack_idx, imp->h_unpack_idx.data(), sizeof(int) * imp->n_unpack, cudaMemcpyHostToDevice));
  FB2_TRY(fb2_import_build_add_plan(imp));
  for (auto& v : imp->h_pack_idx) v += 1;
  for (auto& v : imp->h_unpack_idx) v += 1;
  FB2_TRY(fb2_dalloc(&imp->d_weight, (size_t)n));
  std::vector<double> w((size_t)n, 1.0);
  if (weight) w.assign(weight, weight + n);
  else for (i64 i = 0; i < imp->n_pack; ++i) w[imp->h_pack_idx[i] - 1] = 0.0;   // packed (sent to the owner) = not owned here
  if (n) FB2_CUDA(fb2_memcpy(imp->d_weight, w.data(), sizeof(double) * (size_t)n, cudaMemcpyHostToDevice));
  const i64 nbuf = std::max(imp->n_pack, imp->n_unpack);
  FB2_TRY(fb2_dalloc(&imp->d_snd_buf, (size_t)nbuf)); FB2_TRY(fb2_dalloc(&imp->d_rcv_buf, (size_t)nbuf));
  s->imp = imp;
  // rows from the first interface DOF onwards are treated as interface rows by the overlapped par apply: exact when the
  // numbering is interior-first (dof_import.f90:391-421), merely less overlap otherwise
  s->n_interior = min_itf; s->n_interface = n - min_itf;
  return 0;
}

// ------------------------------------------------------------------------------------------
// matrix assembled by the reference: csr_sparse_matrix_t (irp, ja, val), csr_sparse_matrix.f90:44-50
// ------------------------------------------------------------------------------------------
__global__ void k_irp32_to_64(i64 n1, const int* __restrict__ in, i64* out) {
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n1) out[i] = (i64)in[i] - 1;
}
__global__ void k_check_csr(i64 n, const i64* __restrict__ irp, const int* __restrict__ ja, i64 ncols, int* bad) {
  i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  if (irp[r + 1] < irp[r]) { *bad = 1; return; }
  for (i64 k = irp[r]; k < irp[r + 1]; ++k)
    if (ja[k] < 1 || ja[k] > ncols || (k > irp[r] && ja[k] <= ja[k - 1])) { *bad = 2; return; }
}

// space may be NULL (serial, the matrix owns a bare space) or a space whose num_free_dofs == num_rows (par: with the
// dof_import attached).  irp: Fortran irp(1) = 1, int32 as in the reference; ja 1-based ascending per row.
extern "C" int fb2_csr_create_from_host(fb2_matrix_t* out, fb2_space_t space, int64_t num_rows, const int32_t* irp,
                                        const int32_t* ja, const double* val, int symmetric_storage) {
  FB2_CHECK(out && num_rows >= 0 && irp && irp[0] == 1, "fb2_csr_create_from_host: bad argument (irp(1) must be 1)");
  FB2_CHECK(!space || space->n_free == num_rows, "fb2_csr_create_from_host: the space has a different number of DOFs");
  fb2_matrix_s* A = new fb2_matrix_s();
  if (!space) { FB2_TRY(fb2_fe_space_create_bare(&space, num_rows)); A->owns_space = 1; }
  A->s = space; A->sym = symmetric_storage ? 1 : 0; A->n = num_rows; A->nnz = (i64)irp[num_rows] - 1;
  A->map_state = -1;   // no structured scatter map: batches are scattered by searching ja
  cudaStream_t st = fb2_stream();
  const i64 n = A->n;
  FB2_TRY(fb2_dalloc(&A->d_irp, (size_t)n + 72));
  FB2_CUDA(cudaMemsetAsync(A->d_irp, 0, sizeof(i64) * ((size_t)n + 72), st));
  FB2_TRY(fb2_dalloc(&A->d_ja, (size_t)A->nnz + 16)); FB2_TRY(fb2_dalloc(&A->d_val, (size_t)A->nnz + 16));
  FB2_CUDA(cudaMemsetAsync(A->d_ja + A->nnz, 0, 16 * sizeof(int), st));
  FB2_CUDA(cudaMemsetAsync(A->d_val + A->nnz, 0, 16 * sizeof(double), st));
  int* d_irp32;
  FB2_TRY(fb2_dalloc(&d_irp32, (size_t)n + 1));
  FB2_CUDA(fb2_memcpy(d_irp32, irp, sizeof(int) * (size_t)(n + 1), cudaMemcpyHostToDevice));
  k_irp32_to_64<<<(unsigned)((n + 1 + 255) / 256), 256, 0, st>>>(n + 1, d_irp32, A->d_irp);
  // the tile table pads with the last offset
  if (A->nnz) FB2_CUDA(fb2_memcpy(A->d_ja, ja, sizeof(int) * (size_t)A->nnz, cudaMemcpyHostToDevice));
  if (A->nnz && val) FB2_CUDA(fb2_memcpy(A->d_val, val, sizeof(double) * (size_t)A->nnz, cudaMemcpyHostToDevice));
  else FB2_CUDA(cudaMemsetAsync(A->d_val, 0, sizeof(double) * (size_t)A->nnz, st));
  int *d_bad, h_bad = 0;
  FB2_TRY(fb2_dalloc(&d_bad, 1));
  FB2_CUDA(cudaMemsetAsync(d_bad, 0, sizeof(int), st));
  if (n) // the local matrix of the reference is square over the local DOFs (every vector has n entries): columns beyond n would make
  // the SpMV kernels gather past the end of x
  k_check_csr<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, A->d_irp, A->d_ja, n, d_bad);
  FB2_CUDA(fb2_memcpy(&h_bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost));
  cudaFree(d_bad); cudaFree(d_irp32);
  if (h_bad) { fb2_csr_destroy(A); FB2_CHECK(false, "fb2_csr_create_from_host: irp not monotone or ja not strictly ascending / out of range"); }
  *out = A;
  return 0;
}

// csr initialize_values (base_sparse_matrix.f90 initialize_values / init): every stored entry <- alpha
__global__ void k_fill_val(double* v, double a, i64 n) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) v[i] = a;
}
extern "C" int fb2_csr_initialize_values(fb2_matrix_t A, double alpha) {
  FB2_CHECK(A, "null matrix");
  if (A->nnz) k_fill_val<<<fb2_num_sms() * 8, 256, 0, fb2_stream()>>>(A->d_val, alpha, A->nnz);
  FB2_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------
// element-wise assembler fed in batches: sparse_assembler_t%assembly_matrix / assembly_array
// (sparse_assembler.f90:58-126,154-206) -> csr insert / array add with the skip rules of
// coo append (base_sparse_matrix.f90:3469-3513): ids < 1 or > n are skipped, ia > ja skipped in symmetric storage.
// With fixed_dof_values != NULL the strong-Dirichlet elimination of fe_cell_iterator%assembly
// (sbm_fe_cell_iterator.i90:1610-1637) is applied first: elvec -= elmat(:,j) * g(-id_j) for negative column ids.
// ------------------------------------------------------------------------------------------
__global__ void k_assemble_batch(int ncell, int nd, const int* __restrict__ dofs, const double* __restrict__ elmat,
                                 const double* __restrict__ elvec, const double* __restrict__ fixed, i64 num_fixed, i64 n, int sym,
                                 const i64* __restrict__ irp, const int* __restrict__ ja, double* val, double* rhs, int* missing) {
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int c = blockIdx.x * warps + w; c < ncell; c += gridDim.x * warps) {
    const int* g = dofs + (size_t)c * nd;
    const double* E = elmat ? elmat + (size_t)c * nd * nd : nullptr;
    if (E)
      for (int e = lane; e < nd * nd; e += 32) {
        const int i = e % nd, j = e / nd;            // column-major elmat(idof, jdof): test function = row
        const int gi = g[i], gj = g[j];
        if (gi < 1 || gi > n) continue;
        const double v = E[e];
        if (gj < 0) {
          if (fixed && rhs) {
            if (-(i64)gj > num_fixed) *missing = 2;           // Dirichlet id beyond the fixed values handed over
            else atomicAdd(&rhs[gi - 1], -v * fixed[-gj - 1]);
          }
          continue;
        }
        if (gj < 1 || gj > n || (sym && gi > gj)) continue;
        const i64 b = irp[gi - 1], end = irp[gi];
        i64 lo = b, hi = end;
        while (lo < hi) { i64 mid = (lo + hi) >> 1; if (ja[mid] < gj) lo = mid + 1; else hi = mid; }
        if (lo < end && ja[lo] == gj) atomicAdd(&val[lo], v); else *missing = 1;
      }
    if (elvec && rhs)
      for (int i = lane; i < nd; i += 32) {
        const int gi = g[i];
        if (gi >= 1 && gi <= n) atomicAdd(&rhs[gi - 1], elvec[(size_t)c * nd + i]);
      }
  }
}

extern "C" int fb2_assemble_batch(fb2_matrix_t A, fb2_vector_t rhs, int32_t num_cells, int32_t num_dofs_per_cell,
                                  const int32_t* fe_dofs, const double* elmat, const double* elvec,
                                  const double* fixed_dof_values, int64_t num_fixed) {
  FB2_CHECK(A && num_cells >= 0 && num_dofs_per_cell > 0 && fe_dofs, "fb2_assemble_batch: bad argument");
  FB2_CHECK(!rhs || rhs->n == A->n, "fb2_assemble_batch: rhs size mismatch");
  if (num_cells == 0) return 0;
  const size_t nd = (size_t)num_dofs_per_cell, nc = (size_t)num_cells;
  int *d_dofs = nullptr, *d_missing = nullptr, h_missing = 0;
  double *d_mat = nullptr, *d_vec = nullptr, *d_fix = nullptr;
  cudaStream_t st = fb2_stream();
  FB2_TRY(fb2_dalloc(&d_dofs, nc * nd)); FB2_TRY(fb2_dalloc(&d_missing, 1));
  FB2_CUDA(cudaMemsetAsync(d_missing, 0, sizeof(int), st));
  FB2_CUDA(fb2_memcpy(d_dofs, fe_dofs, sizeof(int) * nc * nd, cudaMemcpyHostToDevice));
  if (elmat) { FB2_TRY(fb2_dalloc(&d_mat, nc * nd * nd)); FB2_CUDA(fb2_memcpy(d_mat, elmat, sizeof(double) * nc * nd * nd, cudaMemcpyHostToDevice)); }
  if (elvec) { FB2_TRY(fb2_dalloc(&d_vec, nc * nd)); FB2_CUDA(fb2_memcpy(d_vec, elvec, sizeof(double) * nc * nd, cudaMemcpyHostToDevice)); }
  if (fixed_dof_values && num_fixed > 0) {
    FB2_TRY(fb2_dalloc(&d_fix, (size_t)num_fixed));
    FB2_CUDA(fb2_memcpy(d_fix, fixed_dof_values, sizeof(double) * (size_t)num_fixed, cudaMemcpyHostToDevice));
  }
  int blocks = (num_cells + 7) / 8;
  if (blocks > fb2_num_sms() * 8) blocks = fb2_num_sms() * 8;
  k_assemble_batch<<<blocks, 256, 0, st>>>(num_cells, num_dofs_per_cell, d_dofs, d_mat, d_vec, d_fix, d_fix ? num_fixed : 0, A->n, A->sym, A->d_irp,
                                           A->d_ja, A->d_val, rhs ? rhs->d : nullptr, d_missing);
  FB2_CUDA(cudaGetLastError());
  FB2_CUDA(fb2_memcpy(&h_missing, d_missing, sizeof(int), cudaMemcpyDeviceToHost));
  cudaFree(d_dofs); cudaFree(d_missing); cudaFree(d_mat); cudaFree(d_vec); cudaFree(d_fix);
  FB2_CHECK(h_missing != 2, "fb2_assemble_batch: a strong-Dirichlet id is beyond num_fixed");
  FB2_CHECK(h_missing == 0, "fb2_assemble_batch: an element entry has no slot in the CSR graph");
  return 0;
}
