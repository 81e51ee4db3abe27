// FE spaces numbered by the reference itself on UNSTRUCTURED conforming hex meshes: the host uploads
// serial_fe_space_t's cell-wise DOF lists (ptr_dofs_x_fe / lst_dofs_gids, sbm_serial_fe_space.i90:2623-2875) and the
// vertex coordinates of every cell; the device builds the CSR graph from those lists the way the reference does --
// every (row, column) pair a cell generates, sorted and compressed (base_sparse_matrix.f90:4479-4628,
// csr_sparse_matrix.f90:416-479) -- and integrates with the real per-cell Jacobians of the trilinear cell map
// (sbm_cell_map.i90:272-296).  No lattice and no scatter tables here: slots are found by binary search in the row.
#include <algorithm>
#include <cub/cub.cuh>
#include "fb2_internal.h"

extern "C" int fb2_fe_space_create_from_host(fb2_space_t* out, int32_t num_cells, int32_t num_local_cells, int order,
                                             int num_components, const int32_t* lst_dofs_gids, const double* cell_coords,
                                             int64_t num_free_dofs, int64_t num_dirichlet_dofs) {
  FB2_CHECK(out && lst_dofs_gids && cell_coords, "fb2_fe_space_create_from_host: null argument");
  FB2_CHECK(order >= 1 && order <= 4, "order must be 1..4");
  FB2_CHECK(num_components == 1 || num_components == 3, "num_components must be 1 or 3");
  FB2_CHECK(num_cells >= 0 && num_local_cells >= 0 && num_local_cells <= num_cells, "bad cell counts");
  FB2_CHECK(num_free_dofs >= 0 && num_free_dofs < 0x7fffffffLL && num_dirichlet_dofs >= 0, "bad DOF counts");
  fb2_space_s* s = new fb2_space_s();
  memset(&s->m, 0, sizeof(s->m));
  s->dir_mask = 0; s->part_id = 0;
  s->m.n_cells = num_cells; s->m.n_local = num_local_cells;
  int rc = fb2_space_setup_reference_fe(s, order, num_components);
  if (rc) { delete s; return rc; }
  const size_t total = (size_t)num_cells * s->rt.ndofs;
  // the lists must be consistent with the counts: free ids 1..n_free (ghost-only DOFs may exceed n_free, they are skipped by
  // the assembly exactly like the structured par spaces), Dirichlet ids -1..-n_dirichlet
  i64 max_dir = 0;
  for (size_t i = 0; i < total; ++i)
    if (lst_dofs_gids[i] < 0) max_dir = std::max<i64>(max_dir, -(i64)lst_dofs_gids[i]);
  if (max_dir > num_dirichlet_dofs) {
    fb2_fe_space_destroy(s);
    fb2_set_error("fb2_fe_space_create_from_host: Dirichlet id -%lld exceeds num_dirichlet_dofs", max_dir);
    return 2;
  }
  s->n_free = num_free_dofs; s->n_dirichlet = num_dirichlet_dofs; s->n_interior = num_free_dofs; s->n_interface = 0;
  FB2_TRY(fb2_dalloc(&s->d_gids, total));
  FB2_TRY(fb2_dalloc(&s->d_coords, (size_t)num_cells * 24));
  if (total) FB2_CUDA(fb2_memcpy(s->d_gids, lst_dofs_gids, sizeof(int) * total, cudaMemcpyHostToDevice));
  if (num_cells) FB2_CUDA(fb2_memcpy(s->d_coords, cell_coords, sizeof(double) * (size_t)num_cells * 24, cudaMemcpyHostToDevice));
  FB2_TRY(fb2_dalloc(&s->d_fixed, (size_t)s->n_dirichlet));
  FB2_CUDA(cudaMemsetAsync(s->d_fixed, 0, sizeof(double) * (s->n_dirichlet ? s->n_dirichlet : 1), fb2_stream()));
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  *out = s;
  return 0;
}

// ------------------------------------------------------------------------------------------
// graph: one 64-bit key (row << 32 | column) per stored pair of every LOCAL cell, radix sort, unique, row offsets by search
// ------------------------------------------------------------------------------------------
__global__ void k_pair_keys(const int* __restrict__ gids, i64 n_local, int nd, i64 n_free, int sym,
                            unsigned long long* __restrict__ keys) {
  const i64 total = n_local * nd * nd;
  for (i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (i64)gridDim.x * blockDim.x) {
    i64 c = t / (nd * nd);
    int e = (int)(t % (nd * nd)), i = e / nd, j = e % nd;
    int gi = gids[c * nd + i], gj = gids[c * nd + j];
    bool keep = gi > 0 && gi <= n_free && gj > 0 && gj <= n_free && (!sym || gj >= gi);
    keys[t] = keep ? (((unsigned long long)(unsigned)(gi - 1)) << 32 | (unsigned)gj) : ~0ull;
  }
}
__global__ void k_row_offsets(const unsigned long long* __restrict__ keys, i64 nkeys, i64 n, i64* __restrict__ irp) {
  i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (r > n) return;
  unsigned long long target = (unsigned long long)r << 32;   // first key of row r (columns are >= 1)
  i64 lo = 0, hi = nkeys;
  while (lo < hi) {
    i64 mid = (lo + hi) >> 1;
    if (keys[mid] < target) lo = mid + 1; else hi = mid;
  }
  irp[r] = lo;
}
__global__ void k_key_columns(const unsigned long long* __restrict__ keys, i64 nnz, int* __restrict__ ja) {
  i64 k = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nnz) ja[k] = (int)(keys[k] & 0xffffffffull);
}

int fb2_csr_graph_from_gids(fb2_matrix_s* A) {
  fb2_space_s* s = A->s;
  cudaStream_t st = fb2_stream();
  const i64 n = A->n;
  const int nd = s->rt.ndofs;
  const i64 nkeys = (i64)s->m.n_local * nd * nd;
  FB2_CHECK(nkeys < ((i64)1 << 33), "too many (row, column) pairs for the one-pass graph build of a host-built space");
  FB2_TRY(fb2_dalloc(&A->d_irp, (size_t)n + 72));
  FB2_CUDA(cudaMemsetAsync(A->d_irp, 0, sizeof(i64) * ((size_t)n + 72), st));
  unsigned long long *k0 = nullptr, *k1 = nullptr, *ku = nullptr;
  i64* d_num = nullptr;
  FB2_TRY(fb2_dalloc(&k0, (size_t)nkeys + 1)); FB2_TRY(fb2_dalloc(&k1, (size_t)nkeys + 1)); FB2_TRY(fb2_dalloc(&d_num, 1));
  i64 n_unique = 0;
  if (nkeys > 0) {
    k_pair_keys<<<fb2_num_sms() * 16, 256, 0, st>>>(s->d_gids, s->m.n_local, nd, n, A->sym, k0);
    FB2_CUDA(cudaGetLastError());
    cub::DoubleBuffer<unsigned long long> db(k0, k1);
    void* tmp = nullptr; size_t bytes = 0;
    int end_bit = 32 + 32;
    FB2_CUDA(cub::DeviceRadixSort::SortKeys(tmp, bytes, db, nkeys, 0, end_bit, st));
    FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
    cudaError_t e = cub::DeviceRadixSort::SortKeys(tmp, bytes, db, nkeys, 0, end_bit, st);
    cudaStreamSynchronize(st);
    cudaFree(tmp);
    FB2_CUDA(e);
    unsigned long long* sorted = db.Current();
    ku = db.Alternate();
    tmp = nullptr; bytes = 0;
    FB2_CUDA(cub::DeviceSelect::Unique(tmp, bytes, sorted, ku, d_num, nkeys, st));
    FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
    e = cub::DeviceSelect::Unique(tmp, bytes, sorted, ku, d_num, nkeys, st);
    cudaStreamSynchronize(st);
    cudaFree(tmp);
    FB2_CUDA(e);
    FB2_CUDA(fb2_memcpy(&n_unique, d_num, sizeof(i64), cudaMemcpyDeviceToHost));
    // the discarded pairs all carry the largest key: it is the last unique entry when present
    unsigned long long last = 0;
    if (n_unique > 0) FB2_CUDA(fb2_memcpy(&last, ku + n_unique - 1, sizeof(last), cudaMemcpyDeviceToHost));
    if (n_unique > 0 && last == ~0ull) --n_unique;
  }
  A->nnz = n_unique;
  FB2_TRY(fb2_dalloc(&A->d_ja, (size_t)A->nnz + 16));
  FB2_TRY(fb2_dalloc(&A->d_val, (size_t)A->nnz + 16));
  FB2_CUDA(cudaMemsetAsync(A->d_ja, 0, sizeof(int) * ((size_t)A->nnz + 16), st));
  FB2_CUDA(cudaMemsetAsync(A->d_val, 0, sizeof(double) * ((size_t)A->nnz + 16), st));
  if (n_unique > 0) {
    k_row_offsets<<<(unsigned)((n + 1 + 255) / 256), 256, 0, st>>>(ku, n_unique, n, A->d_irp);
    k_key_columns<<<(unsigned)((n_unique + 255) / 256), 256, 0, st>>>(ku, n_unique, A->d_ja);
    FB2_CUDA(cudaGetLastError());
  }
  FB2_CUDA(cudaStreamSynchronize(st));
  cudaFree(k0); cudaFree(k1); cudaFree(d_num);
  return 0;
}
