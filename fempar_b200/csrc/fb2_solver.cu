// CSR SpMV, BLAS-1 vector kernels, Jacobi and the preconditioned CG iteration on the device.
// Reference semantics: csr_sparse_matrix.f90:545-672 (matvec), serial_scalar_array.f90:282-419,556-588
// (dot/axpby/nrm2/copy/scal/entrywise_*), par_scalar_array.f90:303-350 (weighted dot + all-reduce),
// jacobi_preconditioner.f90:356-412, cg.f90:101-238,279-407.
#include <math.h>
#include <stdlib.h>
#include "fb2_internal.h"

// ------------------------------------------------------------------------------------------
// deterministic reductions: per-block partials (fixed grid), then one block sums them in order
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}
// sums `v` over the block in a fixed order; result valid on thread 0
__device__ __forceinline__ double block_sum(double v, double* s_red /*[32]*/) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) s_red[w] = v;
  __syncthreads();
  double r = 0.0;
  if (threadIdx.x == 0)
    for (int i = 0; i < nw; ++i) r += s_red[i];
  return r;
}

// out[k] = sum_i partials[k*stride + i], i < nparts, fixed order; one block
__global__ void k_reduce_partials(const double* __restrict__ partials, int nparts, int stride, int nvals, double* out) {
  __shared__ double s_red[32];
  for (int k = 0; k < nvals; ++k) {
    double v = 0.0;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) v += partials[k * stride + i];
    double r = block_sum(v, s_red);
    if (threadIdx.x == 0) out[k] = r;
    __syncthreads();
  }
}

__global__ void k_dot(const double* __restrict__ x, const double* __restrict__ y, const double* __restrict__ w, i64 n,
                      double* partials) {
  __shared__ double s_red[32];
  double acc = 0.0;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x)
    acc += (w ? w[i] : 1.0) * x[i] * y[i];
  double r = block_sum(acc, s_red);
  if (threadIdx.x == 0) partials[blockIdx.x] = r;
}

static int reduce_grid(i64 n) {
  i64 b = (n + 1023) / 1024;
  i64 cap = (i64)fb2_num_sms() * 8;
  if (cap > FB2_MAX_PARTIALS) cap = FB2_MAX_PARTIALS;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// d_result (device) = sum w x y over the local entries; the caller all-reduces when par
int fb2_dot_device(const double* x, const double* y, const double* w, i64 n, double* d_result) {
  int g = reduce_grid(n);
  double* part = fb2_scratch_partials();
  k_dot<<<g, 256, 0, fb2_stream()>>>(x, y, w, n, part);
  k_reduce_partials<<<1, 256, 0, fb2_stream()>>>(part, g, 0, 1, d_result);
  FB2_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------
// SpMV: warp per row, 4 independent 32-wide passes in flight per warp, coalesced val/ja streams
// ------------------------------------------------------------------------------------------
struct CgState {
  double rz, App, rr, rz_new, pp, r2z, tol, tol_extra, rtol, atol, alpha, beta, err, err_extra;   // rr..r2z adjacent (one reduction)
  int iter, converged, done, maxit, crit, icg;
};

template <bool FUSE_DOT>
__global__ void __launch_bounds__(256) k_spmv(i64 n, const i64* __restrict__ irp, const int* __restrict__ ja,
                                              const double* __restrict__ val, const double* __restrict__ x,
                                              double* __restrict__ y, double* partials, const CgState* st) {
  __shared__ double s_red[32];
  if (st && st->done) return;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, warps = blockDim.x >> 5;
  double dacc = 0.0;
  for (i64 row = (i64)blockIdx.x * warps + w; row < n; row += (i64)gridDim.x * warps) {
    const i64 start = irp[row], end = irp[row + 1];
    double acc = 0.0;
    for (i64 e = start + lane; e < end; e += 128) {
      const i64 e1 = e + 32, e2 = e + 64, e3 = e + 96;
      double v0 = val[e];
      int c0 = ja[e];
      double v1 = 0.0, v2 = 0.0, v3 = 0.0;
      int c1 = 1, c2 = 1, c3 = 1;
      if (e1 < end) { v1 = val[e1]; c1 = ja[e1]; }
      if (e2 < end) { v2 = val[e2]; c2 = ja[e2]; }
      if (e3 < end) { v3 = val[e3]; c3 = ja[e3]; }
      acc += v0 * x[c0 - 1];
      acc += v1 * x[c1 - 1];
      acc += v2 * x[c2 - 1];
      acc += v3 * x[c3 - 1];
    }
    acc = warp_sum(acc);
    if (lane == 0) {
      y[row] = acc;
      if (FUSE_DOT) dacc += acc * x[row];
    }
  }
  if (FUSE_DOT) {
    double r = block_sum(dacc, s_red);
    if (threadIdx.x == 0) partials[blockIdx.x] = r;
  }
}

// symmetric storage (upper triangle, diagonal first in each row): y = (U + U^T - D) x, csr_sparse_matrix.f90:597-623
__global__ void k_spmv_sym(i64 n, const i64* __restrict__ irp, const int* __restrict__ ja, const double* __restrict__ val,
                           const double* __restrict__ x, double* __restrict__ y) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, warps = blockDim.x >> 5;
  for (i64 row = (i64)blockIdx.x * warps + w; row < n; row += (i64)gridDim.x * warps) {
    const i64 start = irp[row], end = irp[row + 1];
    const double xr = x[row];
    double acc = 0.0;
    for (i64 e = start + lane; e < end; e += 32) {
      int c = ja[e] - 1;
      double v = val[e];
      acc += v * x[c];
      if (c != row) atomicAdd(&y[c], v * xr);
    }
    acc = warp_sum(acc);
    if (lane == 0) atomicAdd(&y[row], acc);
  }
}

static int g_spmv_variant = 0;   // 0: TMA-streamed tiles (fb2_spmv.cu), 1: warp-per-row
// par apply: overlap the halo exchange with the interior rows (FB2_PAR_OVERLAP=0 restores the sequential order);
// interior tiles per short-lived CTA (FB2_PAR_CHUNK)
static int g_par_overlap = getenv("FB2_PAR_OVERLAP") ? atoi(getenv("FB2_PAR_OVERLAP")) : 1;
static int g_par_chunk = getenv("FB2_PAR_CHUNK") ? atoi(getenv("FB2_PAR_CHUNK")) : 64;
extern "C" int fb2_set_spmv_variant(int v) {
  // 2 = streamed tiles with dictionary-compressed column indices, for matrices whose first SpMV comes after this call
  fb2_spmv_set_compress(v == 2);
  fb2_spmv_set_classes(v != 3);
  g_spmv_variant = (v == 2 || v == 3) ? 0 : v;   // 3: streamed tiles reading ja for every tile (fb2_spmv.cu)
  return 0;
}

static int spmv_grid(i64 n, int warps_per_block) {
  i64 b = (n + warps_per_block - 1) / warps_per_block;
  i64 cap = (i64)fb2_num_sms() * (2048 / (warps_per_block * 32));
  if (cap > FB2_MAX_PARTIALS) cap = FB2_MAX_PARTIALS;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// y = A x on the local rows (no halo exchange)
static int spmv_local(fb2_matrix_t A, const double* x, double* y, bool fuse_dot, const CgState* st, int* grid_out) {
  cudaStream_t s = fb2_stream();
  if (A->sym) {
    FB2_CUDA(cudaMemsetAsync(y, 0, sizeof(double) * (size_t)A->n, s));
    k_spmv_sym<<<spmv_grid(A->n, 8), 256, 0, s>>>(A->n, A->d_irp, A->d_ja, A->d_val, x, y);
    if (grid_out) *grid_out = 0;
  } else {
    if (g_spmv_variant == 0) {
      int g = 0;
      FB2_TRY(fb2_spmv_stream(A, x, y, fuse_dot, st ? &st->done : nullptr, &g));
      if (grid_out) *grid_out = g;
      return 0;
    }
    int g = spmv_grid(A->n, 8);
    if (fuse_dot) k_spmv<true><<<g, 256, 0, s>>>(A->n, A->d_irp, A->d_ja, A->d_val, x, y, fb2_scratch_partials(), st);
    else k_spmv<false><<<g, 256, 0, s>>>(A->n, A->d_irp, A->d_ja, A->d_val, x, y, nullptr, st);
    if (grid_out) *grid_out = g;
  }
  FB2_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int fb2_spmv(fb2_matrix_t A, fb2_vector_t x, fb2_vector_t y) {
  FB2_CHECK(A && x && y && x->n == A->n && y->n == A->n, "fb2_spmv: size mismatch");
  cudaStream_t st = fb2_stream();
  cudaEvent_t e0, e1, e2;
  cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
  cudaEventRecord(e0, st);
  FB2_TRY(spmv_local(A, x->d, y->d, false, nullptr, nullptr));
  cudaEventRecord(e1, st);
  if (A->s->imp && fb2_comm_active()) FB2_TRY(fb2_import_exchange(A->s->imp, y->d));
  cudaEventRecord(e2, st);
  FB2_CUDA(cudaStreamSynchronize(st));
  float ms = 0, ms2 = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventElapsedTime(&ms2, e1, e2);
  fb2_record_phase("spmv", ms);
  fb2_record_phase("halo", ms2);
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
  return 0;
}

// ------------------------------------------------------------------------------------------
// vector kernels
// ------------------------------------------------------------------------------------------
__global__ void k_init(double* x, double a, i64 n) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) x[i] = a;
}
__global__ void k_scal(double* d, double a, const double* s, i64 n) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) d[i] = a * s[i];
}
// op1 = alpha*op2 + beta*op1 (serial_scalar_array.f90:366-402: beta==0 -> copy+scal, no read of op1)
__global__ void k_axpby(double* op1, double a, const double* op2, double b, i64 n) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x)
    op1[i] = (b == 0.0) ? a * op2[i] : a * op2[i] + b * op1[i];
}
__global__ void k_ew_product(double* z, const double* x, const double* y, i64 n) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) z[i] = x[i] * y[i];
}
__global__ void k_ew_invert(double* v, i64 n) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) v[i] = 1.0 / v[i];
}
static int vec_grid(i64 n) {
  i64 b = (n + 255) / 256, cap = (i64)fb2_num_sms() * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

extern "C" int fb2_vec_create_n(fb2_vector_t* out, int64_t n) {
  fb2_vector_s* v = new fb2_vector_s();
  v->n = n;
  FB2_TRY(fb2_dalloc(&v->d, (size_t)n));
  FB2_CUDA(cudaMemsetAsync(v->d, 0, sizeof(double) * (size_t)(n ? n : 1), fb2_stream()));
  *out = v;
  return 0;
}
extern "C" int fb2_vec_create(fb2_vector_t* out, fb2_space_t s) {
  FB2_TRY(fb2_vec_create_n(out, s->n_free));
  (*out)->s = s;
  return 0;
}
extern "C" int fb2_vec_destroy(fb2_vector_t v) {
  if (!v) return 0;
  cudaFree(v->d);
  delete v;
  return 0;
}
extern "C" int fb2_vec_size(fb2_vector_t v, int64_t* n) { *n = v->n; return 0; }
extern "C" void* fb2_vec_device_ptr(fb2_vector_t v) { return v->d; }
extern "C" int fb2_vec_set(fb2_vector_t v, const double* h) {
  FB2_CUDA(cudaMemcpyAsync(v->d, h, sizeof(double) * (size_t)v->n, cudaMemcpyHostToDevice, fb2_stream()));
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  return 0;
}
extern "C" int fb2_vec_get(fb2_vector_t v, double* h) {
  FB2_CUDA(cudaMemcpyAsync(h, v->d, sizeof(double) * (size_t)v->n, cudaMemcpyDeviceToHost, fb2_stream()));
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  return 0;
}
extern "C" int fb2_vec_init(fb2_vector_t v, double a) {
  k_init<<<vec_grid(v->n), 256, 0, fb2_stream()>>>(v->d, a, v->n);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int fb2_vec_copy(fb2_vector_t dst, fb2_vector_t src) {
  FB2_CHECK(dst->n == src->n, "size mismatch");
  FB2_CUDA(cudaMemcpyAsync(dst->d, src->d, sizeof(double) * (size_t)dst->n, cudaMemcpyDeviceToDevice, fb2_stream()));
  return 0;
}
extern "C" int fb2_vec_scal(fb2_vector_t dst, double a, fb2_vector_t src) {
  FB2_CHECK(dst->n == src->n, "size mismatch");
  k_scal<<<vec_grid(dst->n), 256, 0, fb2_stream()>>>(dst->d, a, src->d, dst->n);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int fb2_vec_axpby(fb2_vector_t op1, double alpha, fb2_vector_t op2, double beta) {
  FB2_CHECK(op1->n == op2->n, "size mismatch");
  k_axpby<<<vec_grid(op1->n), 256, 0, fb2_stream()>>>(op1->d, alpha, op2->d, beta, op1->n);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
static const double* weight_of(fb2_vector_t v) {
  return (v->s && v->s->imp && fb2_comm_active()) ? fb2_import_weight(v->s->imp) : nullptr;
}
extern "C" int fb2_vec_dot(fb2_vector_t x, fb2_vector_t y, double* result) {
  FB2_CHECK(x->n == y->n, "size mismatch");
  FB2_CHECK(!(weight_of(x) && fb2_comm_active() && fb2_comm_parts_per_rank() > 1),
            "fb2_vec_dot on one of several sub-domains per rank: use fb2_vec_dot_multi");
  double* d_res = fb2_scratch_partials() + FB2_MAX_PARTIALS * 4;
  FB2_TRY(fb2_dot_device(x->d, y->d, weight_of(x), x->n, d_res));
  if (fb2_comm_active()) FB2_TRY(fb2_allreduce_sum(d_res, 1));
  FB2_CUDA(cudaMemcpyAsync(result, d_res, sizeof(double), cudaMemcpyDeviceToHost, fb2_stream()));
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  return 0;
}
// vector_t%local_dot (vector.f90:41, par_scalar_array.f90:303-326): the weighted dot of this part without the all-reduce
extern "C" int fb2_vec_local_dot(fb2_vector_t x, fb2_vector_t y, double* result) {
  FB2_CHECK(x && y && result && x->n == y->n, "size mismatch");
  double* d_res = fb2_scratch_partials() + FB2_MAX_PARTIALS * 4;
  const double* w = (x->s && x->s->imp) ? fb2_import_weight(x->s->imp) : nullptr;
  FB2_TRY(fb2_dot_device(x->d, y->d, w, x->n, d_res));
  FB2_CUDA(fb2_memcpy(result, d_res, sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}
// vector_t%nullify_non_owned_dofs (par_scalar_array.f90): x_i <- 0 where this part does not own DOF i (weight 0)
__global__ void k_nullify_non_owned(double* x, const double* __restrict__ w, i64 n) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x)
    if (w[i] == 0.0) x[i] = 0.0;
}
extern "C" int fb2_vec_nullify_non_owned_dofs(fb2_vector_t x) {
  FB2_CHECK(x, "null vector");
  if (x->s && x->s->imp) k_nullify_non_owned<<<vec_grid(x->n), 256, 0, fb2_stream()>>>(x->d, fb2_import_weight(x->s->imp), x->n);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
// vector_t%extract_subvector / insert_subvector on an index subset (vector.f90:161-187; indices 1-based like the reference)
__global__ void k_gather_idx(const double* __restrict__ x, const int* __restrict__ idx, int m, double* out) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < m) out[k] = x[idx[k] - 1];
}
__global__ void k_scatter_idx(double* x, const int* __restrict__ idx, int m, const double* __restrict__ in) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < m) x[idx[k] - 1] = in[k];
}
static int subvector(fb2_vector_t v, int32_t m, const int32_t* indices, double* values, bool extract) {
  FB2_CHECK(v && m >= 0 && (m == 0 || (indices && values)), "bad argument");
  if (m == 0) return 0;
  for (int k = 0; k < m; ++k) FB2_CHECK(indices[k] >= 1 && indices[k] <= v->n, "subvector index out of range");
  int* d_idx; double* d_val;
  FB2_TRY(fb2_dalloc(&d_idx, (size_t)m)); FB2_TRY(fb2_dalloc(&d_val, (size_t)m));
  FB2_CUDA(fb2_memcpy(d_idx, indices, sizeof(int) * (size_t)m, cudaMemcpyHostToDevice));
  if (extract) {
    k_gather_idx<<<(m + 255) / 256, 256, 0, fb2_stream()>>>(v->d, d_idx, m, d_val);
    FB2_CUDA(fb2_memcpy(values, d_val, sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost));
  } else {
    FB2_CUDA(fb2_memcpy(d_val, values, sizeof(double) * (size_t)m, cudaMemcpyHostToDevice));
    k_scatter_idx<<<(m + 255) / 256, 256, 0, fb2_stream()>>>(v->d, d_idx, m, d_val);
    FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  }
  cudaFree(d_idx); cudaFree(d_val);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int fb2_vec_extract_subvector(fb2_vector_t v, int32_t size_indices, const int32_t* indices, double* values) {
  return subvector(v, size_indices, indices, values, true);
}
extern "C" int fb2_vec_insert_subvector(fb2_vector_t v, int32_t size_indices, const int32_t* indices, const double* values) {
  return subvector(v, size_indices, indices, const_cast<double*>(values), false);
}
extern "C" int fb2_vec_nrm2(fb2_vector_t x, double* result) {
  FB2_TRY(fb2_vec_dot(x, x, result));
  *result = sqrt(*result);
  return 0;
}
extern "C" int fb2_vec_entrywise_product(fb2_vector_t z, fb2_vector_t x, fb2_vector_t y) {
  FB2_CHECK(z->n == x->n && z->n == y->n, "size mismatch");
  k_ew_product<<<vec_grid(z->n), 256, 0, fb2_stream()>>>(z->d, x->d, y->d, z->n);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int fb2_vec_entrywise_invert(fb2_vector_t v) {
  k_ew_invert<<<vec_grid(v->n), 256, 0, fb2_stream()>>>(v->d, v->n);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int fb2_vec_comm(fb2_vector_t v) {
  if (v->s && v->s->imp && fb2_comm_active()) FB2_TRY(fb2_import_exchange(v->s->imp, v->d));
  return 0;
}

// ------------------------------------------------------------------------------------------
// Jacobi
// ------------------------------------------------------------------------------------------
__global__ void k_extract_diag(i64 n, const i64* __restrict__ irp, const int* __restrict__ ja, const double* __restrict__ val,
                               double* d) {
  i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  i64 lo = irp[r], hi = irp[r + 1];
  const i64 end = hi;
  int col = (int)(r + 1);
  while (lo < hi) {
    i64 mid = (lo + hi) >> 1;
    if (ja[mid] < col) lo = mid + 1; else hi = mid;
  }
  d[r] = (lo < end && ja[lo] == col) ? val[lo] : 0.0;
}
extern "C" int fb2_extract_diagonal(fb2_matrix_t A, fb2_vector_t d) {
  FB2_CHECK(d->n == A->n, "size mismatch");
  if (A->n) k_extract_diag<<<(unsigned)((A->n + 255) / 256), 256, 0, fb2_stream()>>>(A->n, A->d_irp, A->d_ja, A->d_val, d->d);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int fb2_jacobi_setup(fb2_matrix_t A, fb2_vector_t dinv) {
  FB2_TRY(fb2_extract_diagonal(A, dinv));
  if (A->s->imp && fb2_comm_active()) FB2_TRY(fb2_import_exchange(A->s->imp, dinv->d));
  FB2_TRY(fb2_vec_entrywise_invert(dinv));
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  return 0;
}

// ------------------------------------------------------------------------------------------
// fused PCG kernels (scalars stay on the device; every kernel is a no-op once st->done is set,
// so the host may enqueue iterations ahead and still report the reference's exact iteration count)
// ------------------------------------------------------------------------------------------
// x += alpha p ; r -= alpha Ap ; partials of <r,r>_w and <r, dinv r>_w
__global__ void __launch_bounds__(256) k_cg_update_xr(i64 n, CgState* st, const double* __restrict__ p,
                                                      const double* __restrict__ Ap, const double* __restrict__ dinv,
                                                      const double* __restrict__ w, double* __restrict__ x,
                                                      double* __restrict__ r, double* partials, int stride) {
  __shared__ double s_red[32];
  if (st->done) return;
  const double App = st->App;
  const double alpha = (App != 0.0) ? st->rz / App : 0.0;   // cg.f90:195-199
  const bool need_pp = st->crit >= FB2_STOP_DELTA_RHS, icg = st->icg != 0;
  double a_rr = 0.0, a_rz = 0.0, a_pp = 0.0, a_r2z = 0.0;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) {
    const double pi = p[i], r_old = r[i];
    x[i] += alpha * pi;
    double ri = r_old - alpha * Ap[i];
    r[i] = ri;
    double wi = w ? w[i] : 1.0;
    const double zi = dinv ? dinv[i] * ri : ri;      // z = inv(M) r of the NEXT direction update
    a_rr += wi * ri * ri;
    a_rz += wi * ri * zi;
    a_pp += wi * pi * pi;
    a_r2z += wi * r_old * zi;                        // <r2, z> of icg.f90:208-222 (r2 = residual before the update)
  }
  double s1 = block_sum(a_rr, s_red);
  double s2 = block_sum(a_rz, s_red);
  double s3 = need_pp ? block_sum(a_pp, s_red) : 0.0;
  double s4 = icg ? block_sum(a_r2z, s_red) : 0.0;
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = s1; partials[stride + blockIdx.x] = s2; partials[2 * stride + blockIdx.x] = s3;
    partials[3 * stride + blockIdx.x] = s4;
  }
}
// convergence bookkeeping: cg.f90:208-226 and update_convergence_data_and_evaluate_stopping_criteria :331-407 (all ten criteria)
__global__ void k_cg_scalars(CgState* st, double* hist, double* hist_extra) {
  if (st->done) return;
  st->iter += 1;
  const int crit = st->crit;
  const double alpha = (st->App != 0.0) ? st->rz / st->App : 0.0;
  st->alpha = alpha;
  double e;
  if (crit == FB2_STOP_RES_RES || crit == FB2_STOP_RES_RHS) e = sqrt(st->rr);                 // || r(i) ||
  else if (crit == FB2_STOP_RES_NRMGIVEN_RHS_NRMGIVEN || crit == FB2_STOP_RES_NRMGIVEN_RES_NRMGIVEN)
    e = sqrt(st->rz);   // r_nrm_M as the reference passes it: still sqrt(<r,z>) of the PREVIOUS residual (cg.f90:208 vs :228-231)
  else {
    e = alpha * sqrt(st->pp);                                                                 // || dx(i) || = alpha ||p||
    if (st->iter == 1 && (crit == FB2_STOP_DELTA_DELTA || crit == FB2_STOP_DELTA_DELTA_AND_RES_RES ||
                          crit == FB2_STOP_DELTA_DELTA_AND_RES_RHS))
      st->tol = st->rtol * e + st->atol;                                                      // cg.f90:349-357
  }
  st->err = e;
  if (hist) hist[st->iter - 1] = e;
  bool conv = e <= st->tol;
  if (conv && crit >= FB2_STOP_DELTA_RHS_AND_RES_RES) {                                       // 2nd criterion, cg.f90:391-403
    st->err_extra = sqrt(st->rr);
    conv = st->err_extra <= st->tol_extra;
  }
  // the 2nd criterion's estimate as the reference's per-iteration line prints it: updated only when the 1st one is met
  if (hist_extra) hist_extra[st->iter - 1] = st->err_extra;
  st->converged = conv ? 1 : 0;
  if (conv || st->iter >= st->maxit) { st->done = 1; return; }
  st->beta = st->icg ? (1.0 / st->rz) * (st->rz_new - st->r2z) : st->rz_new / st->rz;   // icg.f90:216-222 / cg.f90:220-226
  st->rz = st->rz_new;
}
// p = dinv r + beta p
__global__ void __launch_bounds__(256) k_cg_update_p(i64 n, const CgState* st, const double* __restrict__ r,
                                                     const double* __restrict__ dinv, double* __restrict__ p) {
  if (st->done) return;
  const double beta = st->beta;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x)
    p[i] = (dinv ? dinv[i] * r[i] : r[i]) + beta * p[i];
}
__global__ void k_cg_init_state(CgState* st, double rz, double tol, double tol_extra, double rtol, double atol, int maxit, int crit,
                                int icg) {
  st->icg = icg; st->r2z = 0;
  st->rz = rz; st->App = 0; st->rr = 0; st->rz_new = 0; st->pp = 0; st->tol = tol; st->tol_extra = tol_extra; st->rtol = rtol;
  st->atol = atol; st->alpha = 0; st->beta = 0; st->err = 0; st->err_extra = 0;
  st->iter = 0; st->converged = 0; st->done = 0; st->maxit = maxit; st->crit = crit;
}

// Per-rank timeline of the overlapped par apply (FB2_PAR_TIMELINE=1): events after the interface tiles, after the halo exchange
// (comm stream), after the interior tiles and after the wait for the exchange, for every iteration of fb2_pcg_iterate[_multi];
// averaged and published as phases par_itf_ms / par_exchange_ms / par_interior_ms / par_exposed_wait_ms / par_rest_ms.
struct ParTimeline {
  bool on = false;
  std::vector<cudaEvent_t> ev;     // 6 per iteration: start, itf_done, halo_done (comm stream), interior_done, wait_done, iter_end
  int iters = 0;
  void begin(int n_iterations) {
    on = n_iterations > 0 && getenv("FB2_PAR_TIMELINE") && atoi(getenv("FB2_PAR_TIMELINE")) != 0;
    if (!on) return;
    for (cudaEvent_t e : ev) cudaEventDestroy(e);
    ev.assign((size_t)6 * n_iterations, nullptr);
    for (auto& e : ev) cudaEventCreate(&e);
    iters = 0;
  }
  cudaEvent_t at(int k) { return on && (size_t)(6 * iters + k) < ev.size() ? ev[6 * iters + k] : nullptr; }
  void mark(int k, cudaStream_t st) {
    cudaEvent_t e = at(k);
    if (!e) return;
    if (k == 0) marked = true;
    if (marked) cudaEventRecord(e, st);      // iterations that do not take the overlapped branch are not counted
  }
  bool marked = false;
  void next() { if (on && marked) ++iters; marked = false; }
  void publish() {
    if (!on || iters == 0) return;
    double acc[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < iters; ++i) {
      float t;
      cudaEvent_t* e = &ev[6 * i];
      cudaEventElapsedTime(&t, e[0], e[1]); acc[0] += t;     // interface tiles
      cudaEventElapsedTime(&t, e[1], e[2]); acc[1] += t;     // halo exchange (pack, send/recv x2, unpack) on the comm stream
      cudaEventElapsedTime(&t, e[1], e[3]); acc[2] += t;     // interior tiles
      cudaEventElapsedTime(&t, e[3], e[4]); acc[3] += t;     // exchange not hidden behind the interior tiles
      cudaEventElapsedTime(&t, e[4], e[5]); acc[4] += t;     // dots, all-reduces, vector updates
    }
    const char* names[5] = {"par_itf_ms", "par_exchange_ms", "par_interior_ms", "par_exposed_wait_ms", "par_rest_ms"};
    for (int k = 0; k < 5; ++k) fb2_record_phase(names[k], acc[k] / iters);
    on = false;
  }
};
static ParTimeline g_tl;

static int g_cg_icg = 0;   // 1 while fb2_ils_solve runs the ICG variant (icg.f90) through the fused CG kernels
void fb2_cg_set_icg(int on) { g_cg_icg = on; }

struct CgWork {
  i64 n = 0;
  double *r = nullptr, *p = nullptr, *Ap = nullptr, *hist = nullptr;
  CgState* st = nullptr;
  int hist_cap = 0;
};
// Solver state lives in the matrix handle (re-entrant across handles): Krylov work vectors, the device-resident
// CgState, the convergence history, the user's initial solution and the staging vectors of the host-buffer entry point.
// Two operators of different size can be solved alternately without re-allocation or cross-talk.
struct fb2_solver_ws {
  CgWork w;
  double* x0 = nullptr;
  i64 x0_n = 0;
  bool x0_on = false;
  fb2_vector_t vb = nullptr, vx = nullptr;
  double last[4] = {0, 0, 0, 0};   // rhs_convergence_test, rhs_extra_convergence_test, error estimates of the last solve
  int last_iters = 0, last_tracked = 0;
};
static fb2_solver_ws* ws_of(fb2_matrix_t A) {
  if (!A->ws) A->ws = new fb2_solver_ws();
  return A->ws;
}
static CgWork& work_of(fb2_matrix_t A) { return ws_of(A)->w; }
void fb2_solver_ws_free(fb2_matrix_s* A) {
  if (!A->ws) return;
  CgWork& W = A->ws->w;
  cudaFree(W.r); cudaFree(W.p); cudaFree(W.Ap); cudaFree(W.hist); cudaFree(W.st); cudaFree(A->ws->x0);
  if (A->ws->vb) { fb2_vec_destroy(A->ws->vb); fb2_vec_destroy(A->ws->vx); }
  delete A->ws;
  A->ws = nullptr;
}
static int cg_workspace(fb2_matrix_t A, int hist_len) {
  CgWork& W = work_of(A);
  const i64 n = A->n;
  if (W.n != n) {
    cudaFree(W.r); cudaFree(W.p); cudaFree(W.Ap);
    FB2_TRY(fb2_dalloc(&W.r, (size_t)n)); FB2_TRY(fb2_dalloc(&W.p, (size_t)n)); FB2_TRY(fb2_dalloc(&W.Ap, (size_t)n));
    W.n = n;
  }
  if (!W.st) FB2_TRY(fb2_dalloc(&W.st, 1));
  if (hist_len > W.hist_cap) {
    cudaFree(W.hist);
    FB2_TRY(fb2_dalloc(&W.hist, 2 * (size_t)hist_len));   // error estimates of the 1st criterion, then of the 2nd
    W.hist_cap = hist_len;
  }
  return 0;
}
// the work vectors belong to the matrix handles (released by fb2_csr_destroy, or earlier through this call)
extern "C" int fb2_csr_free_solver_workspace(fb2_matrix_t A) {
  if (A) fb2_solver_ws_free(A);
  return 0;
}
extern "C" int fb2_pcg_free_workspace(void) { return 0; }   // kept for ABI compatibility: nothing is process-global any more

// one PCG iteration enqueued on the stream
static int cg_enqueue_iteration(fb2_matrix_t A, const double* dinv, const double* w, double* x, bool par, double* hist,
                                cudaEvent_t ev_spmv0, cudaEvent_t ev_spmv1) {
  cudaStream_t s = fb2_stream();
  CgWork& W = work_of(A);
  double* part = fb2_scratch_partials();
  const i64 n = A->n;
  int g = 0;
  if (ev_spmv0) cudaEventRecord(ev_spmv0, s);
  if (!par && !A->sym) {
    FB2_TRY(spmv_local(A, W.p, W.Ap, true, W.st, &g));                       // Ap = A p, partial <Ap,p> fused
    if (ev_spmv1) cudaEventRecord(ev_spmv1, s);
    k_reduce_partials<<<1, 1024, 0, s>>>(part, g, 0, 1, &W.st->App);
  } else if (par && !A->sym && g_spmv_variant == 0 && g_par_overlap) {
    // Overlapped par apply (par_sparse_matrix.f90:739-762): interface rows are numbered last, so
    //   (1) the tiles holding interface rows run first,
    //   (2) their halo exchange (pack / NCCL send-recv / gather-add / owner broadcast) runs on the high-priority comm stream
    //   (3) while the interior tiles run as short-lived CTAs on the compute stream with <Ap,p> fused (weights are 1 there),
    //   (4) then the weighted <Ap,p> of the interface rows is added and the partials are summed in a fixed order.
    i64 t_itf = 0, row_itf = 0;
    FB2_TRY(fb2_spmv_interface_split(A, A->s->n_interior, &t_itf, &row_itf));
    static cudaEvent_t ev_itf = nullptr, ev_halo = nullptr;
    if (!ev_itf) { cudaEventCreateWithFlags(&ev_itf, cudaEventDisableTiming); cudaEventCreateWithFlags(&ev_halo, cudaEventDisableTiming); }
    cudaStream_t cs = fb2_comm_stream();
    g_tl.mark(0, s);
    FB2_TRY(fb2_spmv_stream_range(A, W.p, W.Ap, false, &W.st->done, t_itf, A->spmv_ntiles, 0, nullptr, nullptr));
    FB2_CUDA(cudaEventRecord(ev_itf, s));
    g_tl.mark(1, s);
    FB2_CUDA(cudaStreamWaitEvent(cs, ev_itf, 0));
    FB2_TRY(fb2_import_exchange_on(A->s->imp, W.Ap, cs));
    FB2_CUDA(cudaEventRecord(ev_halo, cs));
    g_tl.mark(2, cs);
    FB2_TRY(fb2_spmv_stream_range(A, W.p, W.Ap, true, &W.st->done, 0, t_itf, g_par_chunk, part, &g));
    g_tl.mark(3, s);
    FB2_CUDA(cudaStreamWaitEvent(s, ev_halo, 0));
    g_tl.mark(4, s);
    if (ev_spmv1) cudaEventRecord(ev_spmv1, s);
    int gd = 0;
    if (row_itf < n) {
      gd = reduce_grid(n - row_itf);
      if (gd > 64) gd = 64;
      k_dot<<<gd, 256, 0, s>>>(W.Ap + row_itf, W.p + row_itf, w + row_itf, n - row_itf, part + g);
    }
    k_reduce_partials<<<1, 1024, 0, s>>>(part, g + gd, 0, 1, &W.st->App);
    FB2_TRY(fb2_allreduce_sum(&W.st->App, 1));
  } else {
    FB2_TRY(spmv_local(A, W.p, W.Ap, false, A->sym ? nullptr : W.st, &g));
    if (par) FB2_TRY(fb2_import_exchange(A->s->imp, W.Ap));                  // par_sparse_matrix.f90:757-758
    if (ev_spmv1) cudaEventRecord(ev_spmv1, s);                              // par: the timed apply includes the halo exchange
    FB2_TRY(fb2_dot_device(W.Ap, W.p, w, n, &W.st->App));
    if (par) FB2_TRY(fb2_allreduce_sum(&W.st->App, 1));
  }
  int g2 = reduce_grid(n);
  k_cg_update_xr<<<g2, 256, 0, s>>>(n, W.st, W.p, W.Ap, dinv, w, x, W.r, part, FB2_MAX_PARTIALS);
  k_reduce_partials<<<1, 256, 0, s>>>(part, g2, FB2_MAX_PARTIALS, 4, &W.st->rr);   // rr, rz_new, pp, r2z are adjacent
  if (par) FB2_TRY(fb2_allreduce_sum(&W.st->rr, 4));
  k_cg_scalars<<<1, 1, 0, s>>>(W.st, hist, hist ? hist + W.hist_cap : nullptr);
  k_cg_update_p<<<vec_grid(n), 256, 0, s>>>(n, W.st, W.r, dinv, W.p);
  g_tl.mark(5, s); g_tl.next();
  FB2_CUDA(cudaGetLastError());
  return 0;
}

// initial solution of the next solves on this operator (base_iterative_linear_solver_t%set_initial_solution,
// base_iterative_linear_solver.f90:340-360: the solver keeps its own copy; default zero, :281-282).  Keyed on the matrix
// handle; NULL switches back to the zero default.
extern "C" int fb2_csr_set_initial_solution(fb2_matrix_t A, fb2_vector_t x0) {
  FB2_CHECK(A, "null matrix");
  fb2_solver_ws* S = ws_of(A);
  if (!x0) { S->x0_on = false; return 0; }
  // a vector that does not belong to range(A) is ignored, as the reference does (with a warning, :349-351)
  if (x0->n != A->n) { S->x0_on = false; return 0; }
  if (S->x0_n != x0->n) {
    cudaFree(S->x0); S->x0 = nullptr; S->x0_n = 0;
    FB2_TRY(fb2_dalloc(&S->x0, (size_t)x0->n));
    S->x0_n = x0->n;
  }
  FB2_CUDA(cudaMemcpyAsync(S->x0, x0->d, sizeof(double) * (size_t)x0->n, cudaMemcpyDeviceToDevice, fb2_stream()));
  S->x0_on = true;
  return 0;
}
bool fb2_ils_has_initial_solution(fb2_matrix_t A) { return A->ws && A->ws->x0_on && A->ws->x0_n == A->n; }
int fb2_ils_load_initial_solution(fb2_matrix_t A, double* x) {
  const i64 n = A->n;
  if (fb2_ils_has_initial_solution(A))
    FB2_CUDA(cudaMemcpyAsync(x, A->ws->x0, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, fb2_stream()));
  else k_init<<<vec_grid(n), 256, 0, fb2_stream()>>>(x, 0.0, n);
  FB2_CUDA(cudaGetLastError());
  return 0;
}

// PCG set-up from the zero initial solution in ONE pass over b (and dinv): x <- 0, r <- b, p <- M^-1 b, and the block partials
// of <r,p>, <r,r> (= <b,b>) -- same thread/block mapping and summation order as k_dot, so the sums are the bits the separate
// kernels gave.  partials: [2][gridDim.x]
__global__ void k_cg_setup_zero_x0(const double* __restrict__ b, const double* __restrict__ dinv, const double* __restrict__ w,
                                   i64 n, double* __restrict__ x, double* __restrict__ r, double* __restrict__ p, double* partials) {
  __shared__ double s_red[32];
  double a_rp = 0.0, a_rr = 0.0;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) {
    const double bi = b[i], pi = dinv ? dinv[i] * bi : bi, wi = w ? w[i] : 1.0;
    x[i] = 0.0; r[i] = bi; p[i] = pi;
    a_rp += wi * bi * pi;
    a_rr += wi * bi * bi;
  }
  const double s0 = block_sum(a_rp, s_red);
  __syncthreads();
  const double s1 = block_sum(a_rr, s_red);
  if (threadIdx.x == 0) { partials[blockIdx.x] = s0; partials[gridDim.x + blockIdx.x] = s1; }
}

static int cg_setup(fb2_matrix_t A, const double* dinv, const double* w, const double* b, double* x, bool par, double rtol,
                    double atol, int maxit, int stop) {
  cudaStream_t s = fb2_stream();
  CgWork& W = work_of(A);
  const i64 n = A->n;
  // x <- 0 ; r = A x ; r = b - r   (cg.f90:142,158-162)
  // The initial solution is zero (base_iterative_linear_solver.f90:281-282), so r = b - A*0 = b bit for bit: the
  // reference's initial SpMV and its halo exchange are skipped.
  double* d_tmp = fb2_scratch_partials() + FB2_MAX_PARTIALS * 4;
  if (fb2_ils_has_initial_solution(A)) {   // user-selected initial solution: the full r = b - A x (cg.f90:158-162)
    FB2_TRY(fb2_ils_load_initial_solution(A, x));
    FB2_TRY(fb2_op_apply(A, x, W.r));
    FB2_TRY(fb2_dev_axpby(W.r, 1.0, b, -1.0, n));
    // z = M^-1 r ; p = z ; r_z = <r,z>
    if (dinv) k_ew_product<<<vec_grid(n), 256, 0, s>>>(W.p, dinv, W.r, n);
    else FB2_CUDA(cudaMemcpyAsync(W.p, W.r, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, s));
    FB2_TRY(fb2_dot_device(W.r, W.p, w, n, d_tmp));          // r_z
    FB2_TRY(fb2_dot_device(W.r, W.r, w, n, d_tmp + 1));      // ||r(0)||^2
    FB2_TRY(fb2_dot_device(b, b, w, n, d_tmp + 2));          // ||b||^2
  } else {
    // zero initial solution: one fused pass (x, r, p, <r,p>, <r,r> = <b,b>) instead of 3 fills and 3 dots
    const int g = reduce_grid(n);
    double* part = fb2_scratch_partials();
    k_cg_setup_zero_x0<<<g, 256, 0, s>>>(b, dinv, w, n, x, W.r, W.p, part);
    k_reduce_partials<<<1, 256, 0, s>>>(part, g, g, 2, d_tmp);
    FB2_CUDA(cudaMemcpyAsync(d_tmp + 2, d_tmp + 1, sizeof(double), cudaMemcpyDeviceToDevice, s));   // ||b||^2 = ||r(0)||^2
    FB2_CUDA(cudaGetLastError());
  }
  if (stop == FB2_STOP_RES_NRMGIVEN_RHS_NRMGIVEN) {        // |b|_inv(M): r = inv(M) b ; b.r   (cg.f90:144-154), Ap is free here
    if (dinv) k_ew_product<<<vec_grid(n), 256, 0, s>>>(W.Ap, dinv, b, n);
    FB2_TRY(fb2_dot_device(b, dinv ? W.Ap : b, w, n, d_tmp + 3));
  } else FB2_CUDA(cudaMemsetAsync(d_tmp + 3, 0, sizeof(double), s));
  if (par) FB2_TRY(fb2_allreduce_sum(d_tmp, 4));
  double h[4];
  FB2_CUDA(cudaMemcpyAsync(h, d_tmp, 4 * sizeof(double), cudaMemcpyDeviceToHost, s));
  FB2_CUDA(cudaStreamSynchronize(s));
  // init_convergence_data, cg.f90:279-321
  const double nr0 = sqrt(h[1]), nb = sqrt(h[2]);
  double tol = 0.0, tol_extra = 0.0;
  switch (stop) {
    case FB2_STOP_DELTA_RHS: case FB2_STOP_RES_RHS: tol = rtol * nb + atol; break;
    case FB2_STOP_DELTA_RHS_AND_RES_RHS: case FB2_STOP_DELTA_DELTA_AND_RES_RHS: tol = tol_extra = rtol * nb + atol; break;
    case FB2_STOP_RES_RES: tol = rtol * nr0 + atol; break;
    case FB2_STOP_DELTA_DELTA_AND_RES_RES: tol = tol_extra = rtol * nr0 + atol; break;
    case FB2_STOP_DELTA_RHS_AND_RES_RES: tol = rtol * nb + atol; tol_extra = rtol * nr0 + atol; break;
    case FB2_STOP_RES_NRMGIVEN_RHS_NRMGIVEN: tol = rtol * sqrt(h[3]) + atol; break;
    case FB2_STOP_RES_NRMGIVEN_RES_NRMGIVEN: tol = rtol * sqrt(h[0]) + atol; break;
    default: break;   // delta_delta: set at the first iteration
  }
  k_cg_init_state<<<1, 1, 0, s>>>(W.st, h[0], tol, tol_extra, rtol, atol, maxit, stop, g_cg_icg);
  FB2_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int fb2_pcg_solve(fb2_matrix_t A, fb2_vector_t dinv, fb2_vector_t b, fb2_vector_t x, const fb2_ils_params_t* prm,
                             int* num_iterations, int* did_converge, double* error_estimate, double* history) {
  FB2_CHECK(A && b && x && prm, "null argument");
  FB2_CHECK(b->n == A->n && x->n == A->n && (!dinv || dinv->n == A->n), "size mismatch");
  FB2_CHECK(prm->stopping_criterium >= FB2_STOP_RES_RES && prm->stopping_criterium <= FB2_STOP_DELTA_DELTA_AND_RES_RHS,
            "unknown stopping criterium (cg_supports_stopping_criteria, cg.f90:240-255)");
  const int maxit = prm->max_num_iterations;
  const bool par = A->s->imp && fb2_comm_active();
  const double* w = par ? fb2_import_weight(A->s->imp) : nullptr;
  const bool track = history != nullptr;
  FB2_TRY(cg_workspace(A, track ? maxit : 0));
  CgWork& W = work_of(A);
  cudaStream_t s = fb2_stream();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, s);
  FB2_TRY(cg_setup(A, dinv ? dinv->d : nullptr, w, b->d, x->d, par, prm->rtol, prm->atol, maxit, prm->stopping_criterium));
  int every = prm->check_every > 0 ? prm->check_every : 8;
  CgState h;
  memset(&h, 0, sizeof(h));
  int launched = 0;
  while (launched < maxit) {
    int chunk = every;
    if (launched + chunk > maxit) chunk = maxit - launched;
    for (int k = 0; k < chunk; ++k)
      FB2_TRY(cg_enqueue_iteration(A, dinv ? dinv->d : nullptr, w, x->d, par, track ? W.hist : nullptr, nullptr, nullptr));
    launched += chunk;
    FB2_CUDA(cudaMemcpyAsync(&h, W.st, sizeof(CgState), cudaMemcpyDeviceToHost, s));
    FB2_CUDA(cudaStreamSynchronize(s));
    if (h.done) break;
  }
  cudaEventRecord(e1, s);
  FB2_CUDA(cudaStreamSynchronize(s));
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  fb2_record_phase("pcg", ms);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  *num_iterations = h.iter;
  *did_converge = h.converged;
  *error_estimate = h.err;
  if (track && h.iter > 0) FB2_CUDA(fb2_memcpy(history, W.hist, sizeof(double) * h.iter, cudaMemcpyDeviceToHost));
  fb2_solver_ws* S = ws_of(A);
  S->last[0] = h.tol; S->last[1] = h.tol_extra; S->last[2] = h.err; S->last[3] = h.err_extra;
  S->last_iters = h.iter; S->last_tracked = track ? 1 : 0;
  return 0;
}

// what print_convergence_history_{header,new_line,footer} need beyond the history itself
// (base_iterative_linear_solver.f90:584-676): out[4] = rhs_convergence_test, rhs_extra_convergence_test,
// error_estimate_convergence_test, error_estimate_extra_convergence_test of the last solve on this operator; history_extra
// (may be NULL; needs a solve that tracked the history) = the 2nd criterion's estimate per iteration
extern "C" int fb2_csr_get_convergence_data(fb2_matrix_t A, double out[4], double* history_extra) {
  FB2_CHECK(A && out, "null argument");
  fb2_solver_ws* S = ws_of(A);
  for (int k = 0; k < 4; ++k) out[k] = S->last[k];
  if (history_extra && S->last_iters > 0) {
    FB2_CHECK(S->last_tracked && S->w.hist, "the last solve did not track its convergence history");
    FB2_CUDA(fb2_memcpy(history_extra, S->w.hist + S->w.hist_cap, sizeof(double) * S->last_iters, cudaMemcpyDeviceToHost));
  }
  return 0;
}
void fb2_solver_record_convergence(fb2_matrix_s* A, double tol, double err, int iters) {
  fb2_solver_ws* S = ws_of(A);
  S->last[0] = tol; S->last[1] = 0.0; S->last[2] = err; S->last[3] = 0.0; S->last_iters = iters; S->last_tracked = 0;
}

extern "C" int fb2_pcg_solve_host(fb2_matrix_t A, fb2_vector_t dinv, const double* b_host, double* x_host,
                                  const fb2_ils_params_t* prm, int* num_iterations, int* did_converge, double* error_estimate,
                                  double* history) {
  FB2_CHECK(A && b_host && x_host, "null argument");
  fb2_vector_t &vb = ws_of(A)->vb, &vx = ws_of(A)->vx;   // staging vectors owned by the matrix handle
  if (!vb || vb->n != A->n) {
    if (vb) { fb2_vec_destroy(vb); fb2_vec_destroy(vx); }
    FB2_TRY(fb2_vec_create_n(&vb, A->n));
    FB2_TRY(fb2_vec_create_n(&vx, A->n));
    vb->s = A->s; vx->s = A->s;
  }
  FB2_TRY(fb2_vec_set(vb, b_host));
  FB2_TRY(fb2_pcg_solve(A, dinv, vb, vx, prm, num_iterations, did_converge, error_estimate, history));
  FB2_TRY(fb2_vec_get(vx, x_host));
  return 0;
}

extern "C" int fb2_pcg_iterate(fb2_matrix_t A, fb2_vector_t dinv, fb2_vector_t b, fb2_vector_t x, int n_iterations,
                               double* ms_total, double* ms_spmv) {
  FB2_CHECK(A && b && x && n_iterations > 0, "bad argument");
  const bool par = A->s->imp && fb2_comm_active();
  const double* w = par ? fb2_import_weight(A->s->imp) : nullptr;
  FB2_TRY(cg_workspace(A, 0));
  cudaStream_t s = fb2_stream();
  // tolerance -1: never converges, so exactly n_iterations full iterations run
  FB2_TRY(cg_setup(A, dinv ? dinv->d : nullptr, w, b->d, x->d, par, 0.0, -1.0, n_iterations + 1, FB2_STOP_RES_RES));
  std::vector<cudaEvent_t> ev(2 * (size_t)n_iterations + 2);
  for (auto& e : ev) cudaEventCreate(&e);
  g_tl.begin(par ? n_iterations : 0);
  cudaEventRecord(ev[2 * n_iterations], s);
  for (int k = 0; k < n_iterations; ++k)
    FB2_TRY(cg_enqueue_iteration(A, dinv ? dinv->d : nullptr, w, x->d, par, nullptr, ev[2 * k], ev[2 * k + 1]));
  cudaEventRecord(ev[2 * n_iterations + 1], s);
  FB2_CUDA(cudaStreamSynchronize(s));
  FB2_CUDA(cudaStreamSynchronize(fb2_comm_stream()));
  g_tl.publish();
  float ms = 0, acc = 0;
  cudaEventElapsedTime(&ms, ev[2 * n_iterations], ev[2 * n_iterations + 1]);
  for (int k = 0; k < n_iterations; ++k) {
    float m1 = 0;
    cudaEventElapsedTime(&m1, ev[2 * k], ev[2 * k + 1]);
    acc += m1;
  }
  for (auto& e : ev) cudaEventDestroy(e);
  if (ms_total) *ms_total = ms;
  if (ms_spmv) *ms_spmv = acc;
  return 0;
}

// ------------------------------------------------------------------------------------------
// primitives shared with the host-driven Krylov solvers of fb2_krylov.cu (operator_t%apply, vector_t TBPs on raw
// device arrays; all enqueue on the library stream)
// ------------------------------------------------------------------------------------------
bool fb2_matrix_is_par(fb2_matrix_t A) { return A->s && A->s->imp && fb2_comm_active(); }
const double* fb2_matrix_weight(fb2_matrix_t A) { return fb2_matrix_is_par(A) ? fb2_import_weight(A->s->imp) : nullptr; }
// y = A x  (+ comm() in par: par_sparse_matrix.f90:739-762)
int fb2_op_apply(fb2_matrix_t A, const double* x, double* y) {
  FB2_TRY(spmv_local(A, x, y, false, nullptr, nullptr));
  if (fb2_matrix_is_par(A)) FB2_TRY(fb2_import_exchange(A->s->imp, y));
  return 0;
}
// z = inv(M) r with M = Jacobi (dinv) or identity
int fb2_dev_apply_M(const double* dinv, const double* r, double* z, i64 n) {
  if (dinv) k_ew_product<<<vec_grid(n), 256, 0, fb2_stream()>>>(z, dinv, r, n);
  else FB2_CUDA(cudaMemcpyAsync(z, r, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, fb2_stream()));
  FB2_CUDA(cudaGetLastError());
  return 0;
}
int fb2_dev_axpby(double* op1, double a, const double* op2, double b, i64 n) {
  k_axpby<<<vec_grid(n), 256, 0, fb2_stream()>>>(op1, a, op2, b, n);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
int fb2_dev_scal(double* dst, double a, const double* src, i64 n) {
  k_scal<<<vec_grid(n), 256, 0, fb2_stream()>>>(dst, a, src, n);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
int fb2_dev_init(double* x, double a, i64 n) {
  k_init<<<vec_grid(n), 256, 0, fb2_stream()>>>(x, a, n);
  FB2_CUDA(cudaGetLastError());
  return 0;
}
// out[i] = <xs[i], ys[i]>_w for i < k, all-reduced in par, copied to the host (one synchronisation)
int fb2_dots_host(fb2_matrix_t A, int k, const double* const* xs, const double* const* ys, i64 n, double* out) {
  FB2_CHECK(k >= 1 && k <= 1024, "fb2_dots_host: bad count");
  double* d_tmp = fb2_scratch_partials() + FB2_MAX_PARTIALS * 4;
  const double* w = fb2_matrix_weight(A);
  for (int i = 0; i < k; ++i) FB2_TRY(fb2_dot_device(xs[i], ys[i], w, n, d_tmp + i));
  if (fb2_matrix_is_par(A)) FB2_TRY(fb2_allreduce_sum(d_tmp, k));
  FB2_CUDA(fb2_memcpy(out, d_tmp, sizeof(double) * k, cudaMemcpyDeviceToHost));
  return 0;
}

// ==========================================================================================
// Several sub-domains per process (fb2_comm_set_parts_per_rank): the par_* vector / operator / PCG entry points over
// the list of co-resident parts.  Same kernels and the same device-resident CgState as the single-part path; every part
// writes its partial sums into its own segment of the scratch buffer and one reduction sums all segments in a fixed order.
// ==========================================================================================
struct MultiCtx {
  int np = 0;
  std::vector<fb2_matrix_s*> A;
  std::vector<fb2_import_s*> imp;
  std::vector<int> part;
  std::vector<const double*> w;
};
static int multi_ctx(int np, fb2_matrix_t* A, MultiCtx& C) {
  FB2_CHECK(np >= 1 && A, "multi-part call without parts");
  C.np = np; C.A.assign(A, A + np); C.imp.resize(np); C.part.resize(np); C.w.resize(np);
  for (int i = 0; i < np; ++i) {
    FB2_CHECK(A[i] && A[i]->s && A[i]->s->imp, "multi-part call: every matrix needs a par space (part_id > 0)");
    FB2_CHECK(!A[i]->sym, "multi-part PCG uses full storage");
    C.imp[i] = A[i]->s->imp; C.part[i] = A[i]->s->part_id;
    C.w[i] = fb2_import_weight(C.imp[i]);     // also without a communicator: co-resident parts replicate interface DOFs
  }
  return 0;
}
static int exchange_multi(const MultiCtx& C, double* const* xs, cudaStream_t s) {
  return fb2_import_exchange_multi(C.np, C.imp.data(), C.part.data(), xs, s);
}

extern "C" int fb2_vec_comm_multi(int np, fb2_vector_t* v) {
  FB2_CHECK(np >= 1 && v, "bad argument");
  std::vector<fb2_import_s*> imp(np); std::vector<int> part(np); std::vector<double*> xs(np);
  for (int i = 0; i < np; ++i) {
    FB2_CHECK(v[i] && v[i]->s && v[i]->s->imp, "fb2_vec_comm_multi: vectors must belong to par spaces");
    imp[i] = v[i]->s->imp; part[i] = v[i]->s->part_id; xs[i] = v[i]->d;
  }
  return fb2_import_exchange_multi(np, imp.data(), part.data(), xs.data(), fb2_stream());
}
// sum over the co-resident parts of the weighted local dots, then the all-reduce (par_scalar_array.f90:303-350)
extern "C" int fb2_vec_dot_multi(int np, fb2_vector_t* x, fb2_vector_t* y, double* result) {
  FB2_CHECK(np >= 1 && np <= 64 && x && y, "bad argument");
  double* d_tmp = fb2_scratch_partials() + FB2_MAX_PARTIALS * 4;
  for (int i = 0; i < np; ++i) {
    FB2_CHECK(x[i]->n == y[i]->n, "size mismatch");
    FB2_CHECK(x[i]->s && x[i]->s->imp, "fb2_vec_dot_multi: vectors must belong to par spaces");
    FB2_TRY(fb2_dot_device(x[i]->d, y[i]->d, fb2_import_weight(x[i]->s->imp), x[i]->n, d_tmp + 1 + i));
  }
  k_reduce_partials<<<1, 256, 0, fb2_stream()>>>(d_tmp + 1, np, 0, 1, d_tmp);
  if (fb2_comm_active()) FB2_TRY(fb2_allreduce_sum(d_tmp, 1));
  FB2_CUDA(fb2_memcpy(result, d_tmp, sizeof(double), cudaMemcpyDeviceToHost));
  return 0;
}
extern "C" int fb2_spmv_multi(int np, fb2_matrix_t* A, fb2_vector_t* x, fb2_vector_t* y) {
  MultiCtx C;
  FB2_TRY(multi_ctx(np, A, C));
  std::vector<double*> ys(np);
  for (int i = 0; i < np; ++i) {
    FB2_CHECK(x[i]->n == A[i]->n && y[i]->n == A[i]->n, "size mismatch");
    FB2_TRY(spmv_local(A[i], x[i]->d, y[i]->d, false, nullptr, nullptr));
    ys[i] = y[i]->d;
  }
  FB2_TRY(exchange_multi(C, ys.data(), fb2_stream()));
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  return 0;
}
extern "C" int fb2_jacobi_setup_multi(int np, fb2_matrix_t* A, fb2_vector_t* dinv) {
  MultiCtx C;
  FB2_TRY(multi_ctx(np, A, C));
  std::vector<double*> ds(np);
  for (int i = 0; i < np; ++i) { FB2_TRY(fb2_extract_diagonal(A[i], dinv[i])); ds[i] = dinv[i]->d; }
  FB2_TRY(exchange_multi(C, ds.data(), fb2_stream()));
  for (int i = 0; i < np; ++i) FB2_TRY(fb2_vec_entrywise_invert(dinv[i]));
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  return 0;
}

// the co-resident parts use the work vectors of their own matrices; the shared device state and history are those of the
// first matrix of the list
struct MultiWork {
  std::vector<CgWork*> parts;
  CgState* st = nullptr;
  double* hist = nullptr;
};
static int multi_workspace(const MultiCtx& C, int hist_len, MultiWork& M) {
  M.parts.resize(C.np);
  for (int i = 0; i < C.np; ++i) {
    FB2_CHECK(!fb2_ils_has_initial_solution(C.A[i]), "the multi-part PCG starts from zero: clear the initial solution of the parts");
    FB2_TRY(cg_workspace(C.A[i], i == 0 ? hist_len : 0));
    M.parts[i] = &work_of(C.A[i]);
  }
  M.st = M.parts[0]->st; M.hist = M.parts[0]->hist;
  return 0;
}

static int multi_cg_setup(const MultiCtx& C, MultiWork& M, fb2_vector_t* dinv, fb2_vector_t* b, fb2_vector_t* x, double rtol,
                          double atol, int maxit, int stop) {
  cudaStream_t s = fb2_stream();
  const int np = C.np;
  double* d_tmp = fb2_scratch_partials() + FB2_MAX_PARTIALS * 4;      // [4][np] partial dots, then 4 sums at d_tmp + 4 np
  for (int i = 0; i < np; ++i) {
    CgWork& W = *M.parts[i];
    const i64 n = W.n;
    const double* di = dinv ? dinv[i]->d : nullptr;
    k_init<<<vec_grid(n), 256, 0, s>>>(x[i]->d, 0.0, n);
    FB2_CUDA(cudaMemcpyAsync(W.r, b[i]->d, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, s));
    if (di) k_ew_product<<<vec_grid(n), 256, 0, s>>>(W.p, di, W.r, n);
    else FB2_CUDA(cudaMemcpyAsync(W.p, W.r, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, s));
    FB2_TRY(fb2_dot_device(W.r, W.p, C.w[i], n, d_tmp + 0 * np + i));
    FB2_TRY(fb2_dot_device(W.r, W.r, C.w[i], n, d_tmp + 1 * np + i));
    FB2_TRY(fb2_dot_device(b[i]->d, b[i]->d, C.w[i], n, d_tmp + 2 * np + i));
    if (stop == FB2_STOP_RES_NRMGIVEN_RHS_NRMGIVEN) {
      if (di) k_ew_product<<<vec_grid(n), 256, 0, s>>>(W.Ap, di, b[i]->d, n);
      FB2_TRY(fb2_dot_device(b[i]->d, di ? W.Ap : b[i]->d, C.w[i], n, d_tmp + 3 * np + i));
    } else FB2_CUDA(cudaMemsetAsync(d_tmp + 3 * np + i, 0, sizeof(double), s));
  }
  k_reduce_partials<<<1, 256, 0, s>>>(d_tmp, np, np, 4, d_tmp + 4 * np);
  if (fb2_comm_active()) FB2_TRY(fb2_allreduce_sum(d_tmp + 4 * np, 4));
  double h[4];
  FB2_CUDA(fb2_memcpy(h, d_tmp + 4 * np, 4 * sizeof(double), cudaMemcpyDeviceToHost));
  const double nr0 = sqrt(h[1]), nb = sqrt(h[2]);
  double tol = 0.0, tol_extra = 0.0;
  switch (stop) {
    case FB2_STOP_DELTA_RHS: case FB2_STOP_RES_RHS: tol = rtol * nb + atol; break;
    case FB2_STOP_DELTA_RHS_AND_RES_RHS: case FB2_STOP_DELTA_DELTA_AND_RES_RHS: tol = tol_extra = rtol * nb + atol; break;
    case FB2_STOP_RES_RES: tol = rtol * nr0 + atol; break;
    case FB2_STOP_DELTA_DELTA_AND_RES_RES: tol = tol_extra = rtol * nr0 + atol; break;
    case FB2_STOP_DELTA_RHS_AND_RES_RES: tol = rtol * nb + atol; tol_extra = rtol * nr0 + atol; break;
    case FB2_STOP_RES_NRMGIVEN_RHS_NRMGIVEN: tol = rtol * sqrt(h[3]) + atol; break;
    case FB2_STOP_RES_NRMGIVEN_RES_NRMGIVEN: tol = rtol * sqrt(h[0]) + atol; break;
    default: break;
  }
  k_cg_init_state<<<1, 1, 0, s>>>(M.st, h[0], tol, tol_extra, rtol, atol, maxit, stop, g_cg_icg);
  FB2_CUDA(cudaGetLastError());
  return 0;
}

// one PCG iteration over all co-resident parts (overlapped par apply, see cg_enqueue_iteration)
static int multi_cg_iteration(const MultiCtx& C, MultiWork& M, fb2_vector_t* dinv, fb2_vector_t* x, double* hist, cudaEvent_t ev0,
                              cudaEvent_t ev1) {
  cudaStream_t s = fb2_stream(), cs = fb2_comm_stream();
  const int np = C.np;
  double* part = fb2_scratch_partials();
  static cudaEvent_t ev_itf = nullptr, ev_halo = nullptr;
  if (!ev_itf) { cudaEventCreateWithFlags(&ev_itf, cudaEventDisableTiming); cudaEventCreateWithFlags(&ev_halo, cudaEventDisableTiming); }
  if (ev0) cudaEventRecord(ev0, s);
  g_tl.mark(0, s);
  std::vector<i64> t_itf(np), row_itf(np);
  std::vector<double*> aps(np);
  for (int i = 0; i < np; ++i) {
    CgWork& W = *M.parts[i];
    FB2_TRY(fb2_spmv_interface_split(C.A[i], C.A[i]->s->n_interior, &t_itf[i], &row_itf[i]));
    FB2_TRY(fb2_spmv_stream_range(C.A[i], W.p, W.Ap, false, &M.st->done, t_itf[i], C.A[i]->spmv_ntiles, 0, nullptr, nullptr));
    aps[i] = W.Ap;
  }
  FB2_CUDA(cudaEventRecord(ev_itf, s));
  g_tl.mark(1, s);
  FB2_CUDA(cudaStreamWaitEvent(cs, ev_itf, 0));
  FB2_TRY(exchange_multi(C, aps.data(), cs));
  FB2_CUDA(cudaEventRecord(ev_halo, cs));
  g_tl.mark(2, cs);
  int off = 0;
  for (int i = 0; i < np; ++i) {
    CgWork& W = *M.parts[i];
    int g = 0;
    FB2_TRY(fb2_spmv_stream_range(C.A[i], W.p, W.Ap, true, &M.st->done, 0, t_itf[i], g_par_chunk, part + off, &g));
    off += g;
  }
  g_tl.mark(3, s);
  FB2_CUDA(cudaStreamWaitEvent(s, ev_halo, 0));
  g_tl.mark(4, s);
  if (ev1) cudaEventRecord(ev1, s);
  for (int i = 0; i < np; ++i) {
    CgWork& W = *M.parts[i];
    const i64 n = W.n, r0 = row_itf[i];
    if (r0 < n) {
      int gd = reduce_grid(n - r0);
      if (gd > 64) gd = 64;
      k_dot<<<gd, 256, 0, s>>>(W.Ap + r0, W.p + r0, C.w[i] ? C.w[i] + r0 : nullptr, n - r0, part + off);
      off += gd;
    }
  }
  FB2_CHECK(off <= FB2_MAX_PARTIALS, "too many partial sums for the multi-part PCG");
  k_reduce_partials<<<1, 1024, 0, s>>>(part, off, 0, 1, &M.st->App);
  if (fb2_comm_active()) FB2_TRY(fb2_allreduce_sum(&M.st->App, 1));
  int off2 = 0;
  for (int i = 0; i < np; ++i) {
    CgWork& W = *M.parts[i];
    const int g2 = reduce_grid(W.n);
    k_cg_update_xr<<<g2, 256, 0, s>>>(W.n, M.st, W.p, W.Ap, dinv ? dinv[i]->d : nullptr, C.w[i], x[i]->d, W.r, part + off2, FB2_MAX_PARTIALS);
    off2 += g2;
  }
  k_reduce_partials<<<1, 256, 0, s>>>(part, off2, FB2_MAX_PARTIALS, 4, &M.st->rr);
  if (fb2_comm_active()) FB2_TRY(fb2_allreduce_sum(&M.st->rr, 4));
  k_cg_scalars<<<1, 1, 0, s>>>(M.st, hist, hist ? hist + M.parts[0]->hist_cap : nullptr);
  for (int i = 0; i < np; ++i) {
    CgWork& W = *M.parts[i];
    k_cg_update_p<<<vec_grid(W.n), 256, 0, s>>>(W.n, M.st, W.r, dinv ? dinv[i]->d : nullptr, W.p);
  }
  g_tl.mark(5, s); g_tl.next();
  FB2_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int fb2_pcg_solve_multi(int np, fb2_matrix_t* A, fb2_vector_t* dinv, fb2_vector_t* b, fb2_vector_t* x,
                                   const fb2_ils_params_t* prm, int* num_iterations, int* did_converge, double* error_estimate,
                                   double* history) {
  FB2_CHECK(b && x && prm, "null argument");
  MultiCtx C;
  FB2_TRY(multi_ctx(np, A, C));
  for (int i = 0; i < np; ++i) FB2_CHECK(b[i]->n == A[i]->n && x[i]->n == A[i]->n && (!dinv || dinv[i]->n == A[i]->n), "size mismatch");
  FB2_CHECK(prm->stopping_criterium >= FB2_STOP_RES_RES && prm->stopping_criterium <= FB2_STOP_DELTA_DELTA_AND_RES_RHS, "unknown stopping criterium");
  const int maxit = prm->max_num_iterations;
  const bool track = history != nullptr;
  MultiWork M;
  FB2_TRY(multi_workspace(C, track ? maxit : 0, M));
  FB2_TRY(multi_cg_setup(C, M, dinv, b, x, prm->rtol, prm->atol, maxit, prm->stopping_criterium));
  const int every = prm->check_every > 0 ? prm->check_every : 8;
  CgState h;
  memset(&h, 0, sizeof(h));
  int launched = 0;
  while (launched < maxit) {
    int chunk = every;
    if (launched + chunk > maxit) chunk = maxit - launched;
    for (int k = 0; k < chunk; ++k) FB2_TRY(multi_cg_iteration(C, M, dinv, x, track ? M.hist : nullptr, nullptr, nullptr));
    launched += chunk;
    FB2_CUDA(fb2_memcpy(&h, M.st, sizeof(CgState), cudaMemcpyDeviceToHost));
    if (h.done) break;
  }
  *num_iterations = h.iter; *did_converge = h.converged; *error_estimate = h.err;
  if (track && h.iter > 0) FB2_CUDA(fb2_memcpy(history, M.hist, sizeof(double) * h.iter, cudaMemcpyDeviceToHost));
  fb2_solver_ws* S = ws_of(C.A[0]);      // convergence data of the collective solve: kept with the first part
  S->last[0] = h.tol; S->last[1] = h.tol_extra; S->last[2] = h.err; S->last[3] = h.err_extra;
  S->last_iters = h.iter; S->last_tracked = track ? 1 : 0;
  return 0;
}

extern "C" int fb2_pcg_iterate_multi(int np, fb2_matrix_t* A, fb2_vector_t* dinv, fb2_vector_t* b, fb2_vector_t* x,
                                     int n_iterations, double* ms_total, double* ms_spmv) {
  FB2_CHECK(b && x && n_iterations > 0, "bad argument");
  MultiCtx C;
  FB2_TRY(multi_ctx(np, A, C));
  MultiWork M;
  FB2_TRY(multi_workspace(C, 0, M));
  cudaStream_t s = fb2_stream();
  FB2_TRY(multi_cg_setup(C, M, dinv, b, x, 0.0, -1.0, n_iterations + 1, FB2_STOP_RES_RES));
  std::vector<cudaEvent_t> ev(2 * (size_t)n_iterations + 2);
  for (auto& e : ev) cudaEventCreate(&e);
  g_tl.begin(n_iterations);
  cudaEventRecord(ev[2 * n_iterations], s);
  for (int k = 0; k < n_iterations; ++k) FB2_TRY(multi_cg_iteration(C, M, dinv, x, nullptr, ev[2 * k], ev[2 * k + 1]));
  cudaEventRecord(ev[2 * n_iterations + 1], s);
  FB2_CUDA(cudaStreamSynchronize(s));
  FB2_CUDA(cudaStreamSynchronize(fb2_comm_stream()));
  g_tl.publish();
  float ms = 0, acc = 0;
  cudaEventElapsedTime(&ms, ev[2 * n_iterations], ev[2 * n_iterations + 1]);
  for (int k = 0; k < n_iterations; ++k) { float m1 = 0; cudaEventElapsedTime(&m1, ev[2 * k], ev[2 * k + 1]); acc += m1; }
  for (auto& e : ev) cudaEventDestroy(e);
  if (ms_total) *ms_total = ms;
  if (ms_spmv) *ms_spmv = acc;
  return 0;
}
