// CSR SpMV for sm_100a: row tiles streamed through shared memory with TMA bulk copies
// (cp.async.bulk + mbarrier, SASS UBLKCP), products formed in place, per-row sub-warp reduction,
// optional fused <y,x> partial (the <Ap,p> of CG).  Replaces csr matvec (csr_sparse_matrix.f90:545-568).
//
// Data layout in HBM: val (f64) and ja (int32, 1-based) are two contiguous streams; a tile of R
// consecutive rows is the contiguous element range [irp[r0], irp[r0+R]) of both.  Algorithmic bytes
// per row tile: 12 B per non-zero + 24 B per row (irp, y, x).
#include "fb2_internal.h"

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra LAB_WAIT;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

struct SpmvState { double rz, App, rr, rz_new, tol, alpha, beta, err; int iter, converged, done, maxit; };

#define SPMV_THREADS 256
#define SPMV_STAGES 2

// one block = persistent worker over tiles t = blockIdx.x, blockIdx.x + gridDim.x, ...
template <bool FUSE_DOT>
__global__ void __launch_bounds__(SPMV_THREADS) k_spmv_stream(i64 n, const i64* __restrict__ irp, const int* __restrict__ ja,
                                                             const double* __restrict__ val, const double* __restrict__ x,
                                                             double* __restrict__ y, double* partials, const int* done_flag,
                                                             int R, int cap, i64 ntiles) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t bar[SPMV_STAGES];
  __shared__ i64 s_rng[SPMV_STAGES][2];   // e0a, e1 of the tile in each stage
  __shared__ double s_red[32];
  if (done_flag && *done_flag) return;
  const int tid = threadIdx.x;
  const size_t stage_bytes = (size_t)cap * 12;
  auto sval = [&](int s) { return (double*)(smem_raw + s * stage_bytes); };
  auto sja = [&](int s) { return (int*)(smem_raw + s * stage_bytes + (size_t)cap * 8); };
  i64* s_irp = (i64*)(smem_raw + SPMV_STAGES * stage_bytes);   // [R+1]

  if (tid == 0) {
    for (int s = 0; s < SPMV_STAGES; ++s) mbar_init(&bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  auto issue = [&](i64 t, int s) {   // thread 0 only
    i64 r0 = t * R, r1 = r0 + R;
    if (r1 > n) r1 = n;
    i64 e0 = irp[r0], e1 = irp[r1];
    i64 e0a = e0 & ~(i64)3;
    uint32_t cnt4 = (uint32_t)((e1 - e0a + 3) & ~(i64)3);
    s_rng[s][0] = e0a; s_rng[s][1] = e1;
    mbar_expect_tx(&bar[s], cnt4 * 12u);
    if (cnt4) {
      bulk_g2s(sval(s), val + e0a, cnt4 * 8u, &bar[s]);
      bulk_g2s(sja(s), ja + e0a, cnt4 * 4u, &bar[s]);
    }
  };

  if (tid == 0) {
    for (int s = 0; s < SPMV_STAGES; ++s) {
      i64 t = (i64)blockIdx.x + (i64)s * gridDim.x;
      if (t < ntiles) issue(t, s);
    }
  }
  const int TPR = SPMV_THREADS / R;   // threads per row (R divides 256)
  double dacc = 0.0;
  int it = 0;
  for (i64 t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
    const int s = it % SPMV_STAGES;
    const uint32_t parity = (it / SPMV_STAGES) & 1;
    const i64 r0 = t * R;
    const int nr = (int)((r0 + R <= n) ? R : (n - r0));
    if (tid <= nr) s_irp[tid] = irp[r0 + tid];
    mbar_wait(&bar[s], parity);
    __syncthreads();
    const i64 e0a = s_rng[s][0];
    const int cnt = (int)(s_rng[s][1] - e0a);
    double* sv = sval(s);
    const int* sj = sja(s);
    // products in place: sv[e] *= x[ja[e]-1]
    for (int e = tid; e < cnt; e += 4 * SPMV_THREADS) {
      int e1 = e + SPMV_THREADS, e2 = e + 2 * SPMV_THREADS, e3 = e + 3 * SPMV_THREADS;
      double x0 = x[sj[e] - 1], x1 = 0.0, x2 = 0.0, x3 = 0.0;
      if (e1 < cnt) x1 = x[sj[e1] - 1];
      if (e2 < cnt) x2 = x[sj[e2] - 1];
      if (e3 < cnt) x3 = x[sj[e3] - 1];
      sv[e] *= x0;
      if (e1 < cnt) sv[e1] *= x1;
      if (e2 < cnt) sv[e2] *= x2;
      if (e3 < cnt) sv[e3] *= x3;
    }
    __syncthreads();
    // per-row reduction: TPR consecutive threads per row
    {
      const int row = tid / TPR, sub = tid - row * TPR;
      double acc = 0.0;
      if (row < nr) {
        const int b = (int)(s_irp[row] - e0a), e = (int)(s_irp[row + 1] - e0a);
        for (int k = b + sub; k < e; k += TPR) acc += sv[k];
      }
      for (int o = TPR >> 1; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o, TPR);
      if (sub == 0 && row < nr) {
        y[r0 + row] = acc;
        if (FUSE_DOT) dacc += acc * x[r0 + row];
      }
    }
    __syncthreads();
    if (tid == 0) {
      i64 tn = t + (i64)SPMV_STAGES * gridDim.x;
      if (tn < ntiles) {
        fence_proxy_async();   // generic-proxy writes to this stage are ordered before the async-proxy refill
        issue(tn, s);
      }
    }
  }
  if (FUSE_DOT) {
    // deterministic block sum
    const int lane = tid & 31, w = tid >> 5;
    for (int o = 16; o > 0; o >>= 1) dacc += __shfl_down_sync(0xffffffffu, dacc, o);
    if (lane == 0) s_red[w] = dacc;
    __syncthreads();
    if (tid == 0) {
      double r = 0.0;
      for (int i = 0; i < SPMV_THREADS / 32; ++i) r += s_red[i];
      partials[blockIdx.x] = r;
    }
  }
}

__global__ void k_tile_max(i64 n, const i64* __restrict__ irp, int R, i64 ntiles, int* out_max) {
  i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  int v = 0;
  if (t < ntiles) {
    i64 r0 = t * R, r1 = r0 + R;
    if (r1 > n) r1 = n;
    v = (int)(irp[r1] - (irp[r0] & ~(i64)3));
  }
  for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_down_sync(0xffffffffu, v, o));
  if ((threadIdx.x & 31) == 0 && v > 0) atomicMax(out_max, v);
}

// choose the tile geometry once per matrix
int fb2_spmv_prepare(fb2_matrix_s* A) {
  if (A->spmv_cap > 0 || A->n == 0) return 0;
  const int p = A->s->rt.p, ncomp = A->s->rt.ncomp;
  int maxrow = (2 * p + 1) * (2 * p + 1) * (2 * p + 1) * ncomp;
  int R = 32;
  while (R > 8 && (i64)R * maxrow > 6144) R >>= 1;    // keep a stage below ~72 KB
  if (maxrow <= 27) R = 64;
  i64 ntiles = (A->n + R - 1) / R;
  int* d_max;
  FB2_TRY(fb2_dalloc(&d_max, 1));
  FB2_CUDA(cudaMemsetAsync(d_max, 0, sizeof(int), fb2_stream()));
  k_tile_max<<<(unsigned)((ntiles + 255) / 256), 256, 0, fb2_stream()>>>(A->n, A->d_irp, R, ntiles, d_max);
  int h = 0;
  FB2_CUDA(cudaMemcpyAsync(&h, d_max, sizeof(int), cudaMemcpyDeviceToHost, fb2_stream()));
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  cudaFree(d_max);
  int cap = ((h + 3 + 7) / 8) * 8 + 8;
  size_t smem = (size_t)SPMV_STAGES * cap * 12 + (R + 1) * sizeof(i64) + 16;
  FB2_CHECK(smem <= 200 * 1024, "SpMV tile does not fit in shared memory");
  A->spmv_R = R; A->spmv_cap = cap; A->spmv_ntiles = ntiles; A->spmv_smem = smem;
  FB2_CUDA(cudaFuncSetAttribute(k_spmv_stream<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  FB2_CUDA(cudaFuncSetAttribute(k_spmv_stream<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int bps = 0;
  FB2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_spmv_stream<true>, SPMV_THREADS, smem));
  if (bps < 1) bps = 1;
  i64 grid = (i64)fb2_num_sms() * bps;
  if (grid > ntiles) grid = ntiles;
  if (grid > FB2_MAX_PARTIALS) grid = FB2_MAX_PARTIALS;
  A->spmv_grid = (int)grid;
  return 0;
}

int fb2_spmv_stream(fb2_matrix_s* A, const double* x, double* y, bool fuse_dot, const int* done_flag, int* grid_out) {
  FB2_TRY(fb2_spmv_prepare(A));
  if (A->n == 0) { if (grid_out) *grid_out = 0; return 0; }
  cudaStream_t st = fb2_stream();
  if (fuse_dot)
    k_spmv_stream<true><<<A->spmv_grid, SPMV_THREADS, A->spmv_smem, st>>>(A->n, A->d_irp, A->d_ja, A->d_val, x, y,
                                                                          fb2_scratch_partials(), done_flag, A->spmv_R,
                                                                          A->spmv_cap, A->spmv_ntiles);
  else
    k_spmv_stream<false><<<A->spmv_grid, SPMV_THREADS, A->spmv_smem, st>>>(A->n, A->d_irp, A->d_ja, A->d_val, x, y, nullptr,
                                                                           done_flag, A->spmv_R, A->spmv_cap, A->spmv_ntiles);
  FB2_CUDA(cudaGetLastError());
  if (grid_out) *grid_out = A->spmv_grid;
  return 0;
}
