// CSR SpMV for sm_100a: row tiles streamed through shared memory with TMA bulk copies
// (cp.async.bulk + mbarrier, SASS UBLKCP), products formed in place, per-row sub-warp reduction,
// optional fused <y,x> partial (the <Ap,p> of CG).  Replaces csr matvec (csr_sparse_matrix.f90:545-568).
//
// Data layout in HBM: val (f64) and ja (int32, 1-based) are two contiguous streams; a tile of R
// consecutive rows is the contiguous element range [irp[r0], irp[r0+R]) of both.  Algorithmic bytes
// per row tile: 12 B per non-zero + 24 B per row (irp, y, x).
#include <stdlib.h>
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

#define SPMV_CONSUMERS 256
#define SPMV_THREADS (SPMV_CONSUMERS + 32)
#define SPMV_MAX_STAGES 8

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Persistent block over tiles t = blockIdx.x, blockIdx.x + gridDim.x, ...  Warp 8 is the TMA producer
// (one elected lane): it runs STAGES tiles ahead, waiting on the per-stage "empty" mbarrier and arming
// the "full" one with the byte count of the three bulk copies (val, ja, irp slice).  The 8 consumer
// warps never synchronise with each other: each owns R/8 rows of every tile, gathers x, reduces its
// rows inside TPR-lane groups and releases the stage with one arrive per warp.
template <bool FUSE_DOT>
__global__ void __launch_bounds__(SPMV_THREADS) k_spmv_stream(i64 n, const i64* __restrict__ irp, const int* __restrict__ ja,
                                                             const double* __restrict__ val, const double* __restrict__ x,
                                                             double* __restrict__ y, double* partials, const int* done_flag,
                                                             int R, int cap, i64 ntiles, int STAGES) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full_bar[SPMV_MAX_STAGES], empty_bar[SPMV_MAX_STAGES];
  __shared__ i64 s_e0a[SPMV_MAX_STAGES];
  __shared__ double s_red[32];
  if (done_flag && *done_flag) return;
  const int tid = threadIdx.x;
  const size_t irp_bytes = (size_t)(R + 2) * 8;
  const size_t stage_bytes = (size_t)cap * 12 + irp_bytes;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], SPMV_CONSUMERS / 32); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (tid >= SPMV_CONSUMERS) {
    // ---------------- producer warp ----------------
    if (tid == SPMV_CONSUMERS) {
      i64 t = blockIdx.x;
      i64 e0 = 0, e1 = 0;
      if (t < ntiles) { i64 r0 = t * R, r1 = min(r0 + R, n); e0 = irp[r0]; e1 = irp[r1]; }
      for (int it = 0; t < ntiles; t += gridDim.x, ++it) {
        const int s = it % STAGES;
        // prefetch the element range of the next tile before blocking on the empty barrier
        i64 tn = t + gridDim.x, ne0 = 0, ne1 = 0;
        if (tn < ntiles) { i64 r0 = tn * R, r1 = min(r0 + R, n); ne0 = irp[r0]; ne1 = irp[r1]; }
        mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
        unsigned char* base = smem_raw + s * stage_bytes;
        const i64 e0a = e0 & ~(i64)3;
        const uint32_t cnt4 = (uint32_t)((e1 - e0a + 3) & ~(i64)3);
        s_e0a[s] = e0a;
        mbar_expect_tx(&full_bar[s], cnt4 * 12u + (uint32_t)irp_bytes);
        bulk_g2s(base + (size_t)cap * 12, irp + t * R, (uint32_t)irp_bytes, &full_bar[s]);
        if (cnt4) {
          bulk_g2s(base, val + e0a, cnt4 * 8u, &full_bar[s]);
          bulk_g2s(base + (size_t)cap * 8, ja + e0a, cnt4 * 4u, &full_bar[s]);
        }
        e0 = ne0; e1 = ne1;
      }
    }
    return;
  }
  // ---------------- consumer warps ----------------
  const int TPR = SPMV_CONSUMERS / R;   // lanes per row: 4..32, power of two
  const int row = tid / TPR, sub = tid - row * TPR;
  const int lane = tid & 31, w = tid >> 5;
  double dacc = 0.0;
  int it = 0;
  for (i64 t = blockIdx.x; t < ntiles; t += gridDim.x, ++it) {
    const int s = it % STAGES;
    const i64 r0 = t * R;
    mbar_wait(&full_bar[s], (it / STAGES) & 1);
    const unsigned char* base = smem_raw + s * stage_bytes;
    const double* sv = (const double*)base;
    const int* sj = (const int*)(base + (size_t)cap * 8);
    const i64* s_irp = (const i64*)(base + (size_t)cap * 12);
    const i64 e0a = s_e0a[s];
    double acc = 0.0;
    const bool live = r0 + row < n;
    if (live) {
      const int b = (int)(s_irp[row] - e0a), e = (int)(s_irp[row + 1] - e0a);
      int k = b + sub;
      for (; k + 3 * TPR < e; k += 4 * TPR) {
        const double x0 = x[sj[k] - 1], x1 = x[sj[k + TPR] - 1], x2 = x[sj[k + 2 * TPR] - 1], x3 = x[sj[k + 3 * TPR] - 1];
        acc += sv[k] * x0;
        acc += sv[k + TPR] * x1;
        acc += sv[k + 2 * TPR] * x2;
        acc += sv[k + 3 * TPR] * x3;
      }
      for (; k < e; k += TPR) acc += sv[k] * x[sj[k] - 1];
    }
    for (int o = TPR >> 1; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o, TPR);
    if (sub == 0 && live) {
      y[r0 + row] = acc;
      if (FUSE_DOT) dacc += acc * x[r0 + row];
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty_bar[s]);
  }
  if (FUSE_DOT) {
    // deterministic block sum over the consumer warps (named barrier 1: the producer warp has exited)
    for (int o = 16; o > 0; o >>= 1) dacc += __shfl_down_sync(0xffffffffu, dacc, o);
    if (lane == 0) s_red[w] = dacc;
    asm volatile("bar.sync 1, %0;" ::"r"(SPMV_CONSUMERS) : "memory");
    if (tid == 0) {
      double r = 0.0;
      for (int i = 0; i < SPMV_CONSUMERS / 32; ++i) r += s_red[i];
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
  // Tile geometry from the B200 sweep (profiles/sweep_spmv_r1.txt): 2 stages of ~25 KB and 4 resident blocks
  // per SM beat deeper pipelines; take the largest R whose fullest tile stays below ~2.3k non-zeros.
  int R = 64, STAGES = 2, h = 0;
  const bool fixedR = getenv("FB2_SPMV_R") != nullptr;
  if (fixedR) R = atoi(getenv("FB2_SPMV_R"));
  if (getenv("FB2_SPMV_STAGES")) STAGES = atoi(getenv("FB2_SPMV_STAGES"));
  FB2_CHECK(R >= 8 && R <= 64 && (R & (R - 1)) == 0 && STAGES >= 2 && STAGES <= SPMV_MAX_STAGES, "bad SpMV tile parameters");
  (void)maxrow;
  int* d_max;
  FB2_TRY(fb2_dalloc(&d_max, 1));
  i64 ntiles = 0;
  for (;;) {
    ntiles = (A->n + R - 1) / R;
    FB2_CUDA(cudaMemsetAsync(d_max, 0, sizeof(int), fb2_stream()));
    k_tile_max<<<(unsigned)((ntiles + 255) / 256), 256, 0, fb2_stream()>>>(A->n, A->d_irp, R, ntiles, d_max);
    FB2_CUDA(cudaMemcpyAsync(&h, d_max, sizeof(int), cudaMemcpyDeviceToHost, fb2_stream()));
    FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
    if (fixedR || R == 8 || h <= 2304) break;
    R >>= 1;
  }
  cudaFree(d_max);
  int cap = ((h + 3 + 7) / 8) * 8 + 8;
  size_t stage = (size_t)cap * 12 + (size_t)(R + 2) * 8;
  while (STAGES > 2 && stage * STAGES > 200 * 1024) --STAGES;
  size_t smem = stage * STAGES;
  FB2_CHECK(smem <= 200 * 1024, "SpMV tile does not fit in shared memory");
  A->spmv_R = R; A->spmv_cap = cap; A->spmv_ntiles = ntiles; A->spmv_smem = smem; A->spmv_stages = STAGES;
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
                                                                          A->spmv_cap, A->spmv_ntiles, A->spmv_stages);
  else
    k_spmv_stream<false><<<A->spmv_grid, SPMV_THREADS, A->spmv_smem, st>>>(A->n, A->d_irp, A->d_ja, A->d_val, x, y, nullptr,
                                                                           done_flag, A->spmv_R, A->spmv_cap, A->spmv_ntiles, A->spmv_stages);
  FB2_CUDA(cudaGetLastError());
  if (grid_out) *grid_out = A->spmv_grid;
  return 0;
}
