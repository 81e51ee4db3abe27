// fb2_vec_dot_no_alloc: a dot product that allocates nothing -- for callers that cannot call cudaMalloc (code under CUDA graph
// capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// It is a variant of the two-kernel dot behind fb2_vec_dot (k_dot + k_reduce_partials, fb2_solver.cu), whose block
// partials live in the library's cudaMalloc'ed scratch; here they live in a file-scope __device__ array.
//
// Scratch: g_no_alloc_scratch, FB2_NO_ALLOC_SCRATCH_DOUBLES doubles (512 KiB + 256 B) of device memory on EVERY GPU that loads
// the library, held until the process exits.  Block b reduces the elements [b * 4096, (b + 1) * 4096) to one partial, so the
// array bounds the vector length: n <= FB2_NO_ALLOC_MAX_N = 65536 * 4096 = 2^28 elements; longer vectors are rejected.
// Every slot that a call reads it has written before in the same call (partials [0, nblocks), then the result slot), so nothing
// has to start at zero and nothing is cleared; the array otherwise keeps the last call's contents.  ALL calls share the array:
// calls that overlap on the GPU (other streams, other host threads with their own stream) race on it.  The library's own stream
// serialises the calls made through this entry point.
//
// Host code must not pass the array's NAME to a kernel (in host code it is a host-side stand-in, not a device address); its
// device address is looked up once per device with cudaGetSymbolAddress, with that device current.
#include <stdint.h>
#include "fb2_internal.h"

#define FB2_NO_ALLOC_CHUNK 4096                     // elements per block partial
#define FB2_NO_ALLOC_MAX_PARTIALS 65536
#define FB2_NO_ALLOC_MAX_N ((i64)FB2_NO_ALLOC_MAX_PARTIALS * FB2_NO_ALLOC_CHUNK)   // = include/fempar_b200.h's value
static_assert((long long)FB2_NO_ALLOC_MAX_PARTIALS * FB2_NO_ALLOC_CHUNK == 268435456LL, "header and kernel disagree on the maximum");
#define FB2_NO_ALLOC_SCRATCH_DOUBLES (FB2_NO_ALLOC_MAX_PARTIALS + 32)   // partials, then the result slot (its own 256-byte line)

static __device__ __align__(256) double g_no_alloc_scratch[FB2_NO_ALLOC_SCRATCH_DOUBLES];

// block b: sum of x_i y_i over its chunk, fixed order (thread t takes elements t, t + 256, ...; shuffle tree; warp sums in
// shared memory) -> partial[b].  Deterministic: the result depends on n only, not on the grid.
__global__ void __launch_bounds__(256) k_dot_chunks(const double* __restrict__ x, const double* __restrict__ y, i64 n,
                                                    double* __restrict__ partial) {
  __shared__ double s_red[8];
  const i64 b0 = (i64)blockIdx.x * FB2_NO_ALLOC_CHUNK;
  const i64 b1 = min(b0 + FB2_NO_ALLOC_CHUNK, n);
  double acc = 0.0;
#pragma unroll 4
  for (i64 i = b0 + threadIdx.x; i < b1; i += 256) acc = fma(x[i], y[i], acc);
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double r = 0.0;
    for (int w = 0; w < 8; ++w) r += s_red[w];
    partial[blockIdx.x] = r;
  }
}
// one block: partial[0 .. m) -> *result, fixed order
__global__ void __launch_bounds__(1024) k_sum_partials(const double* __restrict__ partial, int m, double* result) {
  __shared__ double s_red[32];
  double acc = 0.0;
  for (int i = threadIdx.x; i < m; i += 1024) acc += partial[i];
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    double r = s_red[threadIdx.x];
    for (int o = 16; o > 0; o >>= 1) r += __shfl_down_sync(0xffffffffu, r, o);
    if (threadIdx.x == 0) *result = r;
  }
}

// device address of the scratch array, per device (looked up with that device current)
#define FB2_NO_ALLOC_MAX_DEVICES 64
static double* g_scratch_addr[FB2_NO_ALLOC_MAX_DEVICES] = {nullptr};
static int no_alloc_scratch(double** out) {
  int dev = -1;
  FB2_CUDA(cudaGetDevice(&dev));
  FB2_CHECK(dev >= 0 && dev < FB2_NO_ALLOC_MAX_DEVICES, "fb2_vec_dot_no_alloc: device ordinal beyond 63");
  if (!g_scratch_addr[dev]) {
    double* p = nullptr;
    FB2_CUDA(cudaGetSymbolAddress((void**)&p, g_no_alloc_scratch));
    FB2_CHECK(p != nullptr, "fb2_vec_dot_no_alloc: no device address for the scratch array");
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return 0;
}

// result_host <- sum_i x_dev[i] * y_dev[i], i < n (plain local dot: no sub-domain weights, no all-reduce).
// x_dev, y_dev: device pointers (e.g. fb2_vec_device_ptr); result_host: host pointer.  Blocks until the result is there.
// Calls no cudaMalloc / cudaFree / allocating routine.  Returns 2 with fb2_last_error_string() set when n > FB2_NO_ALLOC_MAX_N.
extern "C" int fb2_vec_dot_no_alloc(const double* x_dev, const double* y_dev, int64_t n, double* result_host) {
  FB2_CHECK(result_host != nullptr, "fb2_vec_dot_no_alloc: null result");
  FB2_CHECK(n >= 0, "fb2_vec_dot_no_alloc: negative length");
  FB2_CHECK(n <= FB2_NO_ALLOC_MAX_N,
            "fb2_vec_dot_no_alloc: vector longer than the static scratch allows (n > 268435456 = 65536 partials x 4096 elements); "
            "use fb2_vec_dot");
  if (n == 0) { *result_host = 0.0; return 0; }
  FB2_CHECK(x_dev != nullptr && y_dev != nullptr, "fb2_vec_dot_no_alloc: null vector");
  double* scratch = nullptr;
  FB2_TRY(no_alloc_scratch(&scratch));
  double* d_result = scratch + FB2_NO_ALLOC_MAX_PARTIALS;
  const int nblocks = (int)((n + FB2_NO_ALLOC_CHUNK - 1) / FB2_NO_ALLOC_CHUNK);
  cudaStream_t st = fb2_stream();
  k_dot_chunks<<<nblocks, 256, 0, st>>>(x_dev, y_dev, (i64)n, scratch);
  k_sum_partials<<<1, 1024, 0, st>>>(scratch, nblocks, d_result);
  FB2_CUDA(cudaGetLastError());
  FB2_CUDA(cudaMemcpyAsync(result_host, d_result, sizeof(double), cudaMemcpyDeviceToHost, st));
  FB2_CUDA(cudaStreamSynchronize(st));
  return 0;
}
