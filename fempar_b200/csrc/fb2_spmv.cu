// CSR SpMV for sm_100a: row tiles streamed through shared memory with TMA bulk copies
// (cp.async.bulk + mbarrier, SASS UBLKCP), products formed in place, per-row sub-warp reduction,
// optional fused <y,x> partial (the <Ap,p> of CG).  Replaces csr matvec (csr_sparse_matrix.f90:545-568).
//
// Data layout in HBM: val (f64) and ja (int32, 1-based) are two contiguous streams; a tile of R
// consecutive rows is the contiguous element range [irp[r0], irp[r0+R]) of both.  Algorithmic bytes
// per row tile: 12 B per non-zero + 24 B per row (irp, y, x).
#include <stdlib.h>
#include <algorithm>
#include <vector>
#include <cub/cub.cuh>
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
// the same copy with an L2 eviction policy: the once-read streams (val, ja) are marked evict_first so that they do not push the
// re-used data (x segments, class index tables: L2-resident across neighbouring tiles) out of the 126 MB L2
__device__ __forceinline__ void bulk_g2s_hint(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                   smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
               : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
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
// wait with a suspend-time hint: the thread may be parked by the hardware instead of spinning through the issue slots
__device__ __forceinline__ void mbar_wait_parked(uint64_t* bar, uint32_t parity, uint32_t hint_ns) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "LAB_WAITP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n"
      "@P1 bra DONEP;\n"
      "bra LAB_WAITP;\n"
      "DONEP:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity), "r"(hint_ns)
      : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

#define SPMV_MAX_STAGES 8
#define SPMV_DS 512   // compressed variant: distinct columns of a tile staged in shared memory (larger dictionaries stay global)

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- pattern classes (index-free tiles) ----
// On a structured part most row tiles are translated copies of a few patterns: the columns of tile t are r0 + delta_k with the
// same deltas for every tile of its class (DOF ids are affine in the lattice position away from the boundary).  For such a
// tile neither ja nor x gathers are needed: the class descriptor lists the <= 31 contiguous column segments [r0 + delta,
// r0 + delta + len) the tile touches -- the producer warp bulk-copies them from x (L2-resident: neighbouring tiles share them)
// into an "x window" in the stage -- and a uint16 window position per entry (class constant, copied from an L2-resident
// table).  HBM traffic of a class tile: 8 B per non-zero instead of 12 B, and the consumers read shared memory only.
// The class of a tile is never assumed: fb2_spmv_build_classes compares every tile entry by entry with the representative
// of its class and checks the segment bounds; everything else streams ja as before.  Products are formed in the same
// order from the same operands, so y is bit-identical to the ja path (tests: test_midsize_spmv_variants_bitwise).
struct SpmvClassDesc {
  int nseg, idx_bytes, x_bytes, pad;
  int delta[32];   // first column of the segment relative to the first row of the tile (even, so 16-byte aligned in x)
  int len[32];     // doubles, even
  int off[32];     // position of the segment in the window, doubles
};
#define SPMV_NOCLASS 0xFF
// 0xFE in the tile table: a tile with its own descriptor and index table (no class mates, but its columns still form <= 31
// segments of x: index-free like a class tile, 10 B per non-zero); its descriptor index is in tile_priv[]
#define SPMV_PRIVATE 0xFE
#define SPMV_NONE (-1)
// transposed rows: stagger (entries) of chunk g in the stage, per stream.  val: 8 x 8 B = half of the 32 banks, so that the four
// 8-lane groups of a warp pair up into two conflict-free wavefronts; ja: 8 x 4 B and class index table: 16 x 2 B = 8 banks each
#define SPMV_STAG_VAL 8
#define SPMV_STAG_JA 8
#define SPMV_STAG_IDX 16
#define SPMV_WAIT_NS_DEFAULT 0u
#define SPMV_L2HINT_DEFAULT 1

// Tried and rejected in round 2 (profiles/r2_spmv_history.txt): consumers that stream the rows of a warp's range 32 entries at a
// time (contiguous shared-memory reads, one xor tree per finished row).  It halves the shared-memory wavefronts per non-zero
// but needs 107 instead of 54 instructions per 32 non-zeros (issue slots 82 % busy): 24.9 ms against 15.6 ms at config 2.
// Also tried and dropped (same file, round 2): a per-class row schedule that deals rows to lane groups by length (62 % -> 95 %
// lane utilisation, but equal-type rows of different chunks are 512 entries apart = the same banks: 4.0 instead of 2.5
// wavefronts per LDS.64; 15.9 vs 16.2 ms with schedule, -5 % for Q3) and 16 / 4 lanes per row (18.8 / 17.5 ms).
// Persistent block over tiles t = blockIdx.x, blockIdx.x + gridDim.x, ...  Warp 8 is the TMA producer
// (one elected lane): it runs STAGES tiles ahead, waiting on the per-stage "empty" mbarrier and arming
// the "full" one with the byte count of the three bulk copies (val, ja, irp slice).  The 8 consumer
// warps never synchronise with each other: each owns R/8 rows of every tile, gathers x, reduces its
// rows inside TPR-lane groups and releases the stage with one arrive per warp.
// COMP: the column indices are streamed as uint16 positions in a per-tile dictionary of distinct columns (built once by
// fb2_spmv_compress): 10 B + ~0.5 B of dictionary per non-zero instead of 12 B.  `ja` then points at the uint16 stream, the
// element range of a tile is aligned to 8 entries (16-byte bulk copies), and the products are formed in the same order, so y
// is bit-identical to the uncompressed kernel.
// NC consumer threads (256 or 128) + the producer warp.  With pattern-class tiles the consumers read shared memory only and the
// per-tile bookkeeping of every warp (barrier wait, row bounds, reduction, release: ~150 instructions, half of the stall
// samples in the ncu capture of round 2) is what limits them: 4 warps with 16 entries per lane and tile instead of 8 with 8.
template <bool FUSE_DOT, bool COMP, int NC>
__global__ void __launch_bounds__(NC + 32, 4) k_spmv_stream(i64 n, const i64* __restrict__ irp, const void* __restrict__ ja_any,
                                                             const double* __restrict__ val, const double* __restrict__ x,
                                                             double* __restrict__ y, double* partials, const int* done_flag,
                                                             int R, int cap, i64 ntiles, int STAGES,
                                                             const i64* __restrict__ tile_row, const unsigned char* __restrict__ tile_nrows,
                                                             int chunk, const int* __restrict__ dict, const i64* __restrict__ dict_ptr,
                                                             int dcap, const unsigned char* __restrict__ tile_class,
                                                             const SpmvClassDesc* __restrict__ cls_desc,
                                                             const unsigned short* __restrict__ cls_idx, int tr, uint32_t wait_hint,
                                                             const int* __restrict__ tile_priv, int l2hint) {
  constexpr int IB = COMP ? 2 : 4;                       // bytes per streamed column index
  constexpr i64 AL = COMP ? 7 : 3;                       // element-range alignment mask
  const int* ja = (const int*)ja_any;
  const unsigned short* ja16 = (const unsigned short*)ja_any;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ uint64_t full_bar[SPMV_MAX_STAGES], empty_bar[SPMV_MAX_STAGES];
  __shared__ i64 s_e0a[SPMV_MAX_STAGES], s_r0[SPMV_MAX_STAGES];
  __shared__ int s_nr[SPMV_MAX_STAGES], s_nd[SPMV_MAX_STAGES], s_cls[SPMV_MAX_STAGES], s_xoff[SPMV_MAX_STAGES], s_lastcls[SPMV_MAX_STAGES];
  __shared__ i64 s_d0[SPMV_MAX_STAGES];
  __shared__ double s_red[32];
  if (done_flag && *done_flag) return;
  const int tid = threadIdx.x;
  const size_t irp_bytes = (size_t)(R + 2) * 8;
  // COMP: one region of SPMV_DS doubles per stage: the dictionary lands in its upper half (TMA), is read into registers, and
  // the x values gathered through it then overwrite the region in place (x gathered once per distinct column of the tile)
  const size_t xs_off = (size_t)cap * (8 + IB);
  const size_t dict_off = xs_off + (size_t)SPMV_DS * 4;
  const size_t irp_off = xs_off + (COMP ? (size_t)SPMV_DS * 8 : 0);
  (void)dcap;
  const size_t stage_bytes = irp_off + irp_bytes;
  // tiles of this CTA: persistent (chunk == 0): blockIdx.x, blockIdx.x + gridDim.x, ... ; chunked: `chunk` consecutive tiles
  const i64 t_first = chunk ? (i64)blockIdx.x * chunk : (i64)blockIdx.x;
  const i64 t_step = chunk ? 1 : (i64)gridDim.x;
  const i64 t_end = chunk ? min(t_first + chunk, ntiles) : ntiles;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], NC / 32); s_lastcls[s] = SPMV_NONE; }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  if (tid >= NC) {
    // ---------------- producer warp ----------------
    // Every lane keeps the header of the tile; lane 0 waits for the stage, arms the barrier and issues the irp copy.  The
    // entries are copied in `nch` chunks (1, or one per lane group of a consumer warp when the rows are dealt transposed, see
    // the consumer side): lane g copies chunk g of val and of ja (or of the class index table); for a class tile lane l also
    // copies x segment l - 1.
    const int plane = tid - NC;
    const uint64_t pol_stream = l2_policy_evict_first(), pol_keep = l2_policy_evict_last();
    const int TPRP = (R <= 16 ? 128 : 256) / R;
    const int nch = tr ? 32 / TPRP : 1;                  // chunks per tile
    const int CW = tr ? R / nch : R;                     // rows per chunk
    const i64 CA = tr ? (i64)7 : AL;                     // chunk alignment (entries, relative to the aligned tile start)
    i64 t = t_first;
    i64 e0 = 0, r0 = 0, d0 = 0, d1 = 0, cb = 0, ce = 0;
    int nr = 0, cls = SPMV_NONE;
    auto load_tile = [&](i64 tt, i64& R0, int& NR, i64& E0, i64& CB, i64& CE, i64& D0, i64& D1, int& CLS) {
      R0 = tile_row[tt]; NR = tile_nrows[tt]; E0 = irp[R0];
      CB = CE = 0;
      if (plane < nch) { CB = irp[R0 + min(plane * CW, NR)]; CE = irp[R0 + min((plane + 1) * CW, NR)]; }
      if (COMP) { D0 = dict_ptr[tt]; D1 = dict_ptr[tt + 1]; }
      CLS = SPMV_NONE;
      if (!COMP && tile_class) {
        const int c8 = tile_class[tt];
        CLS = c8 == SPMV_NOCLASS ? SPMV_NONE : c8 == SPMV_PRIVATE ? tile_priv[tt] : c8;
      }
    };
    if (t < t_end) load_tile(t, r0, nr, e0, cb, ce, d0, d1, cls);
    // stage index and ring phase advance incrementally (STAGES is a run-time value: no division per tile)
    int s = 0;
    uint32_t ph = 0;
    for (; t < t_end; t += t_step) {
      unsigned char* base = smem_raw + s * stage_bytes;
      i64 ne0 = 0, nr0 = 0, nd0 = 0, nd1 = 0, ncb = 0, nce = 0;
      int nnr = 0, ncls = SPMV_NONE;
      // prefetch the header of the next tile before blocking on the empty barrier
      const i64 tn = t + t_step;
      if (tn < t_end) load_tile(tn, nr0, nnr, ne0, ncb, nce, nd0, nd1, ncls);
      const i64 e0a = e0 & ~AL;
      const i64 ca = e0a + ((cb - e0a) & ~CA);                                    // aligned start of this lane's chunk
      const uint32_t cn = (plane < nch && ce > cb) ? (uint32_t)((ce - ca + CA) & ~CA) : 0u;   // its entries
      const uint32_t cnt_all = __reduce_add_sync(0xffffffffu, cn);
      const bool is_cls = !COMP && cls != SPMV_NONE;
      // consecutive tiles of an x-row of cells share their class: the index table this stage received two tiles ago is still
      // there (only index / ja copies write that region) and is not copied again
      const bool have_idx = is_cls && s_lastcls[s] == cls;
      __syncwarp();
      if (plane == 0) {
        mbar_wait_parked(&empty_bar[s], ph ^ 1u, 20000u);
        // a dictionary that does not fit the region is not staged: the consumers then go through it in global memory
        const uint32_t dbytes = (COMP && d1 - d0 <= SPMV_DS) ? (uint32_t)(d1 - d0) * 4u : 0u;
        s_e0a[s] = e0a; s_r0[s] = r0; s_nr[s] = nr; s_nd[s] = (int)(d1 - d0); s_d0[s] = d0; s_cls[s] = cls;
        s_lastcls[s] = is_cls ? cls : SPMV_NONE;
        if (is_cls) {
          const SpmvClassDesc& D = cls_desc[cls];
          s_xoff[s] = D.idx_bytes;
          const uint32_t ib = have_idx ? 0u : (tr ? cnt_all * 2u : (uint32_t)D.idx_bytes);
          mbar_expect_tx(&full_bar[s], cnt_all * 8u + ib + (uint32_t)D.x_bytes + (uint32_t)irp_bytes);
          if (!have_idx && !tr) bulk_g2s(base + (size_t)cap * 8, cls_idx + (size_t)cls * cap, (uint32_t)D.idx_bytes, &full_bar[s]);
        } else {
          mbar_expect_tx(&full_bar[s], cnt_all * (uint32_t)(8 + IB) + dbytes + (uint32_t)irp_bytes);
          if (COMP && dbytes) bulk_g2s(base + dict_off, dict + d0, dbytes, &full_bar[s]);
        }
        bulk_g2s(base + irp_off, irp + r0, (uint32_t)irp_bytes, &full_bar[s]);
      }
      __syncwarp();   // the barrier is armed and the stage is free: the other lanes may copy into it
      if (cn) {
        // chunk g lands SPMV_STAG_* x g entries behind its natural place (transposed rows only): the lane groups of a consumer
        // warp then read different shared-memory banks although their rows start a multiple of 128 bytes apart
        const size_t de = (size_t)(ca - e0a);
        const size_t g = tr ? (size_t)plane : 0;
        if (l2hint & 1) bulk_g2s_hint(base + (de + SPMV_STAG_VAL * g) * 8, val + ca, cn * 8u, &full_bar[s], pol_stream);
        else bulk_g2s(base + (de + SPMV_STAG_VAL * g) * 8, val + ca, cn * 8u, &full_bar[s]);
        if (is_cls) {
          if (!have_idx && tr)
            bulk_g2s(base + (size_t)cap * 8 + (de + SPMV_STAG_IDX * g) * 2, cls_idx + (size_t)cls * cap + de, cn * 2u, &full_bar[s]);
        } else if (COMP) bulk_g2s(base + (size_t)cap * 8 + de * 2, ja16 + ca, cn * 2u, &full_bar[s]);
        else if (l2hint & 1) bulk_g2s_hint(base + (size_t)cap * 8 + (de + SPMV_STAG_JA * g) * 4, ja + ca, cn * 4u, &full_bar[s], pol_stream);
        else bulk_g2s(base + (size_t)cap * 8 + (de + SPMV_STAG_JA * g) * 4, ja + ca, cn * 4u, &full_bar[s]);
      }
      if (is_cls) {
        // x window of a class tile: lane l copies segment l - 1
        const SpmvClassDesc& D = cls_desc[cls];
        if (plane >= 1 && plane <= D.nseg) {
          const int g = plane - 1;
          if (l2hint & 2)
            bulk_g2s_hint(base + (size_t)cap * 8 + (size_t)D.idx_bytes + (size_t)D.off[g] * 8, x + (r0 + D.delta[g]),
                          (uint32_t)D.len[g] * 8u, &full_bar[s], pol_keep);
          else
            bulk_g2s(base + (size_t)cap * 8 + (size_t)D.idx_bytes + (size_t)D.off[g] * 8, x + (r0 + D.delta[g]), (uint32_t)D.len[g] * 8u,
                     &full_bar[s]);
        }
      }
      e0 = ne0; r0 = nr0; nr = nnr; d0 = nd0; d1 = nd1; cls = ncls; cb = ncb; ce = nce;
      if (++s == STAGES) { s = 0; ph ^= 1u; }
    }
    return;
  }
  // ---------------- consumer warps ----------------
  // lanes per row: a function of the tile height only (8 for R = 32 and 16, 16 for R = 8, 4 for R = 64), so that the summation
  // order of a row -- and y, bit for bit -- is the same for every kernel variant and consumer count (groups beyond R idle)
  const int TPR = (R <= 16 ? 128 : 256) / R;
  const int lane = tid & 31, w = tid >> 5;
  // tr (transposed rows): lane group g of warp w owns row g * (NC / 32) + w instead of w * (32 / TPR) + g.  Where the row
  // lengths repeat with the period of a cell's created DOFs (8 for Q2) the groups of a warp then hold rows of equal length
  // and finish together; chunk g of the tile (the rows of group g) is staged SPMV_STAG_* x g entries behind its natural place
  // With fewer lane groups than rows the warps go over the tile in passes (NC = 128 and R = 32: two).
  const int GPW = 32 / TPR, rows_pp = (NC / 32) * GPW, NP = (R + rows_pp - 1) / rows_pp;
  const int grp = tr ? lane / TPR : 0;
  const int row0 = tr ? grp * (R / GPW) + w : tid / TPR;
  const int row_step = tr ? NC / 32 : rows_pp;
  const int sub = tid & (TPR - 1);
  double dacc = 0.0;
  int s = 0;
  uint32_t ph = 0;
  for (i64 t = t_first; t < t_end; t += t_step) {
    if (wait_hint) mbar_wait_parked(&full_bar[s], ph, wait_hint);
    else mbar_wait(&full_bar[s], ph);
    const i64 r0 = s_r0[s];
    const unsigned char* base = smem_raw + s * stage_bytes;
    const double* sv = (const double*)base + SPMV_STAG_VAL * grp;
    const int* sj = (const int*)(base + (size_t)cap * 8) + SPMV_STAG_JA * grp;
    const unsigned short* sj16 = (const unsigned short*)(base + (size_t)cap * 8) + (COMP ? 0 : SPMV_STAG_IDX * grp);
    const int* sd = (const int*)(base + dict_off);
    const i64* s_irp = (const i64*)(base + irp_off);
    // COMP: the consumer warps gather x once per DISTINCT column of the tile into shared memory (7-9x fewer gathers than
    // entries); the inner loop then reads shared memory only.  One named barrier per tile; the per-stage buffer is safe to
    // rewrite STAGES tiles later because every warp has passed the barriers of the tiles in between.
    double* xs = (double*)(base + xs_off);
    bool big = false;
    const int* dict_g = nullptr;
    if (COMP) {
      const int nd = s_nd[s];
      big = nd > SPMV_DS;
      dict_g = dict + s_d0[s];
      if (!big) {
        static_assert(!COMP || SPMV_DS <= 2 * NC, "two dictionary entries per consumer thread");
        const int c0 = tid < nd ? sd[tid] : -1, c1 = tid + NC < nd ? sd[tid + NC] : -1;
        asm volatile("bar.sync 2, %0;" ::"r"(NC) : "memory");   // every entry is in a register before the overwrite
        if (c0 >= 0) xs[tid] = x[c0];
        if (c1 >= 0) xs[tid + NC] = x[c1];
        asm volatile("bar.sync 2, %0;" ::"r"(NC) : "memory");
      }
    }
    // class tile: uint16 window positions where ja would be, the x window right behind them
    const bool in_class = !COMP && tile_class && s_cls[s] != SPMV_NONE;
    const double* sxw = (const double*)(base + (size_t)cap * 8 + (in_class ? s_xoff[s] : 0));
    // x value of the k-th staged entry
    auto xat = [&](int k) -> double {
      if (COMP) return big ? x[dict_g[sj16[k]]] : xs[sj16[k]];
      else return in_class ? sxw[sj16[k]] : x[sj[k] - 1];
    };
    const i64 e0a = s_e0a[s];
    const int nr_t = s_nr[s];
    for (int pass = 0, row = row0; pass < NP; ++pass, row += row_step) {
    double acc = 0.0;
    const bool live = row < nr_t;
    // the row's own x for the fused <y,x>: issued before the entry loop so that its latency is hidden behind it
    double xrow = 0.0;
    if (FUSE_DOT && sub == 0 && live) xrow = x[r0 + row];
    if (live && in_class) {
      // shared memory only: a plain strided loop (no gather latency to batch against)
      const int b = (int)(s_irp[row] - e0a), e = (int)(s_irp[row + 1] - e0a);
#pragma unroll 4
      for (int k = b + sub; k < e; k += TPR) acc += sv[k] * sxw[sj16[k]];
    } else if (live) {
      const int b = (int)(s_irp[row] - e0a), e = (int)(s_irp[row + 1] - e0a);
      int k = b + sub;
      for (; k + 3 * TPR < e; k += 4 * TPR) {
        const double x0 = xat(k), x1 = xat(k + TPR), x2 = xat(k + 2 * TPR), x3 = xat(k + 3 * TPR);
        acc += sv[k] * x0;
        acc += sv[k + TPR] * x1;
        acc += sv[k + 2 * TPR] * x2;
        acc += sv[k + 3 * TPR] * x3;
      }
      if (k < e) {
        // tail of up to 3 entries per lane as ONE masked batch, so that its gathers are in flight together
        const int k1 = k + TPR, k2 = k + 2 * TPR;
        const bool p1 = k1 < e, p2 = k2 < e;
        const double x0 = xat(k), x1 = p1 ? xat(k1) : 0.0, x2 = p2 ? xat(k2) : 0.0;
        acc += sv[k] * x0;
        if (p1) acc += sv[k1] * x1;
        if (p2) acc += sv[k2] * x2;
      }
    }
    if (TPR == 8) {   // the usual group width, unrolled
      acc += __shfl_down_sync(0xffffffffu, acc, 4, 8);
      acc += __shfl_down_sync(0xffffffffu, acc, 2, 8);
      acc += __shfl_down_sync(0xffffffffu, acc, 1, 8);
    } else {
      for (int o = TPR >> 1; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o, TPR);
    }
    if (sub == 0 && live) {
      y[r0 + row] = acc;
      if (FUSE_DOT) dacc += acc * xrow;
    }
    }   // passes
    // COMP: the consumers wrote xs[] through the generic proxy into a region the producer's next bulk copy (async proxy)
    // overwrites: order the two proxies before releasing the stage
    if (COMP) fence_proxy_async();
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty_bar[s]);
    if (++s == STAGES) { s = 0; ph ^= 1u; }
  }
  if (FUSE_DOT) {
    // deterministic block sum over the consumer warps (named barrier 1: the producer warp has exited)
    for (int o = 16; o > 0; o >>= 1) dacc += __shfl_down_sync(0xffffffffu, dacc, o);
    if (lane == 0) s_red[w] = dacc;
    asm volatile("bar.sync 1, %0;" ::"r"(NC) : "memory");
    if (tid == 0) {
      double r = 0.0;
      for (int i = 0; i < NC / 32; ++i) r += s_red[i];
      partials[blockIdx.x] = r;
    }
  }
}

// ---- tile table: fixed tiles of R rows, split in halves (down to 2 rows, so that every tile starts on an even row and its
// irp slice is 16-byte aligned for the bulk copy) until the aligned element range fits the stage buffer ----
__device__ __forceinline__ i64 tile_count4(const i64* __restrict__ irp, i64 r0, i64 r1) {
  return (irp[r1] - (irp[r0] & ~(i64)3) + 3) & ~(i64)3;
}
__global__ void k_tile_pieces(i64 n, const i64* __restrict__ irp, int R, int cap, i64 nfixed, int* pieces, int* too_long) {
  i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t > nfixed) return;
  if (t == nfixed) { pieces[t] = 0; return; }
  const i64 r0 = t * R, r1 = min(r0 + R, n);
  int np = 1;
  for (int rows = R; rows >= 2; rows >>= 1, np <<= 1) {
    bool fits = true;
    for (i64 a = r0; a < r1; a += rows) fits &= tile_count4(irp, a, min(a + rows, r1)) <= cap;
    if (fits) { pieces[t] = (int)((r1 - r0 + rows - 1) / rows); return; }
  }
  pieces[t] = (int)((r1 - r0 + 1) / 2);
  *too_long = 1;
}
__global__ void k_tile_fill(i64 n, const i64* __restrict__ irp, int R, int cap, i64 nfixed, const int* __restrict__ off,
                            i64* tile_row, unsigned char* tile_nrows) {
  i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nfixed) return;
  const i64 r0 = t * R, r1 = min(r0 + R, n);
  const int np = off[t + 1] - off[t];
  int rows = R;
  while ((r1 - r0 + rows - 1) / rows < np) rows >>= 1;
  int o = off[t];
  for (i64 a = r0; a < r1; a += rows, ++o) { tile_row[o] = a; tile_nrows[o] = (unsigned char)(min(a + rows, r1) - a); }
}

// ---- compressed column indices: per tile the sorted distinct columns (dictionary) + a uint16 position per entry ----
#define DICT_SORT_MAX 4096
template <bool FILL>
__global__ void __launch_bounds__(256) k_tile_dict(const i64* __restrict__ irp, const int* __restrict__ ja,
                                                   const i64* __restrict__ tile_row, const unsigned char* __restrict__ tile_nrows,
                                                   i64* cnt, int* maxcnt, const i64* __restrict__ dict_ptr, int* dict,
                                                   unsigned short* idx16) {
  __shared__ int sb[DICT_SORT_MAX], su[DICT_SORT_MAX];
  __shared__ int s_scan[256];
  const int tid = threadIdx.x;
  const i64 t = blockIdx.x;
  const i64 r0 = tile_row[t];
  const i64 e0 = irp[r0], e1 = irp[r0 + tile_nrows[t]];
  const int m = (int)(e1 - e0);
  int np2 = 256;
  while (np2 < m) np2 <<= 1;
  for (int i = tid; i < np2; i += 256) sb[i] = i < m ? ja[e0 + i] - 1 : 0x7fffffff;
  __syncthreads();
  for (int k = 2; k <= np2; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < np2; i += 256) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const int a = sb[i], b = sb[ixj];
          if ((a > b) == ((i & k) == 0)) { sb[i] = b; sb[ixj] = a; }
        }
      }
      __syncthreads();
    }
  // ordered compaction of the distinct values: each thread owns a contiguous chunk
  const int per = np2 / 256, lo = tid * per;
  int heads = 0;
  for (int i = lo; i < lo + per; ++i) heads += (i < m && (i == 0 || sb[i] != sb[i - 1])) ? 1 : 0;
  s_scan[tid] = heads;
  __syncthreads();
  for (int o = 1; o < 256; o <<= 1) {            // Hillis-Steele inclusive scan
    const int v = tid >= o ? s_scan[tid - o] : 0;
    __syncthreads();
    s_scan[tid] += v;
    __syncthreads();
  }
  const int nu = s_scan[255];
  int pos = s_scan[tid] - heads;
  for (int i = lo; i < lo + per; ++i)
    if (i < m && (i == 0 || sb[i] != sb[i - 1])) su[pos++] = sb[i];
  __syncthreads();
  const int nu4 = (nu + 3) & ~3;
  if (!FILL) {
    if (tid == 0) { cnt[t] = nu4; atomicMax(maxcnt, nu4); }
    return;
  }
  const i64 d0 = dict_ptr[t];
  for (int i = tid; i < nu4; i += 256) dict[d0 + i] = su[i < nu ? i : nu - 1];
  for (int i = tid; i < m; i += 256) {
    const int c = ja[e0 + i] - 1;
    int a = 0, b = nu;
    while (a < b) { const int mid = (a + b) >> 1; if (su[mid] < c) a = mid + 1; else b = mid; }
    idx16[e0 + i] = (unsigned short)a;
  }
}

// Off by default: at parity with the plain int32 stream on B200, not faster (Q2 256^3: 19.46 vs 19.04 ms per SpMV although
// 11 % fewer bytes move, profiles/r1_spmv_compress_probe.txt) -- the gather phase and its two barriers per tile cost what the
// bytes save while only two stages are in flight.  Kept behind fb2_set_spmv_variant(2) / FB2_SPMV_COMPRESS=1 as the starting
// point for a deeper ring of the small streams.
static int g_compress = getenv("FB2_SPMV_COMPRESS") ? atoi(getenv("FB2_SPMV_COMPRESS")) : 0;
void fb2_spmv_set_compress(bool on) { g_compress = on ? 1 : 0; }
static int spmv_compress(fb2_matrix_s* A, int capbuf, int R, int STAGES) {
  const int enabled = g_compress;
  A->spmv_comp = -1;
  const i64 nt = A->spmv_ntiles;
  if (!enabled || nt == 0 || capbuf > DICT_SORT_MAX || nt > 0x7fffffffLL) return 0;
  cudaStream_t st = fb2_stream();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  i64* cnt; int* d_max;
  FB2_TRY(fb2_dalloc(&cnt, (size_t)nt + 1)); FB2_TRY(fb2_dalloc(&d_max, 1));
  FB2_TRY(fb2_dalloc(&A->d_dict_ptr, (size_t)nt + 1));
  FB2_CUDA(cudaMemsetAsync(cnt + nt, 0, sizeof(i64), st));
  FB2_CUDA(cudaMemsetAsync(d_max, 0, sizeof(int), st));
  k_tile_dict<false><<<(unsigned)nt, 256, 0, st>>>(A->d_irp, A->d_ja, A->d_tile_row, A->d_tile_nrows, cnt, d_max, nullptr, nullptr, nullptr);
  FB2_CUDA(cudaGetLastError());
  {
    void* tmp = nullptr; size_t bytes = 0;
    FB2_CUDA(cub::DeviceScan::ExclusiveSum(tmp, bytes, cnt, A->d_dict_ptr, nt + 1, st));
    FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
    cudaError_t e = cub::DeviceScan::ExclusiveSum(tmp, bytes, cnt, A->d_dict_ptr, nt + 1, st);
    cudaStreamSynchronize(st);
    cudaFree(tmp);
    FB2_CUDA(e);
  }
  i64 total = 0; int dcap = 0;
  FB2_CUDA(fb2_memcpy(&total, A->d_dict_ptr + nt, sizeof(i64), cudaMemcpyDeviceToHost));
  FB2_CUDA(fb2_memcpy(&dcap, d_max, sizeof(int), cudaMemcpyDeviceToHost));
  cudaFree(cnt); cudaFree(d_max);
  const size_t smem = ((size_t)capbuf * 10 + (size_t)SPMV_DS * 8 + (size_t)(R + 2) * 8) * STAGES;
  // worth it only when the dictionaries are small next to the 2 B per entry they save, and the stage still fits
  bool ok = dcap > 0 && smem <= 200 * 1024 && (double)total * 4.0 < 0.5 * 2.0 * (double)A->nnz;
  if (ok) {
    // the extra streams are optional: when HBM is short the uncompressed kernel stays in use
    if (cudaMalloc((void**)&A->d_idx16, sizeof(unsigned short) * ((size_t)A->nnz + 16)) != cudaSuccess) { cudaGetLastError(); A->d_idx16 = nullptr; ok = false; }
    else if (cudaMalloc((void**)&A->d_dict, sizeof(int) * ((size_t)total + 4)) != cudaSuccess) { cudaGetLastError(); A->d_dict = nullptr; ok = false; }
  }
  if (!ok) {
    cudaFree(A->d_idx16); cudaFree(A->d_dict_ptr); A->d_idx16 = nullptr; A->d_dict_ptr = nullptr;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
  }
  FB2_CUDA(cudaMemsetAsync(A->d_idx16, 0, sizeof(unsigned short) * ((size_t)A->nnz + 16), st));
  FB2_CUDA(cudaMemsetAsync(A->d_dict, 0, sizeof(int) * ((size_t)total + 4), st));
  k_tile_dict<true><<<(unsigned)nt, 256, 0, st>>>(A->d_irp, A->d_ja, A->d_tile_row, A->d_tile_nrows, nullptr, nullptr, A->d_dict_ptr, A->d_dict, A->d_idx16);
  FB2_CUDA(cudaGetLastError());
  A->spmv_dcap = dcap; A->spmv_smem_comp = smem;
  // (the attribute is per kernel, not per matrix: always the largest value any matrix may ask for)
  FB2_CUDA(cudaFuncSetAttribute(k_spmv_stream<true, true, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  FB2_CUDA(cudaFuncSetAttribute(k_spmv_stream<false, true, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  cudaEventRecord(e1, st);
  FB2_CUDA(cudaStreamSynchronize(st));
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  fb2_record_phase("spmv_compress", ms);
  fb2_record_phase("spmv_dict_entries", (double)total);
  fb2_record_phase("spmv_dict_cap", (double)dcap);
  A->spmv_comp = 1;
  return 0;
}


// ---- pattern classes: signature hash per tile, the most frequent signatures become classes, every tile is then compared
// entry by entry with the representative of its class ----
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
// warp per tile: order-independent sum of position-keyed hashes of (row lengths, column deltas), seeded by (rows, shift)
__global__ void k_tile_hash(i64 ntiles, const i64* __restrict__ irp, const int* __restrict__ ja, const i64* __restrict__ tile_row,
                            const unsigned char* __restrict__ tile_nrows, unsigned long long* hash, int* tile_id) {
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (i64 t = (i64)blockIdx.x * warps + w; t < ntiles; t += (i64)gridDim.x * warps) {
    const i64 r0 = tile_row[t];
    const int nr = tile_nrows[t];
    const i64 e0 = irp[r0], e1 = irp[r0 + nr];
    unsigned long long h = 0;
    for (int r = lane; r < nr; r += 32) h += mix64(((unsigned long long)(0x40000000u + r) << 32) ^ (unsigned long long)(irp[r0 + r + 1] - irp[r0 + r]));
    for (i64 k = e0 + lane; k < e1; k += 32)
      h += mix64(((unsigned long long)(k - e0) << 32) ^ (unsigned long long)(unsigned)(ja[k] - 1 - (int)r0));
    for (int o = 16; o > 0; o >>= 1) h += __shfl_down_sync(0xffffffffu, h, o);
    if (lane == 0) {
      hash[t] = mix64(h ^ ((unsigned long long)nr << 8) ^ (unsigned long long)(e0 & 3));
      tile_id[t] = (int)t;
    }
  }
}

#define SPMV_CLS_MAX 250
#define SPMV_SEG_GAP 8
// block per class: segments of the representative tile's distinct columns and the window position of every entry
__global__ void __launch_bounds__(256) k_class_build(const i64* __restrict__ irp, const int* __restrict__ ja,
                                                     const i64* __restrict__ tile_row, const unsigned char* __restrict__ tile_nrows,
                                                     const int* __restrict__ rep, int cap, int idx_extra, int x_align, SpmvClassDesc* desc,
                                                     unsigned short* cls_idx, int* valid) {
  __shared__ int sb[DICT_SORT_MAX], su[DICT_SORT_MAX];
  __shared__ int s_scan[256];
  __shared__ SpmvClassDesc D;
  __shared__ int s_ok;
  const int tid = threadIdx.x, c = blockIdx.x;
  const i64 t = rep[c];
  const i64 r0 = tile_row[t];
  const i64 e0 = irp[r0], e1 = irp[r0 + tile_nrows[t]];
  const int m = (int)(e1 - e0), shift = (int)(e0 & 3);
  int np2 = 256;
  while (np2 < m) np2 <<= 1;
  for (int i = tid; i < np2; i += 256) sb[i] = i < m ? ja[e0 + i] - 1 - (int)r0 : 0x7fffffff;
  __syncthreads();
  for (int k = 2; k <= np2; k <<= 1)
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < np2; i += 256) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const int a = sb[i], b = sb[ixj];
          if ((a > b) == ((i & k) == 0)) { sb[i] = b; sb[ixj] = a; }
        }
      }
      __syncthreads();
    }
  const int per = np2 / 256, lo = tid * per;
  int heads = 0;
  for (int i = lo; i < lo + per; ++i) heads += (i < m && (i == 0 || sb[i] != sb[i - 1])) ? 1 : 0;
  s_scan[tid] = heads;
  __syncthreads();
  for (int o = 1; o < 256; o <<= 1) {
    const int v = tid >= o ? s_scan[tid - o] : 0;
    __syncthreads();
    s_scan[tid] += v;
    __syncthreads();
  }
  const int nu = s_scan[255];
  int pos = s_scan[tid] - heads;
  for (int i = lo; i < lo + per; ++i)
    if (i < m && (i == 0 || sb[i] != sb[i - 1])) su[pos++] = sb[i];
  __syncthreads();
  if (tid == 0) {
    // greedy runs of the sorted distinct deltas; starts rounded down and lengths rounded up to even (16-byte bulk copies).
    // x_align: every segment starts at a window position congruent to its first column modulo 16 doubles (= the 32 banks), so
    // that window position == column (mod 16) across ALL segments.  The four lane groups of a consumer warp read columns that
    // differ by multiples of 8 (the same kind of row in consecutive cells) and the 8 lanes of a group read 8 consecutive
    // entries, whose columns are distinct modulo 8 even where they straddle two segments: with packed segments those residues
    // were scrambled (ncu: 3.5 instead of 2 wavefronts per x-window LDS.64).  Falls back to packed segments when the padding
    // does not fit the stage.
    int nseg = 0, total = 0, extent = 0, ok = m > 0;
    // idx_extra: room behind the table for the staggered chunks of the transposed row layout (the x window starts behind it)
    const int idx_n = (shift + m + 7) & ~7;
    const int xw_cap = (cap * 4 - (idx_n + idx_extra) * 2) / 8;
    for (int mode = x_align ? 1 : 0; mode >= 0; --mode) {
      nseg = 0; total = 0; extent = 0; ok = m > 0;
      int i = 0;
      while (ok && i < nu) {
        const int start = su[i] & ~1;
        int last = su[i];
        ++i;
        while (i < nu && su[i] - last <= SPMV_SEG_GAP) { last = su[i]; ++i; }
        const int len = (last - start + 2) & ~1;
        if (nseg >= 31) { ok = 0; break; }
        int off = extent;
        if (mode) off += (((start - off) % 16) + 16) % 16;
        D.delta[nseg] = start; D.len[nseg] = len; D.off[nseg] = off;
        extent = off + len; total += len; ++nseg;
      }
      if (extent > xw_cap || extent > 0xFFFF || idx_n + idx_extra > cap) ok = 0;
      if (ok || nseg >= 31) break;
    }
    D.nseg = nseg; D.idx_bytes = (idx_n + idx_extra) * 2; D.x_bytes = total * 8; D.pad = extent;
    for (int g = nseg; g < 32; ++g) { D.delta[g] = 0; D.len[g] = 0; D.off[g] = 0; }
    s_ok = ok;
  }
  __syncthreads();
  unsigned short* out = cls_idx + (size_t)c * cap;
  for (int i = tid; i < cap; i += 256) {
    unsigned short v = 0;
    const int k = i - shift;
    if (s_ok && k >= 0 && k < m) {
      const int d = ja[e0 + k] - 1 - (int)r0;
      int g = 0;
      while (g + 1 < D.nseg && D.delta[g + 1] <= d) ++g;
      v = (unsigned short)(D.off[g] + d - D.delta[g]);
    }
    out[i] = v;
  }
  if (tid < (int)(sizeof(SpmvClassDesc) / sizeof(int))) ((int*)&desc[c])[tid] = ((const int*)&D)[tid];
  if (tid == 0) valid[c] = s_ok;
}

// warp per tile: exact comparison with the representative of the class with the same hash, and segment bounds
__global__ void k_tile_classify(i64 ntiles, i64 n, const i64* __restrict__ irp, const int* __restrict__ ja,
                                const i64* __restrict__ tile_row, const unsigned char* __restrict__ tile_nrows,
                                const unsigned long long* __restrict__ hash, int ncls, const unsigned long long* __restrict__ cls_hash,
                                const int* __restrict__ rep, const int* __restrict__ valid, const SpmvClassDesc* __restrict__ desc,
                                unsigned char* tile_class, unsigned long long* n_in_class) {
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (i64 t = (i64)blockIdx.x * warps + w; t < ntiles; t += (i64)gridDim.x * warps) {
    const unsigned long long h = hash[t];
    int c = -1;
    for (int k = lane; k < ncls; k += 32) if (cls_hash[k] == h && valid[k]) c = k;
    for (int o = 16; o > 0; o >>= 1) c = max(c, __shfl_xor_sync(0xffffffffu, c, o));
    bool same = c >= 0;
    if (same) {
      const i64 tr = rep[c], r0 = tile_row[t], q0 = tile_row[tr];
      const int nr = tile_nrows[t];
      const i64 e0 = irp[r0], e1 = irp[r0 + nr], f0 = irp[q0], f1 = irp[q0 + tile_nrows[tr]];
      same = nr == tile_nrows[tr] && (e1 - e0) == (f1 - f0) && (e0 & 3) == (f0 & 3);
      if (same) {
        for (int r = lane; r < nr; r += 32) same = same && (irp[r0 + r + 1] - irp[r0 + r]) == (irp[q0 + r + 1] - irp[q0 + r]);
        const int m = (int)(e1 - e0);
        for (int k = lane; k < m; k += 32) same = same && (ja[e0 + k] - (int)r0) == (ja[f0 + k] - (int)q0);
        const SpmvClassDesc& D = desc[c];
        if (lane < D.nseg) same = same && r0 + D.delta[lane] >= 0 && r0 + D.delta[lane] + D.len[lane] <= n;
      }
    }
    same = __all_sync(0xffffffffu, same);
    if (lane == 0) {
      tile_class[t] = same ? (unsigned char)c : (unsigned char)SPMV_NOCLASS;
      if (same) atomicAdd(n_in_class, 1ULL);
    }
  }
}

// tiles left without class, in tile order of discovery (the descriptor index of a private tile carries no meaning)
__global__ void k_list_unclassified(i64 ntiles, const unsigned char* __restrict__ tile_class, int* list, int* count, int limit) {
  const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ntiles || tile_class[t] != SPMV_NOCLASS) return;
  const int k = atomicAdd(count, 1);
  if (k < limit) list[k] = (int)t;
}
// a private descriptor is used when it was built (valid) and every x segment lies inside the vector
__global__ void k_private_assign(int n_priv, const int* __restrict__ list, const int* __restrict__ valid,
                                 const SpmvClassDesc* __restrict__ desc, const i64* __restrict__ tile_row, i64 n, int first,
                                 unsigned char* tile_class, int* tile_priv, unsigned long long* n_used) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_priv) return;
  const int t = list[k];
  bool ok = valid[k] != 0;
  if (ok) {
    const SpmvClassDesc& D = desc[k];
    const i64 r0 = tile_row[t];
    for (int g = 0; g < D.nseg; ++g) ok = ok && r0 + D.delta[g] >= 0 && r0 + D.delta[g] + D.len[g] <= n;
  }
  if (ok) { tile_class[t] = (unsigned char)SPMV_PRIVATE; tile_priv[t] = first + k; atomicAdd(n_used, 1ULL); }
}

static int g_spmv_classes = getenv("FB2_SPMV_CLASSES") ? atoi(getenv("FB2_SPMV_CLASSES")) : 1;
void fb2_spmv_set_classes(bool on) { g_spmv_classes = on ? 1 : 0; }

static int spmv_build_classes(fb2_matrix_s* A) {
  A->spmv_cls = -1;
  const i64 nt = A->spmv_ntiles;
  if (!g_spmv_classes || nt < 64 || nt > 0x7fffffffLL || A->spmv_cap > DICT_SORT_MAX) return 0;
  cudaStream_t st = fb2_stream();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  unsigned long long *hash, *hash_s, *uniq;
  int *tid, *tid_s, *counts, *d_nruns;
  FB2_TRY(fb2_dalloc(&hash, (size_t)nt)); FB2_TRY(fb2_dalloc(&hash_s, (size_t)nt)); FB2_TRY(fb2_dalloc(&uniq, (size_t)nt));
  FB2_TRY(fb2_dalloc(&tid, (size_t)nt)); FB2_TRY(fb2_dalloc(&tid_s, (size_t)nt)); FB2_TRY(fb2_dalloc(&counts, (size_t)nt));
  FB2_TRY(fb2_dalloc(&d_nruns, 1));
  k_tile_hash<<<fb2_num_sms() * 8, 256, 0, st>>>(nt, A->d_irp, A->d_ja, A->d_tile_row, A->d_tile_nrows, hash, tid);
  FB2_CUDA(cudaGetLastError());
  {
    void* tmp = nullptr; size_t bytes = 0;
    FB2_CUDA(cub::DeviceRadixSort::SortPairs(tmp, bytes, hash, hash_s, tid, tid_s, (int)nt, 0, 64, st));
    FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
    FB2_CUDA(cub::DeviceRadixSort::SortPairs(tmp, bytes, hash, hash_s, tid, tid_s, (int)nt, 0, 64, st));
    FB2_CUDA(cudaStreamSynchronize(st));
    cudaFree(tmp);
    tmp = nullptr; bytes = 0;
    FB2_CUDA(cub::DeviceRunLengthEncode::Encode(tmp, bytes, hash_s, uniq, counts, d_nruns, (int)nt, st));
    FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
    FB2_CUDA(cub::DeviceRunLengthEncode::Encode(tmp, bytes, hash_s, uniq, counts, d_nruns, (int)nt, st));
    FB2_CUDA(cudaStreamSynchronize(st));
    cudaFree(tmp);
  }
  int nruns = 0;
  FB2_CUDA(fb2_memcpy(&nruns, d_nruns, sizeof(int), cudaMemcpyDeviceToHost));
  std::vector<unsigned long long> h_uniq((size_t)nruns);
  std::vector<int> h_cnt((size_t)nruns);
  if (nruns) {
    FB2_CUDA(fb2_memcpy(h_uniq.data(), uniq, sizeof(unsigned long long) * (size_t)nruns, cudaMemcpyDeviceToHost));
    FB2_CUDA(fb2_memcpy(h_cnt.data(), counts, sizeof(int) * (size_t)nruns, cudaMemcpyDeviceToHost));
  }
  // the most frequent signatures (at least 16 tiles each), at most SPMV_CLS_MAX
  std::vector<i64> first((size_t)nruns);
  i64 acc = 0;
  for (int k = 0; k < nruns; ++k) { first[k] = acc; acc += h_cnt[k]; }
  std::vector<int> order;
  for (int k = 0; k < nruns; ++k) if (h_cnt[k] >= 16) order.push_back(k);
  std::sort(order.begin(), order.end(), [&](int a, int b) { return h_cnt[a] != h_cnt[b] ? h_cnt[a] > h_cnt[b] : a < b; });
  if ((int)order.size() > SPMV_CLS_MAX) order.resize(SPMV_CLS_MAX);
  const int ncls = (int)order.size();
  int rc = 0;
  if (ncls > 0) {
    std::vector<unsigned long long> h_ch((size_t)ncls);
    std::vector<int> h_rep((size_t)ncls);
    std::vector<int> h_first((size_t)ncls);
    for (int c = 0; c < ncls; ++c) { h_ch[c] = h_uniq[order[c]]; h_first[c] = (int)first[order[c]]; }
    // representative = first tile of the run in the sorted order
    std::vector<int> h_tid_first((size_t)ncls);
    for (int c = 0; c < ncls; ++c) FB2_CUDA(fb2_memcpy(&h_rep[c], tid_s + h_first[c], sizeof(int), cudaMemcpyDeviceToHost));
    unsigned long long* d_ch;
    int *d_rep, *d_valid;
    unsigned long long* d_nin;
    FB2_TRY(fb2_dalloc(&d_ch, (size_t)ncls)); FB2_TRY(fb2_dalloc(&d_rep, (size_t)ncls)); FB2_TRY(fb2_dalloc(&d_valid, (size_t)ncls));
    FB2_TRY(fb2_dalloc(&d_nin, 1));
    FB2_TRY(fb2_dalloc(&A->d_cls_desc, (size_t)ncls * sizeof(SpmvClassDesc)));
    FB2_TRY(fb2_dalloc(&A->d_cls_idx, (size_t)ncls * A->spmv_cap + 8));
    FB2_TRY(fb2_dalloc(&A->d_tile_class, (size_t)nt));
    FB2_CUDA(fb2_memcpy(d_ch, h_ch.data(), sizeof(unsigned long long) * (size_t)ncls, cudaMemcpyHostToDevice));
    FB2_CUDA(fb2_memcpy(d_rep, h_rep.data(), sizeof(int) * (size_t)ncls, cudaMemcpyHostToDevice));
    FB2_CUDA(cudaMemsetAsync(d_nin, 0, sizeof(unsigned long long), st));
    const int tpr = (A->spmv_R <= 16 ? 128 : 256) / A->spmv_R;
    const int idx_extra = A->spmv_tr ? SPMV_STAG_IDX * (32 / tpr - 1) : 0;
    const int x_align = getenv("FB2_SPMV_XALIGN") ? atoi(getenv("FB2_SPMV_XALIGN")) : 0;   // bank-preserving x windows: measured, no gain (profiles/r2_spmv_history.txt)
    k_class_build<<<ncls, 256, 0, st>>>(A->d_irp, A->d_ja, A->d_tile_row, A->d_tile_nrows, d_rep, A->spmv_cap, idx_extra, x_align,
                                        (SpmvClassDesc*)A->d_cls_desc, A->d_cls_idx, d_valid);
    k_tile_classify<<<fb2_num_sms() * 8, 256, 0, st>>>(nt, A->n, A->d_irp, A->d_ja, A->d_tile_row, A->d_tile_nrows, hash, ncls, d_ch,
                                                        d_rep, d_valid, (const SpmvClassDesc*)A->d_cls_desc, A->d_tile_class, d_nin);
    unsigned long long h_nin = 0;
    FB2_CUDA(cudaMemcpyAsync(&h_nin, d_nin, sizeof(h_nin), cudaMemcpyDeviceToHost, st));
    FB2_CUDA(cudaStreamSynchronize(st));
    if (cudaGetLastError() != cudaSuccess) rc = 1;
    A->spmv_ncls = ncls; A->spmv_tiles_in_class = (i64)h_nin;
    if (rc == 0 && h_nin * 4 >= (unsigned long long)nt) A->spmv_cls = 1;   // worth the extra branch: a quarter of the tiles or more
    // Private descriptors for the tiles without class mates (next to sub-domain interfaces and mesh boundaries their columns
    // still form a few segments of x): the same index-free consumer path, 10 B per non-zero instead of 12 B + gathers.
    // Only when classes are in use and the rest is small (<= a quarter of the tiles, <= 1 GiB of tables).
    A->spmv_tiles_private = 0;
    const i64 n_rest = nt - (i64)h_nin;
    static const int use_private = getenv("FB2_SPMV_PRIVATE") ? atoi(getenv("FB2_SPMV_PRIVATE")) : 1;
    if (A->spmv_cls == 1 && use_private && n_rest > 0 && n_rest * 4 <= nt &&
        (size_t)n_rest * ((size_t)A->spmv_cap * 2 + sizeof(SpmvClassDesc)) <= ((size_t)1 << 30)) {
      int *d_list, *d_cnt, *d_valid2;
      unsigned long long* d_used;
      FB2_TRY(fb2_dalloc(&d_list, (size_t)n_rest)); FB2_TRY(fb2_dalloc(&d_cnt, 1)); FB2_TRY(fb2_dalloc(&d_valid2, (size_t)n_rest));
      FB2_TRY(fb2_dalloc(&d_used, 1));
      FB2_CUDA(cudaMemsetAsync(d_cnt, 0, sizeof(int), st));
      FB2_CUDA(cudaMemsetAsync(d_used, 0, sizeof(unsigned long long), st));
      k_list_unclassified<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(nt, A->d_tile_class, d_list, d_cnt, (int)n_rest);
      // grow the descriptor and index tables: classes first, private tiles behind them
      void* desc2; unsigned short* idx2;
      FB2_TRY(fb2_dalloc((char**)&desc2, ((size_t)ncls + n_rest) * sizeof(SpmvClassDesc)));
      FB2_TRY(fb2_dalloc(&idx2, ((size_t)ncls + n_rest) * A->spmv_cap + 8));
      FB2_CUDA(cudaMemcpyAsync(desc2, A->d_cls_desc, (size_t)ncls * sizeof(SpmvClassDesc), cudaMemcpyDeviceToDevice, st));
      FB2_CUDA(cudaMemcpyAsync(idx2, A->d_cls_idx, (size_t)ncls * A->spmv_cap * sizeof(unsigned short), cudaMemcpyDeviceToDevice, st));
      FB2_TRY(fb2_dalloc(&A->d_tile_priv, (size_t)nt));
      FB2_CUDA(cudaMemsetAsync(A->d_tile_priv, 0xff, sizeof(int) * (size_t)nt, st));
      SpmvClassDesc* pdesc = (SpmvClassDesc*)desc2 + ncls;
      k_class_build<<<(unsigned)n_rest, 256, 0, st>>>(A->d_irp, A->d_ja, A->d_tile_row, A->d_tile_nrows, d_list, A->spmv_cap, idx_extra, x_align,
                                                     pdesc, idx2 + (size_t)ncls * A->spmv_cap, d_valid2);
      k_private_assign<<<(unsigned)((n_rest + 255) / 256), 256, 0, st>>>((int)n_rest, d_list, d_valid2, pdesc, A->d_tile_row, A->n, ncls,
                                                                          A->d_tile_class, A->d_tile_priv, d_used);
      unsigned long long h_used = 0;
      FB2_CUDA(cudaMemcpyAsync(&h_used, d_used, sizeof(h_used), cudaMemcpyDeviceToHost, st));
      FB2_CUDA(cudaStreamSynchronize(st));
      if (cudaGetLastError() != cudaSuccess) rc = 1;
      cudaFree(A->d_cls_desc); cudaFree(A->d_cls_idx);
      A->d_cls_desc = desc2; A->d_cls_idx = idx2;
      A->spmv_tiles_private = (i64)h_used;
      cudaFree(d_list); cudaFree(d_cnt); cudaFree(d_valid2); cudaFree(d_used);
    }
    cudaFree(d_ch); cudaFree(d_rep); cudaFree(d_valid); cudaFree(d_nin);
  }
  cudaFree(hash); cudaFree(hash_s); cudaFree(uniq); cudaFree(tid); cudaFree(tid_s); cudaFree(counts); cudaFree(d_nruns);
  cudaEventRecord(e1, st);
  FB2_CUDA(cudaStreamSynchronize(st));
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  fb2_record_phase("spmv_classes_ms", ms);
  fb2_record_phase("spmv_classes", (double)A->spmv_ncls);
  fb2_record_phase("spmv_tiles_in_class", (double)A->spmv_tiles_in_class);
  fb2_record_phase("spmv_tiles_private", (double)A->spmv_tiles_private);
  fb2_record_phase("spmv_signatures", (double)nruns);
  if (rc) { fb2_set_error("SpMV class construction failed"); return 1; }
  return 0;
}

// lane-group iterations of the consumer warps over the fixed tiles for the two row maps: a warp advances its G rows together,
// so a tile costs sum over warps of max over the warp's rows of ceil(len / TPR).  [0] natural (warp w: rows w G .. w G + G - 1),
// [1] transposed (rows w, w + W, ...)
__global__ void k_row_map_cost(i64 n, const i64* __restrict__ irp, int R, int TPR, int W, int G, i64 nfixed,
                               unsigned long long* cost) {
  const i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long c0 = 0, c1 = 0;
  if (t < nfixed) {
    const i64 r0 = t * R;
    const int NP = R / (W * G);
    for (int p = 0; p < NP; ++p)
      for (int w = 0; w < W; ++w) {
        int m0 = 0, m1 = 0;
        for (int g = 0; g < G; ++g) {
          const i64 a = r0 + (p * W + w) * G + g, b = r0 + g * (R / G) + p * W + w;
          if (a < n) m0 = max(m0, (int)((irp[a + 1] - irp[a] + TPR - 1) / TPR));
          if (b < n) m1 = max(m1, (int)((irp[b + 1] - irp[b] + TPR - 1) / TPR));
        }
        c0 += m0; c1 += m1;
      }
  }
  for (int o = 16; o > 0; o >>= 1) { c0 += __shfl_down_sync(0xffffffffu, c0, o); c1 += __shfl_down_sync(0xffffffffu, c1, o); }
  if ((threadIdx.x & 31) == 0) { atomicAdd(&cost[0], c0); atomicAdd(&cost[1], c1); }
}

// choose the tile geometry once per matrix
int fb2_spmv_prepare(fb2_matrix_s* A) {
  if (A->spmv_cap > 0 || A->n == 0) return 0;
  cudaStream_t st = fb2_stream();
  // Geometry from the B200 sweep (profiles/sweep_spmv_r1.txt): 2 stages of <= 28 KB keep 4 CTAs resident per SM, which beat
  // deeper pipelines.  R = the power of two whose average tile fills ~90 % of a stage; tiles that overflow are split.
  int cap = 2304, STAGES = 2;
  const double avg_row = (double)A->nnz / (double)A->n;
  int R = 64;
  while (R > 8 && R * avg_row > 0.9 * cap) R >>= 1;
  if (getenv("FB2_SPMV_R")) R = atoi(getenv("FB2_SPMV_R"));
  if (getenv("FB2_SPMV_STAGES")) STAGES = atoi(getenv("FB2_SPMV_STAGES"));
  if (getenv("FB2_SPMV_CAP")) cap = atoi(getenv("FB2_SPMV_CAP"));
  FB2_CHECK(R >= 8 && R <= 64 && (R & (R - 1)) == 0 && STAGES >= 2 && STAGES <= SPMV_MAX_STAGES && cap % 8 == 0, "bad SpMV tile parameters");
  const i64 nfixed = (A->n + R - 1) / R;
  // consumer threads: 8 lanes per row measured best (profiles/r2_spmv_history.txt): 256 consumer threads for 32-row tiles, 128 for
  // <= 16-row tiles (Q3: 13.5 vs 14.1 ms; 16 lanes per row 18.8-19.2 ms, 4 lanes 17.5-18.7 ms at config 2)
  int nc = R <= 16 ? 128 : 256;
  if (getenv("FB2_SPMV_CONSUMERS")) nc = atoi(getenv("FB2_SPMV_CONSUMERS"));
  if (g_compress) nc = 256;   // the dictionary variant stages two entries per consumer thread
  FB2_CHECK(nc == 128 || nc == 256, "bad FB2_SPMV_CONSUMERS");
  A->spmv_nc = nc;
  // row map: transposed when it saves >= 10 % of the lane-group iterations (rows of equal length meet in one warp);
  // FB2_SPMV_TRANSPOSE=0/1 overrides.  Needs 8 lanes per row and whole passes of the lane groups over the tile.
  {
    const int tpr = (R <= 16 ? 128 : 256) / R, W = nc / 32, G = 32 / tpr;
    int tr = 0;
    if (tpr == 8 && R % (W * G) == 0 && !g_compress) {
      unsigned long long *d_cost, h_cost[2] = {0, 0};
      FB2_TRY(fb2_dalloc(&d_cost, 2));
      FB2_CUDA(cudaMemsetAsync(d_cost, 0, 2 * sizeof(unsigned long long), st));
      k_row_map_cost<<<(unsigned)((nfixed + 255) / 256), 256, 0, st>>>(A->n, A->d_irp, R, tpr, W, G, nfixed, d_cost);
      FB2_CUDA(cudaMemcpyAsync(h_cost, d_cost, sizeof(h_cost), cudaMemcpyDeviceToHost, st));
      FB2_CUDA(cudaStreamSynchronize(st));
      cudaFree(d_cost);
      tr = (double)h_cost[1] < 0.9 * (double)h_cost[0] ? 1 : 0;
      if (getenv("FB2_SPMV_TRANSPOSE")) tr = atoi(getenv("FB2_SPMV_TRANSPOSE")) ? 1 : 0;
      fb2_record_phase("spmv_row_map_cost_natural", (double)h_cost[0]);
      fb2_record_phase("spmv_row_map_cost_transposed", (double)h_cost[1]);
    }
    A->spmv_tr = tr;
    fb2_record_phase("spmv_transposed_rows", tr);
  }
  // transposed rows: the chunks are staged up to 3 x 8 entries behind their place and aligned to 8 entries
  const int cap_fit = A->spmv_tr ? cap - 32 : cap;
  int *pieces, *off, *d_flag, h_flag = 0, total = 0;
  FB2_TRY(fb2_dalloc(&pieces, (size_t)nfixed + 1)); FB2_TRY(fb2_dalloc(&off, (size_t)nfixed + 1)); FB2_TRY(fb2_dalloc(&d_flag, 1));
  FB2_CUDA(cudaMemsetAsync(d_flag, 0, sizeof(int), st));
  k_tile_pieces<<<(unsigned)((nfixed + 256) / 256), 256, 0, st>>>(A->n, A->d_irp, R, cap_fit, nfixed, pieces, d_flag);
  {
    void* tmp = nullptr; size_t bytes = 0;
    FB2_CUDA(cub::DeviceScan::ExclusiveSum(tmp, bytes, pieces, off, nfixed + 1, st));
    FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
    FB2_CUDA(cub::DeviceScan::ExclusiveSum(tmp, bytes, pieces, off, nfixed + 1, st));
    FB2_CUDA(cudaStreamSynchronize(st));
    cudaFree(tmp);
  }
  FB2_CUDA(fb2_memcpy(&total, off + nfixed, sizeof(int), cudaMemcpyDeviceToHost));
  FB2_CUDA(fb2_memcpy(&h_flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost));
  FB2_CHECK(h_flag == 0, "a pair of rows exceeds the SpMV stage buffer: use fb2_set_spmv_variant(1)");
  FB2_TRY(fb2_dalloc(&A->d_tile_row, (size_t)total)); FB2_TRY(fb2_dalloc(&A->d_tile_nrows, (size_t)total));
  k_tile_fill<<<(unsigned)((nfixed + 255) / 256), 256, 0, st>>>(A->n, A->d_irp, R, cap_fit, nfixed, off, A->d_tile_row, A->d_tile_nrows);
  FB2_CUDA(cudaStreamSynchronize(st));
  cudaFree(pieces); cudaFree(off); cudaFree(d_flag);
  const i64 ntiles = total;
  const int capbuf = cap + 8;
  size_t stage = (size_t)capbuf * 12 + (size_t)(R + 2) * 8;
  size_t smem = stage * STAGES;
  FB2_CHECK(smem <= 200 * 1024, "SpMV tile does not fit in shared memory");
  A->spmv_R = R; A->spmv_cap = capbuf; A->spmv_ntiles = ntiles; A->spmv_smem = smem; A->spmv_stages = STAGES;
  // the attribute is per kernel, not per matrix (two matrices with different R must not lower each other's limit)
  FB2_CUDA(cudaFuncSetAttribute(k_spmv_stream<true, false, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  FB2_CUDA(cudaFuncSetAttribute(k_spmv_stream<false, false, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  FB2_CUDA(cudaFuncSetAttribute(k_spmv_stream<true, false, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  FB2_CUDA(cudaFuncSetAttribute(k_spmv_stream<false, false, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  int bps = 0;
  if (nc == 128) FB2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_spmv_stream<true, false, 128>, nc + 32, smem));
  else FB2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_spmv_stream<true, false, 256>, nc + 32, smem));
  if (bps < 1) bps = 1;
  i64 grid = (i64)fb2_num_sms() * bps;
  if (grid > ntiles) grid = ntiles;
  if (grid > FB2_MAX_PARTIALS) grid = FB2_MAX_PARTIALS;
  A->spmv_grid = (int)grid;
  fb2_record_phase("spmv_R", R); fb2_record_phase("spmv_tiles", (double)ntiles); fb2_record_phase("spmv_ctas_per_sm", bps);
  FB2_TRY(spmv_compress(A, capbuf, R, STAGES));
  if (A->spmv_comp != 1) FB2_TRY(spmv_build_classes(A));
  return 0;
}

// drops the tile table and the pattern classes: the next apply prepares again (with the FB2_SPMV_* environment of that
// moment).  Used by the tuning sweeps under profiles/.
extern "C" int fb2_csr_spmv_reset(fb2_matrix_t A) {
  FB2_CHECK(A, "null matrix");
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  cudaFree(A->d_tile_row); cudaFree(A->d_tile_nrows); cudaFree(A->d_tile_class); cudaFree(A->d_cls_desc); cudaFree(A->d_cls_idx); cudaFree(A->d_tile_priv);
  cudaFree(A->d_idx16); cudaFree(A->d_dict); cudaFree(A->d_dict_ptr);
  A->d_tile_row = nullptr; A->d_tile_nrows = nullptr; A->d_tile_class = nullptr; A->d_cls_desc = nullptr; A->d_cls_idx = nullptr; A->d_tile_priv = nullptr;
  A->d_idx16 = nullptr; A->d_dict = nullptr; A->d_dict_ptr = nullptr;
  A->spmv_R = A->spmv_cap = A->spmv_grid = A->spmv_stages = 0; A->spmv_ntiles = 0; A->spmv_smem = 0;
  A->spmv_nc = 0; A->spmv_tr = 0; A->spmv_comp = 0; A->spmv_cls = 0; A->spmv_ncls = 0; A->spmv_tiles_in_class = 0; A->spmv_tiles_private = 0; A->spmv_t_itf = -1; A->spmv_row_itf = -1;
  return 0;
}

int fb2_spmv_stream_range(fb2_matrix_s* A, const double* x, double* y, bool fuse_dot, const int* done_flag, i64 t0, i64 t1,
                          int chunk, double* partials, int* grid_out) {
  FB2_TRY(fb2_spmv_prepare(A));
  const i64 nt = t1 - t0;
  if (A->n == 0 || nt <= 0) { if (grid_out) *grid_out = 0; return 0; }
  cudaStream_t st = fb2_stream();
  if (chunk && (nt + chunk - 1) / chunk > FB2_MAX_PARTIALS / 2) chunk = (int)((nt + FB2_MAX_PARTIALS / 2 - 1) / (FB2_MAX_PARTIALS / 2));
  i64 grid = chunk ? (nt + chunk - 1) / chunk : (nt < A->spmv_grid ? nt : (i64)A->spmv_grid);
  // class tiles bulk-copy segments of x: they need a 16-byte aligned x (always true for library vectors, not for views)
  const unsigned char* tcls = (A->spmv_cls == 1 && (((uintptr_t)x) & 15) == 0) ? A->d_tile_class + t0 : nullptr;
  // L2 eviction hints of the bulk copies: bit 0 = val / ja streams evict_first, bit 1 = x segments evict_last (FB2_SPMV_L2HINT)
  const int l2hint = getenv("FB2_SPMV_L2HINT") ? atoi(getenv("FB2_SPMV_L2HINT")) : SPMV_L2HINT_DEFAULT;
  // suspend-time hint (ns) of the consumers' wait for a full stage; 0 = plain try_wait spin
  const uint32_t wait_hint = getenv("FB2_SPMV_WAIT_NS") ? (uint32_t)atoi(getenv("FB2_SPMV_WAIT_NS")) : SPMV_WAIT_NS_DEFAULT;
#define SPMV_LAUNCH_NC(FD, CP, NCT, JA, SMEM, DICT, DPTR, DCAP)                                                                   \
  k_spmv_stream<FD, CP, NCT><<<(unsigned)grid, NCT + 32, SMEM, st>>>(A->n, A->d_irp, JA, A->d_val, x, y, FD ? partials : nullptr, \
                                                                    done_flag, A->spmv_R, A->spmv_cap, nt, A->spmv_stages,   \
                                                                    A->d_tile_row + t0, A->d_tile_nrows + t0, chunk, DICT, DPTR, DCAP, \
                                                                    CP ? nullptr : tcls, (const SpmvClassDesc*)A->d_cls_desc, A->d_cls_idx, CP ? 0 : A->spmv_tr, wait_hint, A->d_tile_priv ? A->d_tile_priv + t0 : nullptr, l2hint)
#define SPMV_LAUNCH(FD, CP, JA, SMEM, DICT, DPTR, DCAP)                                                   \
  do {                                                                                                    \
    if (!CP && A->spmv_nc == 128) SPMV_LAUNCH_NC(FD, false, 128, JA, SMEM, DICT, DPTR, DCAP);             \
    else SPMV_LAUNCH_NC(FD, CP, 256, JA, SMEM, DICT, DPTR, DCAP);                                         \
  } while (0)
  if (A->spmv_comp == 1) {
    if (fuse_dot) SPMV_LAUNCH(true, true, A->d_idx16, A->spmv_smem_comp, A->d_dict, A->d_dict_ptr + t0, A->spmv_dcap);
    else SPMV_LAUNCH(false, true, A->d_idx16, A->spmv_smem_comp, A->d_dict, A->d_dict_ptr + t0, A->spmv_dcap);
  } else {
    if (fuse_dot) SPMV_LAUNCH(true, false, A->d_ja, A->spmv_smem, nullptr, nullptr, 0);
    else SPMV_LAUNCH(false, false, A->d_ja, A->spmv_smem, nullptr, nullptr, 0);
  }
#undef SPMV_LAUNCH
#undef SPMV_LAUNCH_NC
  FB2_CUDA(cudaGetLastError());
  if (grid_out) *grid_out = (int)grid;
  return 0;
}

int fb2_spmv_stream(fb2_matrix_s* A, const double* x, double* y, bool fuse_dot, const int* done_flag, int* grid_out) {
  FB2_TRY(fb2_spmv_prepare(A));
  // short-lived CTAs over 32 consecutive tiles beat persistent striding CTAs (dynamic balance, contiguous streams per CTA):
  // profiles/r1_spmv_chunk_sweep.txt
  const int chunk = getenv("FB2_SPMV_CHUNK") ? atoi(getenv("FB2_SPMV_CHUNK")) : 32;
  return fb2_spmv_stream_range(A, x, y, fuse_dot, done_flag, 0, A->spmv_ntiles, chunk, fb2_scratch_partials(), grid_out);
}

// par overlap: the tile table is in row order and interface rows are numbered last (dof_import.f90:391-421), so the tiles
// from the one holding row n_interior onwards are the "interface" launch
int fb2_spmv_interface_split(fb2_matrix_s* A, i64 n_interior, i64* t_itf, i64* row_itf) {
  FB2_TRY(fb2_spmv_prepare(A));
  if (A->spmv_t_itf < 0) {
    std::vector<i64> rows((size_t)A->spmv_ntiles);
    if (A->spmv_ntiles) FB2_CUDA(fb2_memcpy(rows.data(), A->d_tile_row, sizeof(i64) * rows.size(), cudaMemcpyDeviceToHost));
    // last tile whose first row is <= n_interior (it holds row n_interior unless n_interior == n)
    i64 lo = 0, hi = A->spmv_ntiles;
    while (lo < hi) { i64 mid = (lo + hi) >> 1; if (rows[mid] <= n_interior) lo = mid + 1; else hi = mid; }
    i64 t = lo > 0 ? lo - 1 : 0;
    if (n_interior >= A->n) t = A->spmv_ntiles;
    A->spmv_t_itf = t;
    A->spmv_row_itf = t < A->spmv_ntiles ? rows[t] : A->n;
  }
  *t_itf = A->spmv_t_itf; *row_itf = A->spmv_row_itf;
  return 0;
}
