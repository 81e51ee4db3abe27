// Row-owner ("gather") assembly for the scalar Laplacian on structured parts, sm_100a.
//
// The scatter kernel (fb2_assemble_fast.cu) adds every element-matrix entry into CSR with an L2 atomic: 729 RED.ADD.F64 per Q2
// cell on top of a memset of the whole value array, 3.3x the algorithmic DRAM traffic (profiles/r1_final_ncu_asm_q2_128.txt).
// Here every CSR row is WRITTEN ONCE, completely, by plain coalesced stores, with no memset and no atomics:
//
//  * FEMPAR numbers the DOFs a cell creates consecutively (sbm_serial_fe_space.i90:2623-2875: own-cell DOFs first, then the
//    n-faces the cell is the first to touch).  Away from the boundary a cell creates the P^3 nodes with local index 1..P per
//    direction, so the rows of these nodes are ONE contiguous chunk of val (Q2: 8 rows, 512 entries, 4 KB).
//  * A warp owns such a cell ("owner cell").  The rows of its nodes receive contributions from the (P+1)^3 pairs
//    (cell c+d, local row i'), d in {0,1}^3: exactly one pair per lane (Q2: 27).  Lane u evaluates row i' of the element
//    matrix of cell c+d by sum factorisation (same arithmetic as the scatter kernel, fb2_sumfact.cuh).
//  * Rows of one node are spread over up to 8 lanes whose column sets overlap on the planes through the node: three
//    shuffle passes (one per direction) hand the overlapping plane of the d_k = 1 lane to its d_k = 0 partner.  After that
//    every chunk entry lives in exactly one lane, is placed through a per-class table into a shared-memory image of the chunk
//    and the image is streamed out with coalesced 256-byte stores.  Deterministic: no atomics anywhere.
//  * The geometry factors D(q) = w detJ J^-1 J^-T of a cell are needed by 8 owner cells.  A CTA therefore walks a pencil of
//    4 x 4 owner-cell columns along x and keeps the D of two x-slabs of 5 x 5 cells in shared memory (each cell's geometry is
//    evaluated 25/16 times instead of 8).
//  * Which cells qualify is not assumed but VERIFIED on the device against irp/ja/lst_dofs_gids for every cell (k_rows_verify),
//    as for the scatter map: a cell that fails (boundary layers, sub-domain interfaces after the interior-then-interface
//    renumbering) leaves its rows to the scatter kernel, which then visits only the cells touching such rows and skips
//    the rows owned here (FastArgs::cell_list / row_fast).
// Replaces, like the scatter kernel, the triple loop of the user plugin + fe_cell_iterator%assembly
// (tutorial_01_..._modules.f90:114-128, sbm_fe_cell_iterator.i90:1610-1637, base_sparse_matrix.f90:3469-3513).
#include <cub/cub.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <stdlib.h>
#include "fb2_sumfact.cuh"

namespace {

constexpr int TY = 4, TZ = 4;            // owner-cell columns per CTA (pencil cross-section)
constexpr int GY = TY + 1, GZ = TZ + 1;  // geometry cells per x-slab

__host__ __device__ constexpr int rows_chunk(int P) {   // entries of the chunk: per direction P-1 nodes with P+1 columns, 1 with 2P+1
  int w = (P - 1) * (P + 1) + (2 * P + 1);
  return w * w * w;
}

struct RowsArgs {
  CellGeom g;
  const int* perm;            // total-box lex -> local cell id (nullptr: identity)
  const i64* chunk_base;      // [n_local] irp of the first row of the owner cell's chunk, -1: not handled here
  const int* gids;            // lst_dofs_gids (source term only: row ids of the chunk)
  const unsigned short* plan; // [NN][NN] chunk-relative slot of (unit, column), 0xFFFF: handed to the partner lane
  const double *qw, *gphi, *gdphi;
  const double* src_qp;       // user samples f(x_q) [local cell][q] or nullptr
  double* val;
  double* rhs;
  FuncPrm fn;
  int nseg, seg_len;          // x-segments per pencil, owner cells per segment
  int npy, npz;               // pencils per direction
  unsigned char stage_swz[32];     // chunk-image swizzle, see h_stage_swz (zeros = identity, FB2_ASM_SWIZZLE=0)
  unsigned char row_of_node[27];   // chunk row (0..P^3-1) of the owner cell's node (a1-1) + P (a2-1) + P^2 (a3-1)
};

// unit type per direction, t in [0, P]: t < P: node a = t + 1 of the owner cell, row i' = a of the owner cell itself (d = 0);
// t = P: node a = P seen from the next cell (d = 1) where it is local node 0
template <int P> __device__ __forceinline__ int unit_d(int t) { return t == P ? 1 : 0; }
template <int P> __device__ __forceinline__ int unit_i(int t) { return t == P ? 0 : t + 1; }
template <int P> __device__ __forceinline__ int unit_a(int t) { return t == P ? P : t + 1; }

// ------------------------------------------------------------------------------------------
// the gather kernel (Q2: 27 units = 27 lanes, one owner cell per warp pass)
// ------------------------------------------------------------------------------------------
// Chunk image in shared memory: slot s lives at s ^ stage_swz[s >> 4], a permutation inside every aligned group of 16 doubles
// (= the 16 eight-byte banks; table passed in RowsArgs::stage_swz), so the coalesced read-out of 32 consecutive slots stays conflict-free whatever the table holds.
// The table was searched offline against the scatter of the interior Q2 plan (27 lanes store column j of their row: 112
// wavefronts per owner cell with the identity, 76 with this table, 51 if conflict-free; ncu had 19 % of the kernel's
// shared-memory wavefronts on that store).  Any table is correct; this one is only tuned for the interior plan.
static const unsigned char h_stage_swz[32] = {4, 8, 1, 15, 0, 2, 13, 14, 1, 4, 4, 6, 4, 1, 3, 2,
                                              4, 8, 1, 9, 5, 12, 13, 12, 4, 9, 7, 9, 1, 10, 11, 0};

template <int P, bool SRC, bool ZSKIP>
__global__ void __launch_bounds__(256, 2) k_assemble_rows(RowsArgs R) {
  static_assert(P == 2, "row-owner kernel: Q2 (27 units per owner cell = one warp pass)");
  constexpr int N1 = P + 1, NN = N1 * N1 * N1, NQ = NN, CH = rows_chunk(P), WARPS = 8;
  // D(q) of a geometry cell: NQ x 6 doubles.  Strides (in doubles) chosen so that the 8 cells a warp reads at one q fall into
  // disjoint bank groups for 16-byte loads: z-neighbour 4, y-neighbour 8, x-slab 16 (mod 32 words)
  constexpr int SC = NQ * 6;                                   // 162 doubles = 324 words = 4 (mod 32)
  constexpr int SB = GZ * SC + 10;                             // 820 doubles = 1640 words = 8 (mod 32)
  constexpr int SA = GY * SB + 4;                              // 4104 doubles = 8208 words = 16 (mod 32)
  static_assert((2 * SC) % 32 == 4 && (2 * SB) % 32 == 8 && (2 * SA) % 32 == 16, "bank layout");
  extern __shared__ __align__(16) double smem[];
  double* s_D = smem;                                          // [2][SA]
  double* s_stage = s_D + 2 * SA;                              // [WARPS][CH]
  double* s_fw = s_stage + WARPS * CH;                         // [2][GY*GZ][NQ] (SRC only)
  // strides 26 / 10 doubles (= 20 words mod 32): lanes are consecutive quadrature points and read 16 bytes each, so the 8
  // lanes of a quarter warp cover the 32 banks once; with strides of 24 / 8 doubles they met on two bank groups (ncu: 15.8
  // wavefronts per LDS.128 in jacobian_at, 35 % of all shared-memory wavefronts of the kernel)
  double* s_gd = s_fw + (SRC ? 2 * GY * GZ * NQ : 0);          // [NQ][26]
  double* s_gp = s_gd + NQ * 26;                               // [NQ][10] (SRC only, analytic f)
  double* s_qw = s_gp + (SRC ? NQ * 10 : 0);                    // [NQ]
  double* s_B = s_qw + NQ;                                     // [N1*N1]
  double* s_C = s_B + N1 * N1;
  unsigned short* s_plan = (unsigned short*)(s_C + N1 * N1);   // [NN columns][32 lanes]: swizzled image position of (unit, column)
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  for (int e = tid; e < NQ * 24; e += blockDim.x) s_gd[(e / 24) * 26 + e % 24] = R.gdphi[e];
  if (SRC) for (int e = tid; e < NQ * 8; e += blockDim.x) s_gp[(e / 8) * 10 + e % 8] = R.gphi[e];
  for (int e = tid; e < NQ; e += blockDim.x) s_qw[e] = R.qw[e];
  if (tid < N1 * N1) { s_B[tid] = c_B1[P - 1][tid]; s_C[tid] = c_C1[P - 1][tid]; }
  static_assert(CH <= 512, "stage_swz covers 32 groups of 16 slots");
  for (int e = tid; e < NN * NN; e += blockDim.x) {   // transposed: the lanes of a warp read consecutive uint16 (one wavefront)
    const unsigned short slot = R.plan[e];
    s_plan[(e % NN) * 32 + e / NN] = slot == 0xFFFF ? slot : (unsigned short)(slot ^ R.stage_swz[slot >> 4]);
  }

  // pencil and x-segment of this CTA (local-box cell coordinates)
  int b = blockIdx.x;
  const int seg = b % R.nseg; b /= R.nseg;
  const int py = b % R.npy, pz = b / R.npy;
  const int y0 = py * TY, z0 = pz * TZ;
  const int x_begin = seg * R.seg_len, x_end = min(x_begin + R.seg_len, R.g.nloc[0]);
  double dN[3], rN[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) { dN[k] = (double)R.g.N[k]; rN[k] = 1.0 / dN[k]; }

  // geometry of one x-slab: thread <-> (cell (gy, gz) of the slab, q)
  auto slab_geometry = [&](int xs) {
    double* Dslab = s_D + (xs & 1) * SA;
    for (int item = tid; item < GY * GZ * NQ; item += blockDim.x) {
      const int q = item % NQ, cell = item / NQ, gy = cell % GY, gz = cell / GY;
      const int cl[3] = {xs, y0 + gy, z0 + gz};
      if (cl[0] >= R.g.nloc[0] || cl[1] >= R.g.nloc[1] || cl[2] >= R.g.nloc[2]) continue;   // never read by a verified owner cell
      // vertex coordinates (first + ijk + {0,1}) / N (uniform_hex_mesh_generator.f90:608-635), correctly rounded quotient
      // through the reciprocal and one residual correction (as in the scatter kernel)
      double lohi[3][2];
#pragma unroll
      for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int bb = 0; bb < 2; ++bb) {
          const double a = (double)(R.g.first[k] + R.g.L[k] + cl[k] + bb);
          const double q0 = a * rN[k];
          lohi[k][bb] = fma(fma(-q0, dN[k], a), rN[k], q0);
        }
      double X[8][3];
#pragma unroll
      for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int k = 0; k < 3; ++k) X[a][k] = lohi[k][(a >> k) & 1];
      double inv[3][3], det;
      jacobian_at(X, s_gd + q * 26, inv, det);
      const double wd = det * s_qw[q];
      double* D = Dslab + gy * SB + gz * SC + q * 6;
      D[0] = wd * (inv[0][0] * inv[0][0] + inv[1][0] * inv[1][0] + inv[2][0] * inv[2][0]);
      D[1] = wd * (inv[0][0] * inv[0][1] + inv[1][0] * inv[1][1] + inv[2][0] * inv[2][1]);
      D[2] = wd * (inv[0][0] * inv[0][2] + inv[1][0] * inv[1][2] + inv[2][0] * inv[2][2]);
      D[3] = wd * (inv[0][1] * inv[0][1] + inv[1][1] * inv[1][1] + inv[2][1] * inv[2][1]);
      D[4] = wd * (inv[0][1] * inv[0][2] + inv[1][1] * inv[1][2] + inv[2][1] * inv[2][2]);
      D[5] = wd * (inv[0][2] * inv[0][2] + inv[1][2] * inv[1][2] + inv[2][2] * inv[2][2]);
      if (SRC) {
        double f;
        if (R.src_qp) {   // user samples, indexed by the local cell id
          const int lex = (cl[0] + R.g.L[0]) + R.g.ntot[0] * ((cl[1] + R.g.L[1]) + R.g.ntot[1] * (cl[2] + R.g.L[2]));
          const int c = R.perm ? R.perm[lex] : lex;
          f = R.src_qp[(i64)c * NQ + q];
        } else {
          double xq[3] = {0, 0, 0};
#pragma unroll
          for (int a = 0; a < 8; ++a)
#pragma unroll
            for (int k = 0; k < 3; ++k) xq[k] += s_gp[q * 10 + a] * X[a][k];
          f = func_f(R.fn, xq);
        }
        s_fw[((xs & 1) * GY * GZ + cell) * NQ + q] = wd * f;
      }
    }
  };

  // lane <-> unit
  const int t0 = lane % N1, t1 = (lane / N1) % N1, t2 = lane / (N1 * N1);
  const bool unit_on = lane < NN;
  const int d0 = unit_d<P>(t0), d1 = unit_d<P>(t1), d2 = unit_on ? unit_d<P>(t2) : 0;
  const int irow = unit_i<P>(t0) + N1 * (unit_i<P>(t1) + N1 * (unit_on ? unit_i<P>(t2) : 0));   // row of the element matrix
  const int node = (unit_a<P>(t0) - 1) + P * ((unit_a<P>(t1) - 1) + P * ((unit_on ? unit_a<P>(t2) : 1) - 1));
  const bool rcv0 = unit_on && t0 == P - 1, rcv1 = unit_on && t1 == P - 1, rcv2 = unit_on && t2 == P - 1;
  const bool owner_lane = unit_on && d0 == 0 && d1 == 0 && d2 == 0;
  double* stage = s_stage + w * CH;
  const unsigned short* myplan = s_plan + (unit_on ? lane : 0);

  __syncthreads();
  slab_geometry(x_begin);
  for (int xs = x_begin; xs < x_end; ++xs) {
    // owner-cell data of this step: issued before the geometry so that the loads are back when the contraction starts
    i64 cbase[2];
    int crow0[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int idx = w * 2 + h, oy = idx % TY, oz = idx / TY;
      const int cl1 = y0 + oy, cl2 = z0 + oz;
      cbase[h] = -1; crow0[h] = 0;
      if (cl1 < R.g.nloc[1] && cl2 < R.g.nloc[2]) {
        const int lex = (xs + R.g.L[0]) + R.g.ntot[0] * ((cl1 + R.g.L[1]) + R.g.ntot[1] * (cl2 + R.g.L[2]));
        const int c = R.perm ? R.perm[lex] : lex;
        cbase[h] = R.chunk_base[c];
        if (SRC && cbase[h] >= 0) crow0[h] = R.gids[(i64)c * NN + (1 + N1 + N1 * N1)];   // first DOF the cell created
      }
    }
    slab_geometry(xs + 1);
    __syncthreads();
#pragma unroll 1
    for (int h = 0; h < 2; ++h) {
      if (cbase[h] < 0) continue;                        // warp-uniform
      const int idx = w * 2 + h, oy = idx % TY, oz = idx / TY;
      double acc[NN];
      double sacc = 0.0;
      if (unit_on) {
        const double* Dc = s_D + ((xs + d0) & 1) * SA + (oy + d1) * SB + (oz + d2) * SC;
        column_sumfact<P, ZSKIP>([&](int q, double g0, double g1, double g2, double& u0, double& u1, double& u2) {
          const double2 Da = *(const double2*)(Dc + q * 6), Db = *(const double2*)(Dc + q * 6 + 2),
                        Dd = *(const double2*)(Dc + q * 6 + 4);
          u0 = Da.x * g0 + Da.y * g1 + Db.x * g2;
          u1 = Da.y * g0 + Db.y * g1 + Dd.x * g2;
          u2 = Db.x * g0 + Dd.x * g1 + Dd.y * g2;
        }, s_B, s_C, irow, acc);
        if (SRC) {
          const double* fw = s_fw + ((((xs + d0) & 1) * GY * GZ) + (oy + d1) + GY * (oz + d2)) * NQ;
          const int i1 = irow % N1, i2 = (irow / N1) % N1, i3 = irow / (N1 * N1);
#pragma unroll 1
          for (int q = 0; q < NQ; ++q) {
            const int q1 = q % N1, q2 = (q / N1) % N1, q3 = q / (N1 * N1);
            sacc += fw[q] * (s_B[q1 * N1 + i1] * s_B[q2 * N1 + i2] * s_B[q3 * N1 + i3]);
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < NN; ++i) acc[i] = 0.0;
      }
      // overlap reduction: direction k, the d_k = 1 lane hands its plane j_k = 0 to the lane N1^k below (plane j_k = P)
#pragma unroll
      for (int jb = 0; jb < N1; ++jb)
#pragma unroll
        for (int ja = 0; ja < N1; ++ja) {
          const double v = __shfl_down_sync(0xffffffffu, acc[0 + N1 * (ja + N1 * jb)], 1);
          if (rcv0) acc[P + N1 * (ja + N1 * jb)] += v;
        }
#pragma unroll
      for (int jb = 0; jb < N1; ++jb)
#pragma unroll
        for (int ja = 0; ja < N1; ++ja) {
          const double v = __shfl_down_sync(0xffffffffu, acc[ja + N1 * (0 + N1 * jb)], N1);
          if (rcv1) acc[ja + N1 * (P + N1 * jb)] += v;
        }
#pragma unroll
      for (int jb = 0; jb < N1; ++jb)
#pragma unroll
        for (int ja = 0; ja < N1; ++ja) {
          const double v = __shfl_down_sync(0xffffffffu, acc[ja + N1 * (jb + N1 * 0)], N1 * N1);
          if (rcv2) acc[ja + N1 * (jb + N1 * P)] += v;
        }
      if (SRC) {
        double v = __shfl_down_sync(0xffffffffu, sacc, 1);
        if (rcv0) sacc += v;
        v = __shfl_down_sync(0xffffffffu, sacc, N1);
        if (rcv1) sacc += v;
        v = __shfl_down_sync(0xffffffffu, sacc, N1 * N1);
        if (rcv2) sacc += v;
        if (owner_lane) R.rhs[crow0[h] - 1 + R.row_of_node[node]] = sacc;
      }
      // place the live entries into the chunk image, then stream the image out
      if (unit_on) {
#pragma unroll
        for (int j = 0; j < NN; ++j) {
          const unsigned short slot = myplan[j * 32];
          if (slot != 0xFFFF) stage[slot] = acc[j];
        }
      }
      __syncwarp();
      double* dst = R.val + cbase[h];
#pragma unroll
      for (int e = 0; e < CH; e += 32) dst[e + lane] = stage[(e + lane) ^ R.stage_swz[(e + lane) >> 4]];
      __syncwarp();
    }
    __syncthreads();      // the next step's geometry overwrites the slab of this step
  }
}

// ------------------------------------------------------------------------------------------
// plan (from one deep-interior representative cell) and per-cell verification
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int local_cell_at(const CellGeom& g, const int* __restrict__ perm, const int* cl /*local-box*/) {
  const int lex = (cl[0] + g.L[0]) + g.ntot[0] * ((cl[1] + g.L[1]) + g.ntot[1] * (cl[2] + g.L[2]));
  return perm ? perm[lex] : lex;
}

// one warp: unit u = lane.  row_of_node[node] = chunk row of the owner cell's node; plan[u][j] = chunk-relative slot
template <int P>
__global__ void k_rows_plan(CellGeom g, const int* __restrict__ perm, int rep_lex_local /*local-box lexicographic index*/,
                            const int* __restrict__ gids, const i64* __restrict__ irp, const int* __restrict__ ja, i64 n_free,
                            unsigned short* plan, int* row_of_node, int* ok) {
  constexpr int N1 = P + 1, NN = N1 * N1 * N1, CH = rows_chunk(P), NR = P * P * P;
  const int lane = threadIdx.x;
  if (lane >= NN) return;
  const int cl[3] = {rep_lex_local % g.nloc[0], (rep_lex_local / g.nloc[0]) % g.nloc[1], rep_lex_local / (g.nloc[0] * g.nloc[1])};
  const int c = local_cell_at(g, perm, cl);
  const int g0 = gids[(i64)c * NN + (1 + N1 + N1 * N1)];
  const int t[3] = {lane % N1, (lane / N1) % N1, lane / (N1 * N1)};
  int cd[3], irow = 0, node = 0, mul = 1, mulp = 1;
  bool good = g0 > 0 && g0 + NR - 1 <= n_free;
  for (int k = 0; k < 3; ++k) {
    cd[k] = cl[k] + unit_d<P>(t[k]);
    good = good && cd[k] < g.nloc[k];
    irow += mul * unit_i<P>(t[k]); node += mulp * (unit_a<P>(t[k]) - 1);
    mul *= N1; mulp *= P;
  }
  if (good) {
    const int c2 = local_cell_at(g, perm, cd);
    const int* gl = gids + (i64)c2 * NN;
    const int grow = gl[irow], r = grow - g0;
    good = r >= 0 && r < NR;
    if (good) {
      if (unit_d<P>(t[0]) + unit_d<P>(t[1]) + unit_d<P>(t[2]) == 0) row_of_node[node] = r;
      const i64 cb = irp[g0 - 1];
      for (int j = 0; j < NN; ++j) {
        const int jk[3] = {j % N1, (j / N1) % N1, j / (N1 * N1)};
        bool live = true;
        for (int k = 0; k < 3; ++k) live = live && !(unit_d<P>(t[k]) == 1 && jk[k] == 0);
        unsigned short slot = 0xFFFF;
        if (live) {
          const int gc = gl[j];
          good = good && gc > 0 && gc <= n_free;
          if (good) {
            const i64 s = find_slot(ja, irp[grow - 1], irp[grow], gc);
            good = s < irp[grow] && ja[s] == gc && s - cb < CH;
            if (good) slot = (unsigned short)(s - cb);
          }
        }
        plan[lane * NN + j] = slot;
      }
    }
  }
  if (!good) atomicExch(ok, 0);
}

// warp per local cell: does the plan reproduce this cell's chunk?  chunk_base[c] = irp of its first row, or -1
template <int P>
__global__ void k_rows_verify(CellGeom g, const int* __restrict__ perm, int n_local, const int* __restrict__ gids,
                              const i64* __restrict__ irp, const int* __restrict__ ja, i64 n_free,
                              const unsigned short* __restrict__ plan, const int* __restrict__ row_of_node, i64* chunk_base,
                              unsigned* row_fast, unsigned long long* n_fast) {
  constexpr int N1 = P + 1, NN = N1 * N1 * N1, CH = rows_chunk(P), NR = P * P * P;
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t[3] = {lane % N1, (lane / N1) % N1, lane / (N1 * N1)};
  for (int li = blockIdx.x * warps + w; li < n_local; li += gridDim.x * warps) {
    const int cl[3] = {li % g.nloc[0], (li / g.nloc[0]) % g.nloc[1], li / (g.nloc[0] * g.nloc[1])};
    const int c = local_cell_at(g, perm, cl);
    const int g0 = gids[(i64)c * NN + (1 + N1 + N1 * N1)];
    bool good = g0 > 0 && (i64)g0 + NR - 1 <= n_free;
    i64 cb = -1;
    if (good) { cb = irp[g0 - 1]; good = irp[g0 - 1 + NR] - cb == CH; }
    if (good && lane < NN) {
      int cd[3], irow = 0, node = 0, mul = 1, mulp = 1;
      for (int k = 0; k < 3; ++k) {
        cd[k] = cl[k] + unit_d<P>(t[k]);
        good = good && cd[k] < g.nloc[k];
        irow += mul * unit_i<P>(t[k]); node += mulp * (unit_a<P>(t[k]) - 1);
        mul *= N1; mulp *= P;
      }
      if (good) {
        const int c2 = local_cell_at(g, perm, cd);
        const int* gl = gids + (i64)c2 * NN;
        const int grow = gl[irow];
        good = grow == g0 + row_of_node[node];
        if (good) {
          const i64 rb = irp[grow - 1], re = irp[grow];
          for (int j = 0; j < NN; ++j) {
            const unsigned short slot = plan[lane * NN + j];
            if (slot == 0xFFFF) continue;
            const int gc = gl[j];
            const i64 s = cb + slot;
            good = good && gc > 0 && gc <= n_free && s >= rb && s < re && ja[s] == gc;
          }
        }
      }
    }
    good = __all_sync(0xffffffffu, good);
    if (lane == 0) {
      chunk_base[c] = good ? cb : -1;
      if (good) {
        atomicAdd(n_fast, 1ULL);
        for (int r = 0; r < NR; ++r) atomicOr(&row_fast[(g0 - 1 + r) >> 5], 1u << ((g0 - 1 + r) & 31));
      }
    }
    __syncwarp();
  }
}

// flags for the compaction of the slow lists
__global__ void k_rows_flag_rows(i64 n, const unsigned* __restrict__ row_fast, unsigned char* flag) {
  i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) flag[r] = !((row_fast[r >> 5] >> (r & 31)) & 1u);
}
__global__ void k_rows_flag_cells(CellGeom g, const int* __restrict__ perm, int n_local, int nd, const int* __restrict__ gids,
                                  i64 n_free, const unsigned* __restrict__ row_fast, unsigned char* flag, int* lex_out) {
  int li = blockIdx.x * blockDim.x + threadIdx.x;
  if (li >= n_local) return;
  const int cl[3] = {li % g.nloc[0], (li / g.nloc[0]) % g.nloc[1], li / (g.nloc[0] * g.nloc[1])};
  const int lex = (cl[0] + g.L[0]) + g.ntot[0] * ((cl[1] + g.L[1]) + g.ntot[1] * (cl[2] + g.L[2]));
  const int c = perm ? perm[lex] : lex;
  bool slow = false;
  for (int l = 0; l < nd; ++l) {
    const int gi = gids[(i64)c * nd + l];
    if (gi > 0 && gi <= n_free && !((row_fast[(gi - 1) >> 5] >> ((gi - 1) & 31)) & 1u)) { slow = true; break; }
  }
  flag[li] = slow;
  lex_out[li] = lex;
}
// warp per slow row: zero its values
__global__ void k_rows_zero(i64 n_slow, const i64* __restrict__ rows, const i64* __restrict__ irp, double* val) {
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (i64 k = (i64)blockIdx.x * warps + w; k < n_slow; k += (i64)gridDim.x * warps) {
    const i64 r = rows[k], b = irp[r], e = irp[r + 1];
    for (i64 s = b + lane; s < e; s += 32) val[s] = 0.0;
  }
}

template <int P>
size_t rows_smem_bytes(bool src) {
  constexpr int N1 = P + 1, NN = N1 * N1 * N1, NQ = NN, CH = rows_chunk(P);
  constexpr int SC = NQ * 6, SB = GZ * SC + 10, SA = GY * SB + 4;
  size_t d = 2 * (size_t)SA + 8 * (size_t)CH + (src ? 2 * GY * GZ * NQ + NQ * 10 : 0) + NQ * 26 + NQ + 2 * N1 * N1;
  return d * sizeof(double) + (size_t)NN * 32 * sizeof(unsigned short) + 16;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
// set-up (paid once per matrix, like the scatter map): returns with A->rows_state = 1 (usable) or -1
int fb2_rows_prepare(fb2_matrix_s* A, int kind) {
  if (A->rows_state != 0) return 0;
  fb2_space_s* s = A->s;
  A->rows_state = -1;
  static const int off = getenv("FB2_ASM_ROWS") ? atoi(getenv("FB2_ASM_ROWS")) == 0 : 0;
  if (off || kind != FB2_KIND_POISSON || s->rt.p != 2 || s->rt.ncomp != 1 || A->sym || s->d_coords || !s->d_gids) return 0;
  const int n_local = s->m.n_local;
  for (int k = 0; k < 3; ++k) if (s->m.nloc[k] < 8) return 0;   // no deep-interior representative
  if (n_local == 0 || A->n < 64) return 0;
  constexpr int P = 2, NN = 27;
  cudaStream_t st = fb2_stream();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  CellGeom g = make_geom(s);
  int *d_ron, *d_ok;
  unsigned long long* d_nfast;
  FB2_TRY(fb2_dalloc(&A->d_rows_plan, (size_t)NN * NN));
  FB2_TRY(fb2_dalloc(&d_ron, 32)); FB2_TRY(fb2_dalloc(&d_ok, 1)); FB2_TRY(fb2_dalloc(&d_nfast, 1));
  int one = 1;
  FB2_CUDA(cudaMemcpyAsync(d_ok, &one, sizeof(int), cudaMemcpyHostToDevice, st));
  FB2_CUDA(cudaMemsetAsync(d_ron, 0, 32 * sizeof(int), st));
  FB2_CUDA(cudaMemsetAsync(d_nfast, 0, sizeof(unsigned long long), st));
  const int rc[3] = {s->m.nloc[0] / 2, s->m.nloc[1] / 2, s->m.nloc[2] / 2};     // a cell in the middle of the local box
  const int rep = rc[0] + s->m.nloc[0] * (rc[1] + s->m.nloc[1] * rc[2]);
  k_rows_plan<P><<<1, 32, 0, st>>>(g, s->d_perm, rep, s->d_gids, A->d_irp, A->d_ja, A->n, A->d_rows_plan, d_ron, d_ok);
  int h_ok = 0, h_ron[32];
  FB2_CUDA(cudaMemcpyAsync(&h_ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost, st));
  FB2_CUDA(cudaMemcpyAsync(h_ron, d_ron, sizeof(h_ron), cudaMemcpyDeviceToHost, st));
  FB2_CUDA(cudaStreamSynchronize(st));
  FB2_CUDA(cudaGetLastError());
  int rc_all = 0;
  if (h_ok) {
    for (int k = 0; k < P * P * P; ++k) A->rows_row_of_node[k] = (unsigned char)h_ron[k];
    const size_t words = (size_t)((A->n + 31) / 32) + 1;
    FB2_TRY(fb2_dalloc(&A->d_chunk_base, (size_t)n_local));
    FB2_TRY(fb2_dalloc(&A->d_row_fast, words));
    FB2_CUDA(cudaMemsetAsync(A->d_row_fast, 0, words * sizeof(unsigned), st));
    k_rows_verify<P><<<fb2_num_sms() * 8, 256, 0, st>>>(g, s->d_perm, n_local, s->d_gids, A->d_irp, A->d_ja, A->n, A->d_rows_plan,
                                                        d_ron, A->d_chunk_base, A->d_row_fast, d_nfast);
    unsigned long long h_nfast = 0;
    FB2_CUDA(cudaMemcpyAsync(&h_nfast, d_nfast, sizeof(h_nfast), cudaMemcpyDeviceToHost, st));
    FB2_CUDA(cudaStreamSynchronize(st));
    FB2_CUDA(cudaGetLastError());
    A->n_fast_cells = (i64)h_nfast;
    if (h_nfast * 2 >= (unsigned long long)n_local) {     // worth it: at least half of the cells qualify
      // slow rows and the cells touching them, both in ascending order (cub select keeps the order)
      unsigned char* flag;
      int* lex_all;
      const i64 nmax = A->n > n_local ? A->n : n_local;
      FB2_TRY(fb2_dalloc(&flag, (size_t)nmax));
      FB2_TRY(fb2_dalloc(&lex_all, (size_t)n_local));
      i64* d_cnt;
      FB2_TRY(fb2_dalloc(&d_cnt, 2));
      // rows
      k_rows_flag_rows<<<(unsigned)((A->n + 255) / 256), 256, 0, st>>>(A->n, A->d_row_fast, flag);
      i64 n_slow_rows_max = A->n - (i64)h_nfast * (P * P * P);
      FB2_TRY(fb2_dalloc(&A->d_slow_rows, (size_t)n_slow_rows_max + 1));
      {
        void* tmp = nullptr; size_t bytes = 0;
        thrust::counting_iterator<i64> it(0);
        FB2_CUDA(cub::DeviceSelect::Flagged(tmp, bytes, it, flag, A->d_slow_rows, d_cnt, A->n, st));
        FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
        FB2_CUDA(cub::DeviceSelect::Flagged(tmp, bytes, it, flag, A->d_slow_rows, d_cnt, A->n, st));
        FB2_CUDA(cudaStreamSynchronize(st));
        cudaFree(tmp);
      }
      // cells
      k_rows_flag_cells<<<(n_local + 255) / 256, 256, 0, st>>>(g, s->d_perm, n_local, s->rt.ndofs, s->d_gids, A->n, A->d_row_fast,
                                                              flag, lex_all);
      int* slow_tmp;
      FB2_TRY(fb2_dalloc(&slow_tmp, (size_t)n_local));
      {
        void* tmp = nullptr; size_t bytes = 0;
        FB2_CUDA(cub::DeviceSelect::Flagged(tmp, bytes, lex_all, flag, slow_tmp, d_cnt + 1, n_local, st));
        FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
        FB2_CUDA(cub::DeviceSelect::Flagged(tmp, bytes, lex_all, flag, slow_tmp, d_cnt + 1, n_local, st));
        FB2_CUDA(cudaStreamSynchronize(st));
        cudaFree(tmp);
      }
      i64 h_cnt[2];
      FB2_CUDA(fb2_memcpy(h_cnt, d_cnt, sizeof(h_cnt), cudaMemcpyDeviceToHost));
      A->n_slow_rows = h_cnt[0]; A->n_slow_cells = (int)h_cnt[1];
      FB2_TRY(fb2_dalloc(&A->d_slow_cells, (size_t)A->n_slow_cells));
      FB2_CUDA(fb2_memcpy(A->d_slow_cells, slow_tmp, sizeof(int) * (size_t)A->n_slow_cells, cudaMemcpyDeviceToDevice));
      cudaFree(slow_tmp); cudaFree(flag); cudaFree(lex_all); cudaFree(d_cnt);
      A->rows_state = 1;
    }
  }
  cudaFree(d_ron); cudaFree(d_ok); cudaFree(d_nfast);
  cudaEventRecord(e1, st);
  cudaStreamSynchronize(st);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  fb2_record_phase("gather_plan", ms);
  fb2_record_phase("gather_fast_cells", (double)A->n_fast_cells);
  fb2_record_phase("gather_slow_cells", (double)A->n_slow_cells);
  fb2_record_phase("gather_slow_rows", (double)A->n_slow_rows);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  FB2_CUDA(cudaGetLastError());
  return rc_all;
}

// zero the rows the scatter kernel accumulates into (instead of a memset of all of val)
int fb2_rows_zero_slow(fb2_matrix_s* A) {
  if (A->n_slow_rows == 0) return 0;
  i64 blocks = (A->n_slow_rows + 7) / 8, cap = (i64)fb2_num_sms() * 16;
  k_rows_zero<<<(unsigned)(blocks < cap ? blocks : cap), 256, 0, fb2_stream()>>>(A->n_slow_rows, A->d_slow_rows, A->d_irp, A->d_val);
  FB2_CUDA(cudaGetLastError());
  return 0;
}

// the gather kernel over every verified owner cell
int fb2_rows_launch(fb2_matrix_s* A, const AsmArgs& a) {
  constexpr int P = 2;
  fb2_space_s* s = A->s;
  FB2_TRY(load_const_tables(s));
  const bool src = a.kind == FB2_KIND_POISSON && (a.fn.kind == FB2_FUNC_SHARP_WAVE || a.src_qp != nullptr);
  const bool zskip = tables_have_centre_zeros(s);
  RowsArgs R;
  R.g = a.g; R.perm = a.perm; R.chunk_base = A->d_chunk_base; R.gids = a.gids; R.plan = A->d_rows_plan;
  static const int use_swz = getenv("FB2_ASM_SWIZZLE") ? atoi(getenv("FB2_ASM_SWIZZLE")) : 1;
  for (int k = 0; k < 32; ++k) R.stage_swz[k] = use_swz ? h_stage_swz[k] : 0;
  R.qw = a.qw; R.gphi = a.gphi; R.gdphi = a.gdphi; R.src_qp = a.src_qp; R.val = a.val; R.rhs = a.rhs; R.fn = a.fn;
  for (int k = 0; k < 27; ++k) R.row_of_node[k] = k < P * P * P ? A->rows_row_of_node[k] : 0;
  R.npy = (s->m.nloc[1] + TY - 1) / TY; R.npz = (s->m.nloc[2] + TZ - 1) / TZ;
  const i64 npencil = (i64)R.npy * R.npz;
  // enough CTAs for ~8 waves of 2 CTAs per SM, segments of at least 8 cells
  i64 nseg = (8LL * 2 * fb2_num_sms() + npencil - 1) / npencil;
  const int max_seg = (s->m.nloc[0] + 7) / 8;
  if (nseg > max_seg) nseg = max_seg;
  if (nseg < 1) nseg = 1;
  R.seg_len = (int)((s->m.nloc[0] + nseg - 1) / nseg);
  R.nseg = (s->m.nloc[0] + R.seg_len - 1) / R.seg_len;
  const size_t smem = rows_smem_bytes<P>(src);
  const i64 grid = npencil * R.nseg;
  FB2_CHECK(grid < 0x7fffffffLL, "too many pencils");
#define FB2_ROWS_LAUNCH(S, Z)                                                                                      \
  do {                                                                                                             \
    FB2_CUDA(cudaFuncSetAttribute(k_assemble_rows<P, S, Z>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    k_assemble_rows<P, S, Z><<<(unsigned)grid, 256, smem, fb2_stream()>>>(R);                                       \
  } while (0)
  if (src) { if (zskip) FB2_ROWS_LAUNCH(true, true); else FB2_ROWS_LAUNCH(true, false); }
  else { if (zskip) FB2_ROWS_LAUNCH(false, true); else FB2_ROWS_LAUNCH(false, false); }
#undef FB2_ROWS_LAUNCH
  FB2_CUDA(cudaGetLastError());
  return 0;
}
