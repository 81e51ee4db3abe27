// Fast assembly path for the Laplacian-type operators (Poisson, vector Laplacian) on sm_100a.
//
// (1) Scatter map ("cell-local -> CSR-slot map", north_star): slot(c,i,j) = irp[gid(c,i)] + tab[cell_tab[c]][i][j].
//     An explicit map would cost nd^2 * 4 B per cell (49 GB at Q2 256^3); on a structured part the relative
//     slots repeat for every cell with the same position class (distance to the ends of the local cell box,
//     clipped at 3, per direction), so the map is stored as <= 4096 class tables of nd^2 uint16 plus explicit
//     tables for the cells whose class table does not verify (e.g. next to a sub-domain interface after the
//     interior-then-interface renumbering).  Every cell is verified on the device against ja.
// (2) Integration: one warp per cell, lane j owns column j of the element matrix,
//     E[i][j] = sum_q grad_i(q) . (w detJ J^-1 J^-T)(q) grad_j(q), evaluated by SUM FACTORISATION over the three
//     tensor directions with the 1-D basis tables as constant-bank operands of DFMA (no table traffic in the
//     inner loop, 2.7x fewer flops than the dense q-loop); replaces the triple loop of the user plugin
//     (tutorial_01_..._modules.f90:114-128, par_test_poisson_discrete_integration.f90:117-131).
// (3) Scatter: RED.ADD.F64 (atomic variant) or plain read-modify-write on one colour class (colored variant)
//     into val[slot]; strong Dirichlet columns go to the RHS (sbm_fe_cell_iterator.i90:1610-1637).
#include <stdlib.h>
#include "fb2_sumfact.cuh"

// ------------------------------------------------------------------------------------------
// scatter map
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int cell_class_key(const CellGeom& g, const int* cl /*local-box coordinates*/) {
  int key = 0, mul = 1;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    // clip at 3: next to a sub-domain interface the interface cells are numbered last, so the cell that creates the DOFs of
    // the shared n-faces sits one layer further in than at a Dirichlet boundary (measured: clip 2 left ~all cells irregular)
    int lo = min(cl[k], 3), hi = min(g.nloc[k] - 1 - cl[k], 3);
    key += mul * (lo * 4 + hi);
    mul *= 16;
  }
  return key;
}

__global__ void k_map_keys(CellGeom g, const int* __restrict__ perm, int n_local, int* cell_tab, int* rep /*[4096]*/) {
  int ci = blockIdx.x * blockDim.x + threadIdx.x;
  if (ci >= n_local) return;
  int cl[3] = {ci % g.nloc[0], (ci / g.nloc[0]) % g.nloc[1], ci / (g.nloc[0] * g.nloc[1])};
  int lex = (cl[0] + g.L[0]) + g.ntot[0] * ((cl[1] + g.L[1]) + g.ntot[1] * (cl[2] + g.L[2]));
  int c = perm ? perm[lex] : lex;
  int key = cell_class_key(g, cl);
  cell_tab[c] = key;
  // the interior key is shared by almost every cell: look before the atomic (an out-of-date larger value only costs an atomic)
  if (((volatile int*)rep)[key] > c) atomicMin(&rep[key], c);
}
__global__ void k_map_remap(int n_local, const int* __restrict__ class_of_key, int* cell_tab) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < n_local) cell_tab[c] = class_of_key[cell_tab[c]];
}

// one block per table: relative slot of every (i,j) pair of the listed cell
__global__ void k_map_build(const int* __restrict__ cells, int nd, const int* __restrict__ gids, const i64* __restrict__ irp,
                            const int* __restrict__ ja, i64 n_free, int sym, unsigned short* tabs /*first table of the list*/) {
  const int c = cells[blockIdx.x];
  const int* g = gids + (i64)c * nd;
  unsigned short* tab = tabs + (size_t)blockIdx.x * nd * nd;
  for (int e = threadIdx.x; e < nd * nd; e += blockDim.x) {
    int i = e / nd, j = e - i * nd;
    int gi = g[i], gj = g[j];
    unsigned short rel = 0xFFFF;
    if (gi > 0 && gi <= n_free && gj > 0 && gj <= n_free && !(sym && gi > gj)) {
      i64 b = irp[gi - 1], end = irp[gi];
      i64 s = find_slot(ja, b, end, gj);
      if (s < end && ja[s] == gj && s - b < 0xFFFF) rel = (unsigned short)(s - b);
    }
    tab[e] = rel;
  }
}

// warp per cell: does the table assigned to the cell reproduce the slots of this cell?
// (list != nullptr: only the listed cells -- total-box lexicographic ids -- are verified: the cells the scatter kernel will visit)
__global__ void k_map_verify(int n_local, int nd, const int* __restrict__ gids, const i64* __restrict__ irp,
                             const int* __restrict__ ja, i64 n_free, int sym, const int* __restrict__ cell_tab,
                             const unsigned short* __restrict__ tabs, int* bad_list, int* n_bad, const int* __restrict__ list,
                             const int* __restrict__ perm, const unsigned* __restrict__ row_fast) {
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int k = blockIdx.x * warps + w; k < n_local; k += gridDim.x * warps) {
    const int c = list ? (perm ? perm[list[k]] : list[k]) : k;
    const int* g = gids + (i64)c * nd;
    const unsigned short* tab = tabs + (size_t)cell_tab[c] * nd * nd;
    bool bad = false;
    for (int e = lane; e < nd * nd; e += 32) {
      int i = e / nd, j = e - i * nd;
      int gi = g[i], gj = g[j];
      bool valid = gi > 0 && gi <= n_free && gj > 0 && gj <= n_free && !(sym && gi > gj);
      // rows the gather kernel writes are never scattered through this table (same filter as k_assemble_fast)
      if (row_fast && gi > 0 && gi <= n_free && ((row_fast[(gi - 1) >> 5] >> ((gi - 1) & 31)) & 1u)) continue;
      unsigned short rel = tab[e];
      if (valid) {
        i64 s = irp[gi - 1] + rel;
        bad |= (rel == 0xFFFF) || s >= irp[gi] || ja[s] != gj;
      } else bad |= (rel != 0xFFFF);
    }
    bad = __any_sync(0xffffffffu, bad);
    if (bad && lane == 0) bad_list[atomicAdd(n_bad, 1)] = c;
    __syncwarp();
  }
}
__global__ void k_map_assign(const int* __restrict__ bad_list, int n_bad, int first_tab, int* cell_tab) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n_bad) cell_tab[bad_list ? bad_list[k] : k] = first_tab + k;   // bad_list == nullptr: identity (every cell)
}

// only_cells / n_only: when the row-owner kernel owns most rows, only the cells the scatter kernel visits need a verified table
int fb2_build_scatter_map(fb2_matrix_s* A, const int* only_cells, int n_only) {
  if (A->map_state == 1 && A->map_partial && !only_cells) {
    // a map verified for the scatter cells only is now needed for every cell (e.g. the colored or scatter-only variant): rebuild
    cudaFree(A->d_cell_tab); cudaFree(A->d_tabs);
    A->d_cell_tab = nullptr; A->d_tabs = nullptr; A->map_state = 0; A->map_partial = 0;
  }
  if (A->map_state != 0) return 0;
  A->map_partial = only_cells ? 1 : 0;
  fb2_space_s* s = A->s;
  const int nd = s->rt.ndofs, n_local = s->m.n_local;
  cudaStream_t st = fb2_stream();
  if (n_local == 0 || A->n == 0) { A->map_state = -1; return 0; }
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  if (s->d_coords) {   // host-numbered space: no position classes, one explicit table per cell (as for the "bad" cells below)
    const size_t elems = (size_t)nd * nd;
    if ((size_t)n_local * elems * 2 > ((size_t)4 << 30)) A->map_state = -1;
    else {
      int* d_ids;
      FB2_TRY(fb2_dalloc(&A->d_cell_tab, (size_t)n_local));
      FB2_TRY(fb2_dalloc(&d_ids, (size_t)n_local));
      FB2_TRY(fb2_dalloc(&A->d_tabs, elems * n_local));
      k_map_assign<<<(n_local + 255) / 256, 256, 0, st>>>(nullptr, n_local, 0, A->d_cell_tab);
      FB2_CUDA(cudaMemcpyAsync(d_ids, A->d_cell_tab, sizeof(int) * (size_t)n_local, cudaMemcpyDeviceToDevice, st));
      k_map_build<<<n_local, 256, 0, st>>>(d_ids, nd, s->d_gids, A->d_irp, A->d_ja, A->n, A->sym, A->d_tabs);
      FB2_CUDA(cudaGetLastError());
      FB2_CUDA(cudaStreamSynchronize(st));
      cudaFree(d_ids);
      A->n_class_tabs = 0; A->n_tabs = n_local; A->common_tab = -1; A->map_state = 1;
    }
    cudaEventRecord(e1, st);
    cudaStreamSynchronize(st);
    float ms_h = 0;
    cudaEventElapsedTime(&ms_h, e0, e1);
    fb2_record_phase("scatter_map", ms_h);
    fb2_record_phase("scatter_map_class_tables", 0.0);
    fb2_record_phase("scatter_map_tables", (double)A->n_tabs);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
  }
  CellGeom g = make_geom(s);
  int *d_rep, *d_class_of_key, *d_bad, *d_nbad, *d_reps;
  FB2_TRY(fb2_dalloc(&A->d_cell_tab, (size_t)n_local));
  FB2_TRY(fb2_dalloc(&d_rep, 4096)); FB2_TRY(fb2_dalloc(&d_class_of_key, 4096)); FB2_TRY(fb2_dalloc(&d_reps, 4096));
  FB2_TRY(fb2_dalloc(&d_bad, (size_t)n_local)); FB2_TRY(fb2_dalloc(&d_nbad, 1));
  FB2_CUDA(cudaMemsetAsync(d_rep, 0x7f, 4096 * sizeof(int), st));
  FB2_CUDA(cudaMemsetAsync(d_nbad, 0, sizeof(int), st));
  k_map_keys<<<(n_local + 255) / 256, 256, 0, st>>>(g, s->d_perm, n_local, A->d_cell_tab, d_rep);
  int h_rep[4096], h_class[4096], h_reps[4096], ncls = 0;
  FB2_CUDA(cudaMemcpyAsync(h_rep, d_rep, sizeof(h_rep), cudaMemcpyDeviceToHost, st));
  FB2_CUDA(cudaStreamSynchronize(st));
  for (int k = 0; k < 4096; ++k) {
    h_class[k] = -1;
    if (h_rep[k] < n_local) { h_class[k] = ncls; h_reps[ncls++] = h_rep[k]; }
  }
  A->common_tab = h_class[(3 * 4 + 3) * (1 + 16 + 256)];   // lo = hi = 3 (clipped) in x, y, z: the interior class
  FB2_CUDA(cudaMemcpyAsync(d_class_of_key, h_class, sizeof(h_class), cudaMemcpyHostToDevice, st));
  FB2_CUDA(cudaMemcpyAsync(d_reps, h_reps, sizeof(int) * ncls, cudaMemcpyHostToDevice, st));
  k_map_remap<<<(n_local + 255) / 256, 256, 0, st>>>(n_local, d_class_of_key, A->d_cell_tab);
  unsigned short* d_ctabs;
  const size_t tab_elems = (size_t)nd * nd;
  FB2_TRY(fb2_dalloc(&d_ctabs, tab_elems * ncls));
  k_map_build<<<ncls, 256, 0, st>>>(d_reps, nd, s->d_gids, A->d_irp, A->d_ja, A->n, A->sym, d_ctabs);
  k_map_verify<<<fb2_num_sms() * 8, 256, 0, st>>>(only_cells ? n_only : n_local, nd, s->d_gids, A->d_irp, A->d_ja, A->n, A->sym,
                                                   A->d_cell_tab, d_ctabs, d_bad, d_nbad, only_cells, s->d_perm,
                                                   only_cells ? A->d_row_fast : nullptr);
  int nbad = 0;
  FB2_CUDA(cudaMemcpyAsync(&nbad, d_nbad, sizeof(int), cudaMemcpyDeviceToHost, st));
  FB2_CUDA(cudaStreamSynchronize(st));
  FB2_CUDA(cudaGetLastError());
  int rc = 0;
  if ((size_t)nbad * tab_elems * 2 > ((size_t)4 << 30)) {
    // too irregular for a compressed map: the assembly falls back to searching ja (generic kernel)
    cudaFree(d_ctabs); cudaFree(A->d_cell_tab); A->d_cell_tab = nullptr;
    A->map_state = -1;
  } else {
    A->n_class_tabs = ncls; A->n_tabs = ncls + nbad;
    if (nbad == 0) A->d_tabs = d_ctabs;
    else {
      rc = fb2_dalloc(&A->d_tabs, tab_elems * A->n_tabs);
      if (rc == 0) {
        cudaMemcpyAsync(A->d_tabs, d_ctabs, tab_elems * ncls * 2, cudaMemcpyDeviceToDevice, st);
        k_map_build<<<nbad, 256, 0, st>>>(d_bad, nd, s->d_gids, A->d_irp, A->d_ja, A->n, A->sym, A->d_tabs + tab_elems * ncls);
        k_map_assign<<<(nbad + 255) / 256, 256, 0, st>>>(d_bad, nbad, ncls, A->d_cell_tab);
        cudaStreamSynchronize(st);
      }
      cudaFree(d_ctabs);
    }
    A->map_state = 1;
  }
  cudaEventRecord(e1, st);
  cudaStreamSynchronize(st);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  fb2_record_phase("scatter_map", ms);
  fb2_record_phase("scatter_map_class_tables", (double)A->n_class_tabs);
  fb2_record_phase("scatter_map_tables", (double)A->n_tabs);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(d_rep); cudaFree(d_class_of_key); cudaFree(d_bad); cudaFree(d_nbad); cudaFree(d_reps);
  FB2_CUDA(cudaGetLastError());
  return rc;
}

// ------------------------------------------------------------------------------------------
// integration + scatter
// ------------------------------------------------------------------------------------------
struct FastArgs {
  AsmArgs a;
  const int* cell_tab;
  const unsigned short* tabs;
  int common_tab;   // table id of the interior position class (-1: none); cached in shared memory by the kernel
  int tile;         // (x,y) tile edge of the cell traversal of the atomic variant; 0 = lexicographic
  // row-owner hybrid (fb2_assemble_rows.cu): only the listed cells are visited (total-box lexicographic ids, nullptr = all
  // local cells) and rows whose bit is set in row_fast are left alone (the gather kernel writes them completely)
  const int* cell_list;
  const unsigned* row_fast;
};

// Per-cell input data (cell id through perm, DOF ids, row starts, Dirichlet values, table id) is software-pipelined three
// cells ahead in registers: the ncu source view of the previous version showed 32 % of the stall samples as long-scoreboard
// waits on exactly these dependent loads (perm -> gids -> irp / fixed) and on the per-entry scatter-table loads; the table of
// the most frequent position class (interior cells) is now kept in shared memory, as are the geometry-basis gradients.
// ELAST (NCOMP = 3 only): isotropic linear elasticity sigma = lambda tr(eps) I + 2 mu eps
// (par_test_pb_bddc_linear_elasticity/.../sbm_irreducible_discrete_integration.i90:93-125):
//   E[(i,a),(j,b)] = sum_q w detJ [ lambda G_i^a G_j^b + mu G_i^b G_j^a + delta_ab mu G_i . G_j ],  G = J^-1-mapped gradient,
// i.e. for column (j,b) and row component a the same sum-factorised contraction with the flux
//   t_k = w detJ [ lambda B_ak G_j^b + mu B_bk G_j^a + delta_ab mu sum_c B_ck G_j^c ],  B_rk = d xi_k / d x_r.
// HOSTGEO: host-numbered space (fb2_unstructured.cu): cells in the caller's order, vertex coordinates read from memory, one
// explicit scatter table per cell; a separate instantiation so that the structured kernels keep their register budget.
template <int P, int NCOMP, bool ATOMIC, bool ELAST, bool HOSTGEO = false>
__global__ void __launch_bounds__(128, (P == 3 ? 2 : (ELAST ? 2 : 3))) k_assemble_fast(FastArgs F, int cells_in_launch) {
  static_assert(!ELAST || NCOMP == 3, "elasticity needs 3 components");
  static_assert(!HOSTGEO || ATOMIC, "host-numbered spaces scatter with atomics");
  constexpr int DW = ELAST ? 10 : 6;                 // per quadrature point: B (9) + w detJ, or the symmetric D (6)
  constexpr int N1 = P + 1, NN = N1 * N1 * N1, NQ = NN, ND = NN * NCOMP;
  constexpr int CPW = NN <= 16 ? 32 / NN : 1;        // cells per warp (Q1: 4)
  constexpr int CPL = (NN + 31) / 32;                // columns per lane (Q3: 2)
  constexpr int G = CPL * NCOMP;                     // DOFs handled by one lane in its "lane as DOF" role
  constexpr int WARPS = 4;
  constexpr bool TABCACHE = ND * ND * 2 <= 16384;    // scalar Q1-Q3 and vector Q1/Q2
  const AsmArgs& A = F.a;
  __shared__ double s_B[N1 * N1], s_C[N1 * N1];      // 1-D tables for the lane-dependent factors
  // trilinear geometry gradients at the quadrature points (stride 26 doubles = 20 words mod 32: lanes are consecutive q
  // and read 16 bytes each; a stride of 24 doubles put them on two bank groups -- 15.8 wavefronts per LDS.128, ncu round 2), weights
  __shared__ __align__(16) double s_gd[NQ * 26];
  __shared__ double s_qw[NQ];
  __shared__ double s_D[WARPS * CPW][NQ][DW];
  __shared__ double s_fw[WARPS * CPW][NQ];
  __shared__ double s_gfix[WARPS * CPW][ND];
  __shared__ i64 s_base[WARPS * CPW][ND];
  __shared__ int s_gid[WARPS * CPW][ND];
  constexpr bool TABSTAGE = ND * ND * 2 <= 2048;     // per-cell staging of non-interior tables: Q1, Q2 scalar, vector Q1
  __shared__ unsigned short s_tabc[TABCACHE ? ND * ND : 2];
  __shared__ unsigned short s_tabw[TABSTAGE ? WARPS * CPW : 1][TABSTAGE ? ND * ND : 2];
  int staged = -1;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int sub = CPW > 1 ? lane / NN : 0, j0 = CPW > 1 ? lane - sub * NN : lane;
  const bool lane_on = lane < CPW * NN || CPL > 1;
  const int slot = w * CPW + (lane_on ? sub : 0);
  if (threadIdx.x < N1 * N1) { s_B[threadIdx.x] = c_B1[P - 1][threadIdx.x]; s_C[threadIdx.x] = c_C1[P - 1][threadIdx.x]; }
  for (int e = threadIdx.x; e < NQ * 24; e += blockDim.x) s_gd[(e / 24) * 26 + e % 24] = A.gdphi[e];
  for (int e = threadIdx.x; e < NQ; e += blockDim.x) s_qw[e] = A.qw[e];
  const bool have_cache = TABCACHE && F.common_tab >= 0;
  if (have_cache) {
    const unsigned short* src = F.tabs + (size_t)F.common_tab * (ND * ND);
    for (int e = threadIdx.x; e < ND * ND; e += blockDim.x) s_tabc[e] = src[e];
  }
  __syncthreads();
  int cdim[3], coff[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    coff[k] = A.color < 0 ? 0 : (A.color >> k) & 1;
    cdim[k] = A.color < 0 ? A.g.nloc[k] : (A.g.nloc[k] - coff[k] + 1) / 2;
  }
  const bool has_source = A.kind == FB2_KIND_POISSON && (A.fn.kind == FB2_FUNC_SHARP_WAVE || A.src_qp != nullptr);
  const int stride = gridDim.x * WARPS * CPW;
  double dN[3], rN[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) { dN[k] = (double)A.g.N[k]; rN[k] = 1.0 / dN[k]; }
  // total-box position and lexicographic index of the ci-th cell of this launch
  // Traversal order of the atomic variant: (x,y) tiles of T x T cell columns, each walked plane by plane in z, so that the
  // matrix rows a cell shares with its z-neighbour are still in L2 when that neighbour scatters (a whole z-plane of config 2
  // is 382 MB of values, three times the L2; a 32 x 32 tile plane is 6 MB).  Ragged last tiles are handled in closed form.
  const int T = F.tile;
  auto cell_pos = [&](int ci, int* ijk) {
    int r = ci;
    if (F.cell_list) {
      const int lex = F.cell_list[ci];
      ijk[0] = lex % A.g.ntot[0];
      const int rr = lex / A.g.ntot[0];
      ijk[1] = rr % A.g.ntot[1]; ijk[2] = rr / A.g.ntot[1];
      return lex;
    }
    if (A.color < 0 && T > 0) {
      const int nx = cdim[0], ny = cdim[1], nz = cdim[2];
      const int row_cells = nx * T * nz;                    // every tile row but the last is T cells high
      const int ty = r / row_cells;
      r -= ty * row_cells;
      const int wy = min(T, ny - ty * T);
      const int tile_cells = T * wy * nz;                   // every tile of the row but the last is T cells wide
      const int tx = r / tile_cells;
      r -= tx * tile_cells;
      const int wx = min(T, nx - tx * T);
      const int plane = wx * wy;
      ijk[2] = r / plane; r -= ijk[2] * plane;
      const int yy = r / wx;
      ijk[1] = ty * T + yy; ijk[0] = tx * T + (r - yy * wx);
    } else {
      ijk[0] = r % cdim[0]; r /= cdim[0];
      ijk[1] = r % cdim[1]; ijk[2] = r / cdim[1];
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) ijk[k] = (A.color < 0 ? ijk[k] : 2 * ijk[k] + coff[k]) + A.g.L[k];
    return ijk[0] + A.g.ntot[0] * (ijk[1] + A.g.ntot[1] * ijk[2]);
  };
  auto cell_id = [&](int ci) {            // stage A: local cell id (one dependent load through perm in the par numbering)
    if (!lane_on || ci >= cells_in_launch) return -1;
    if constexpr (HOSTGEO) return ci;
    int ijk[3];
    const int lex = cell_pos(ci, ijk);
    return A.perm ? A.perm[lex] : lex;
  };
  auto load_gids = [&](int c, int* gid, int& tabid) {   // stage B
#pragma unroll
    for (int jj = 0; jj < CPL; ++jj)
#pragma unroll
      for (int cc = 0; cc < NCOMP; ++cc) {
        const int q = j0 + 32 * jj;
        gid[jj * NCOMP + cc] = (c >= 0 && q < NQ) ? A.gids[(i64)c * ND + cc * NN + q] : 0;
      }
    tabid = c >= 0 ? F.cell_tab[c] : 0;
  };
  auto load_rows = [&](const int* gid, i64* base, double* fix) {   // stage C
#pragma unroll
    for (int g = 0; g < G; ++g) {
      bool mine = gid[g] > 0 && gid[g] <= A.n_free;
      if (mine && F.row_fast) mine = !((F.row_fast[(gid[g] - 1) >> 5] >> ((gid[g] - 1) & 31)) & 1u);
      base[g] = mine ? A.irp[gid[g] - 1] : -1;
      fix[g] = gid[g] < 0 ? A.fixed[-gid[g] - 1] : 0.0;
    }
  };
  // ---- pipeline prologue ----
  const int ci_first = (blockIdx.x * WARPS + w) * CPW + sub;
  int gidC[G], gidN[G], tabC = 0, tabN = 0, cNN;
  i64 baseC[G];
  double fixC[G];
  {
    const int c0 = cell_id(ci_first), c1 = cell_id(ci_first + stride);
    cNN = cell_id(ci_first + 2 * stride);
    load_gids(c0, gidC, tabC);
    load_gids(c1, gidN, tabN);
    load_rows(gidC, baseC, fixC);
  }
  for (int ci0 = (blockIdx.x * WARPS + w) * CPW; ci0 < cells_in_launch; ci0 += stride) {
    const int ci = ci0 + sub;
    const bool on = lane_on && ci < cells_in_launch;
    const int tab_cur = tabC;
    bool any_dir = false;
    // ---- hand the current cell's DOF data to the warp, then advance the pipeline by one cell ----
#pragma unroll
    for (int jj = 0; jj < CPL; ++jj)
#pragma unroll
      for (int cc = 0; cc < NCOMP; ++cc) {
        const int q = j0 + 32 * jj, g = jj * NCOMP + cc;
        if (on && q < NQ) {
          const int l = cc * NN + q;
          s_gid[slot][l] = gidC[g];
          s_gfix[slot][l] = fixC[g];
          s_base[slot][l] = baseC[g];
          any_dir |= gidC[g] < 0;
        }
      }
    load_rows(gidN, baseC, fixC);                       // rows / Dirichlet values of cell k+1
#pragma unroll
    for (int g = 0; g < G; ++g) gidC[g] = gidN[g];
    tabC = tabN;
    load_gids(cNN, gidN, tabN);                         // DOF ids and table id of cell k+2
    cNN = cell_id(ci + 3 * stride);                     // cell id of cell k+3
    if (on) {
      int ijk[3];
      int lex_cur = ci;
      if constexpr (!HOSTGEO) lex_cur = cell_pos(ci, ijk);
      // vertex coordinates (first + ijk + {0,1}) / N (uniform_hex_mesh_generator.f90:608-635): the quotient is formed with the
      // precomputed reciprocal and one residual correction (q0 = a*y; r = fma(-q0, N, a); q = fma(r, y, q0)), which gives the
      // correctly rounded a/N in 3 DFMA-class operations instead of the ~25-instruction IEEE division sequence per coordinate
      double X[8][3];
      if constexpr (HOSTGEO) cell_vertices_from(A.coords, ci, X);
      else if constexpr (P >= 3) cell_vertices(A.g, ijk, X);   // Q3 is at the 255-register limit: the two extra invariants spill
      else {
        double lohi[3][2];
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
          for (int b = 0; b < 2; ++b) {
            const double a = (double)(A.g.first[k] + ijk[k] + b);
            const double q0 = a * rN[k];
            lohi[k][b] = fma(fma(-q0, dN[k], a), rN[k], q0);
          }
#pragma unroll
        for (int a = 0; a < 8; ++a)
#pragma unroll
          for (int k = 0; k < 3; ++k) X[a][k] = lohi[k][(a >> k) & 1];
      }
#pragma unroll
      for (int jj = 0; jj < CPL; ++jj) {
        const int q = j0 + 32 * jj;
        if (q < NQ) {
          // --- lane as quadrature point: D(q) = w detJ J^-1 J^-T (sbm_cell_map.i90:272-296), source term ---
          double inv[3][3], det;
          jacobian_at(X, s_gd + q * 26, inv, det);
          const double wd = det * s_qw[q];
          double* D = s_D[slot][q];
          if constexpr (ELAST) {
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int k = 0; k < 3; ++k) D[r * 3 + k] = inv[r][k];
            D[9] = wd;
          } else {
            D[0] = wd * (inv[0][0] * inv[0][0] + inv[1][0] * inv[1][0] + inv[2][0] * inv[2][0]);
            D[1] = wd * (inv[0][0] * inv[0][1] + inv[1][0] * inv[1][1] + inv[2][0] * inv[2][1]);
            D[2] = wd * (inv[0][0] * inv[0][2] + inv[1][0] * inv[1][2] + inv[2][0] * inv[2][2]);
            D[3] = wd * (inv[0][1] * inv[0][1] + inv[1][1] * inv[1][1] + inv[2][1] * inv[2][1]);
            D[4] = wd * (inv[0][1] * inv[0][2] + inv[1][1] * inv[1][2] + inv[2][1] * inv[2][2]);
            D[5] = wd * (inv[0][2] * inv[0][2] + inv[1][2] * inv[1][2] + inv[2][2] * inv[2][2]);
          }
          if (has_source) {
            if (A.src_qp) {   // user samples, indexed by the local cell id
              const int c_cur = HOSTGEO ? ci : (A.perm ? A.perm[lex_cur] : lex_cur);
              s_fw[slot][q] = wd * A.src_qp[(i64)c_cur * NQ + q];
            } else {
              double xq[3] = {0, 0, 0};
#pragma unroll
              for (int a = 0; a < 8; ++a)
#pragma unroll
                for (int k = 0; k < 3; ++k) xq[k] += A.gphi[q * 8 + a] * X[a][k];
              s_fw[slot][q] = wd * func_f(A.fn, xq);
            }
          }
        }
      }
    }
    const bool use_common = have_cache && tab_cur == F.common_tab;
    if constexpr (TABSTAGE) {
      if (on && !use_common && staged != tab_cur) {      // the lanes of this cell stage its (boundary-class or explicit) table
        const unsigned short* src = F.tabs + (size_t)tab_cur * (ND * ND);
        constexpr int NL = CPW > 1 ? NN : (NN < 32 ? NN : 32);
        for (int e = j0; e < ND * ND; e += NL) s_tabw[slot][e] = src[e];
        staged = tab_cur;
      }
    }
    any_dir = __any_sync(0xffffffffu, any_dir);
    __syncwarp();
    if (on) {
      // scatter table in shared memory: the block-wide copy of the interior position class or this cell's staged copy
      const unsigned short* tab;
      if constexpr (TABSTAGE) tab = use_common ? (const unsigned short*)s_tabc : (const unsigned short*)s_tabw[slot];
      else tab = use_common ? (const unsigned short*)s_tabc : F.tabs + (size_t)tab_cur * (ND * ND);
#pragma unroll 1
      for (int jj = 0; jj < CPL; ++jj) {
        const int j = j0 + 32 * jj;
        if (j >= NN) break;
        if constexpr (!ELAST) {
          double acc[NN];
          const double (*Dc)[DW] = s_D[slot];
          column_sumfact<P>([&](int q, double g0, double g1, double g2, double& t0, double& t1, double& t2) {
            const double* Dq = Dc[q];
            t0 = Dq[0] * g0 + Dq[1] * g1 + Dq[2] * g2;
            t1 = Dq[1] * g0 + Dq[3] * g1 + Dq[4] * g2;
            t2 = Dq[2] * g0 + Dq[4] * g1 + Dq[5] * g2;
          }, s_B, s_C, j, acc);
          // --- element vector of row j: source term and strong Dirichlet columns (E is symmetric: E[j][i] = acc[i]) ---
          double ev[NCOMP];
#pragma unroll
          for (int cc = 0; cc < NCOMP; ++cc) ev[cc] = 0.0;
          if (has_source) {
            const int j1 = j % N1, j2 = (j / N1) % N1, j3 = j / (N1 * N1);
            double sacc = 0.0;
#pragma unroll 1
            for (int q = 0; q < NQ; ++q) {
              const int q1 = q % N1, q2 = (q / N1) % N1, q3 = q / (N1 * N1);
              sacc += s_fw[slot][q] * (s_B[q1 * N1 + j1] * s_B[q2 * N1 + j2] * s_B[q3 * N1 + j3]);
            }
            ev[0] = sacc;
          }
          if (any_dir) {
#pragma unroll
            for (int cc = 0; cc < NCOMP; ++cc) {
              double sacc = 0.0;
#pragma unroll
              for (int i = 0; i < NN; ++i) sacc += acc[i] * s_gfix[slot][cc * NN + i];
              ev[cc] -= sacc;
            }
          }
          // --- scatter through the precomputed map ---
#pragma unroll
          for (int cc = 0; cc < NCOMP; ++cc) {
#pragma unroll
            for (int i = 0; i < NN; ++i) {
              const int li = cc * NN + i;
              const unsigned short rel = tab[li * ND + cc * NN + j];
              if (rel != 0xFFFF && s_base[slot][li] >= 0) {
                double* dst = A.val + (s_base[slot][li] + rel);
                if (ATOMIC) atomicAdd(dst, acc[i]); else *dst += acc[i];
              }
            }
            const int g = s_gid[slot][cc * NN + j];
            if (g > 0 && g <= A.n_free && ev[cc] != 0.0 && s_base[slot][cc * NN + j] >= 0) {
              if (ATOMIC) atomicAdd(&A.rhs[g - 1], ev[cc]); else A.rhs[g - 1] += ev[cc];
            }
          }
        } else {
          const double lam = A.lambda, mu = A.mu;
          const double (*Bc)[DW] = s_D[slot];
#pragma unroll 1
          for (int b = 0; b < 3; ++b) {                 // column (j, b)
            double evb = 0.0;
#pragma unroll 1
            for (int a = 0; a < 3; ++a) {               // rows (i, a)
              double acc[NN];
              column_sumfact<P>([&](int q, double g0, double g1, double g2, double& t0, double& t1, double& t2) {
                const double* Bq = Bc[q];
                const double G0 = Bq[0] * g0 + Bq[1] * g1 + Bq[2] * g2;      // physical gradient of phi_j
                const double G1 = Bq[3] * g0 + Bq[4] * g1 + Bq[5] * g2;
                const double G2 = Bq[6] * g0 + Bq[7] * g1 + Bq[8] * g2;
                const double Ga = a == 0 ? G0 : (a == 1 ? G1 : G2), Gb = b == 0 ? G0 : (b == 1 ? G1 : G2);
                const double lG = Bq[9] * lam * Gb, mG = Bq[9] * mu * Ga, md = a == b ? Bq[9] * mu : 0.0;
                const double* Ba = Bq + 3 * a;
                const double* Bb = Bq + 3 * b;
                t0 = Ba[0] * lG + Bb[0] * mG + md * (Bq[0] * G0 + Bq[3] * G1 + Bq[6] * G2);
                t1 = Ba[1] * lG + Bb[1] * mG + md * (Bq[1] * G0 + Bq[4] * G1 + Bq[7] * G2);
                t2 = Ba[2] * lG + Bb[2] * mG + md * (Bq[2] * G0 + Bq[5] * G1 + Bq[8] * G2);
              }, s_B, s_C, j, acc);
              // strong Dirichlet columns through the symmetry E[(j,b),(i,a)] = E[(i,a),(j,b)]
              if (any_dir) {
#pragma unroll
                for (int i = 0; i < NN; ++i) evb -= acc[i] * s_gfix[slot][a * NN + i];
              }
#pragma unroll
              for (int i = 0; i < NN; ++i) {
                const int li = a * NN + i;
                const unsigned short rel = tab[li * ND + b * NN + j];
                if (rel != 0xFFFF && s_base[slot][li] >= 0) {
                  double* dst = A.val + (s_base[slot][li] + rel);
                  if (ATOMIC) atomicAdd(dst, acc[i]); else *dst += acc[i];
                }
              }
            }
            const int g = s_gid[slot][b * NN + j];
            if (g > 0 && g <= A.n_free && evb != 0.0 && s_base[slot][b * NN + j] >= 0) {
              if (ATOMIC) atomicAdd(&A.rhs[g - 1], evb); else A.rhs[g - 1] += evb;
            }
          }
        }
      }
    }
    __syncwarp();
  }
}

template <int P, int NCOMP, bool ELAST>
static int launch_fast_host(const FastArgs& F0, const fb2_space_s* s) {
  constexpr int NN = (P + 1) * (P + 1) * (P + 1), CPW = NN <= 16 ? 32 / NN : 1, WARPS = 4;
  FastArgs F = F0;
  F.a.color = -1; F.tile = 0;
  int bps = 0;
  FB2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_assemble_fast<P, NCOMP, true, ELAST, true>, WARPS * 32, 0));
  if (bps < 1) bps = 1;
  const int ncell = s->m.n_local;
  i64 need = ((i64)ncell + WARPS * CPW - 1) / (WARPS * CPW), cap = (i64)fb2_num_sms() * bps;
  k_assemble_fast<P, NCOMP, true, ELAST, true><<<(int)(need < cap ? need : cap), WARPS * 32, 0, fb2_stream()>>>(F, ncell);
  FB2_CUDA(cudaGetLastError());
  return 0;
}

template <int P, int NCOMP, bool ELAST = false>
static int launch_fast(const FastArgs& F0, const fb2_space_s* s, int scatter_variant, int n_listed = -1) {
  constexpr int NN = (P + 1) * (P + 1) * (P + 1), CPW = NN <= 16 ? 32 / NN : 1, WARPS = 4;
  FastArgs F = F0;
  cudaStream_t st = fb2_stream();
  int bps = 0;
  FB2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_assemble_fast<P, NCOMP, true, ELAST>, WARPS * 32, 0));
  if (bps < 1) bps = 1;
  auto grid_for = [&](int ncell) {
    i64 need = ((i64)ncell + WARPS * CPW - 1) / (WARPS * CPW), cap = (i64)fb2_num_sms() * bps;
    return (int)(need < cap ? need : cap);
  };
  if (scatter_variant == FB2_SCATTER_ATOMIC) {
    F.a.color = -1;
    const int ncell = n_listed >= 0 ? n_listed : s->m.n_local;   // listed cells only (row-owner hybrid) or every local cell
    if (ncell > 0) k_assemble_fast<P, NCOMP, true, ELAST><<<grid_for(ncell), WARPS * 32, 0, st>>>(F, ncell);
  } else {
    for (int color = 0; color < 8; ++color) {
      F.a.color = color;
      int ncell = 1;
      for (int k = 0; k < 3; ++k) ncell *= (s->m.nloc[k] - ((color >> k) & 1) + 1) / 2;
      if (ncell == 0) continue;
      k_assemble_fast<P, NCOMP, false, ELAST><<<grid_for(ncell), WARPS * 32, 0, st>>>(F, ncell);
    }
  }
  FB2_CUDA(cudaGetLastError());
  return 0;
}

// returns 0 when the fast path ran, -1 when it does not apply (the caller uses the generic kernel), >0 on error
int fb2_rows_launch(fb2_matrix_s* A, const AsmArgs& a);   // fb2_assemble_rows.cu
int fb2_assemble_fast(fb2_matrix_s* A, const AsmArgs& args, int scatter_variant, bool rows_path) {
  const fb2_space_s* s = A->s;
  if (s->rt.p > 3) return -1;
  if (args.kind == FB2_KIND_ELASTICITY && (s->rt.ncomp != 3 || s->rt.p > 2)) return -1;   // Q3 elasticity: generic kernel
  // rows_path: the caller built the map for the cells the scatter kernel visits (asking for the full map here would throw that
  // partial map away and verify every cell after all)
  if (!rows_path) FB2_TRY(fb2_build_scatter_map(A, nullptr, 0));
  if (A->map_state != 1) return -1;
  FB2_TRY(load_const_tables(s));
  FastArgs F;
  F.a = args; F.cell_tab = A->d_cell_tab; F.tabs = A->d_tabs; F.common_tab = A->common_tab;
  F.tile = getenv("FB2_ASM_TILE") ? atoi(getenv("FB2_ASM_TILE")) : 32;
  F.cell_list = nullptr; F.row_fast = nullptr;
  if (rows_path) {   // verified owner cells: rows written once by the gather kernel; the rest: scatter over the cells touching them
    FB2_TRY(fb2_rows_launch(A, args));
    F.cell_list = A->d_slow_cells; F.row_fast = A->d_row_fast;
    return launch_fast<2, 1>(F, s, FB2_SCATTER_ATOMIC, A->n_slow_cells);
  }
  if (s->d_coords) {   // host-numbered space
    if (args.kind == FB2_KIND_ELASTICITY) return s->rt.p == 1 ? launch_fast_host<1, 3, true>(F, s) : launch_fast_host<2, 3, true>(F, s);
    if (s->rt.ncomp != 1)   // vector Laplacian
      return s->rt.p == 1 ? launch_fast_host<1, 3, false>(F, s) : s->rt.p == 2 ? launch_fast_host<2, 3, false>(F, s) : -1;
    return s->rt.p == 1 ? launch_fast_host<1, 1, false>(F, s) : s->rt.p == 2 ? launch_fast_host<2, 1, false>(F, s)
                                                                              : launch_fast_host<3, 1, false>(F, s);
  }
  if (args.kind == FB2_KIND_ELASTICITY)
    return s->rt.p == 1 ? launch_fast<1, 3, true>(F, s, scatter_variant) : launch_fast<2, 3, true>(F, s, scatter_variant);
  if (s->rt.p == 1) return s->rt.ncomp == 1 ? launch_fast<1, 1>(F, s, scatter_variant) : launch_fast<1, 3>(F, s, scatter_variant);
  if (s->rt.p == 2) return s->rt.ncomp == 1 ? launch_fast<2, 1>(F, s, scatter_variant) : launch_fast<2, 3>(F, s, scatter_variant);
  return s->rt.ncomp == 1 ? launch_fast<3, 1>(F, s, scatter_variant) : launch_fast<3, 3>(F, s, scatter_variant);
}
