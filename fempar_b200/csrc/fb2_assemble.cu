// Batched per-cell Galerkin integration + scatter into the device-built CSR matrix and RHS
// (replaces the serial cell loop of the user discrete_integration_t plugins and
//  fe_cell_iterator%assembly -> sparse_assembler -> coo append -> sort_and_compress:
//  sbm_fe_affine_operator.i90:15-34, sbm_fe_cell_iterator.i90:1610-1637,1818-1879,
//  base_sparse_matrix.f90:3469-3513).
#include "fb2_internal.h"

#include "fb2_assemble.cuh"

template <bool ATOMIC>
__device__ __forceinline__ void add_entry(const AsmArgs& A, int gi, int gj, double v, double gfix_j, double* elvec_i) {
  if (gi <= 0 || gi > A.n_free) return;
  if (gj < 0) { atomicAdd(elvec_i, -v * gfix_j); return; }   // strong Dirichlet column -> RHS
  if (gj == 0 || gj > A.n_free) return;
  if (A.sym && gi > gj) return;
  i64 slot = find_slot(A.ja, A.irp[gi - 1], A.irp[gi], gj);
  if (ATOMIC) atomicAdd(&A.val[slot], v); else A.val[slot] += v;
}

// Generic kernel: one warp per cell, lanes over the (i,j) node pairs; reference tables in shared memory.
template <bool ATOMIC>
__global__ void __launch_bounds__(128) k_assemble_generic(AsmArgs A, int cells_in_launch) {
  extern __shared__ double sm[];
  const int nq = A.nq, nn = A.nn, nd = A.nd;
  double* s_dphi = sm;                      // [nq][nn][3]
  double* s_phi = s_dphi + nq * nn * 3;     // [nq][nn]
  double* s_qw = s_phi + nq * nn;           // [nq]
  double* s_gphi = s_qw + nq;               // [nq][8]
  double* s_gdphi = s_gphi + nq * 8;        // [nq][8][3]
  double* s_warp = s_gdphi + nq * 24;
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per_warp = nq * 17 + 2 * nd + ((nd + 1) / 2);   // D[6] Jinv[9] wdet fw | gfix elvec | gid(int, nd)
  double* w_D = s_warp + (size_t)w * per_warp;  // [nq][6]
  double* w_Ji = w_D + nq * 6;                  // [nq][9]
  double* w_wd = w_Ji + nq * 9;                 // [nq]
  double* w_fw = w_wd + nq;                     // [nq]
  double* w_gfix = w_fw + nq;                   // [nd]
  double* w_elvec = w_gfix + nd;                // [nd]
  int* w_gid = (int*)(w_elvec + nd);            // [nd]
  for (int i = threadIdx.x; i < nq * nn * 3; i += blockDim.x) s_dphi[i] = A.dphi[i];
  for (int i = threadIdx.x; i < nq * nn; i += blockDim.x) s_phi[i] = A.phi[i];
  for (int i = threadIdx.x; i < nq; i += blockDim.x) s_qw[i] = A.qw[i];
  for (int i = threadIdx.x; i < nq * 8; i += blockDim.x) s_gphi[i] = A.gphi[i];
  for (int i = threadIdx.x; i < nq * 24; i += blockDim.x) s_gdphi[i] = A.gdphi[i];
  __syncthreads();

  // colour classes: cells with the same parity of (i,j,k) share no DOF
  int cdim[3], coff[3];
  for (int k = 0; k < 3; ++k) {
    coff[k] = A.color < 0 ? 0 : (A.color >> k) & 1;
    cdim[k] = A.color < 0 ? A.g.nloc[k] : (A.g.nloc[k] - coff[k] + 1) / 2;
  }
  for (int ci = blockIdx.x * warps + w; ci < cells_in_launch; ci += gridDim.x * warps) {
    int c;
    double X[8][3];
    if (A.coords) {   // host-built space: cells in the caller's order, vertices uploaded
      c = ci;
      cell_vertices_from(A.coords, c, X);
    } else {
      int ijk[3];
      int r = ci;
      ijk[0] = r % cdim[0]; r /= cdim[0];
      ijk[1] = r % cdim[1]; ijk[2] = r / cdim[1];
      for (int k = 0; k < 3; ++k) ijk[k] = (A.color < 0 ? ijk[k] : 2 * ijk[k] + coff[k]) + A.g.L[k];
      int lex = ijk[0] + A.g.ntot[0] * (ijk[1] + A.g.ntot[1] * ijk[2]);
      c = A.perm ? A.perm[lex] : lex;
      cell_vertices(A.g, ijk, X);
    }
    // geometry at the quadrature points (lanes over q)
    for (int q = lane; q < nq; q += 32) {
      double inv[3][3], det;
      jacobian_at(X, s_gdphi + q * 24, inv, det);
      double wd = det * s_qw[q];
      // D = wd * Jinv Jinv^T  where grad_x = Jinv * grad_xi (Jinv[i][j] = d xi_j / d x_i)
      double D[6];
      D[0] = inv[0][0] * inv[0][0] + inv[1][0] * inv[1][0] + inv[2][0] * inv[2][0];
      D[1] = inv[0][0] * inv[0][1] + inv[1][0] * inv[1][1] + inv[2][0] * inv[2][1];
      D[2] = inv[0][0] * inv[0][2] + inv[1][0] * inv[1][2] + inv[2][0] * inv[2][2];
      D[3] = inv[0][1] * inv[0][1] + inv[1][1] * inv[1][1] + inv[2][1] * inv[2][1];
      D[4] = inv[0][1] * inv[0][2] + inv[1][1] * inv[1][2] + inv[2][1] * inv[2][2];
      D[5] = inv[0][2] * inv[0][2] + inv[1][2] * inv[1][2] + inv[2][2] * inv[2][2];
      for (int k = 0; k < 6; ++k) w_D[q * 6 + k] = wd * D[k];
      for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) w_Ji[q * 9 + i * 3 + j] = inv[i][j];
      w_wd[q] = wd;
      double xq[3] = {0, 0, 0};
      for (int a = 0; a < 8; ++a)
        for (int k = 0; k < 3; ++k) xq[k] += s_gphi[q * 8 + a] * X[a][k];
      w_fw[q] = (A.kind == FB2_KIND_POISSON) ? wd * (A.src_qp ? A.src_qp[(i64)c * nq + q] : func_f(A.fn, xq)) : 0.0;
    }
    for (int l = lane; l < nd; l += 32) {
      int g = A.gids[(i64)c * nd + l];
      w_gid[l] = g;
      w_gfix[l] = g < 0 ? A.fixed[-g - 1] : 0.0;
      w_elvec[l] = 0.0;
    }
    __syncwarp();
    // element vector: sum_q f(x_q) w detJ phi_i(q)
    if (A.kind == FB2_KIND_POISSON)
      for (int i = lane; i < nn; i += 32) {
        double s = 0.0;
        for (int q = 0; q < nq; ++q) s += w_fw[q] * s_phi[q * nn + i];
        w_elvec[i] = s;
      }
    __syncwarp();
    // element matrix entries, node pairs (ni, nj)
    for (int e = lane; e < nn * nn; e += 32) {
      int ni = e % nn, nj = e / nn;
      if (A.kind != FB2_KIND_ELASTICITY) {
        double s = 0.0;
        for (int q = 0; q < nq; ++q) {
          const double* gi = s_dphi + (q * nn + ni) * 3;
          const double* gj = s_dphi + (q * nn + nj) * 3;
          const double* D = w_D + q * 6;
          double t0 = D[0] * gj[0] + D[1] * gj[1] + D[2] * gj[2];
          double t1 = D[1] * gj[0] + D[3] * gj[1] + D[4] * gj[2];
          double t2 = D[2] * gj[0] + D[4] * gj[1] + D[5] * gj[2];
          s += gi[0] * t0 + gi[1] * t1 + gi[2] * t2;
        }
        for (int cc = 0; cc < A.ncomp; ++cc) {
          int i = cc * nn + ni, j = cc * nn + nj;
          add_entry<ATOMIC>(A, w_gid[i], w_gid[j], s, w_gfix[j], &w_elvec[i]);
        }
      } else {
        double acc[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
        for (int q = 0; q < nq; ++q) {
          const double* ri = s_dphi + (q * nn + ni) * 3;
          const double* rj = s_dphi + (q * nn + nj) * 3;
          const double* Ji = w_Ji + q * 9;
          double gi[3], gj[3];
          for (int k = 0; k < 3; ++k) {
            gi[k] = Ji[k * 3] * ri[0] + Ji[k * 3 + 1] * ri[1] + Ji[k * 3 + 2] * ri[2];
            gj[k] = Ji[k * 3] * rj[0] + Ji[k * 3 + 1] * rj[1] + Ji[k * 3 + 2] * rj[2];
          }
          double dotg = gi[0] * gj[0] + gi[1] * gj[1] + gi[2] * gj[2], wd = w_wd[q];
          for (int ci2 = 0; ci2 < 3; ++ci2)
            for (int cj = 0; cj < 3; ++cj)
              acc[ci2][cj] += wd * (A.lambda * gi[ci2] * gj[cj] + A.mu * ((ci2 == cj ? dotg : 0.0) + gi[cj] * gj[ci2]));
        }
        for (int ci2 = 0; ci2 < 3; ++ci2)
          for (int cj = 0; cj < 3; ++cj) {
            int i = ci2 * nn + ni, j = cj * nn + nj;
            add_entry<ATOMIC>(A, w_gid[i], w_gid[j], acc[ci2][cj], w_gfix[j], &w_elvec[i]);
          }
      }
    }
    __syncwarp();
    for (int i = lane; i < nd; i += 32) {
      int g = w_gid[i];
      double v = w_elvec[i];
      if (g > 0 && g <= A.n_free && v != 0.0) {
        if (ATOMIC) atomicAdd(&A.rhs[g - 1], v); else A.rhs[g - 1] += v;
      }
    }
    __syncwarp();
  }
}

static size_t generic_smem_bytes(const fb2_space_s* s, int warps) {
  int nq = s->rt.nq, nn = s->rt.nnode, nd = s->rt.ndofs;
  size_t per_warp = (size_t)nq * 17 + 2 * nd + ((nd + 1) / 2);
  return sizeof(double) * ((size_t)nq * nn * 4 + nq + nq * 32 + per_warp * warps);
}

int fb2_assemble_fast(fb2_matrix_s* A, const AsmArgs& args, int scatter_variant, bool rows_path);  // fb2_assemble_fast.cu
static int g_asm_variant = 0;   // 0: fast paths where they apply, 1: always the generic kernel (kept for ncu comparison),
                                // 2: sum-factorised scatter kernel only (no row-owner gather)
int fb2_assembly_variant() { return g_asm_variant; }
extern "C" int fb2_set_assembly_variant(int v) { g_asm_variant = v; return 0; }

extern "C" int fb2_integrate_assemble(fb2_matrix_t A, fb2_vector_t rhs, int kind, int func, const double* func_params,
                                      const double* mat_params, int scatter_variant) {
  FB2_CHECK(A && rhs, "null handle");
  fb2_space_s* s = A->s;
  FB2_CHECK(rhs->n == A->n, "rhs size mismatch");
  FB2_CHECK(kind == FB2_KIND_POISSON ? s->rt.ncomp == 1 : s->rt.ncomp == 3, "operator kind does not match the field type");
  cudaStream_t st = fb2_stream();
  // fe_affine_operator%compute starts from a zeroed matrix/array (sbm_fe_operator.i90:136-137; force_compute zeroes too)
  AsmArgs a;
  a.g = make_geom(s);
  a.p = s->rt.p; a.nn = s->rt.nnode; a.ncomp = s->rt.ncomp; a.nd = s->rt.ndofs; a.nq = s->rt.nq;
  a.perm = s->d_perm; a.gids = s->d_gids; a.coords = s->d_coords; a.fixed = s->d_fixed;
  a.src_qp = kind == FB2_KIND_POISSON ? s->d_src_qp : nullptr;
  a.qw = s->d_qw; a.phi = s->d_phi; a.dphi = s->d_dphi; a.gphi = s->d_gphi; a.gdphi = s->d_gdphi;
  a.irp = A->d_irp; a.ja = A->d_ja; a.val = A->d_val; a.rhs = rhs->d; a.n_free = A->n;
  a.kind = kind; a.sym = A->sym; a.fn = make_func(func, func_params);
  a.lambda = mat_params ? mat_params[0] : 0.0; a.mu = mat_params ? mat_params[1] : 1.0;
  // the scatter map is set-up (like the reference's COO->CSR compression it is paid once): build it before the timed region
  const bool host_built = s->d_coords != nullptr;   // no lattice: generic kernel, slots by binary search, atomic scatter
  FB2_CHECK(s->d_gids != nullptr, "this space has no cell DOF lists (bare space): use fb2_assemble_batch");
  if (host_built) scatter_variant = FB2_SCATTER_ATOMIC;

  // row-owner gather path (scalar Q2 Laplacian on structured parts): plan + verification are set-up as well
  bool rows_path = false;
  if (g_asm_variant == 0 && scatter_variant == FB2_SCATTER_ATOMIC && !host_built) FB2_TRY(fb2_rows_prepare(A, kind));
  // the scatter map is set-up too: with the row-owner plan only the cells left to the scatter kernel need verified tables
  if (g_asm_variant != 1 && s->rt.p <= 3 && !(kind == FB2_KIND_ELASTICITY && s->rt.p > 2)) {
    const bool rows_ok = g_asm_variant == 0 && scatter_variant == FB2_SCATTER_ATOMIC && A->rows_state == 1;
    FB2_TRY(fb2_build_scatter_map(A, rows_ok ? A->d_slow_cells : nullptr, rows_ok ? A->n_slow_cells : 0));
    rows_path = rows_ok && A->map_state == 1;
  }
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  // rows written completely by the gather kernel need no zero fill: only the rows left to the scatter kernel are cleared
  if (rows_path) FB2_TRY(fb2_rows_zero_slow(A));
  else FB2_CUDA(cudaMemsetAsync(A->d_val, 0, sizeof(double) * (size_t)A->nnz, st));
  FB2_CUDA(cudaMemsetAsync(rhs->d, 0, sizeof(double) * (size_t)rhs->n, st));
  const int warps = 4, threads = warps * 32;
  size_t smem = generic_smem_bytes(s, warps);
  FB2_CHECK(smem <= 227 * 1024, "reference tables do not fit in shared memory for this order");
  FB2_CUDA(cudaFuncSetAttribute(k_assemble_generic<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  FB2_CUDA(cudaFuncSetAttribute(k_assemble_generic<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int bps = (int)(227 * 1024 / smem); if (bps < 1) bps = 1; if (bps > 8) bps = 8;
  int fast_rc = g_asm_variant != 1 ? fb2_assemble_fast(A, a, scatter_variant, rows_path) : -1;
  if (fast_rc > 0) return fast_rc;
  if (fast_rc == 0) {
  } else if (scatter_variant == FB2_SCATTER_ATOMIC) {
    a.color = -1;
    int ncell = s->m.n_local;
    int blocks = fb2_num_sms() * bps;
    if ((i64)blocks * warps > ncell) blocks = (ncell + warps - 1) / warps;
    k_assemble_generic<true><<<blocks, threads, smem, st>>>(a, ncell);
  } else {
    for (int color = 0; color < 8; ++color) {
      a.color = color;
      int ncell = 1;
      for (int k = 0; k < 3; ++k) ncell *= (s->m.nloc[k] - ((color >> k) & 1) + 1) / 2;
      if (ncell == 0) continue;
      int blocks = fb2_num_sms() * bps;
      if ((i64)blocks * warps > ncell) blocks = (ncell + warps - 1) / warps;
      k_assemble_generic<false><<<blocks, threads, smem, st>>>(a, ncell);
    }
  }
  FB2_CUDA(cudaGetLastError());
  cudaEventRecord(e1, st);
  FB2_CUDA(cudaStreamSynchronize(st));
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  fb2_record_phase("integrate", ms);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  // par: assemble the RHS across parts (par_sparse_assembler.f90:171-176)
  // (with several sub-domains per process the caller assembles the RHS of all co-resident parts with fb2_vec_comm_multi)
  if (s->imp && fb2_comm_active() && fb2_comm_parts_per_rank() == 1) FB2_TRY(fb2_import_exchange(s->imp, rhs->d));
  return 0;
}

// ------------------------------------------------------------------------------------------
// strong Dirichlet values by nodal interpolation (sbm_serial_fe_space.i90:997-1176)
// ------------------------------------------------------------------------------------------
// physical position of a Lagrange node of cell c (structured: closed form; host-built: trilinear map of the uploaded vertices)
__device__ __forceinline__ void node_position(const CellGeom& g, const double* __restrict__ coords, const int* __restrict__ lex_of,
                                              int c, int node, int p, double* x) {
  int n1 = p + 1;
  int in[3] = {node % n1, (node / n1) % n1, node / (n1 * n1)};
  if (coords) {
    double X[8][3], xi[3] = {(double)in[0] / (double)p, (double)in[1] / (double)p, (double)in[2] / (double)p};
    cell_vertices_from(coords, c, X);
    trilinear_point(X, xi, x);
    return;
  }
  int lex = lex_of ? lex_of[c] : c;
  int ijk[3] = {lex % g.ntot[0], (lex / g.ntot[0]) % g.ntot[1], lex / (g.ntot[0] * g.ntot[1])};
  for (int k = 0; k < 3; ++k) {
    double x0 = (double)(g.first[k] + ijk[k]) / (double)g.N[k], x1 = (double)(g.first[k] + ijk[k] + 1) / (double)g.N[k];
    double sfrac = (double)in[k] / (double)p;
    x[k] = (1.0 - sfrac) * x0 + sfrac * x1;
  }
}

__global__ void k_interp_dirichlet(CellGeom g, const double* __restrict__ coords, int p, int nn, int nd, int n_cells, const int* __restrict__ lex_of,
                                   const int* __restrict__ gids, FuncPrm fn, double* fixed) {
  i64 tid = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (i64)n_cells * nd) return;
  int c = (int)(tid / nd), l = (int)(tid % nd);
  int gid = gids[tid];
  if (gid >= 0) return;
  double x[3];
  node_position(g, coords, lex_of, c, l % nn, p, x);
  fixed[-gid - 1] = func_u(fn, x);
}

extern "C" int fb2_fe_space_interpolate_dirichlet(fb2_space_t s, int func, const double* func_params) {
  if (s->n_dirichlet == 0) return 0;
  i64 total = (i64)s->m.n_cells * s->rt.ndofs;
  k_interp_dirichlet<<<(unsigned)((total + 255) / 256), 256, 0, fb2_stream()>>>(
      make_geom(s), s->d_coords, s->rt.p, s->rt.nnode, s->rt.ndofs, s->m.n_cells, s->d_lex_of, s->d_gids, make_func(func, func_params),
      s->d_fixed);
  FB2_CUDA(cudaGetLastError());
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  return 0;
}

// nodal interpolation of an analytic scalar function into the FREE DOF values of an fe_function
// (serial_fe_space_interpolate_scalar_function, sbm_serial_fe_space.i90:734-819); every component gets u
__global__ void k_interp_free(CellGeom g, const double* __restrict__ coords, int p, int nn, int nd, int n_local, const int* __restrict__ lex_of,
                              const int* __restrict__ gids, i64 n_free, FuncPrm fn, double* x_out) {
  i64 tid = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (i64)n_local * nd) return;
  int c = (int)(tid / nd), l = (int)(tid % nd);
  int gid = gids[tid];
  if (gid <= 0 || gid > n_free) return;
  double x[3];
  node_position(g, coords, lex_of, c, l % nn, p, x);
  x_out[gid - 1] = func_u(fn, x);   // every cell sharing the node writes the same value
}
extern "C" int fb2_fe_space_interpolate(fb2_space_t s, int func, const double* func_params, fb2_vector_t x) {
  FB2_CHECK(s && x && x->n == s->n_free, "fb2_fe_space_interpolate: size mismatch");
  i64 total = (i64)s->m.n_local * s->rt.ndofs;
  if (total > 0)
    k_interp_free<<<(unsigned)((total + 255) / 256), 256, 0, fb2_stream()>>>(make_geom(s), s->d_coords, s->rt.p, s->rt.nnode, s->rt.ndofs,
                                                                            s->m.n_local, s->d_lex_of, s->d_gids, s->n_free,
                                                                            make_func(func, func_params), x->d);
  FB2_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------
// user source term: the integrand seam for an arbitrary f (discrete_integration_t%integrate_galerkin evaluates
// source_term%get_value_space at quadrature_points_coordinates, tutorial_01_..._modules.f90:100-118): the host asks for the
// physical quadrature points of the local cells, evaluates its own function there and hands the samples back
// ------------------------------------------------------------------------------------------
__global__ void k_quad_points(CellGeom g, const double* __restrict__ coords, const int* __restrict__ lex_of, int n_local, int nq,
                              const double* __restrict__ gphi, double* __restrict__ out) {
  i64 tid = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (i64)n_local * nq) return;
  const int c = (int)(tid / nq), q = (int)(tid % nq);
  double X[8][3];
  if (coords) cell_vertices_from(coords, c, X);
  else {
    const int lex = lex_of ? lex_of[c] : c;
    int ijk[3] = {lex % g.ntot[0], (lex / g.ntot[0]) % g.ntot[1], lex / (g.ntot[0] * g.ntot[1])};
    cell_vertices(g, ijk, X);
  }
  double xq[3] = {0, 0, 0};
  for (int a = 0; a < 8; ++a)
    for (int k = 0; k < 3; ++k) xq[k] += gphi[q * 8 + a] * X[a][k];
  for (int k = 0; k < 3; ++k) out[tid * 3 + k] = xq[k];
}
extern "C" int fb2_fe_space_get_quadrature_points(fb2_space_t s, double* host_out) {
  FB2_CHECK(s && host_out && s->d_gphi, "fb2_fe_space_get_quadrature_points: bad argument (bare space?)");
  const i64 total = (i64)s->m.n_local * s->rt.nq;
  if (total == 0) return 0;
  double* d;
  FB2_TRY(fb2_dalloc(&d, (size_t)total * 3));
  k_quad_points<<<(unsigned)((total + 255) / 256), 256, 0, fb2_stream()>>>(make_geom(s), s->d_coords, s->d_lex_of, s->m.n_local,
                                                                          s->rt.nq, s->d_gphi, d);
  FB2_CUDA(cudaGetLastError());
  FB2_CUDA(fb2_memcpy(host_out, d, sizeof(double) * (size_t)total * 3, cudaMemcpyDeviceToHost));
  cudaFree(d);
  return 0;
}
extern "C" int fb2_fe_space_set_source_values(fb2_space_t s, const double* host_in) {
  FB2_CHECK(s && s->d_gphi, "fb2_fe_space_set_source_values: bad argument (bare space?)");
  const size_t total = (size_t)s->m.n_local * s->rt.nq;
  if (!host_in) { cudaFree(s->d_src_qp); s->d_src_qp = nullptr; return 0; }   // back to the analytic selector
  if (!s->d_src_qp) FB2_TRY(fb2_dalloc(&s->d_src_qp, total));
  if (total) FB2_CUDA(fb2_memcpy(s->d_src_qp, host_in, sizeof(double) * total, cudaMemcpyHostToDevice));
  return 0;
}

// ------------------------------------------------------------------------------------------
// error norms of u - u_h for a scalar field (Integration/ErrorNorms/error_norms.f90:160-221,
// sbm_error_norms_scalar.i90:114-232; tutorial_01_..._modules.f90:611-661): warp per cell, lanes over q.
// One pass accumulates the cell contributions of every norm type:
//   out[0] sum e w (mean)   out[1] sum |e| w (l1)   out[2] sum e^2 w (l2, h1)   out[3] sum |e|^p w (lp, w1p)
//   out[4] max |e| (linfty, w1infty)   out[5] sum |grad e|^2 w (h1 semi)   out[6] sum (|grad e|^2)^(p/2) w (w1p semi)
//   out[7] max_k |d_k e| (w1infty semi);   w = detJ * quadrature weight.  Maxima of non-negative doubles are taken on their bit
//   patterns (monotone), sums with FP64 atomics (order-dependent in the last bits, like the reference's MPI reduction).
// ------------------------------------------------------------------------------------------
__global__ void k_error_norms(AsmArgs A, const double* __restrict__ x, int ncells, double expo, double* out /*[8]*/) {
  extern __shared__ double sm[];
  const int nq = A.nq, nn = A.nn;
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double* w_u = sm + (size_t)w * nn;
  double a_mean = 0.0, a_l1 = 0.0, a_l2 = 0.0, a_lp = 0.0, a_max = 0.0, a_h1 = 0.0, a_w1p = 0.0, a_gmax = 0.0;
  for (int ci = blockIdx.x * warps + w; ci < ncells; ci += gridDim.x * warps) {
    int c;
    double X[8][3];
    if (A.coords) {
      c = ci;
      cell_vertices_from(A.coords, c, X);
    } else {
      int ijk[3] = {ci % A.g.nloc[0] + A.g.L[0], (ci / A.g.nloc[0]) % A.g.nloc[1] + A.g.L[1],
                    ci / (A.g.nloc[0] * A.g.nloc[1]) + A.g.L[2]};
      int lex = ijk[0] + A.g.ntot[0] * (ijk[1] + A.g.ntot[1] * ijk[2]);
      c = A.perm ? A.perm[lex] : lex;
      cell_vertices(A.g, ijk, X);
    }
    for (int l = lane; l < nn; l += 32) {
      int g = A.gids[(i64)c * A.nd + l];
      w_u[l] = g > 0 ? x[g - 1] : (g < 0 ? A.fixed[-g - 1] : 0.0);
    }
    __syncwarp();
    for (int q = lane; q < nq; q += 32) {
      double inv[3][3], det;
      jacobian_at(X, A.gdphi + q * 24, inv, det);
      double wd = det * A.qw[q], xq[3] = {0, 0, 0}, uh = 0.0, gr[3] = {0, 0, 0};
      for (int a = 0; a < 8; ++a)
        for (int k = 0; k < 3; ++k) xq[k] += A.gphi[q * 8 + a] * X[a][k];
      for (int a = 0; a < nn; ++a) {
        uh += w_u[a] * A.phi[q * nn + a];
        for (int k = 0; k < 3; ++k) gr[k] += w_u[a] * A.dphi[(q * nn + a) * 3 + k];
      }
      double gu[3], d2 = 0.0;
      func_grad(A.fn, xq, gu);
      for (int i = 0; i < 3; ++i) {
        double gh = inv[i][0] * gr[0] + inv[i][1] * gr[1] + inv[i][2] * gr[2];
        d2 += (gu[i] - gh) * (gu[i] - gh);
        a_gmax = fmax(a_gmax, fabs(gu[i] - gh));
      }
      const double du = func_u(A.fn, xq) - uh;
      a_mean += wd * du; a_l1 += wd * fabs(du); a_l2 += wd * du * du; a_lp += wd * pow(fabs(du), expo);
      a_max = fmax(a_max, fabs(du));
      a_h1 += wd * d2; a_w1p += wd * pow(d2, 0.5 * expo);
    }
    __syncwarp();
  }
  for (int o = 16; o > 0; o >>= 1) {
    a_mean += __shfl_down_sync(0xffffffffu, a_mean, o); a_l1 += __shfl_down_sync(0xffffffffu, a_l1, o);
    a_l2 += __shfl_down_sync(0xffffffffu, a_l2, o); a_lp += __shfl_down_sync(0xffffffffu, a_lp, o);
    a_h1 += __shfl_down_sync(0xffffffffu, a_h1, o); a_w1p += __shfl_down_sync(0xffffffffu, a_w1p, o);
    a_max = fmax(a_max, __shfl_down_sync(0xffffffffu, a_max, o)); a_gmax = fmax(a_gmax, __shfl_down_sync(0xffffffffu, a_gmax, o));
  }
  if (lane == 0) {
    atomicAdd(&out[0], a_mean); atomicAdd(&out[1], a_l1); atomicAdd(&out[2], a_l2); atomicAdd(&out[3], a_lp);
    atomicAdd(&out[5], a_h1); atomicAdd(&out[6], a_w1p);
    atomicMax((unsigned long long*)&out[4], (unsigned long long)__double_as_longlong(a_max));
    atomicMax((unsigned long long*)&out[7], (unsigned long long)__double_as_longlong(a_gmax));
  }
}

int fb2_allreduce_max(double* d_buf, int count);   // fb2_comm.cu

// error_norms_scalar_t%compute for every norm type of error_norms.f90:41-54 that applies to an H1 scalar field.
// norms[11] = mean, l1, l2, lp, linfty, h1_seminorm, h1, w1p_seminorm, w1p, w1infty_seminorm, w1infty; `exponent` as the
// optional argument of compute (<= 0: the reference's default 2.0, error_norms.f90:308-331).  finalize_norm (:174-221) is
// restated as written, including `norm = norm**exponent` for the p-norms.
extern "C" int fb2_error_norms_all(fb2_space_t s, fb2_vector_t x, int func, const double* func_params, double exponent,
                                   double norms[11]) {
  FB2_CHECK(s && x && norms, "null argument");
  FB2_CHECK(s->rt.ncomp == 1, "error norms are implemented for scalar fields");
  FB2_CHECK(s->d_gids != nullptr, "this space has no cell DOF lists (bare space)");
  const double expo = exponent > 0.0 ? exponent : 2.0;
  AsmArgs a;
  a.g = make_geom(s);
  a.p = s->rt.p; a.nn = s->rt.nnode; a.ncomp = 1; a.nd = s->rt.ndofs; a.nq = s->rt.nq;
  a.perm = s->d_perm; a.gids = s->d_gids; a.coords = s->d_coords; a.fixed = s->d_fixed; a.src_qp = nullptr;
  a.qw = s->d_qw; a.phi = s->d_phi; a.dphi = s->d_dphi; a.gphi = s->d_gphi; a.gdphi = s->d_gdphi;
  a.fn = make_func(func, func_params);
  double* d_out = fb2_scratch_partials();
  cudaStream_t st = fb2_stream();
  FB2_CUDA(cudaMemsetAsync(d_out, 0, 8 * sizeof(double), st));
  int warps = 4, blocks = fb2_num_sms() * 8;
  k_error_norms<<<blocks, warps * 32, warps * a.nn * sizeof(double), st>>>(a, x->d, s->m.n_local, expo, d_out);
  FB2_CUDA(cudaGetLastError());
  double h[8];
  if (fb2_comm_active()) {
    // local cells partition the mesh: l1_sum of the sums, l1_max of the maxima (error_norms.f90:188-207); the two maxima are
    // moved next to each other for one max-reduction
    double* d_mx = d_out + 8;
    FB2_CUDA(cudaMemcpyAsync(d_mx, d_out + 4, sizeof(double), cudaMemcpyDeviceToDevice, st));
    FB2_CUDA(cudaMemcpyAsync(d_mx + 1, d_out + 7, sizeof(double), cudaMemcpyDeviceToDevice, st));
    FB2_TRY(fb2_allreduce_sum(d_out, 8));
    FB2_TRY(fb2_allreduce_max(d_mx, 2));
    FB2_CUDA(cudaMemcpyAsync(d_out + 4, d_mx, sizeof(double), cudaMemcpyDeviceToDevice, st));
    FB2_CUDA(cudaMemcpyAsync(d_out + 7, d_mx + 1, sizeof(double), cudaMemcpyDeviceToDevice, st));
  }
  FB2_CUDA(cudaMemcpyAsync(h, d_out, 8 * sizeof(double), cudaMemcpyDeviceToHost, st));
  FB2_CUDA(cudaStreamSynchronize(st));
  norms[0] = h[0];                          // mean_norm
  norms[1] = h[1];                          // l1_norm
  norms[2] = sqrt(h[2]);                    // l2_norm
  norms[3] = pow(h[3], expo);               // lp_norm        (norm**exponent, as written)
  norms[4] = h[4];                          // linfty_norm
  norms[5] = sqrt(h[5]);                    // h1_seminorm
  norms[6] = sqrt(h[2] + h[5]);             // h1_norm
  norms[7] = pow(h[6], expo);               // w1p_seminorm
  norms[8] = pow(h[3] + h[6], expo);        // w1p_norm
  norms[9] = h[7];                          // w1infty_seminorm
  norms[10] = h[4] + h[7];                  // w1infty_norm   (l1_max of the values + l1_max of the gradients)
  return 0;
}

extern "C" int fb2_error_norms(fb2_space_t s, fb2_vector_t x, int func, const double* func_params, double* h1_seminorm,
                               double* l2_norm) {
  double all[11];
  FB2_TRY(fb2_error_norms_all(s, x, func, func_params, 2.0, all));
  *h1_seminorm = all[5];
  *l2_norm = all[2];
  return 0;
}

// ------------------------------------------------------------------------------------------
// fe_operator_t residual / tangent at an evaluation point: what compute_internal_residual ->
// discrete_integration%integrate_residual and compute_tangent -> integrate_tangent do for a non-linear operator
// (sbm_fe_operator.i90:568-639; integrate_residual of test_transient_poisson_discrete_integration.f90:248-319 is the
// beta = 0 case).  Device integrand: quasi-linear diffusion  -div(k(u) grad u) = f,  k(u) = 1 + beta u^2:
//   residual_i = sum_K int  k(u_h) grad u_h . grad phi_i  -  f phi_i
//   tangent_ij = sum_K int  k(u_h) grad phi_j . grad phi_i  +  k'(u_h) phi_j grad u_h . grad phi_i        (not symmetric)
// u_h = the fe_function of the evaluation point: free DOF values from `u`, strong Dirichlet values from fixed_dof_values
// (fe_cell_function_scalar_t%update, as the reference's u_h%update(fe, fe_function)).  Warp per cell, lanes over the quadrature
// points for the geometry and u_h(q), grad u_h(q); then lanes over the rows (residual) or the (i, j) pairs (tangent).
// ------------------------------------------------------------------------------------------
template <bool TANGENT>
__global__ void __launch_bounds__(128) k_integrate_nonlinear(AsmArgs A, const double* __restrict__ u, double beta, double* residual,
                                                             int cells_in_launch) {
  extern __shared__ double sm[];
  const int nq = A.nq, nn = A.nn;
  double* s_dphi = sm;                      // [nq][nn][3]
  double* s_phi = s_dphi + nq * nn * 3;     // [nq][nn]
  double* s_warp = s_phi + nq * nn;
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per_warp = nq * 15 + nn + ((nn + 1) / 2);   // Jinv[9] wd uq gu[3] fw | u_local | gid(int)
  double* w_Ji = s_warp + (size_t)w * per_warp;   // [nq][9]
  double* w_wd = w_Ji + nq * 9;                   // [nq]
  double* w_uq = w_wd + nq;                       // [nq]
  double* w_gu = w_uq + nq;                       // [nq][3]
  double* w_fw = w_gu + nq * 3;                   // [nq]
  double* w_ul = w_fw + nq;                       // [nn]
  int* w_gid = (int*)(w_ul + nn);                 // [nn]
  for (int i = threadIdx.x; i < nq * nn * 3; i += blockDim.x) s_dphi[i] = A.dphi[i];
  for (int i = threadIdx.x; i < nq * nn; i += blockDim.x) s_phi[i] = A.phi[i];
  __syncthreads();
  for (int ci = blockIdx.x * warps + w; ci < cells_in_launch; ci += gridDim.x * warps) {
    int c;
    double X[8][3];
    if (A.coords) {
      c = ci;
      cell_vertices_from(A.coords, c, X);
    } else {
      int ijk[3] = {ci % A.g.nloc[0] + A.g.L[0], (ci / A.g.nloc[0]) % A.g.nloc[1] + A.g.L[1],
                    ci / (A.g.nloc[0] * A.g.nloc[1]) + A.g.L[2]};
      int lex = ijk[0] + A.g.ntot[0] * (ijk[1] + A.g.ntot[1] * ijk[2]);
      c = A.perm ? A.perm[lex] : lex;
      cell_vertices(A.g, ijk, X);
    }
    for (int l = lane; l < nn; l += 32) {
      const int g = A.gids[(i64)c * A.nd + l];
      w_gid[l] = g;
      w_ul[l] = g > 0 ? u[g - 1] : (g < 0 ? A.fixed[-g - 1] : 0.0);
    }
    __syncwarp();
    for (int q = lane; q < nq; q += 32) {
      double inv[3][3], det;
      jacobian_at(X, A.gdphi + q * 24, inv, det);
      const double wd = det * A.qw[q];
      double uq = 0.0, gr[3] = {0, 0, 0};
      for (int a = 0; a < nn; ++a) {
        uq += w_ul[a] * s_phi[q * nn + a];
        for (int k = 0; k < 3; ++k) gr[k] += w_ul[a] * s_dphi[(q * nn + a) * 3 + k];
      }
      for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) w_Ji[q * 9 + i * 3 + j] = inv[i][j];
        w_gu[q * 3 + i] = inv[i][0] * gr[0] + inv[i][1] * gr[1] + inv[i][2] * gr[2];     // physical gradient of u_h
      }
      w_wd[q] = wd; w_uq[q] = uq;
      double f;
      if (A.src_qp) f = A.src_qp[(i64)c * nq + q];
      else {
        double xq[3] = {0, 0, 0};
        for (int a = 0; a < 8; ++a)
          for (int k = 0; k < 3; ++k) xq[k] += A.gphi[q * 8 + a] * X[a][k];
        f = func_f(A.fn, xq);
      }
      w_fw[q] = wd * f;
    }
    __syncwarp();
    if (!TANGENT) {
      for (int i = lane; i < nn; i += 32) {
        const int gi = w_gid[i];
        if (gi <= 0 || gi > A.n_free) continue;
        double r = 0.0;
        for (int q = 0; q < nq; ++q) {
          const double* ri = s_dphi + (q * nn + i) * 3;
          const double* Ji = w_Ji + q * 9;
          double d = 0.0;
          for (int k = 0; k < 3; ++k) d += (Ji[k * 3] * ri[0] + Ji[k * 3 + 1] * ri[1] + Ji[k * 3 + 2] * ri[2]) * w_gu[q * 3 + k];
          const double uq = w_uq[q];
          r += w_wd[q] * (1.0 + beta * uq * uq) * d - w_fw[q] * s_phi[q * nn + i];
        }
        atomicAdd(&residual[gi - 1], r);
      }
    } else {
      for (int e = lane; e < nn * nn; e += 32) {
        const int ni = e % nn, nj = e / nn;
        const int gi = w_gid[ni], gj = w_gid[nj];
        if (gi <= 0 || gi > A.n_free || gj <= 0 || gj > A.n_free) continue;   // tangent with respect to the free DOFs
        double s = 0.0;
        for (int q = 0; q < nq; ++q) {
          const double* ri = s_dphi + (q * nn + ni) * 3;
          const double* rj = s_dphi + (q * nn + nj) * 3;
          const double* Ji = w_Ji + q * 9;
          double dij = 0.0, diu = 0.0;
          for (int k = 0; k < 3; ++k) {
            const double gik = Ji[k * 3] * ri[0] + Ji[k * 3 + 1] * ri[1] + Ji[k * 3 + 2] * ri[2];
            const double gjk = Ji[k * 3] * rj[0] + Ji[k * 3 + 1] * rj[1] + Ji[k * 3 + 2] * rj[2];
            dij += gik * gjk; diu += gik * w_gu[q * 3 + k];
          }
          const double uq = w_uq[q];
          s += w_wd[q] * ((1.0 + beta * uq * uq) * dij + 2.0 * beta * uq * s_phi[q * nn + nj] * diu);
        }
        const i64 b = A.irp[gi - 1], end = A.irp[gi];
        const i64 slot = find_slot(A.ja, b, end, gj);
        if (slot < end && A.ja[slot] == gj) atomicAdd(&A.val[slot], s);
      }
    }
    __syncwarp();
  }
}

static int nonlinear_launch(fb2_matrix_t A, fb2_vector_t u, int func, const double* func_params, double beta, fb2_vector_t residual,
                            bool tangent) {
  FB2_CHECK(A && u && u->n == A->n, "evaluation point size mismatch");
  fb2_space_s* s = A->s;
  FB2_CHECK(s && s->d_gids && s->rt.ncomp == 1, "the non-linear integrand is defined for scalar spaces with cell DOF lists");
  FB2_CHECK(!A->sym, "the tangent of the quasi-linear operator is not symmetric: full storage required");
  AsmArgs a;
  a.g = make_geom(s);
  a.p = s->rt.p; a.nn = s->rt.nnode; a.ncomp = 1; a.nd = s->rt.ndofs; a.nq = s->rt.nq;
  a.perm = s->d_perm; a.gids = s->d_gids; a.coords = s->d_coords; a.fixed = s->d_fixed; a.src_qp = s->d_src_qp;
  a.qw = s->d_qw; a.phi = s->d_phi; a.dphi = s->d_dphi; a.gphi = s->d_gphi; a.gdphi = s->d_gdphi;
  a.irp = A->d_irp; a.ja = A->d_ja; a.val = A->d_val; a.rhs = nullptr; a.n_free = A->n;
  a.kind = FB2_KIND_POISSON; a.sym = 0; a.color = -1; a.fn = make_func(func, func_params); a.lambda = 0; a.mu = 0;
  cudaStream_t st = fb2_stream();
  const int warps = 4, nq = a.nq, nn = a.nn;
  const size_t per_warp = (size_t)nq * 15 + nn + ((nn + 1) / 2);
  const size_t smem = sizeof(double) * ((size_t)nq * nn * 4 + per_warp * warps);
  FB2_CHECK(smem <= 227 * 1024, "reference tables do not fit in shared memory for this order");
  int blocks = fb2_num_sms() * 4;
  if ((i64)blocks * warps > s->m.n_local) blocks = (s->m.n_local + warps - 1) / warps;
  if (blocks < 1) blocks = 1;
  if (tangent) {
    FB2_CUDA(cudaFuncSetAttribute(k_integrate_nonlinear<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    FB2_CUDA(cudaMemsetAsync(A->d_val, 0, sizeof(double) * (size_t)A->nnz, st));
    k_integrate_nonlinear<true><<<blocks, warps * 32, smem, st>>>(a, u->d, beta, nullptr, s->m.n_local);
  } else {
    FB2_CHECK(residual && residual->n == A->n, "residual size mismatch");
    FB2_CUDA(cudaFuncSetAttribute(k_integrate_nonlinear<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    FB2_CUDA(cudaMemsetAsync(residual->d, 0, sizeof(double) * (size_t)residual->n, st));
    k_integrate_nonlinear<false><<<blocks, warps * 32, smem, st>>>(a, u->d, beta, residual->d, s->m.n_local);
    // par: the residual is assembled across parts like the RHS (par_sparse_assembler.f90:171-176)
    if (s->imp && fb2_comm_active() && fb2_comm_parts_per_rank() == 1) FB2_TRY(fb2_import_exchange(s->imp, residual->d));
  }
  FB2_CUDA(cudaGetLastError());
  FB2_CUDA(cudaStreamSynchronize(st));
  return 0;
}
extern "C" int fb2_integrate_residual(fb2_matrix_t A, fb2_vector_t u, int func, const double* func_params, double beta,
                                      fb2_vector_t residual) {
  return nonlinear_launch(A, u, func, func_params, beta, residual, false);
}
extern "C" int fb2_integrate_tangent(fb2_matrix_t A, fb2_vector_t u, int func, const double* func_params, double beta) {
  return nonlinear_launch(A, u, func, func_params, beta, nullptr, true);
}
