// The reference's other Krylov solvers that share the CG path's primitives, on the device:
//   RICHARDSON richardson.f90:101-205, ICG icg.f90:106-240 (through the fused CG kernels of fb2_solver.cu),
//   RGMRES rgmres.f90:120-356, FGMRES fgmres.f90:120-356, LGMRES lgmres.f90:152-420,
//   orthonormalisation iterative_solver_utils.f90:46-106 (mgsro) and :128-196 (icgsro), Givens :202-235.
// Vectors stay on the device; the O(dkrymax^2) Hessenberg / Givens / triangular-solve arithmetic runs on the host
// exactly as in the reference (it is scalar work), fed by one batched reduction per orthogonalisation pass:
// multidot reads the new vector once per 8 basis vectors, multiaxpy writes it once per 8.
#include <math.h>
#include <algorithm>
#include "fb2_internal.h"

#define MV_CHUNK 8

struct BasisPtrs { const double* v[MV_CHUNK]; };

// partials[j * stride + block] = sum_i w_i V_j[i] q[i], j < nv
__global__ void __launch_bounds__(256) k_multidot(int nv, BasisPtrs V, const double* __restrict__ q, const double* __restrict__ w,
                                                  i64 n, double* partials, int stride) {
  __shared__ double s_red[MV_CHUNK][8];
  double acc[MV_CHUNK];
#pragma unroll
  for (int j = 0; j < MV_CHUNK; ++j) acc[j] = 0.0;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) {
    const double qi = (w ? w[i] : 1.0) * q[i];
#pragma unroll
    for (int j = 0; j < MV_CHUNK; ++j)
      if (j < nv) acc[j] += V.v[j][i] * qi;
  }
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < MV_CHUNK; ++j) {
    double v = acc[j];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) s_red[j][wp] = v;
  }
  __syncthreads();
  if (threadIdx.x < nv) {
    double r = 0.0;
    for (int k = 0; k < 8; ++k) r += s_red[threadIdx.x][k];     // fixed order: deterministic
    partials[threadIdx.x * stride + blockIdx.x] = r;
  }
}
__global__ void k_reduce_rows(const double* __restrict__ partials, int nparts, int stride, int nvals, double* out) {
  // one warp per value, fixed order inside the warp
  const int v = blockIdx.x, lane = threadIdx.x;
  if (v >= nvals) return;
  double a = 0.0;
  for (int i = lane; i < nparts; i += 32) a += partials[v * stride + i];
  for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
  if (lane == 0) out[v] = a;
}
struct Coefs { double c[MV_CHUNK]; };
// q += alpha * sum_j c_j V_j
__global__ void __launch_bounds__(256) k_multiaxpy(int nv, BasisPtrs V, Coefs C, double alpha, double* __restrict__ q, i64 n) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x) {
    double a = 0.0;
#pragma unroll
    for (int j = 0; j < MV_CHUNK; ++j)
      if (j < nv) a += C.c[j] * V.v[j][i];
    q[i] += alpha * a;
  }
}

static int mv_grid(i64 n) {
  i64 b = (n + 1023) / 1024, cap = (i64)fb2_num_sms() * 8;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

// multivector_t%multidot: p[i] = <V_i, q>, i < k  (one all-reduce, one host synchronisation)
static int multidot(fb2_matrix_s* A, int k, double* const* V, const double* q, i64 n, double* p_host) {
  if (k <= 0) return 0;
  FB2_CHECK(k <= 1024, "Krylov basis larger than 1024 vectors");
  double* part = fb2_scratch_partials();
  double* d_out = fb2_scratch_partials() + FB2_MAX_PARTIALS * 4;
  const int g = mv_grid(n);
  const double* w = fb2_matrix_weight(A);
  for (int j0 = 0; j0 < k; j0 += MV_CHUNK) {
    const int nv = std::min(MV_CHUNK, k - j0);
    BasisPtrs B;
    for (int j = 0; j < MV_CHUNK; ++j) B.v[j] = V[j0 + std::min(j, nv - 1)];
    k_multidot<<<g, 256, 0, fb2_stream()>>>(nv, B, q, w, n, part, g);
    k_reduce_rows<<<nv, 32, 0, fb2_stream()>>>(part, g, g, nv, d_out + j0);
  }
  FB2_CUDA(cudaGetLastError());
  if (fb2_matrix_is_par(A)) FB2_TRY(fb2_allreduce_sum(d_out, k));
  FB2_CUDA(fb2_memcpy(p_host, d_out, sizeof(double) * k, cudaMemcpyDeviceToHost));
  return 0;
}
// multivector_t%multiaxpy: q += alpha * sum_i c_i V_i
static int multiaxpy(int k, double* const* V, const double* c, double alpha, double* q, i64 n) {
  for (int j0 = 0; j0 < k; j0 += MV_CHUNK) {
    const int nv = std::min(MV_CHUNK, k - j0);
    BasisPtrs B; Coefs C;
    for (int j = 0; j < MV_CHUNK; ++j) { B.v[j] = V[j0 + std::min(j, nv - 1)]; C.c[j] = j < nv ? c[j0 + j] : 0.0; }
    k_multiaxpy<<<mv_grid(n) * 4, 256, 0, fb2_stream()>>>(nv, B, C, alpha, q, n);
  }
  FB2_CUDA(cudaGetLastError());
  return 0;
}
static int dot1(fb2_matrix_s* A, const double* x, const double* y, i64 n, double* out) {
  const double* xs[1] = {x}; const double* ys[1] = {y};
  return fb2_dots_host(A, 1, xs, ys, n, out);
}
static int nrm2(fb2_matrix_s* A, const double* x, i64 n, double* out) {
  FB2_TRY(dot1(A, x, x, n, out));
  *out = sqrt(*out);
  return 0;
}

// iterative_solver_utils.f90:46-106.  V[m-1] against V[0..m-2]; hh[0..m-1]; ierrc as in the reference.
static int modified_gs_reorthonorm(fb2_matrix_s* A, int m, double* const* V, i64 n, double* hh, int* ierrc) {
  const double reorth = (double)0.98f, rzero = 0.0, rone = 1.0;
  double* q = V[m - 1];
  double norm;
  FB2_TRY(dot1(A, q, q, n, &norm));
  if (norm <= rzero) { *ierrc = -1; return 0; }
  if (norm > rzero && rone / norm > rzero) *ierrc = 0; else { *ierrc = -2; return 0; }
  double thr = norm * reorth;
  for (int i = 0; i < m - 1; ++i) {
    double fct;
    FB2_TRY(dot1(A, q, V[i], n, &fct));
    hh[i] = fct;
    FB2_TRY(fb2_dev_axpby(q, -fct, V[i], rone, n));
    if (fct * fct > thr) {
      FB2_TRY(dot1(A, q, V[i], n, &fct));
      hh[i] += fct;
      FB2_TRY(fb2_dev_axpby(q, -fct, V[i], rone, n));
    }
    norm -= hh[i] * hh[i];
    if (norm < rzero) norm = rzero;
    thr = norm * reorth;
  }
  FB2_TRY(nrm2(A, q, n, &hh[m - 1]));
  if (hh[m - 1] <= rzero) { *ierrc = -3; return 0; }
  FB2_TRY(fb2_dev_scal(q, rone / hh[m - 1], q, n));
  return 0;
}

// iterative_solver_utils.f90:128-196 (iterated classical Gram-Schmidt)
static int iterative_gs_reorthonorm(fb2_matrix_s* A, int k, double* const* Q, i64 n, double* s, int* ierrc) {
  const double alpha = 0.5, rzero = 0.0, rone = 1.0;
  for (int i = 0; i < k; ++i) s[i] = 0.0;
  double* q = Q[k - 1];
  double delta_prev, delta;
  FB2_TRY(nrm2(A, q, n, &delta_prev));
  if (delta_prev <= rzero) { *ierrc = -1; return 0; }
  if (delta_prev > rzero && rone / delta_prev > rzero) *ierrc = 0; else { *ierrc = -2; return 0; }
  std::vector<double> p((size_t)std::max(k - 1, 1));
  for (;;) {
    FB2_TRY(multidot(A, k - 1, Q, q, n, p.data()));
    FB2_TRY(multiaxpy(k - 1, Q, p.data(), -1.0, q, n));
    for (int i = 0; i < k - 1; ++i) s[i] += p[i];
    FB2_TRY(nrm2(A, q, n, &delta));
    if (delta > delta_prev * alpha || delta == rzero) break;
    delta_prev = delta;
  }
  s[k - 1] = delta;
  if (s[k - 1] <= rzero) { *ierrc = -3; return 0; }
  FB2_TRY(fb2_dev_scal(q, rone / s[k - 1], q, n));
  return 0;
}

// iterative_solver_utils.f90:202-235
static void apply_givens_rotation(double& x, double& y, double& c, double& s) {
  if (x == 0.0 && y == 0.0) { c = 1.0; s = 0.0; }
  else if (fabs(y) > fabs(x)) { double t = x / y; x = sqrt(1.0 + t * t); s = copysign(1.0 / x, y); c = t * s; }
  else if (fabs(y) <= fabs(x)) { double t = y / x; y = sqrt(1.0 + t * t); c = copysign(1.0 / y, x); s = t * c; }
  else { x = 0.0; y = 0.0; c = 1.0; s = 0.0; }
  x = fabs(x * y);
  y = 0.0;
}

// ------------------------------------------------------------------------------------------
// workspace (allocate_workspace / free_workspace of the solvers)
// ------------------------------------------------------------------------------------------
struct KrylovWork {
  std::vector<double*> bufs;
  ~KrylovWork() { for (double* p : bufs) cudaFree(p); }
  int get(double** out, i64 n) {
    double* p = nullptr;
    FB2_TRY(fb2_dalloc(&p, (size_t)n));
    bufs.push_back(p);
    *out = p;
    return 0;
  }
};

// richardson.f90:101-205: x <- x + relaxation * inv(M) (b - A x); supports res_res, res_rhs
static int richardson_solve(fb2_matrix_s* A, const double* dinv, const double* b, double* x, const fb2_ils_params_t* prm,
                            double relaxation, int* iters, int* conv, double* err, double* hist) {
  const i64 n = A->n;
  const int stop = prm->stopping_criterium;
  FB2_CHECK(stop == FB2_STOP_RES_RES || stop == FB2_STOP_RES_RHS, "RICHARDSON supports RES_RES and RES_RHS (richardson.f90:207-214)");
  KrylovWork W;
  double *r, *z;
  FB2_TRY(W.get(&r, n)); FB2_TRY(W.get(&z, n));
  double rhs_norm = 0.0, res_norm, tol;
  if (stop == FB2_STOP_RES_RHS) FB2_TRY(nrm2(A, b, n, &rhs_norm));
  FB2_TRY(fb2_ils_load_initial_solution(A, x));   // zero unless set_initial_solution was called
  FB2_TRY(fb2_op_apply(A, x, r));
  FB2_TRY(fb2_dev_axpby(r, 1.0, b, -1.0, n));
  FB2_TRY(nrm2(A, r, n, &res_norm));
  tol = prm->rtol * (stop == FB2_STOP_RES_RHS ? rhs_norm : res_norm) + prm->atol;
  double e = res_norm;
  bool did = e <= tol;
  int it = 0;
  while (!did && it < prm->max_num_iterations) {
    FB2_TRY(fb2_dev_apply_M(dinv, r, z, n));
    FB2_TRY(fb2_dev_axpby(x, relaxation, z, 1.0, n));
    ++it;
    FB2_TRY(fb2_op_apply(A, x, r));
    FB2_TRY(fb2_dev_axpby(r, 1.0, b, -1.0, n));
    FB2_TRY(nrm2(A, r, n, &e));
    if (hist) hist[it - 1] = e;
    did = e <= tol;
  }
  *iters = it; *conv = did ? 1 : 0; *err = e;
  fb2_solver_record_convergence(A, tol, e, it);
  return 0;
}

// solve hh(1:k,1:k) y = g in place: the non-BLAS branch of rgmres.f90:327-337 (= DTRSV 'U','N','N')
static void back_substitute(const std::vector<double>& hh, int ld, double* g, int k) {
  for (int j = k - 1; j >= 0; --j) {
    g[j] = g[j] / hh[(size_t)j * ld + j];
    for (int i = j - 1; i >= 0; --i) g[i] -= hh[(size_t)j * ld + i] * g[j];
  }
}

// variant 'R' rgmres.f90:120-356, 'F' fgmres.f90:120-356, 'L' lgmres.f90:152-420
static int gmres_solve(fb2_matrix_s* A, const double* dinv, const double* b, double* x, const fb2_ils_params_t* prm, char variant,
                       int dkrymax, int strat, int* iters, int* conv, double* err, double* hist) {
  const i64 n = A->n;
  const int stop = prm->stopping_criterium, maxit = prm->max_num_iterations;
  const bool true_res = variant == 'L' && (stop == FB2_STOP_RES_RES || stop == FB2_STOP_RES_RHS);
  FB2_CHECK(stop == FB2_STOP_RES_NRMGIVEN_RHS_NRMGIVEN || stop == FB2_STOP_RES_NRMGIVEN_RES_NRMGIVEN || true_res,
            "stopping criterium not supported by this GMRES variant (rgmres.f90:358-366, fgmres.f90:364-372, lgmres.f90:477-487)");
  FB2_CHECK(strat == FB2_ORTHO_MGSRO || strat == FB2_ORTHO_ICGSRO, "unknown orthonormalisation strategy");
  int m = dkrymax < maxit ? dkrymax : maxit;      // more basis vectors than iterations can never be used
  if (m < 1) m = 1;
  const int ld = m + 1;                            // hh(dkrymax+1, dkrymax) column-major
  KrylovWork W;
  double *r, *z;
  FB2_TRY(W.get(&r, n)); FB2_TRY(W.get(&z, n));
  std::vector<double*> V((size_t)m + 1), Z(variant == 'F' ? (size_t)m : 0);
  for (auto& p : V) FB2_TRY(W.get(&p, n));
  for (auto& p : Z) FB2_TRY(W.get(&p, n));
  std::vector<double> hh((size_t)ld * m, 0.0), g((size_t)m + 1, 0.0), g_aux((size_t)m + 1, 0.0), cs(2 * (size_t)m, 0.0);
  double rhs_norm = 0.0, res_norm = 0.0, res_2_norm = 0.0, tol = 0.0, e = 0.0;
  // NOTE lgmres.f90:190-192 evaluates ||b|| only for res_nrmgiven_rhs_nrmgiven although res_rhs uses it (:217): here ||b||
  // is evaluated for res_rhs too (the evident intent; the reference reads an unset variable)
  if (stop == FB2_STOP_RES_NRMGIVEN_RHS_NRMGIVEN || stop == FB2_STOP_RES_RHS) FB2_TRY(nrm2(A, b, n, &rhs_norm));
  FB2_TRY(fb2_ils_load_initial_solution(A, x));   // zero unless set_initial_solution was called
  FB2_TRY(fb2_op_apply(A, x, r));
  FB2_TRY(fb2_dev_axpby(r, 1.0, b, -1.0, n));
  FB2_TRY(nrm2(A, r, n, &res_norm));
  res_2_norm = res_norm;
  if (variant == 'L') { FB2_TRY(fb2_dev_apply_M(dinv, r, z, n)); FB2_TRY(nrm2(A, z, n, &res_norm)); }
  if (stop == FB2_STOP_RES_NRMGIVEN_RHS_NRMGIVEN) { tol = prm->rtol * rhs_norm + prm->atol; e = res_norm; }
  else if (stop == FB2_STOP_RES_NRMGIVEN_RES_NRMGIVEN) { tol = prm->rtol * res_norm + prm->atol; e = res_norm; }
  else if (stop == FB2_STOP_RES_RES) { tol = prm->rtol * res_2_norm + prm->atol; e = res_2_norm; }
  else { tol = prm->rtol * rhs_norm + prm->atol; e = res_2_norm; }
  bool exit_loop = e <= tol, have_z = false;
  int it = 0, ierrc = 0;
  while (!exit_loop && it < maxit) {
    if (it != 0) {
      FB2_TRY(fb2_op_apply(A, x, r));
      FB2_TRY(fb2_dev_axpby(r, 1.0, b, -1.0, n));
      if (variant == 'L') { FB2_TRY(fb2_dev_apply_M(dinv, r, z, n)); FB2_TRY(nrm2(A, z, n, &res_norm)); }
      else FB2_TRY(nrm2(A, r, n, &res_norm));
    }
    if (res_norm != 0.0) FB2_TRY(fb2_dev_scal(V[0], 1.0 / res_norm, variant == 'L' ? z : r, n));
    std::fill(g.begin(), g.end(), 0.0);
    g[0] = res_norm;
    int kloc = 0;
    bool null_rank = false;
    while (!exit_loop && it < maxit && kloc < m) {
      ++kloc; ++it;
      if (variant == 'R') { FB2_TRY(fb2_dev_apply_M(dinv, V[kloc - 1], z, n)); FB2_TRY(fb2_op_apply(A, z, V[kloc])); }
      else if (variant == 'F') { FB2_TRY(fb2_dev_apply_M(dinv, V[kloc - 1], Z[kloc - 1], n)); FB2_TRY(fb2_op_apply(A, Z[kloc - 1], V[kloc])); }
      else { FB2_TRY(fb2_op_apply(A, V[kloc - 1], r)); FB2_TRY(fb2_dev_apply_M(dinv, r, V[kloc], n)); }
      double* col = &hh[(size_t)(kloc - 1) * ld];
      if (strat == FB2_ORTHO_MGSRO) FB2_TRY(modified_gs_reorthonorm(A, kloc + 1, V.data(), n, col, &ierrc));
      else FB2_TRY(iterative_gs_reorthonorm(A, kloc + 1, V.data(), n, col, &ierrc));
      if (ierrc < 0) { exit_loop = true; break; }
      for (int j = 0; j < kloc - 1; ++j) {           // previous Givens rotations on the new column
        const double a = col[j], c = cs[2 * j], s = cs[2 * j + 1];
        col[j] = c * a + s * col[j + 1];
        col[j + 1] = c * col[j + 1] - s * a;
      }
      double c, s;
      apply_givens_rotation(col[kloc - 1], col[kloc], c, s);
      cs[2 * (kloc - 1)] = c; cs[2 * (kloc - 1) + 1] = s;
      const double alpha = -s * g[kloc - 1];
      g[kloc - 1] = c * g[kloc - 1];
      g[kloc] = alpha;
      res_norm = fabs(alpha);
      if (true_res) {                                // lgmres.f90:300-335: true residual of the current iterate
        int ka = kloc;
        std::copy(g.begin(), g.begin() + ka, g_aux.begin());
        while (ka > 0 && hh[(size_t)(ka - 1) * ld + ka - 1] == 0.0) --ka;
        if (ka <= 0) { null_rank = true; break; }
        back_substitute(hh, ld, g_aux.data(), ka);
        FB2_CUDA(cudaMemcpyAsync(z, x, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, fb2_stream()));
        FB2_TRY(multiaxpy(ka, V.data(), g_aux.data(), 1.0, z, n));
        FB2_TRY(fb2_op_apply(A, z, r));
        FB2_TRY(fb2_dev_axpby(r, 1.0, b, -1.0, n));
        FB2_TRY(nrm2(A, r, n, &res_2_norm));
        have_z = true;
        e = res_2_norm;
      } else e = res_norm;
      if (hist) hist[it - 1] = e;
      exit_loop = e <= tol;
    }
    if (null_rank || ierrc == -2) break;
    if (true_res) {
      if (have_z) FB2_CUDA(cudaMemcpyAsync(x, z, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, fb2_stream()));
    } else if (kloc > 0) {
      while (kloc > 0 && hh[(size_t)(kloc - 1) * ld + kloc - 1] == 0.0) --kloc;
      if (kloc <= 0) break;
      back_substitute(hh, ld, g.data(), kloc);
      if (variant == 'R') {                          // r <- V y ; z = inv(M) r ; x += z
        FB2_TRY(fb2_dev_init(r, 0.0, n));
        FB2_TRY(multiaxpy(kloc, V.data(), g.data(), 1.0, r, n));
        FB2_TRY(fb2_dev_apply_M(dinv, r, z, n));
        FB2_TRY(fb2_dev_axpby(x, 1.0, z, 1.0, n));
      } else FB2_TRY(multiaxpy(kloc, variant == 'F' ? Z.data() : V.data(), g.data(), 1.0, x, n));
    }
    exit_loop = e <= tol;
  }
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  *iters = it; *conv = e <= tol ? 1 : 0; *err = e;
  fb2_solver_record_convergence(A, tol, e, it);
  return 0;
}

// LU with partial pivoting of the k x k leading block (column-major, leading dimension ld) and solve: what
// dgetrf/dgetrs do in lfom.f90:250-262 (LAPACK is third party; same algorithm, unblocked).  Returns false if singular.
static bool lu_solve(const std::vector<double>& hh, int ld, int k, double* g) {
  std::vector<double> a((size_t)k * k);
  for (int j = 0; j < k; ++j)
    for (int i = 0; i < k; ++i) a[(size_t)j * k + i] = hh[(size_t)j * ld + i];
  for (int c = 0; c < k; ++c) {
    int piv = c;
    for (int i = c + 1; i < k; ++i)
      if (fabs(a[(size_t)c * k + i]) > fabs(a[(size_t)c * k + piv])) piv = i;
    if (a[(size_t)c * k + piv] == 0.0) return false;
    if (piv != c) {
      for (int j = 0; j < k; ++j) std::swap(a[(size_t)j * k + c], a[(size_t)j * k + piv]);
      std::swap(g[c], g[piv]);
    }
    for (int i = c + 1; i < k; ++i) {
      const double l = a[(size_t)c * k + i] / a[(size_t)c * k + c];
      a[(size_t)c * k + i] = l;
      for (int j = c + 1; j < k; ++j) a[(size_t)j * k + i] -= l * a[(size_t)j * k + c];
      g[i] -= l * g[c];
    }
  }
  for (int i = k - 1; i >= 0; --i) {
    for (int j = i + 1; j < k; ++j) g[i] -= a[(size_t)j * k + i] * g[j];
    g[i] /= a[(size_t)i * k + i];
  }
  return true;
}

// lfom.f90:124-330: left-preconditioned full orthogonalisation method; RES_RES / RES_RHS on the true residual.
// Restated as written, including g(1) = the INITIAL preconditioned residual norm after a restart (lfom.f90:247).
static int lfom_solve(fb2_matrix_s* A, const double* dinv, const double* b, double* x, const fb2_ils_params_t* prm, int dkrymax,
                      int strat, int* iters, int* conv, double* err, double* hist) {
  const i64 n = A->n;
  const int stop = prm->stopping_criterium, maxit = prm->max_num_iterations;
  FB2_CHECK(stop == FB2_STOP_RES_RES || stop == FB2_STOP_RES_RHS, "LFOM supports RES_RES and RES_RHS (lfom.f90:345-352)");
  FB2_CHECK(strat == FB2_ORTHO_MGSRO || strat == FB2_ORTHO_ICGSRO, "unknown orthonormalisation strategy");
  int m = dkrymax < maxit ? dkrymax : maxit;
  if (m < 1) m = 1;
  const int ld = m + 1;
  KrylovWork W;
  double *r, *z;
  FB2_TRY(W.get(&r, n)); FB2_TRY(W.get(&z, n));
  std::vector<double*> V((size_t)m + 1);
  for (auto& p : V) FB2_TRY(W.get(&p, n));
  std::vector<double> hh((size_t)ld * m, 0.0), g((size_t)m + 1, 0.0);
  double rhs_norm = 0.0, res_norm = 0.0, res_2_norm = 0.0, res_norm_initial, tol, e;
  if (stop == FB2_STOP_RES_RHS) FB2_TRY(nrm2(A, b, n, &rhs_norm));
  FB2_TRY(fb2_ils_load_initial_solution(A, x));   // zero unless set_initial_solution was called
  FB2_TRY(fb2_op_apply(A, x, r));
  FB2_TRY(fb2_dev_axpby(r, 1.0, b, -1.0, n));
  FB2_TRY(nrm2(A, r, n, &res_2_norm));
  FB2_TRY(fb2_dev_apply_M(dinv, r, z, n));
  FB2_TRY(nrm2(A, z, n, &res_norm));
  res_norm_initial = res_norm;
  tol = prm->rtol * (stop == FB2_STOP_RES_RES ? res_2_norm : rhs_norm) + prm->atol;
  e = res_2_norm;
  bool exit_loop = e <= tol, have_z = false, failed = false;
  int it = 0, ierrc = 0;
  while (!exit_loop && it < maxit && !failed) {
    std::fill(hh.begin(), hh.end(), 0.0);
    if (it != 0) {
      FB2_TRY(fb2_op_apply(A, x, r));
      FB2_TRY(fb2_dev_axpby(r, 1.0, b, -1.0, n));
      FB2_TRY(fb2_dev_apply_M(dinv, r, z, n));
      FB2_TRY(nrm2(A, z, n, &res_norm));
    }
    if (res_norm != 0.0) FB2_TRY(fb2_dev_scal(V[0], 1.0 / res_norm, z, n));
    int kloc = 0;
    while (!exit_loop && it < maxit && kloc < m) {
      ++kloc; ++it;
      FB2_TRY(fb2_op_apply(A, V[kloc - 1], r));
      FB2_TRY(fb2_dev_apply_M(dinv, r, V[kloc], n));
      double* col = &hh[(size_t)(kloc - 1) * ld];
      if (strat == FB2_ORTHO_MGSRO) FB2_TRY(modified_gs_reorthonorm(A, kloc + 1, V.data(), n, col, &ierrc));
      else FB2_TRY(iterative_gs_reorthonorm(A, kloc + 1, V.data(), n, col, &ierrc));
      if (ierrc < 0) { exit_loop = true; break; }
      std::fill(g.begin(), g.end(), 0.0);
      g[0] = res_norm_initial;
      if (!lu_solve(hh, ld, kloc, g.data())) { failed = true; break; }       // dgetrf info /= 0
      FB2_CUDA(cudaMemcpyAsync(z, x, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, fb2_stream()));
      FB2_TRY(multiaxpy(kloc, V.data(), g.data(), 1.0, z, n));
      FB2_TRY(fb2_op_apply(A, z, r));
      FB2_TRY(fb2_dev_axpby(r, 1.0, b, -1.0, n));
      FB2_TRY(nrm2(A, r, n, &res_2_norm));
      have_z = true;
      e = res_2_norm;
      if (hist) hist[it - 1] = e;
      exit_loop = e <= tol;
    }
    if (failed || ierrc == -2) break;
    if (have_z) FB2_CUDA(cudaMemcpyAsync(x, z, sizeof(double) * (size_t)n, cudaMemcpyDeviceToDevice, fb2_stream()));
    exit_loop = e <= tol;
  }
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  *iters = it; *conv = e <= tol ? 1 : 0; *err = e;
  fb2_solver_record_convergence(A, tol, e, it);
  return 0;
}

// minres.f90:361-740: Paige & Saunders' MINRES as adapted in the reference (symmetric, possibly indefinite A; SPD M).
// Only RES_RES is accepted (minres.f90:764-770); the test applied is phibar <= Anorm * ynorm * rtol + atol.
static int minres_solve(fb2_matrix_s* A, const double* dinv, const double* b, double* x, const fb2_ils_params_t* prm,
                        int* iters, int* conv, double* err, double* hist) {
  const i64 n = A->n;
  FB2_CHECK(prm->stopping_criterium == FB2_STOP_RES_RES, "MINRES supports RES_RES only (minres.f90:764-770)");
  const double zero = 0.0, one = 1.0, ten = 10.0, eps = 2.220446049250313e-16;   // epsilon(1.0_rp)
  const size_t bytes = sizeof(double) * (size_t)n;
  cudaStream_t st = fb2_stream();
  KrylovWork W;
  double *r1, *r2, *v1, *v2, *y, *w, *w1, *w2;
  FB2_TRY(W.get(&r1, n)); FB2_TRY(W.get(&r2, n)); FB2_TRY(W.get(&v1, n)); FB2_TRY(W.get(&v2, n));
  FB2_TRY(W.get(&y, n)); FB2_TRY(W.get(&w, n)); FB2_TRY(W.get(&w1, n)); FB2_TRY(W.get(&w2, n));
  int istop = 0, it = 0;
  double Anorm = zero, Acond = zero, ynorm = zero, e = zero, tol = zero;
  FB2_TRY(fb2_ils_load_initial_solution(A, x));   // zero unless set_initial_solution was called
  FB2_TRY(fb2_op_apply(A, x, r1));
  FB2_TRY(fb2_dev_axpby(r1, 1.0, b, -1.0, n));
  FB2_TRY(fb2_dev_apply_M(dinv, r1, v1, n));
  double beta1;
  FB2_TRY(dot1(A, r1, v1, n, &beta1));
  if (beta1 < zero) istop = 8;
  else if (beta1 == zero) istop = 0;
  else {
    beta1 = sqrt(beta1);
    double oldb = zero, beta = beta1, dbar = zero, epsln = zero, qrnorm = beta1, phibar = beta1, rhs1 = beta1, rhs2 = zero;
    double tnorm2 = zero, ynorm2 = zero, cs = -one, sn = zero, gmax = zero, gmin = zero;
    FB2_TRY(fb2_dev_init(w, 0.0, n));
    FB2_TRY(fb2_dev_init(w2, 0.0, n));
    FB2_CUDA(cudaMemcpyAsync(r2, r1, bytes, cudaMemcpyDeviceToDevice, st));
    for (;;) {
      ++it;
      const double s = one / beta;
      FB2_TRY(fb2_dev_scal(v2, s, v1, n));
      FB2_TRY(fb2_op_apply(A, v2, y));
      if (it >= 2) FB2_TRY(fb2_dev_axpby(y, -(beta / oldb), r1, 1.0, n));
      double alpha;
      FB2_TRY(dot1(A, v2, y, n, &alpha));
      FB2_TRY(fb2_dev_axpby(y, -(alpha / beta), r2, 1.0, n));
      std::swap(r1, r2);                                             // r1 = r2
      FB2_CUDA(cudaMemcpyAsync(r2, y, bytes, cudaMemcpyDeviceToDevice, st));   // r2 = y
      FB2_TRY(fb2_dev_apply_M(dinv, r2, v1, n));
      oldb = beta;
      FB2_TRY(dot1(A, r2, v1, n, &beta));
      if (beta < zero) { istop = 6; break; }
      beta = sqrt(beta);
      tnorm2 = tnorm2 + alpha * alpha + oldb * oldb + beta * beta;
      if (it == 1) {
        if (beta / beta1 <= ten * eps) istop = -1;
        gmax = fabs(alpha);
        gmin = gmax;
      }
      const double oldeps = epsln;
      const double delta = cs * dbar + sn * alpha;
      const double gbar = sn * dbar - cs * alpha;
      epsln = sn * beta;
      dbar = -cs * beta;
      const double gamma = sqrt(gbar * gbar + beta * beta);
      cs = gbar / gamma;
      sn = beta / gamma;
      const double phi = cs * phibar;
      phibar = sn * phibar;
      const double denom = one / gamma;
      {                                                              // w1 = w2 ; w2 = w ; w = (v - oldeps w1 - delta w2) * denom
        double* t = w1; w1 = w2; w2 = w; w = t;
        FB2_CUDA(cudaMemcpyAsync(w, v2, bytes, cudaMemcpyDeviceToDevice, st));
        FB2_TRY(fb2_dev_axpby(w, -oldeps, w1, 1.0, n));
        FB2_TRY(fb2_dev_axpby(w, -delta, w2, 1.0, n));
        FB2_TRY(fb2_dev_scal(w, denom, w, n));
      }
      FB2_TRY(fb2_dev_axpby(x, phi, w, 1.0, n));
      gmax = gmax > gamma ? gmax : gamma;
      gmin = gmin < gamma ? gmin : gamma;
      const double z = rhs1 / gamma;
      ynorm2 = z * z + ynorm2;
      rhs1 = rhs2 - delta * z;
      rhs2 = -epsln * z;
      Anorm = sqrt(tnorm2);
      ynorm = sqrt(ynorm2);
      const double epsx = Anorm * ynorm * eps, epsr = Anorm * ynorm * prm->rtol + prm->atol;
      tol = epsr; e = phibar;
      qrnorm = phibar;
      Acond = gmax / gmin;
      if (istop == 0) {
        if (it >= prm->max_num_iterations) istop = 5;
        if (Acond >= 0.1 / eps) istop = 4;
        if (epsx >= beta1) istop = 3;
        if (qrnorm <= epsx) istop = 2;
        if (qrnorm <= epsr) istop = 1;
      }
      if (hist) hist[it - 1] = e;
      if (istop != 0) break;
    }
  }
  fb2_solver_record_convergence(A, tol, e, it);
  FB2_CUDA(cudaStreamSynchronize(st));
  fb2_record_phase("minres_istop", (double)istop);
  fb2_record_phase("minres_Anorm", Anorm);
  fb2_record_phase("minres_Acond", Acond);
  *iters = it; *conv = (istop == 0 || istop == -1 || istop == 1 || istop == 2) ? 1 : 0; *err = e;
  return 0;
}

// iterative_linear_solver_t%apply for every solver type on the device path (factory keys of
// iterative_linear_solver_creational_methods_dictionary.f90:97-106): x = A^-1 b with x0 = 0
extern "C" int fb2_ils_solve(fb2_matrix_t A, fb2_vector_t dinv, fb2_vector_t b, fb2_vector_t x, int ils_type,
                             const fb2_ils_params_t* prm, const fb2_ils_ext_params_t* ext, int* num_iterations,
                             int* did_converge, double* error_estimate, double* history) {
  FB2_CHECK(A && b && x && prm && num_iterations && did_converge && error_estimate, "fb2_ils_solve: null argument");
  FB2_CHECK(b->n == A->n && x->n == A->n && (!dinv || dinv->n == A->n), "fb2_ils_solve: size mismatch");
  fb2_ils_ext_params_t e;
  e.dkrymax = 1000; e.orthonorm_strat = FB2_ORTHO_ICGSRO; e.relaxation = 1.0;     // iterative_solver_parameters.f90:181-183
  if (ext) e = *ext;
  const double* d = dinv ? dinv->d : nullptr;
  int rc = 0;
  switch (ils_type) {
    case FB2_ILS_CG: return fb2_pcg_solve(A, dinv, b, x, prm, num_iterations, did_converge, error_estimate, history);
    case FB2_ILS_ICG:
      fb2_cg_set_icg(1);
      rc = fb2_pcg_solve(A, dinv, b, x, prm, num_iterations, did_converge, error_estimate, history);
      fb2_cg_set_icg(0);
      return rc;
    case FB2_ILS_RICHARDSON:
      return richardson_solve(A, d, b->d, x->d, prm, e.relaxation, num_iterations, did_converge, error_estimate, history);
    case FB2_ILS_RGMRES:
      return gmres_solve(A, d, b->d, x->d, prm, 'R', e.dkrymax, e.orthonorm_strat, num_iterations, did_converge, error_estimate, history);
    case FB2_ILS_FGMRES:
      return gmres_solve(A, d, b->d, x->d, prm, 'F', e.dkrymax, e.orthonorm_strat, num_iterations, did_converge, error_estimate, history);
    case FB2_ILS_LGMRES:
      return gmres_solve(A, d, b->d, x->d, prm, 'L', e.dkrymax, e.orthonorm_strat, num_iterations, did_converge, error_estimate, history);
    case FB2_ILS_LFOM:
      return lfom_solve(A, d, b->d, x->d, prm, e.dkrymax, e.orthonorm_strat, num_iterations, did_converge, error_estimate, history);
    case FB2_ILS_MINRES:
      return minres_solve(A, d, b->d, x->d, prm, num_iterations, did_converge, error_estimate, history);
    default: break;
  }
  FB2_CHECK(false, "fb2_ils_solve: unknown solver type");
  return 2;
}
