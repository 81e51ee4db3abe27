// Shared device helpers of the assembly kernels (fb2_assemble.cu: generic warp-per-cell kernel,
// fb2_assemble_fast.cu: register-blocked contraction + precomputed scatter map).
#pragma once
#include "fb2_internal.h"

struct CellGeom {  // structured-part geometry needed to rebuild vertex coordinates of a cell
  int N[3], ntot[3], nloc[3], L[3], first[3];
  int n_local;
};
static inline CellGeom make_geom(const fb2_space_s* s) {
  CellGeom g;
  for (int k = 0; k < 3; ++k) {
    g.N[k] = s->m.N[k]; g.ntot[k] = s->m.ntot[k]; g.nloc[k] = s->m.nloc[k]; g.L[k] = s->m.L[k]; g.first[k] = s->m.first[k];
  }
  g.n_local = s->m.n_local;
  return g;
}

struct FuncPrm { int kind; double alpha, r, c[3]; };

__device__ __forceinline__ double func_u(const FuncPrm& f, const double* x) {
  if (f.kind == FB2_FUNC_LINEAR) return x[0] + x[1] + x[2];
  if (f.kind == FB2_FUNC_SHARP_WAVE) {
    double t = (x[0] - f.c[0]) * (x[0] - f.c[0]) + (x[1] - f.c[1]) * (x[1] - f.c[1]) + (x[2] - f.c[2]) * (x[2] - f.c[2]);
    return atan(f.alpha * (sqrt(t) - f.r));
  }
  return 0.0;
}
__device__ __forceinline__ void func_grad(const FuncPrm& f, const double* x, double* g) {
  g[0] = g[1] = g[2] = 0.0;
  if (f.kind == FB2_FUNC_LINEAR) { g[0] = g[1] = g[2] = 1.0; return; }
  if (f.kind == FB2_FUNC_SHARP_WAVE) {
    double t1 = (x[0] - f.c[0]) * (x[0] - f.c[0]) + (x[1] - f.c[1]) * (x[1] - f.c[1]) + (x[2] - f.c[2]) * (x[2] - f.c[2]);
    double s = sqrt(t1), t2 = s - f.r, den = s * (f.alpha * f.alpha * t2 * t2 + 1.0);
    for (int k = 0; k < 3; ++k) g[k] = f.alpha * (x[k] - f.c[k]) / den;
  }
}
__device__ __forceinline__ double func_f(const FuncPrm& f, const double* x) {
  if (f.kind != FB2_FUNC_SHARP_WAVE) return 0.0;
  double a = f.alpha, a2 = a * a, a3 = a2 * a, dx2[3], t1 = 0.0;
  for (int k = 0; k < 3; ++k) { dx2[k] = (x[k] - f.c[k]) * (x[k] - f.c[k]); t1 += dx2[k]; }
  double s = sqrt(t1), t2 = s - f.r, t22 = t2 * t2, w = a2 * t22 + 1.0;
  double den1 = s * w, den2 = t1 * s * w, den3 = t1 * w * w, res = 0.0;
  for (int k = 0; k < 3; ++k) res -= a / den1 - (a * dx2[k]) / den2 - (2.0 * a3 * dx2[k] * t2) / den3;
  return res;
}
static inline FuncPrm make_func(int func, const double* prm) {
  FuncPrm f;
  f.kind = func; f.alpha = 200.0; f.r = 0.7; f.c[0] = f.c[1] = f.c[2] = -0.05;
  if (prm) { f.alpha = prm[0]; f.r = prm[1]; f.c[0] = prm[2]; f.c[1] = prm[3]; f.c[2] = prm[4]; }
  return f;
}

// vertex coordinates of the cell at total-box position ijk (unit cube; uniform_hex_mesh_generator.f90:608-635)
__device__ __forceinline__ void cell_vertices(const CellGeom& g, const int* ijk, double X[8][3]) {
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int k = 0; k < 3; ++k) X[a][k] = (double)(g.first[k] + ijk[k] + ((a >> k) & 1)) / (double)g.N[k];
}

// vertex coordinates of a cell of a host-built (unstructured) space
__device__ __forceinline__ void cell_vertices_from(const double* __restrict__ coords, int c, double X[8][3]) {
  const double* p = coords + (i64)c * 24;
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int k = 0; k < 3; ++k) X[a][k] = p[a * 3 + k];
}
// physical position of the reference point xi in [0,1]^3 under the trilinear cell map
__device__ __forceinline__ void trilinear_point(const double X[8][3], const double* xi, double* x) {
  x[0] = x[1] = x[2] = 0.0;
#pragma unroll
  for (int a = 0; a < 8; ++a) {
    double w = ((a & 1) ? xi[0] : 1.0 - xi[0]) * ((a & 2) ? xi[1] : 1.0 - xi[1]) * ((a & 4) ? xi[2] : 1.0 - xi[2]);
#pragma unroll
    for (int k = 0; k < 3; ++k) x[k] += w * X[a][k];
  }
}

// J(q) = sum_a x_a (x) gradN_a(q), inverse, |det| (sbm_cell_map.i90:272-296, sbm_cell_integrator.i90:448-561)
__device__ __forceinline__ void jacobian_at(const double X[8][3], const double* __restrict__ gd /*[8][3]*/, double inv[3][3],
                                            double& det) {
  double J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) J[i][j] += X[a][j] * gd[a * 3 + i];  // J[i][j] = d x_j / d xi_i
  double c00 = J[1][1] * J[2][2] - J[1][2] * J[2][1], c01 = J[1][2] * J[2][0] - J[1][0] * J[2][2],
         c02 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
  det = J[0][0] * c00 + J[0][1] * c01 + J[0][2] * c02;
  double id = 1.0 / det;
  inv[0][0] = c00 * id; inv[1][0] = c01 * id; inv[2][0] = c02 * id;
  inv[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) * id;
  inv[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) * id;
  inv[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) * id;
  inv[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) * id;
  inv[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) * id;
  inv[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) * id;
  det = fabs(det);
}

struct AsmArgs {
  CellGeom g;
  int p, nn, ncomp, nd, nq;
  const int* perm;       // lex -> local cell id (nullptr: identity)
  const int* gids;
  const double* coords;  // host-built spaces: vertex coordinates [cell][8][3]; nullptr = structured closed form
  const double* fixed;
  const double* src_qp;  // user source term at the quadrature points [local cell][q] (nullptr: the analytic selector `fn`)
  const double *qw, *phi, *dphi, *gphi, *gdphi;
  const i64* irp;
  const int* ja;
  double* val;
  double* rhs;
  i64 n_free;
  int kind, sym, color;  // color < 0: atomic scatter over all cells; 0..7: plain read-modify-write on one colour class
  FuncPrm fn;
  double lambda, mu;
};

__device__ __forceinline__ i64 find_slot(const int* __restrict__ ja, i64 lo, i64 hi, int col) {
  while (lo < hi) {
    i64 mid = (lo + hi) >> 1;
    if (ja[mid] < col) lo = mid + 1; else hi = mid;
  }
  return lo;
}

