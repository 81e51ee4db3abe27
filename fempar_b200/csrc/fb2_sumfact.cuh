// Sum-factorised evaluation of one column (= one row, the Laplacian-type element matrices are symmetric) of the element
// matrix, shared by the scatter kernel (fb2_assemble_fast.cu) and the row-owner gather kernel (fb2_assemble_rows.cu).
#pragma once
#include "fb2_assemble.cuh"

// 1-D Lagrange values B[q][i] and derivatives C[q][i] at the 1-D Gauss points, orders 1..3.  Every index below is
// a compile-time constant after unrolling, so these become constant-bank / uniform-register operands of DFMA.
// (static: every translation unit that includes this header owns and loads its copy)
static __constant__ double c_B1[3][16];
static __constant__ double c_C1[3][16];

static int g_const_loaded[4] = {0, 0, 0, 0};
static int load_const_tables(const fb2_space_s* s) {
  const int p = s->rt.p;
  if (g_const_loaded[p]) return 0;
  double b[16] = {0}, c[16] = {0};
  for (size_t k = 0; k < s->h_b1d.size(); ++k) { b[k] = s->h_b1d[k]; c[k] = s->h_c1d[k]; }
  FB2_CUDA(cudaMemcpyToSymbol(c_B1, b, sizeof(b), sizeof(b) * (p - 1)));
  FB2_CUDA(cudaMemcpyToSymbol(c_C1, c, sizeof(c), sizeof(c) * (p - 1)));
  g_const_loaded[p] = 1;
  return 0;
}
// Even orders have a Gauss point AND a Lagrange node at the centre of the reference interval: there every basis function but
// the middle one vanishes exactly and the middle one has an exactly zero derivative.  The products with these exact zeros are
// left out of the unrolled contraction when the host has checked the tables (bit-identical results: fma(0, t, a) == a).
static bool tables_have_centre_zeros(const fb2_space_s* s) {
  const int p = s->rt.p, n1 = p + 1;
  if (p != 2) return false;
  const int qm = 1, im = 1;
  for (int i = 0; i < n1; ++i) {
    if (i != im && s->h_b1d[qm * n1 + i] != 0.0) return false;
    if (i == im && s->h_c1d[qm * n1 + i] != 0.0) return false;
  }
  return true;
}

// Column j of the element matrix by sum factorisation.  With q = (q1,q2,q3), i = (i1,i2,i3) and
// t(q) = w detJ J^-1 J^-T(q) grad_ref phi_j(q):
//   E[i][j] = sum_q  C[q1][i1] B[q2][i2] B[q3][i3] t0 + B[q1][i1] C[q2][i2] B[q3][i3] t1 + B[q1][i1] B[q2][i2] C[q3][i3] t2
// contracted direction by direction: 9 (t) + 9 (x) + 9 (y) + 6 (z) FMA per quadrature point for p = 2, i.e. 891 DFMA
// per column instead of the 2430 of the dense q-loop, and no table traffic in the inner loop.
// tfun(q, g0, g1, g2, t0, t1, t2) maps the reference gradient of phi_j at quadrature point q to the flux t(q).
// ZSKIP (P = 2): B[1][0] = B[1][2] = C[1][1] = 0 exactly -> 774 DFMA per column.
template <int P, bool ZSKIP = false, class TFun>
__device__ __forceinline__ void column_sumfact(TFun tfun, const double* __restrict__ sB,
                                               const double* __restrict__ sC, int j, double* __restrict__ acc) {
  constexpr int N1 = P + 1, NN = N1 * N1 * N1;
  // is the table entry [q][i] an exact zero?  (compile-time after unrolling)
#define FB2_BZ(q, i) (ZSKIP && (q) == 1 && (i) != 1)
#define FB2_CZ(q, i) (ZSKIP && (q) == 1 && (i) == 1)
  const int j1 = j % N1, j2 = (j / N1) % N1, j3 = j / (N1 * N1);
#pragma unroll
  for (int i = 0; i < NN; ++i) acc[i] = 0.0;
#pragma unroll
  for (int q3 = 0; q3 < N1; ++q3) {
    const double b3 = sB[q3 * N1 + j3], c3 = sC[q3 * N1 + j3];
    double w01[N1 * N1], w2[N1 * N1];
#pragma unroll
    for (int k = 0; k < N1 * N1; ++k) { w01[k] = 0.0; w2[k] = 0.0; }
#pragma unroll
    for (int q2 = 0; q2 < N1; ++q2) {
      const double b2 = sB[q2 * N1 + j2], c2 = sC[q2 * N1 + j2];
      const double m0 = b2 * b3, m1 = c2 * b3, m2 = b2 * c3;   // grad_ref phi_j = (c1 m0, b1 m1, b1 m2)
      double a0[N1], a1[N1], a2[N1];
#pragma unroll
      for (int k = 0; k < N1; ++k) { a0[k] = 0.0; a1[k] = 0.0; a2[k] = 0.0; }
#pragma unroll
      for (int q1 = 0; q1 < N1; ++q1) {
        const int q = q1 + N1 * (q2 + N1 * q3);
        const double b1 = sB[q1 * N1 + j1], c1 = sC[q1 * N1 + j1];
        const double g0 = c1 * m0, g1 = b1 * m1, g2 = b1 * m2;
        double t0, t1, t2;
        tfun(q, g0, g1, g2, t0, t1, t2);
#pragma unroll
        for (int i1 = 0; i1 < N1; ++i1) {
          if (!FB2_CZ(q1, i1)) a0[i1] = fma(c_C1[P - 1][q1 * N1 + i1], t0, a0[i1]);
          if (!FB2_BZ(q1, i1)) {
            a1[i1] = fma(c_B1[P - 1][q1 * N1 + i1], t1, a1[i1]);
            a2[i1] = fma(c_B1[P - 1][q1 * N1 + i1], t2, a2[i1]);
          }
        }
      }
#pragma unroll
      for (int i2 = 0; i2 < N1; ++i2)
#pragma unroll
        for (int i1 = 0; i1 < N1; ++i1) {
          double v = w01[i1 + N1 * i2];
          if (!FB2_CZ(q2, i2)) v = fma(c_C1[P - 1][q2 * N1 + i2], a1[i1], v);
          if (!FB2_BZ(q2, i2)) {
            v = fma(c_B1[P - 1][q2 * N1 + i2], a0[i1], v);
            w2[i1 + N1 * i2] = fma(c_B1[P - 1][q2 * N1 + i2], a2[i1], w2[i1 + N1 * i2]);
          }
          w01[i1 + N1 * i2] = v;
        }
    }
#pragma unroll
    for (int i3 = 0; i3 < N1; ++i3)
#pragma unroll
      for (int k = 0; k < N1 * N1; ++k) {
        double v = acc[k + N1 * N1 * i3];
        if (!FB2_CZ(q3, i3)) v = fma(c_C1[P - 1][q3 * N1 + i3], w2[k], v);
        if (!FB2_BZ(q3, i3)) v = fma(c_B1[P - 1][q3 * N1 + i3], w01[k], v);
        acc[k + N1 * N1 * i3] = v;
      }
  }
#undef FB2_BZ
#undef FB2_CZ
}
