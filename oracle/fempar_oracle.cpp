// =====================================================================================
//  fempar_oracle.cpp  --  TEST INFRASTRUCTURE ONLY (parity oracle + timed CPU baseline)
//
//  A from-scratch CPU restatement (C++17, no GPU, no BLAS) of the FEMPAR hot path named
//  by BASELINE.json:north_star.  The reference itself (Fortran 2003/2008 + FPL/FLAP
//  submodules that are absent from /root/reference) cannot be compiled in this image, so
//  this file follows the reference's algorithms step by step and cites, for every
//  function, the reference file:line it restates.  Paths are relative to
//  /root/reference/Sources/Lib unless stated.
//
//  PARITY PIN: validated against the only numeric golden data the reference holds for
//  this path -- Tutorials/tutorial_01_poisson_sharp_circular_wave/convergence_plot/
//  uniform_mesh_error_order_{1,2,4} (tests/test_oracle_golden.py) -- and against the
//  par_test_poisson acceptance criterion (all error norms < 1e-4 for u = x+y+z).
//  The DOF numbering / CSR graph have no fixture upstream: for them this restatement IS
//  the pin, argued by the citations below ("parity unpinned" in the strict sense for the
//  ORDER of the DOFs).
//
//  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
//  legs may load this library.  The product (fempar_b200/) never does.
// =====================================================================================
#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <thread>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace fo {

typedef long long i64;

// static row/cell range split over std::threads (no OpenMP runtime in this image)
template <class F>
static void parallel_for(int n, int nt, F f) {
  if (nt <= 1 || n < 4096) { f(0, n); return; }
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t) {
    int lo = int((long long)n * t / nt), hi = int((long long)n * (t + 1) / nt);
    th.emplace_back([=]() { f(lo, hi); });
  }
  for (auto& t : th) t.join();
}

// ---------------------------------------------------------------------------
// polytope_t (hypercube only): n-faces sorted by (dimension, bit code)
//   Integration/ReferenceFE/sbm_polytope.i90:175-227 (create + sort_n_faces)
//   code = free_bits << d | anchor_bits ; "n_face_dir_is_fixed" returns the FREE bit
//   (sbm_polytope.i90:342-349), "n_face_dir_coordinate" the anchor bit (:352-359).
// ---------------------------------------------------------------------------
struct Polytope {
  int d = 0;
  std::vector<int> code;          // sorted n-face codes; last one is the cell itself
  std::vector<int> index_of_code; // ijk_to_index (1-based, 0 = none)
  void create(int dims) {
    d = dims;
    code.clear();
    for (int fr = 0; fr < (1 << d); ++fr)
      for (int an = 0; an < (1 << d); ++an)
        if ((fr & an) == 0) code.push_back((fr << d) | an);
    std::stable_sort(code.begin(), code.end(), [&](int a, int b) {
      int da = __builtin_popcount(a >> d), db = __builtin_popcount(b >> d);
      if (da != db) return da < db;
      return a < b;
    });
    index_of_code.assign(((1 << d) - 1) * (1 << d) + 1, 0);
    for (size_t i = 0; i < code.size(); ++i) index_of_code[code[i]] = int(i) + 1;
  }
  int num_n_faces() const { return int(code.size()); }
  int free_bits(int f) const { return code[f] >> d; }
  int anchor(int f) const { return code[f] & ((1 << d) - 1); }
  int dim(int f) const { return __builtin_popcount(free_bits(f)); }
  int is_free(int f, int dir) const { return (free_bits(f) >> dir) & 1; }
  int coord(int f, int dir) const { return (anchor(f) >> dir) & 1; }
};

// ijk <-> spatial numbering, Geometry/Tools/uniform_hex_mesh_generator.f90:783-815 (0-based here)
static inline i64 ijk_to_lex(int d, const int* n, const int* ijk) {
  i64 r = 0, prev = 1;
  for (int k = 0; k < d; ++k) { r += prev * ijk[k]; prev *= n[k]; }
  return r;
}
static inline void lex_to_ijk(int d, const int* n, i64 j, int* ijk) {
  for (int k = 0; k < d; ++k) { ijk[k] = int(j % n[k]); j /= n[k]; }
  for (int k = d; k < 3; ++k) ijk[k] = 0;
}

// ---------------------------------------------------------------------------
// uniform_hex_mesh_t%generate_part_local_mesh
//   Geometry/Tools/uniform_hex_mesh_generator.f90:145-647
// ---------------------------------------------------------------------------
struct Mesh {
  int d = 3;
  int N[3] = {1, 1, 1}, P[3] = {1, 1, 1}, pijk[3] = {0, 0, 0};
  bool par = false;
  int part_id = 1;
  int nloc[3] = {1, 1, 1}, ntot[3] = {1, 1, 1}, L[3] = {0, 0, 0}, R[3] = {0, 0, 0}, first[3] = {0, 0, 0};
  int n_local = 0, n_ghost = 0, n_itfc = 0, n_cells = 0;
  Polytope poly;
  int nvef = 0;       // n-faces per cell without the cell itself (26 in 3-D)
  int ntypes = 0;
  std::vector<std::array<int, 3>> tot_x_dir, glob_x_dir;
  std::vector<int> tot_off;    // 0-based first vef id of each type (size ntypes+1)
  std::vector<i64> glob_off;   // 1-based like the reference (num_global_n_faces)
  std::vector<int> type_first_face; // polytope n-face (0-based) defining each type
  int n_vefs = 0, n_vertices = 0;
  std::vector<int> perm, lex_of;   // lex (total box, 0-based) -> local cell id (0-based) and inverse
  std::vector<int> lst_vefs;       // [n_cells * nvef] local vef ids, 0-based
  std::vector<int> vef_bid;        // boundary id per vef: 1..3^d-1 on the boundary, 3^d interior
  std::vector<i64> cells_gid, vefs_gid; // 1-based global ids
  std::vector<int> cells_part;     // 1-based part ids
  std::vector<double> vcoord;      // [n_vertices*3]
  std::vector<char> vef_ghost;     // vef touches no local cell
  double dom[3][2] = {{0, 1}, {0, 1}, {0, 1}};

  void create(int dims, const int* Nc, const int* Pp, int part /*1-based; <=0: serial*/) {
    d = dims;
    for (int k = 0; k < 3; ++k) { N[k] = k < d ? Nc[k] : 1; P[k] = 1; pijk[k] = 0; }
    par = part > 0;
    part_id = par ? part : 1;
    poly.create(d);
    nvef = poly.num_n_faces() - 1;
    if (par) {
      for (int k = 0; k < d; ++k) P[k] = Pp[k];
      lex_to_ijk(d, P, part - 1, pijk);      // :263-268
    }
    for (int k = 0; k < 3; ++k) { L[k] = R[k] = 0; }
    for (int k = 0; k < d; ++k) {            // :269-276 (non periodic)
      L[k] = R[k] = 1;
      if (pijk[k] == 0) L[k] = 0;
      if (pijk[k] == P[k] - 1) R[k] = 0;
    }
    if (!par) for (int k = 0; k < 3; ++k) L[k] = R[k] = 0;
    n_local = 1;
    for (int k = 0; k < 3; ++k) { nloc[k] = 1; ntot[k] = 1; first[k] = 0; }
    for (int k = 0; k < d; ++k) {            // :281-291
      nloc[k] = N[k] / P[k];
      first[k] = pijk[k] * nloc[k];
      n_local *= nloc[k];
      ntot[k] = nloc[k] + L[k] + R[k];
      first[k] -= L[k];
    }
    n_cells = 1;
    for (int k = 0; k < d; ++k) n_cells *= ntot[k];
    n_ghost = n_cells - n_local;
    // number of interface cells, :303-315
    n_itfc = 1;
    for (int k = 0; k < d; ++k) {
      if (nloc[k] > L[k] + R[k]) n_itfc *= (nloc[k] - L[k] - R[k]);
      else { n_itfc = 0; break; }
    }
    n_itfc = n_local - n_itfc;
    // cell permutation interior -> interface (filled backwards) -> ghost (filled backwards), :316-351
    perm.assign(n_cells, -1);
    {
      int ighost = 0, ilocal = 0, iitfc = 0, ijk[3];
      for (int ic = 0; ic < n_cells; ++ic) {
        lex_to_ijk(d, ntot, ic, ijk);
        int p1 = 0; // 1-based like the reference, 0 = unset
        for (int k = 0; k < d; ++k)
          if ((L[k] == 1 && ijk[k] == 0) || (R[k] == 1 && ijk[k] == ntot[k] - 1)) {
            p1 = n_local + n_ghost - ighost; ++ighost; break;
          }
        if (p1 == 0)
          for (int k = 0; k < d; ++k)
            if ((L[k] == 1 && ijk[k] == 1) || (R[k] == 1 && ijk[k] == ntot[k] - 2)) {
              p1 = n_local - iitfc; ++iitfc; break;
            }
        if (p1 == 0) { ++ilocal; p1 = ilocal; }
        perm[ic] = p1 - 1;
      }
      assert(ilocal == n_local - n_itfc && iitfc == n_itfc && ighost == n_ghost);
    }
    lex_of.assign(n_cells, -1);
    for (int ic = 0; ic < n_cells; ++ic) lex_of[perm[ic]] = ic;
    // n-face types, :371-430
    ntypes = 0; tot_x_dir.clear(); glob_x_dir.clear(); type_first_face.clear();
    for (int f = 0; f < poly.num_n_faces(); ++f)
      if (poly.dim(f) < d && poly.anchor(f) == 0) {
        std::array<int, 3> t = {1, 1, 1}, g = {1, 1, 1};
        for (int k = 0; k < d; ++k) {
          g[k] = N[k] + 1 - poly.is_free(f, k);
          t[k] = ntot[k] + 1 - poly.is_free(f, k);
        }
        tot_x_dir.push_back(t); glob_x_dir.push_back(g); type_first_face.push_back(f);
        ++ntypes;
      }
    tot_off.assign(ntypes + 1, 0); glob_off.assign(ntypes + 1, 1);
    for (int t = 0; t < ntypes; ++t) {
      i64 gt = 1; int tt = 1;
      for (int k = 0; k < d; ++k) { gt *= glob_x_dir[t][k]; tt *= tot_x_dir[t][k]; }
      tot_off[t + 1] = tot_off[t] + tt;
      glob_off[t + 1] = glob_off[t] + gt;
    }
    n_vefs = tot_off[ntypes];
    n_vertices = tot_off[1];
    // cell -> vef list, :445-466
    lst_vefs.assign(size_t(n_cells) * nvef, -1);
    for (int ic = 0; ic < n_cells; ++ic) {
      int ijk[3], fijk[3];
      lex_to_ijk(d, ntot, ic, ijk);
      int itype = -1;
      for (int f = 0; f < nvef; ++f) {
        if (poly.anchor(f) == 0) ++itype;
        for (int k = 0; k < d; ++k) fijk[k] = (ijk[k] + poly.coord(f, k)) % tot_x_dir[itype][k];
        lst_vefs[size_t(perm[ic]) * nvef + f] = tot_off[itype] + int(ijk_to_lex(d, tot_x_dir[itype].data(), fijk));
      }
    }
    // cells gids / parts, :468-489
    cells_gid.assign(n_cells, 0); cells_part.assign(n_cells, 1);
    for (int ic = 0; ic < n_cells; ++ic) {
      int ijk[3], mp[3] = {0, 0, 0}, g[3] = {0, 0, 0};
      lex_to_ijk(d, ntot, ic, ijk);
      for (int k = 0; k < d; ++k) {
        if (L[k] == 1 && ijk[k] == 0) mp[k] = pijk[k] - 1;
        else if (R[k] == 1 && ijk[k] == ntot[k] - 1) mp[k] = pijk[k] + 1;
        else mp[k] = pijk[k];
        g[k] = first[k] + ijk[k];
      }
      cells_part[perm[ic]] = 1 + int(ijk_to_lex(d, P, mp));
      cells_gid[perm[ic]] = 1 + ijk_to_lex(d, N, g);
    }
    // vef gids, coordinates, boundary ids, :581-635
    vefs_gid.assign(n_vefs, 0); vef_bid.assign(n_vefs, -1); vcoord.assign(size_t(n_vertices) * 3, 0.0);
    for (int t = 0; t < ntypes; ++t) {
      int f = type_first_face[t];
      for (int v = tot_off[t]; v < tot_off[t + 1]; ++v) {
        int fijk[3];
        lex_to_ijk(d, tot_x_dir[t].data(), v - tot_off[t], fijk);
        for (int k = 0; k < d; ++k) fijk[k] += first[k];
        vefs_gid[v] = glob_off[t] + ijk_to_lex(d, glob_x_dir[t].data(), fijk);
        if (t == 0)
          for (int k = 0; k < d; ++k) {
            double c = double(fijk[k]) / double(N[k]);
            vcoord[size_t(v) * 3 + k] = (dom[k][1] - dom[k][0]) * c + dom[k][0];
          }
        int index = 0;
        for (int k = 0; k < d; ++k) {
          if (poly.is_free(f, k) == 0 && fijk[k] == 0) { /* bit stays 0 */ }
          else if (poly.is_free(f, k) == 0 && fijk[k] == glob_x_dir[t][k] - 1) index |= (1 << k);
          else index |= (1 << (d + k));
        }
        vef_bid[v] = poly.index_of_code[index];
      }
    }
    // a vef is ghost iff no local cell touches it (bst_vef_iterator is_ghost semantics)
    vef_ghost.assign(n_vefs, 1);
    for (int c = 0; c < n_local; ++c)
      for (int f = 0; f < nvef; ++f) vef_ghost[lst_vefs[size_t(c) * nvef + f]] = 0;
  }
  bool vef_at_boundary(int v) const { return vef_bid[v] != poly.num_n_faces(); }
  // vertex coordinates of local cell c (2^d vertices, polytope order = x fastest)
  void cell_vertices(int c, double* X /*[2^d][3]*/) const {
    for (int a = 0; a < (1 << d); ++a) {
      int v = lst_vefs[size_t(c) * nvef + a];
      for (int k = 0; k < 3; ++k) X[a * 3 + k] = vcoord[size_t(v) * 3 + k];
    }
  }
};

// ---------------------------------------------------------------------------
// 1-D Lagrange basis on equispaced nodes of [-1,1]   (sbm_polynomial.i90:60-108)
// Gauss-Legendre rule on [-1,1]                      (sbm_quadrature.i90:767-823)
// ---------------------------------------------------------------------------
static void lagrange_1d(int p, double x, double* val, double* der) {
  std::vector<double> xn(p + 1);
  for (int k = 0; k <= p; ++k) xn[k] = p == 0 ? 0.0 : -1.0 + 2.0 * k / p;
  for (int k = 0; k <= p; ++k) {
    double v = 1.0;
    for (int m = 0; m <= p; ++m) if (m != k) v *= (x - xn[m]) / (xn[k] - xn[m]);
    double dsum = 0.0;
    for (int n = 0; n <= p; ++n) {
      if (n == k) continue;
      double t = 1.0 / (xn[k] - xn[n]);
      for (int m = 0; m <= p; ++m) if (m != k && m != n) t *= (x - xn[m]) / (xn[k] - xn[m]);
      dsum += t;
    }
    val[k] = v; der[k] = dsum;
  }
}
static void gauss_legendre(int n, double* x, double* w) {
  for (int i = 0; i < n; ++i) {
    double z = std::cos(M_PI * (i + 0.75) / (n + 0.5)), pp = 0;
    for (int it = 0; it < 100; ++it) {
      double p1 = 1.0, p2 = 0.0;
      for (int j = 1; j <= n; ++j) { double p3 = p2; p2 = p1; p1 = ((2.0 * j - 1.0) * z * p2 - (j - 1.0) * p3) / j; }
      pp = n * (z * p1 - p2) / (z * z - 1.0);
      double dz = p1 / pp; z -= dz;
      if (std::fabs(dz) < 1e-16) break;
    }
    { double p1 = 1.0, p2 = 0.0;
      for (int j = 1; j <= n; ++j) { double p3 = p2; p2 = p1; p1 = ((2.0 * j - 1.0) * z * p2 - (j - 1.0) * p3) / j; }
      pp = n * (z * p1 - p2) / (z * z - 1.0); }
    x[n - 1 - i] = z; w[n - 1 - i] = 2.0 / ((1.0 - z * z) * pp * pp);
  }
  for (int i = 0; i < n / 2; ++i) { // symmetrise
    double xa = 0.5 * (x[n - 1 - i] - x[i]); x[i] = -xa; x[n - 1 - i] = xa;
    double wa = 0.5 * (w[i] + w[n - 1 - i]); w[i] = w[n - 1 - i] = wa;
  }
  if (n % 2) x[n / 2] = 0.0;
}

// ---------------------------------------------------------------------------
// Reference FE: hex Lagrangian of order p with ncomp components.
//   node / shape index = i + j(p+1) + k(p+1)^2 (sbm_polytope.i90:882-893,
//   sbm_hex_lagrangian_reference_fe.i90:301-316); vector fields = component-major
//   blocks (sbm_lagrangian_reference_fe.i90:1141-1149); own DOFs of an n-face = its
//   interior lattice nodes, lowest free direction fastest (sbm_polytope.i90:731-880),
//   for ncomp>1 [comp1 nodes.., comp2 nodes..] (sbm_lagrangian_reference_fe.i90:1250-1286).
//   Quadrature: p+1 GL points per direction (sbm_lagrangian_reference_fe.i90:1414-1423),
//   x index fastest, weight = product.
// ---------------------------------------------------------------------------
struct RefFE {
  int d, p, ncomp, n1, nnode, ndofs, nq;
  std::vector<std::vector<int>> own;   // per n-face (incl. cell): cell-local dof ids (0-based)
  std::vector<double> qx, qw;          // qx[nq*3], qw[nq]
  std::vector<double> phi;             // [nq][nnode]
  std::vector<double> dphi;            // [nq][nnode][3]  (reference gradients)
  std::vector<double> gphi, gdphi;     // geometry (order 1) basis at the q-points: [nq][2^d], [nq][2^d][3]
  std::vector<double> node_ref;        // reference coordinates of the nodes [nnode*3] in [-1,1]
  std::vector<double> gphi_nodes;      // geometry basis at the FE nodes [nnode][2^d]
  void create(const Polytope& poly, int order, int ncomponents) {
    d = poly.d; p = order; ncomp = ncomponents; n1 = p + 1;
    nnode = 1; for (int k = 0; k < d; ++k) nnode *= n1;
    ndofs = nnode * ncomp;
    own.assign(poly.num_n_faces(), {});
    for (int f = 0; f < poly.num_n_faces(); ++f) {
      std::vector<int> nodes;
      int lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};
      for (int k = 0; k < d; ++k) {
        if (poly.is_free(f, k)) { lo[k] = 1; hi[k] = p - 1; }
        else { lo[k] = hi[k] = poly.coord(f, k) * p; }
      }
      for (int k2 = lo[2]; k2 <= hi[2]; ++k2)
        for (int k1 = lo[1]; k1 <= hi[1]; ++k1)
          for (int k0 = lo[0]; k0 <= hi[0]; ++k0) nodes.push_back(k0 + n1 * (k1 + (d > 2 ? n1 * k2 : 0)));
      for (int c = 0; c < ncomp; ++c)
        for (int nd : nodes) own[f].push_back(c * nnode + nd);
    }
    int nq1 = p + 1;
    std::vector<double> x1(nq1), w1(nq1);
    gauss_legendre(nq1, x1.data(), w1.data());
    nq = 1; for (int k = 0; k < d; ++k) nq *= nq1;
    qx.assign(size_t(nq) * 3, 0.0); qw.assign(nq, 0.0);
    int nqd[3] = {nq1, nq1, d > 2 ? nq1 : 1}, nnd[3] = {n1, n1, d > 2 ? n1 : 1};
    for (int q = 0; q < nq; ++q) {
      int qi[3]; lex_to_ijk(d, nqd, q, qi);
      double w = 1.0;
      for (int k = 0; k < d; ++k) { qx[q * 3 + k] = x1[qi[k]]; w *= w1[qi[k]]; }
      qw[q] = w;
    }
    auto tabulate = [&](int ord, const double* pts, int npts, std::vector<double>& v, std::vector<double>* dv) {
      int m1 = ord + 1, nn = 1; for (int k = 0; k < d; ++k) nn *= m1;
      int mm[3] = {m1, m1, d > 2 ? m1 : 1};
      v.assign(size_t(npts) * nn, 0.0);
      if (dv) dv->assign(size_t(npts) * nn * 3, 0.0);
      std::vector<double> l(3 * m1, 1.0), dl(3 * m1, 0.0);
      for (int q = 0; q < npts; ++q) {
        for (int k = 0; k < d; ++k) lagrange_1d(ord, pts[q * 3 + k], &l[k * m1], &dl[k * m1]);
        for (int a = 0; a < nn; ++a) {
          int ai[3]; lex_to_ijk(d, mm, a, ai);
          double val = 1.0;
          for (int k = 0; k < d; ++k) val *= l[k * m1 + ai[k]];
          v[size_t(q) * nn + a] = val;
          if (dv)
            for (int k = 0; k < d; ++k) {
              double g = 1.0;
              for (int k2 = 0; k2 < d; ++k2) g *= (k2 == k) ? dl[k2 * m1 + ai[k2]] : l[k2 * m1 + ai[k2]];
              (*dv)[(size_t(q) * nn + a) * 3 + k] = g;
            }
        }
      }
    };
    tabulate(p, qx.data(), nq, phi, &dphi);
    tabulate(1, qx.data(), nq, gphi, &gdphi);
    node_ref.assign(size_t(nnode) * 3, 0.0);
    for (int a = 0; a < nnode; ++a) {
      int ai[3]; lex_to_ijk(d, nnd, a, ai);
      for (int k = 0; k < d; ++k) node_ref[a * 3 + k] = -1.0 + 2.0 * ai[k] / p;
    }
    tabulate(1, node_ref.data(), nnode, gphi_nodes, nullptr);
  }
};

// ---------------------------------------------------------------------------
// analytic data ("functions")
// ---------------------------------------------------------------------------
enum { FUNC_ZERO = 0, FUNC_LINEAR = 1, FUNC_SHARP_WAVE = 2 };
struct Func {
  int kind = FUNC_LINEAR, d = 3;
  double alpha = 200.0, r = 0.7, c[3] = {-0.05, -0.05, -0.05};
  // exact solution / boundary function
  double u(const double* x) const {
    if (kind == FUNC_LINEAR) { double s = 0; for (int k = 0; k < d; ++k) s += x[k]; return s; } // par_test_poisson_analytical_functions.f90:71-124
    if (kind == FUNC_SHARP_WAVE) { // tutorial_01_..._modules.f90:462-484
      double t = 0; for (int k = 0; k < d; ++k) t += (x[k] - c[k]) * (x[k] - c[k]);
      return std::atan(alpha * (std::sqrt(t) - r));
    }
    return 0.0;
  }
  void grad_u(const double* x, double* g) const {
    g[0] = g[1] = g[2] = 0.0;
    if (kind == FUNC_LINEAR) { for (int k = 0; k < d; ++k) g[k] = 1.0; return; }
    if (kind == FUNC_SHARP_WAVE) { // tutorial_01_..._modules.f90:486-527
      double t1 = 0; for (int k = 0; k < d; ++k) t1 += (x[k] - c[k]) * (x[k] - c[k]);
      double s = std::sqrt(t1), t2 = s - r, den = s * (alpha * alpha * t2 * t2 + 1.0);
      for (int k = 0; k < d; ++k) g[k] = alpha * (x[k] - c[k]) / den;
    }
  }
  double f(const double* x) const { // source term
    if (kind != FUNC_SHARP_WAVE) return 0.0;
    // tutorial_01_..._modules.f90:395-458
    double a = alpha, a2 = a * a, a3 = a2 * a, t1 = 0, dx2[3];
    for (int k = 0; k < d; ++k) { dx2[k] = (x[k] - c[k]) * (x[k] - c[k]); t1 += dx2[k]; }
    double s = std::sqrt(t1), t2 = s - r, t22 = t2 * t2;
    double den1 = s * (a2 * t22 + 1.0), den2 = std::pow(t1, 1.5) * (a2 * t22 + 1.0),
           den3 = t1 * (a2 * t22 + 1.0) * (a2 * t22 + 1.0);
    double res = 0;
    for (int k = 0; k < d; ++k) res -= a / den1 - (a * dx2[k]) / den2 - (2.0 * a3 * dx2[k] * t2) / den3;
    return res;
  }
};

enum { KIND_POISSON = 0, KIND_VECTOR_LAPLACIAN = 1, KIND_ELASTICITY = 2 };

// ---------------------------------------------------------------------------
// The FE space + linear system of one part (or of the serial problem)
// ---------------------------------------------------------------------------
struct Space {
  Mesh m;
  RefFE fe;
  unsigned dir_mask = 0xFFFFFFFFu;     // bit (b-1) set => boundary id b carries a strong Dirichlet condition
  std::vector<int> gids;               // lst_dofs_gids [n_cells*ndofs], 1-based signed (fe_space.f90:636-637)
  int n_dirichlet = 0, n_free = 0, n_free_ghost = 0;
  // linear system
  bool sym_storage = false;
  std::vector<i64> irp;                // 0-based offsets here, size n+1 (exported 1-based on request)
  std::vector<int> ja;                 // 1-based column ids like the reference
  std::vector<double> val, rhs, fixed;
  std::vector<double> dinv;            // Jacobi inverse diagonal
  // par data (build_par_layer): dof_import_t (LinearAlgebra/Vector/dof_import.f90:49-90); part ids 1-based,
  // ptrs 0-based offsets (reference ptrs = these + 1), pack/unpack_idx 1-based local dof ids like the reference
  std::vector<double> weight;
  std::vector<int> list_rcv, rcv_ptrs, unpack_idx, list_snd, snd_ptrs, pack_idx;
  std::vector<int> perm_old2new;       // 1-based -> 1-based, interior first then interface
  int n_interior = 0, n_interface = 0;
  double t_integrate = 0, t_pattern = 0;
  int pattern_built = -1;   // -1 none, 0 full, 1 symmetric storage
  int nthreads = 1;

  bool vef_dirichlet(int v) const {
    int b = m.vef_bid[v];
    if (b == m.poly.num_n_faces()) return false;
    return (dir_mask >> (b - 1)) & 1u;
  }

  void create(int d, const int* N, const int* P, int part, int order, int ncomp, unsigned mask) {
    m.create(d, N, P, part);
    fe.create(m.poly, order, ncomp);
    dir_mask = mask;
    gids.assign(size_t(m.n_cells) * fe.ndofs, 0);
    setup_strong_dirichlet();
    list_dofs();
    n_interior = n_free; n_interface = 0;
    if (m.par) build_par_layer();
  }

  // serial_fe_space_set_up_strong_dirichlet_bcs  Integration/FESpace/sbm_serial_fe_space.i90:525-651
  //   + ..._on_vef_and_field :653-709, generate_own_dofs_vef_component_wise sbm_fe_cell_iterator.i90:686-732,
  //   fetch_own_dofs_vef_from_source_fe :803-856 (identity permutation on an oriented structured mesh)
  void setup_strong_dirichlet() {
    const int nvef = m.nvef, nd = fe.ndofs;
    n_dirichlet = 0;
    std::vector<int> vis_cell(m.n_vefs, -1), vis_ivef(m.n_vefs, -1);
    for (int c = 0; c < m.n_cells; ++c)          // the cell iterator runs over local AND ghost cells
      for (int f = 0; f < nvef; ++f) {
        int v = m.lst_vefs[size_t(c) * nvef + f];
        if (!vef_dirichlet(v)) continue;         // vef%get_set_id() /= 0 and the condition flags every component
        int prev = n_dirichlet;
        if (vis_cell[v] == -1) {
          for (int l : fe.own[f]) {
            int& g = gids[size_t(c) * nd + l];
            if (g >= 0) { ++n_dirichlet; g = -n_dirichlet; }
          }
        }
        if (vis_cell[v] != -1) {
          int sc = vis_cell[v], sf = vis_ivef[v];
          for (size_t i = 0; i < fe.own[f].size(); ++i)
            gids[size_t(c) * nd + fe.own[f][i]] = gids[size_t(sc) * nd + fe.own[sf][i]];
        }
        if (vis_cell[v] == -1 && prev < n_dirichlet) { vis_cell[v] = c; vis_ivef[v] = f; }
      }
  }

  // serial_fe_space_list_dofs  Integration/FESpace/sbm_serial_fe_space.i90:2623-2875
  //   generate_own_dofs_cell sbm_fe_cell_iterator.i90:410-454 ; generate_own_dofs_vef :507-544
  void list_dofs() {
    const int nvef = m.nvef, nd = fe.ndofs;
    std::vector<int> own_cell(m.n_vefs, -1), own_ivef(m.n_vefs, -1);
    int cur = 0;
    for (int c = 0; c < m.n_cells; ++c) {        // first loop: every cell, local ones first
      bool local = c < m.n_local;
      if (local)
        for (int l : fe.own[nvef]) gids[size_t(c) * nd + l] = ++cur;   // interior DOFs, unconditional
      for (int f = 0; f < nvef; ++f) {
        int v = m.lst_vefs[size_t(c) * nvef + f];
        if (m.vef_ghost[v]) continue;
        if (own_cell[v] == -1) {
          if (local) {
            int prev = cur;
            for (int l : fe.own[f]) {
              int& g = gids[size_t(c) * nd + l];
              if (g >= 0) g = ++cur;
            }
            if (prev < cur) { own_cell[v] = c; own_ivef[v] = f; }
          }
        } else {
          int sc = own_cell[v], sf = own_ivef[v];
          for (size_t i = 0; i < fe.own[f].size(); ++i)
            gids[size_t(c) * nd + fe.own[f][i]] = gids[size_t(sc) * nd + fe.own[sf][i]];
        }
      }
    }
    n_free = cur;
    int gcur = cur;
    for (int c = m.n_local; c < m.n_cells; ++c) { // second loop: ghost cells, :2812-2859
      for (int l : fe.own[nvef]) gids[size_t(c) * nd + l] = ++gcur;
      for (int f = 0; f < nvef; ++f) {
        int v = m.lst_vefs[size_t(c) * nvef + f];
        if (!m.vef_ghost[v]) continue;
        if (own_cell[v] == -1) {
          int prev = gcur;
          for (int l : fe.own[f]) {
            int& g = gids[size_t(c) * nd + l];
            if (g >= 0) g = ++gcur;
          }
          if (prev < gcur) { own_cell[v] = c; own_ivef[v] = f; }
        } else {
          int sc = own_cell[v], sf = own_ivef[v];
          for (size_t i = 0; i < fe.own[f].size(); ++i)
            gids[size_t(c) * nd + fe.own[f][i]] = gids[size_t(sc) * nd + fe.own[sf][i]];
        }
      }
    }
    n_free_ghost = gcur - cur;
  }


  // par_fe_space_t: interface couplings, owner decision, dof_import and the interior-then-interface renumbering.
  //   compute_raw_interface_data_by_continuity   Integration/FESpace/sbm_par_fe_space.i90:1877-2033
  //   raw_interface_data_by_continuity_decide_owner                                      :2737-2851
  //   generate_non_consecutive_dof_ggid                                                  :1103-1118
  //   dof_import_create / fill_first_I_then_G_renumbering / renum_dofs  LinearAlgebra/Vector/dof_import.f90:102-213,365-421
  //   renum_dofs_first_interior_then_interface                          sbm_par_fe_space.i90:449-474
  //   cells around a vef = cells in increasing local id (sbm_base_static_triangulation.i90:697-741);
  //   interface vef = touched by local AND ghost cells (:748-784); itfc vefs visited in increasing vef id (sbm_triangulation.i90:201-218)
  void build_par_layer() {
    const int nvef = m.nvef, nd = fe.ndofs;
    const int mypart = m.part_id;
    int nparts = 1; for (int k = 0; k < m.d; ++k) nparts *= m.P[k];
    // cells around each vef, increasing cell id
    std::vector<int> cptr(m.n_vefs + 1, 0);
    for (int c = 0; c < m.n_cells; ++c) for (int f = 0; f < nvef; ++f) cptr[m.lst_vefs[size_t(c) * nvef + f] + 1]++;
    for (int v = 0; v < m.n_vefs; ++v) cptr[v + 1] += cptr[v];
    std::vector<int> ccell(cptr[m.n_vefs]), cface(cptr[m.n_vefs]), pos(cptr.begin(), cptr.end() - 1);
    for (int c = 0; c < m.n_cells; ++c)
      for (int f = 0; f < nvef; ++f) { int v = m.lst_vefs[size_t(c) * nvef + f]; ccell[pos[v]] = c; cface[pos[v]] = f; pos[v]++; }
    struct Col { i64 part, flag, ggid; int dof, maxpart; };
    std::vector<Col> cols;
    std::vector<int> local_idof(n_free + 1, 0), lst;
    std::vector<int> parts_visited;
    for (int v = 0; v < m.n_vefs; ++v) {
      bool has_local = false, has_ghost = false;
      for (int k = cptr[v]; k < cptr[v + 1]; ++k) { if (ccell[k] < m.n_local) has_local = true; else has_ghost = true; }
      if (!(has_local && has_ghost)) continue;                 // vef%is_at_interface()
      parts_visited.clear(); lst.clear();
      int max_part = mypart; i64 max_ggid = 0;
      for (int k = cptr[v]; k < cptr[v + 1]; ++k) {
        int c = ccell[k], f = cface[k];
        int part = m.cells_part[c]; i64 ggid = m.cells_gid[c];
        for (int l : fe.own[f]) {
          int g = gids[size_t(c) * nd + l];
          if (!(g > 0 && g <= n_free)) continue;                // is_free_dof
          if (ggid >= max_ggid) {
            max_ggid = ggid;
            if (local_idof[g] == 0) lst.push_back(g);
            local_idof[g] = l + 1;
          }
          if (part != mypart && std::find(parts_visited.begin(), parts_visited.end(), part) == parts_visited.end()) {
            parts_visited.push_back(part);
            max_part = std::max(max_part, part);
          }
        }
      }
      for (int g : lst)
        for (int part : parts_visited) {
          Col cc; cc.part = part; cc.flag = i64(parts_visited.size()) + 1; cc.dof = g; cc.maxpart = max_part;
          cc.ggid = (max_ggid - 1) + (i64(local_idof[g] - 1) << 44);   // cell_ggid_shift = 44, field_id = 1 (fe_space_parameters.f90:137-139)
          cols.push_back(cc);
        }
    }
    // sort by (neighbour part, #parts around + 1, non-consecutive global dof id): sort_array_cols_by_row_section, Tools/sort.f90:118-141
    std::stable_sort(cols.begin(), cols.end(), [](const Col& a, const Col& b) {
      if (a.part != b.part) return a.part < b.part;
      if (a.flag != b.flag) return a.flag < b.flag;
      return a.ggid < b.ggid;
    });
    const i64 OWNER = -2, NON_OWNER = -3, UNCOUPLED = -1;
    size_t col = 0;
    while (col < cols.size()) {                                  // :2797-2845
      if (cols[col].flag == 2) {
        size_t start = col;
        while (col < cols.size() && cols[col].flag == 2 && cols[col].part == cols[start].part) ++col;
        size_t size = col - start;
        bool i_am_max = mypart == cols[start].maxpart;
        for (size_t j = start; j < start + size / 2; ++j) cols[j].flag = i_am_max ? OWNER : NON_OWNER;
        for (size_t j = start + size / 2; j < start + size; ++j) cols[j].flag = i_am_max ? NON_OWNER : OWNER;
      } else {
        if (mypart == cols[col].maxpart) cols[col].flag = OWNER;
        else cols[col].flag = (cols[col].part == cols[col].maxpart) ? NON_OWNER : UNCOUPLED;
        ++col;
      }
    }
    // dof_import_create
    weight.assign(n_free, 1.0);
    list_rcv.clear(); list_snd.clear();
    std::vector<int> rcv_vis(nparts + 1, -1), snd_vis(nparts + 1, -1), rcnt, scnt;
    for (const Col& c : cols) {
      if (c.flag == OWNER) { if (rcv_vis[c.part] < 0) { rcv_vis[c.part] = int(list_rcv.size()); list_rcv.push_back(int(c.part)); rcnt.push_back(0); } rcnt[rcv_vis[c.part]]++; }
      else if (c.flag == NON_OWNER) { if (snd_vis[c.part] < 0) { snd_vis[c.part] = int(list_snd.size()); list_snd.push_back(int(c.part)); scnt.push_back(0); } scnt[snd_vis[c.part]]++; }
    }
    rcv_ptrs.assign(list_rcv.size() + 1, 0); snd_ptrs.assign(list_snd.size() + 1, 0);
    for (size_t i = 0; i < list_rcv.size(); ++i) rcv_ptrs[i + 1] = rcv_ptrs[i] + rcnt[i];
    for (size_t i = 0; i < list_snd.size(); ++i) snd_ptrs[i + 1] = snd_ptrs[i] + scnt[i];
    unpack_idx.assign(rcv_ptrs.back(), 0); pack_idx.assign(snd_ptrs.back(), 0);
    std::vector<int> rp(rcv_ptrs.begin(), rcv_ptrs.end()), sp(snd_ptrs.begin(), snd_ptrs.end());
    std::vector<char> visited(n_free + 1, 0);
    n_interface = 0;
    for (const Col& c : cols) {
      if (c.flag == OWNER) unpack_idx[rp[rcv_vis[c.part]]++] = c.dof;
      else if (c.flag == NON_OWNER) { pack_idx[sp[snd_vis[c.part]]++] = c.dof; weight[c.dof - 1] = 0.0; }
      if (!visited[c.dof]) { visited[c.dof] = 1; ++n_interface; }
    }
    n_interior = n_free - n_interface;
    // fill_first_I_then_G_renumbering
    perm_old2new.assign(n_free + 1, -1);
    int cur = n_interior + 1;
    for (int g : pack_idx) if (perm_old2new[g] == -1) perm_old2new[g] = cur++;
    for (int g : unpack_idx) if (perm_old2new[g] == -1) perm_old2new[g] = cur++;
    cur = 1;
    for (int g = 1; g <= n_free; ++g) if (perm_old2new[g] == -1) perm_old2new[g] = cur++;
    // renum_dofs_block (only free dofs: sbm_fe_cell_iterator.i90:907-933) and dof_import%renum_dofs
    for (auto& g : gids) if (g > 0 && g <= n_free) g = perm_old2new[g];
    for (auto& g : pack_idx) g = perm_old2new[g];
    for (auto& g : unpack_idx) g = perm_old2new[g];
    std::vector<double> w2(n_free);
    for (int g = 1; g <= n_free; ++g) w2[perm_old2new[g] - 1] = weight[g - 1];
    weight.swap(w2);
  }

  // physical coordinates of the FE nodes of a cell through the order-1 geometry map
  void node_coords(int c, double* xn /*[nnode][3]*/) const {
    double X[8 * 3]; m.cell_vertices(c, X);
    int nv = 1 << m.d;
    for (int a = 0; a < fe.nnode; ++a)
      for (int k = 0; k < 3; ++k) {
        double s = 0; for (int v = 0; v < nv; ++v) s += fe.gphi_nodes[size_t(a) * nv + v] * X[v * 3 + k];
        xn[a * 3 + k] = s;
      }
  }

  // interpolate_dirichlet_values: fixed_dof_values(-gid) = g(x_node)  sbm_serial_fe_space.i90:997-1176
  void interpolate_dirichlet(const Func& fn) {
    fixed.assign(n_dirichlet, 0.0);
    std::vector<double> xn(size_t(fe.nnode) * 3);
    for (int c = 0; c < m.n_cells; ++c) {
      bool any = false;
      for (int l = 0; l < fe.ndofs; ++l) if (gids[size_t(c) * fe.ndofs + l] < 0) { any = true; break; }
      if (!any) continue;
      node_coords(c, xn.data());
      for (int l = 0; l < fe.ndofs; ++l) {
        int g = gids[size_t(c) * fe.ndofs + l];
        if (g < 0) fixed[-g - 1] = fn.u(&xn[size_t(l % fe.nnode) * 3]);
      }
    }
  }

  // cell_map_update + apply_cell_map_to_interpolation for one cell:
  //   J(q) = sum_a x_a (x) grad N_a(q)  (sbm_cell_map.i90:272-296), invmtx (sbm_cell_integrator.i90:448-561),
  //   grad phi = J^-T grad_ref phi   (sbm_lagrangian_reference_fe.i90:1333-1347)
  void cell_geometry(int c, double* detj /*[nq]*/, double* grad /*[nq][nnode][3]*/, double* xq /*[nq][3]*/) const {
    const int d = m.d, nv = 1 << d;
    double X[8 * 3]; m.cell_vertices(c, X);
    for (int q = 0; q < fe.nq; ++q) {
      double J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
      for (int k = 0; k < 3; ++k) xq[q * 3 + k] = 0.0;
      for (int a = 0; a < nv; ++a) {
        for (int i = 0; i < d; ++i)
          for (int j = 0; j < d; ++j) J[i][j] += X[a * 3 + j] * fe.gdphi[(size_t(q) * nv + a) * 3 + i]; // J(i,j)= d x_j / d xi_i
        for (int k = 0; k < 3; ++k) xq[q * 3 + k] += fe.gphi[size_t(q) * nv + a] * X[a * 3 + k];
      }
      double inv[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, det;
      if (d == 2) {
        det = J[0][0] * J[1][1] - J[0][1] * J[1][0];
        inv[0][0] = J[1][1] / det; inv[1][1] = J[0][0] / det; inv[0][1] = -J[0][1] / det; inv[1][0] = -J[1][0] / det;
      } else {
        double c00 = J[1][1] * J[2][2] - J[1][2] * J[2][1], c01 = J[1][2] * J[2][0] - J[1][0] * J[2][2],
               c02 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
        det = J[0][0] * c00 + J[0][1] * c01 + J[0][2] * c02;
        inv[0][0] = c00 / det; inv[1][0] = c01 / det; inv[2][0] = c02 / det;
        inv[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) / det;
        inv[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) / det;
        inv[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) / det;
        inv[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) / det;
        inv[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) / det;
        inv[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) / det;
      }
      detj[q] = std::fabs(det);
      // with J(i,j) = d x_j/d xi_i : grad_x phi = J^{-1} grad_xi phi
      for (int a = 0; a < fe.nnode; ++a)
        for (int i = 0; i < 3; ++i) {
          double s = 0;
          for (int j = 0; j < d; ++j) s += inv[i][j] * fe.dphi[(size_t(q) * fe.nnode + a) * 3 + j];
          grad[(size_t(q) * fe.nnode + a) * 3 + i] = i < d ? s : 0.0;
        }
    }
  }

  // the user integrands:
  //  Poisson   Tutorials/tutorial_01_.../modules/tutorial_01_..._modules.f90:71-137 and
  //            Tests/Par/par_test_poisson/modules/par_test_poisson_discrete_integration.f90:62-142
  //  vector Laplacian  Tests/Serial/test_poisson/modules/vector_poisson_discrete_integration.f90:62-140
  //  elasticity  Tests/Par/par_test_pb_bddc_linear_elasticity/modules/sbm_irreducible_discrete_integration.i90:93-125
  void element(int c, int kind, const Func& fn, const double* mat /*lambda,mu*/, double* elmat /*col-major nd x nd*/, double* elvec,
               double* detj, double* grad, double* xq) const {
    const int nd = fe.ndofs, nn = fe.nnode;
    cell_geometry(c, detj, grad, xq);
    std::fill(elmat, elmat + size_t(nd) * nd, 0.0);
    std::fill(elvec, elvec + nd, 0.0);
    for (int q = 0; q < fe.nq; ++q) {
      double factor = detj[q] * fe.qw[q];
      const double* g = &grad[size_t(q) * nn * 3];
      if (kind == KIND_POISSON) {
        for (int i = 0; i < nd; ++i)
          for (int j = 0; j < nd; ++j)
            elmat[i + size_t(j) * nd] += factor * (g[j * 3] * g[i * 3] + g[j * 3 + 1] * g[i * 3 + 1] + g[j * 3 + 2] * g[i * 3 + 2]);
        double f = fn.f(&xq[q * 3]);
        for (int i = 0; i < nd; ++i) elvec[i] += factor * f * fe.phi[size_t(q) * nn + i];
      } else if (kind == KIND_VECTOR_LAPLACIAN) {
        for (int i = 0; i < nd; ++i)
          for (int j = 0; j < nd; ++j) {
            int ci = i / nn, cj = j / nn, ni = i % nn, nj = j % nn;
            double dc = (ci == cj) ? (g[nj * 3] * g[ni * 3] + g[nj * 3 + 1] * g[ni * 3 + 1] + g[nj * 3 + 2] * g[ni * 3 + 2]) : 0.0;
            elmat[i + size_t(j) * nd] += factor * dc;
          }
        // source term is zero in vector_poisson_analytical_functions.f90:69-82
      } else { // isotropic linear elasticity: sigma = lambda tr(eps) I + 2 mu eps
        double lambda = mat[0], mu = mat[1];
        for (int i = 0; i < nd; ++i)
          for (int j = 0; j < nd; ++j) {
            int ci = i / nn, cj = j / nn, ni = i % nn, nj = j % nn;
            // eps(phi_j e_cj) : sigma(phi_i e_ci)
            double gi[3] = {g[ni * 3], g[ni * 3 + 1], g[ni * 3 + 2]}, gj[3] = {g[nj * 3], g[nj * 3 + 1], g[nj * 3 + 2]};
            double dotg = gi[0] * gj[0] + gi[1] * gj[1] + gi[2] * gj[2];
            double v = lambda * gi[ci] * gj[cj] + mu * ((ci == cj ? dotg : 0.0) + gi[cj] * gj[ci]);
            elmat[i + size_t(j) * nd] += factor * v;
          }
      }
    }
  }

  // fe_cell_iterator_impose_strong_dirichlet_bcs_conforming  sbm_fe_cell_iterator.i90:1610-1637
  void impose_dirichlet(int c, const double* elmat, double* elvec) const {
    const int nd = fe.ndofs;
    for (int j = 0; j < nd; ++j) {
      int g = gids[size_t(c) * nd + j];
      if (g < 0) {
        double gv = fixed[-g - 1];
        for (int i = 0; i < nd; ++i) elvec[i] -= elmat[i + size_t(j) * nd] * gv;
      }
    }
  }

  // Assembly.  route 0 = reference-faithful COO append -> stable sort -> sum duplicates in insertion order
  //   (base_sparse_matrix.f90:3469-3513, 4479-4628; csr_sparse_matrix.f90:416-479);
  // route 1 = direct: graph first, then accumulate in the same cell order (bitwise identical sums,
  //   needed beyond the int32/COO memory limits).
  // Matrix rows/cols are the LOCAL free dofs 1..n_free; gids > n_free (ghost-only) are dropped
  // (base_sparse_matrix.f90:3469-3513 filters ia/ja outside 1..n), Dirichlet (<1) dropped.
  void assemble(int kind, const Func& fn, const double* mat, bool sym, int route) {
    sym_storage = sym;
    const int nd = fe.ndofs, n = n_free;
    interpolate_dirichlet(fn);
    rhs.assign(n, 0.0);
    std::vector<double> elmat(size_t(nd) * nd), elvec(nd), detj(fe.nq), grad(size_t(fe.nq) * fe.nnode * 3), xq(size_t(fe.nq) * 3);
    if (route == 0) {
      std::vector<int> ci, cj; std::vector<double> cv;
      for (int c = 0; c < m.n_local; ++c) {
        element(c, kind, fn, mat, elmat.data(), elvec.data(), detj.data(), grad.data(), xq.data());
        impose_dirichlet(c, elmat.data(), elvec.data());
        const int* g = &gids[size_t(c) * nd];
        for (int j = 0; j < nd; ++j) {         // coo append: loops j (columns) outer, i inner
          if (g[j] < 1 || g[j] > n) continue;
          for (int i = 0; i < nd; ++i) {
            if (g[i] < 1 || g[i] > n) continue;
            if (sym && g[i] > g[j]) continue;
            ci.push_back(g[i]); cj.push_back(g[j]); cv.push_back(elmat[i + size_t(j) * nd]);
          }
        }
        for (int i = 0; i < nd; ++i) if (g[i] >= 1 && g[i] <= n) rhs[g[i] - 1] += elvec[i]; // serial_scalar_array.f90:247-262
      }
      std::vector<size_t> idx(ci.size());
      std::iota(idx.begin(), idx.end(), size_t(0));
      std::stable_sort(idx.begin(), idx.end(), [&](size_t a, size_t b) {
        if (ci[a] != ci[b]) return ci[a] < ci[b];
        return cj[a] < cj[b];
      });
      irp.assign(n + 1, 0); ja.clear(); val.clear();
      int lr = -1, lc = -1;
      for (size_t k : idx) {
        if (ci[k] == lr && cj[k] == lc) { val.back() += cv[k]; continue; }
        lr = ci[k]; lc = cj[k];
        ja.push_back(lc); val.push_back(cv[k]); irp[lr]++;
      }
      for (int i = 0; i < n; ++i) irp[i + 1] += irp[i];
      pattern_built = -1;
    } else {
      if (pattern_built != int(sym)) build_pattern(sym);
      std::fill(val.begin(), val.end(), 0.0);
      for (int c = 0; c < m.n_local; ++c) {
        element(c, kind, fn, mat, elmat.data(), elvec.data(), detj.data(), grad.data(), xq.data());
        impose_dirichlet(c, elmat.data(), elvec.data());
        const int* g = &gids[size_t(c) * nd];
        for (int i = 0; i < nd; ++i) {
          if (g[i] < 1 || g[i] > n) continue;
          const int* rb = &ja[irp[g[i] - 1]]; const int* re = &ja[irp[g[i]]];
          for (int j = 0; j < nd; ++j) {
            if (g[j] < 1 || g[j] > n) continue;
            if (sym && g[i] > g[j]) continue;
            const int* it = std::lower_bound(rb, re, g[j]);
            val[it - &ja[0]] += elmat[i + size_t(j) * nd];
          }
          rhs[g[i] - 1] += elvec[i];
        }
      }
    }
  }

  void build_pattern(bool sym) {
    sym_storage = sym; pattern_built = int(sym);
    const int nd = fe.ndofs, n = n_free;
    // dof -> cells adjacency
    std::vector<i64> dptr(n + 1, 0);
    for (int c = 0; c < m.n_local; ++c)
      for (int l = 0; l < nd; ++l) { int g = gids[size_t(c) * nd + l]; if (g >= 1 && g <= n) dptr[g]++; }
    for (int i = 0; i < n; ++i) dptr[i + 1] += dptr[i];
    std::vector<int> dcell(dptr[n]); std::vector<i64> pos(dptr.begin(), dptr.end() - 1);
    for (int c = 0; c < m.n_local; ++c)
      for (int l = 0; l < nd; ++l) { int g = gids[size_t(c) * nd + l]; if (g >= 1 && g <= n) dcell[pos[g - 1]++] = c; }
    irp.assign(n + 1, 0);
    std::vector<int> tmp;
    for (int pass = 0; pass < 2; ++pass) {
      for (int r = 0; r < n; ++r) {
        tmp.clear();
        for (i64 k = dptr[r]; k < dptr[r + 1]; ++k) {
          const int* g = &gids[size_t(dcell[k]) * nd];
          for (int l = 0; l < nd; ++l) if (g[l] >= 1 && g[l] <= n && (!sym || g[l] >= r + 1)) tmp.push_back(g[l]);
        }
        std::sort(tmp.begin(), tmp.end());
        tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
        if (pass == 0) irp[r + 1] = irp[r] + i64(tmp.size());
        else std::copy(tmp.begin(), tmp.end(), ja.begin() + irp[r]);
      }
      if (pass == 0) { ja.assign(irp[n], 0); val.assign(irp[n], 0.0); }
    }
  }

  // csr matvec / matvec_symmetric_storage   LinearAlgebra/SparseMatrix/csr_sparse_matrix.f90:545-568,597-623
  void spmv(const double* x, double* y) const {
    const int n = n_free;
    if (!sym_storage) {
      parallel_for(n, nthreads, [&](int lo, int hi) {
        for (int i = lo; i < hi; ++i) {
          double s = 0.0;
          for (i64 k = irp[i]; k < irp[i + 1]; ++k) s += val[k] * x[ja[k] - 1];
          y[i] = s;
        }
      });
    } else {
      std::fill(y, y + n, 0.0);
      for (int i = 0; i < n; ++i) {
        i64 k0 = irp[i];
        if (k0 == irp[i + 1]) continue;
        // the diagonal is the first stored entry of each row
        y[i] += val[k0] * x[ja[k0] - 1];
        for (i64 k = k0 + 1; k < irp[i + 1]; ++k) {
          y[i] += val[k] * x[ja[k] - 1];
          y[ja[k] - 1] += val[k] * x[i];
        }
      }
    }
  }

  // csr extract_diagonal (csr_sparse_matrix.f90:4231-4256) + entrywise_invert (jacobi_preconditioner.f90:371-386)
  void jacobi_setup() {
    const int n = n_free;
    dinv.assign(n, 0.0);
    for (int i = 0; i < n; ++i) {
      const int* rb = &ja[irp[i]]; const int* re = &ja[irp[i + 1]];
      const int* it = std::lower_bound(rb, re, i + 1);
      if (it != re && *it == i + 1) dinv[i] = val[it - &ja[0]];
    }
    for (int i = 0; i < n; ++i) dinv[i] = 1.0 / dinv[i];
  }

  static double dot(const double* a, const double* b, int n) { double s = 0; for (int i = 0; i < n; ++i) s += a[i] * b[i]; return s; }

  // cg_solve_body  LinearAlgebra/IterativeLinearSolver/cg.f90:101-238; init_convergence_data :279-321;
  // update_convergence_data_and_evaluate_stopping_criteria :323-407.  stop (iterative_solver_parameters.f90:37-46):
  //  0 res_res (CG default)  1 res_rhs  2 res_nrmgiven_rhs_nrmgiven  3 res_nrmgiven_res_nrmgiven  4 delta_rhs  5 delta_delta
  //  6 delta_rhs_and_res_res  7 delta_rhs_and_res_rhs  8 delta_delta_and_res_res  9 delta_delta_and_res_rhs
  int cg(int prec /*0 identity, 1 jacobi*/, double rtol, double atol, int maxit, int stop, const double* b, double* x,
         int* iters, int* converged, double* err, double* hist, double* t_spmv_acc) {
    const int n = n_free;
    std::vector<double> r(n), z(n), p(n), Ap(n);
    auto applyM = [&](const double* in, double* out) {
      if (prec == 1) for (int i = 0; i < n; ++i) out[i] = dinv[i] * in[i];
      else std::copy(in, in + n, out);
    };
    auto nrm2 = [&](const double* v) { return std::sqrt(dot(v, v, n)); };
    std::fill(x, x + n, 0.0);                       // x <- initial_solution (zero)
    double b_nrm_M = 0.0;                           // |b|_inv(M), cg.f90:144-154
    if (stop == 2) { applyM(b, r.data()); b_nrm_M = std::sqrt(dot(b, r.data(), n)); }
    spmv(x, r.data());                              // r = A x
    for (int i = 0; i < n; ++i) r[i] = b[i] - r[i]; // r = b - r
    applyM(r.data(), z.data());
    double r_z = dot(r.data(), z.data(), n);
    double r_nrm_M = std::sqrt(r_z);
    p = z;
    // init_convergence_data
    double tol = 0.0, tol_extra = 0.0;
    switch (stop) {
      case 4: case 1: tol = rtol * nrm2(b) + atol; break;
      case 7: case 9: tol = tol_extra = rtol * nrm2(b) + atol; break;
      case 0: tol = rtol * nrm2(r.data()) + atol; break;
      case 8: tol = tol_extra = rtol * nrm2(r.data()) + atol; break;
      case 6: tol = rtol * nrm2(b) + atol; tol_extra = rtol * nrm2(r.data()) + atol; break;
      case 2: tol = rtol * b_nrm_M + atol; break;
      case 3: tol = rtol * r_nrm_M + atol; break;
      default: break;
    }
    int it = 0; bool conv = false; double e = 0;
    for (;;) {
      ++it;
      spmv(p.data(), Ap.data());
      double App = dot(Ap.data(), p.data(), n);
      double alpha = (App != 0.0) ? r_z / App : 0.0;
      for (int i = 0; i < n; ++i) x[i] += alpha * p[i];
      for (int i = 0; i < n; ++i) r[i] -= alpha * Ap[i];
      // update_convergence_data_and_evaluate_stopping_criteria(r, r_nrm_M, alpha, p)
      const bool delta = stop >= 4;
      if (it == 1 && delta) {
        e = alpha * nrm2(p.data());
        if (stop == 5 || stop == 8 || stop == 9) tol = rtol * e + atol;
      }
      if (stop == 0 || stop == 1) e = nrm2(r.data());
      else if (delta) { if (it != 1) e = alpha * nrm2(p.data()); }
      else e = r_nrm_M;                              // NOTE: still the M-norm of the previous residual (updated below, :228-231)
      if (hist) hist[it - 1] = e;
      conv = e <= tol;
      if (stop >= 6 && conv) conv = nrm2(r.data()) <= tol_extra;
      if (conv || it >= maxit) break;
      applyM(r.data(), z.data());
      double beta = 1.0 / r_z;
      r_z = dot(r.data(), z.data(), n);
      beta *= r_z;
      r_nrm_M = std::sqrt(r_z);
      for (int i = 0; i < n; ++i) p[i] = z[i] + beta * p[i];
    }
    *iters = it; *converged = conv ? 1 : 0; *err = e;
    (void)t_spmv_acc;
    return 0;
  }

  // energy-norm error sqrt(sum_K int |grad(u - u_h)|^2)  tutorial_01_..._modules.f90:611-661; also L2
  void errors(const double* x, const Func& fn, double* h1semi, double* l2) const {
    const int nd = fe.ndofs, nn = fe.nnode;
    std::vector<double> detj(fe.nq), grad(size_t(fe.nq) * nn * 3), xq(size_t(fe.nq) * 3), ul(nd);
    double e1 = 0, e0 = 0;
    for (int c = 0; c < m.n_local; ++c) {
      cell_geometry(c, detj.data(), grad.data(), xq.data());
      for (int l = 0; l < nd; ++l) {
        int g = gids[size_t(c) * nd + l];
        ul[l] = g > 0 ? x[g - 1] : fixed[-g - 1];
      }
      double loc = 0;
      for (int q = 0; q < fe.nq; ++q) {
        double factor = detj[q] * fe.qw[q], gu[3], uh = 0, gh[3] = {0, 0, 0};
        fn.grad_u(&xq[q * 3], gu);
        for (int a = 0; a < nn; ++a) {
          uh += ul[a] * fe.phi[size_t(q) * nn + a];
          for (int k = 0; k < 3; ++k) gh[k] += ul[a] * grad[(size_t(q) * nn + a) * 3 + k];
        }
        double dd = 0; for (int k = 0; k < 3; ++k) dd += (gu[k] - gh[k]) * (gu[k] - gh[k]);
        loc += factor * dd;
        double du = fn.u(&xq[q * 3]) - uh; e0 += factor * du * du;
      }
      e1 += loc;
    }
    *h1semi = std::sqrt(e1); *l2 = std::sqrt(e0);
  }

  // error_norms_scalar_t%compute for the eleven norm types of par_test_poisson's check_solution
  // (Tests/Par/par_test_poisson/modules/par_test_poisson_driver.f90:557-589):
  //   cell contributions   Integration/ErrorNorms/sbm_error_norms_scalar.i90:187-232
  //   update_norm          Integration/ErrorNorms/error_norms.f90:160-172 (sum, or max for the *infty norms)
  //   finalize_norm        error_norms.f90:174-221 (restated as written: the p-norms end with norm**exponent)
  //   default exponent 2   error_norms.f90:308-331
  // acc[8] = the un-finalised sums / maxima of THIS part (mean, l1, l2, lp, max|e|, h1, w1p, max|d_k e|), so that a multi-part
  // caller can l1_sum / l1_max them before finalising; out[11] = mean, l1, l2, lp, linfty, h1_semi, h1, w1p_semi, w1p,
  // w1infty_semi, w1infty of this part alone.
  void errors_all(const double* x, const Func& fn, double expo, double* acc, double* out) const {
    const int nd = fe.ndofs, nn = fe.nnode;
    std::vector<double> detj(fe.nq), grad(size_t(fe.nq) * nn * 3), xq(size_t(fe.nq) * 3), ul(nd);
    for (int k = 0; k < 8; ++k) acc[k] = 0.0;
    for (int c = 0; c < m.n_local; ++c) {
      cell_geometry(c, detj.data(), grad.data(), xq.data());
      for (int l = 0; l < nd; ++l) {
        int g = gids[size_t(c) * nd + l];
        ul[l] = g > 0 ? x[g - 1] : fixed[-g - 1];
      }
      double cv[5] = {0, 0, 0, 0, 0}, cg[3] = {0, 0, 0};
      for (int q = 0; q < fe.nq; ++q) {
        const double factor = detj[q] * fe.qw[q];
        double gu[3] = {0, 0, 0}, uh = 0, gh[3] = {0, 0, 0};
        fn.grad_u(&xq[q * 3], gu);
        for (int a = 0; a < nn; ++a) {
          uh += ul[a] * fe.phi[size_t(q) * nn + a];
          for (int k = 0; k < m.d; ++k) gh[k] += ul[a] * grad[(size_t(q) * nn + a) * 3 + k];
        }
        const double e = fn.u(&xq[q * 3]) - uh;
        double g2 = 0;
        for (int k = 0; k < m.d; ++k) { g2 += (gu[k] - gh[k]) * (gu[k] - gh[k]); cg[2] = std::max(cg[2], std::fabs(gu[k] - gh[k])); }
        cv[0] += e * factor; cv[1] += std::fabs(e) * factor; cv[2] += e * e * factor; cv[3] += std::pow(std::fabs(e), expo) * factor;
        cv[4] = std::max(cv[4], std::fabs(e));
        cg[0] += g2 * factor; cg[1] += std::pow(g2, expo / 2.0) * factor;
      }
      acc[0] += cv[0]; acc[1] += cv[1]; acc[2] += cv[2]; acc[3] += cv[3]; acc[4] = std::max(acc[4], cv[4]);
      acc[5] += cg[0]; acc[6] += cg[1]; acc[7] = std::max(acc[7], cg[2]);
    }
    finalize_norms(acc, expo, out);
  }
  static void finalize_norms(const double* a, double expo, double* out) {
    out[0] = a[0]; out[1] = a[1]; out[2] = std::sqrt(a[2]); out[3] = std::pow(a[3], expo); out[4] = a[4];
    out[5] = std::sqrt(a[5]); out[6] = std::sqrt(a[2] + a[5]); out[7] = std::pow(a[6], expo); out[8] = std::pow(a[3] + a[6], expo);
    out[9] = a[7]; out[10] = a[4] + a[7];
  }
};

} // namespace fo

// =====================================================================================
// flat C API for ctypes
// =====================================================================================
using fo::Space; using fo::Func; using fo::i64;
static Func make_func(int d, int kind, const double* prm) {
  Func f; f.kind = kind; f.d = d;
  if (prm) { f.alpha = prm[0]; f.r = prm[1]; f.c[0] = prm[2]; f.c[1] = prm[3]; f.c[2] = prm[4]; }
  return f;
}
extern "C" {
void* fo_create(int d, const int* N, const int* P, int part_id, int order, int ncomp, unsigned dir_mask) {
  Space* s = new Space(); s->create(d, N, P, part_id, order, ncomp, dir_mask); return s;
}
void fo_destroy(void* h) { delete (Space*)h; }
int fo_num_cells(void* h) { return ((Space*)h)->m.n_cells; }
int fo_num_local_cells(void* h) { return ((Space*)h)->m.n_local; }
int fo_num_itfc_cells(void* h) { return ((Space*)h)->m.n_itfc; }
int fo_num_vefs(void* h) { return ((Space*)h)->m.n_vefs; }
int fo_ndofs_per_cell(void* h) { return ((Space*)h)->fe.ndofs; }
int fo_num_free(void* h) { return ((Space*)h)->n_free; }
int fo_num_free_ghost(void* h) { return ((Space*)h)->n_free_ghost; }
int fo_num_dirichlet(void* h) { return ((Space*)h)->n_dirichlet; }
const int* fo_lst_dofs_gids(void* h) { return ((Space*)h)->gids.data(); }
const int* fo_lst_vefs(void* h) { return ((Space*)h)->m.lst_vefs.data(); }
const int* fo_vef_bid(void* h) { return ((Space*)h)->m.vef_bid.data(); }
const i64* fo_cells_gid(void* h) { return ((Space*)h)->m.cells_gid.data(); }
const int* fo_cells_part(void* h) { return ((Space*)h)->m.cells_part.data(); }
const i64* fo_vefs_gid(void* h) { return ((Space*)h)->m.vefs_gid.data(); }
int fo_assemble(void* h, int kind, int func, const double* fprm, const double* mat, int sym, int route) {
  Space* s = (Space*)h; Func f = make_func(s->m.d, func, fprm);
  s->assemble(kind, f, mat, sym != 0, route); return 0;
}
int fo_build_pattern(void* h, int sym) { ((Space*)h)->build_pattern(sym != 0); return 0; }
i64 fo_nnz(void* h) { return i64(((Space*)h)->ja.size()); }
const i64* fo_irp(void* h) { return ((Space*)h)->irp.data(); }      // 0-based offsets (reference irp = this + 1)
const int* fo_ja(void* h) { return ((Space*)h)->ja.data(); }        // 1-based
const double* fo_val(void* h) { return ((Space*)h)->val.data(); }
const double* fo_rhs(void* h) { return ((Space*)h)->rhs.data(); }
const double* fo_fixed(void* h) { return ((Space*)h)->fixed.data(); }
int fo_spmv(void* h, const double* x, double* y) { ((Space*)h)->spmv(x, y); return 0; }
int fo_jacobi_setup(void* h) { ((Space*)h)->jacobi_setup(); return 0; }
const double* fo_dinv(void* h) { return ((Space*)h)->dinv.data(); }
int fo_cg(void* h, int prec, double rtol, double atol, int maxit, int stop, const double* b, double* x, int* iters,
          int* converged, double* err, double* hist) {
  return ((Space*)h)->cg(prec, rtol, atol, maxit, stop, b, x, iters, converged, err, hist, nullptr);
}
int fo_errors(void* h, const double* x, int func, const double* fprm, double* h1semi, double* l2) {
  Space* s = (Space*)h; Func f = make_func(s->m.d, func, fprm);
  s->errors(x, f, h1semi, l2); return 0;
}
int fo_errors_all(void* h, const double* x, int func, const double* fprm, double expo, double* acc8, double* out11) {
  Space* s = (Space*)h; Func f = make_func(s->m.d, func, fprm);
  s->errors_all(x, f, expo > 0 ? expo : 2.0, acc8, out11); return 0;
}
int fo_finalize_norms(const double* acc8, double expo, double* out11) { Space::finalize_norms(acc8, expo > 0 ? expo : 2.0, out11); return 0; }
// element matrix of one cell (column-major nd x nd) and element vector, before Dirichlet elimination
int fo_element(void* h, int cell, int kind, int func, const double* fprm, const double* mat, double* elmat, double* elvec) {
  Space* s = (Space*)h; Func f = make_func(s->m.d, func, fprm);
  std::vector<double> detj(s->fe.nq), grad(size_t(s->fe.nq) * s->fe.nnode * 3), xq(size_t(s->fe.nq) * 3);
  s->element(cell, kind, f, mat, elmat, elvec, detj.data(), grad.data(), xq.data());
  return 0;
}
int fo_ref_tables(void* h, double* qw, double* phi, double* dphi) {
  Space* s = (Space*)h;
  std::copy(s->fe.qw.begin(), s->fe.qw.end(), qw);
  std::copy(s->fe.phi.begin(), s->fe.phi.end(), phi);
  std::copy(s->fe.dphi.begin(), s->fe.dphi.end(), dphi);
  return 0;
}
int fo_num_interior(void* h) { return ((Space*)h)->n_interior; }
int fo_num_interface(void* h) { return ((Space*)h)->n_interface; }
int fo_imp_sizes(void* h, int* out) {
  Space* s = (Space*)h;
  out[0] = int(s->list_rcv.size()); out[1] = int(s->list_snd.size()); out[2] = int(s->unpack_idx.size()); out[3] = int(s->pack_idx.size());
  return 0;
}
const int* fo_imp_ptr(void* h, int which) {
  Space* s = (Space*)h;
  switch (which) {
    case 0: return s->list_rcv.data(); case 1: return s->rcv_ptrs.data(); case 2: return s->unpack_idx.data();
    case 3: return s->list_snd.data(); case 4: return s->snd_ptrs.data(); case 5: return s->pack_idx.data();
    case 6: return s->perm_old2new.data();
  }
  return nullptr;
}
const double* fo_weight(void* h) { return ((Space*)h)->weight.data(); }
int fo_nq(void* h) { return ((Space*)h)->fe.nq; }
int fo_hw_threads() { int n = int(std::thread::hardware_concurrency()); return n > 0 ? n : 1; }
int fo_set_threads(void* h, int nt) { ((Space*)h)->nthreads = nt > 0 ? nt : 1; return 0; }
int fo_num_threads(void* h) { return ((Space*)h)->nthreads; }
// -------------------------------------------------------------------------------------
// The reference's MPI path with every rank a host thread (BASELINE.md 3 variant 2):
// one Space per part, sub-assembled local matrices, and per PCG iteration
//   par_sparse_matrix apply  = local matvec + comm      LinearAlgebra/SparseMatrix/par_sparse_matrix.f90:739-762
//   par_scalar_array%comm    = 2-stage neighbour exchange LinearAlgebra/Vector/par_scalar_array.f90:498-533
//                              (pack_rp / unpack_rp, Tools/mpi_context.f90:522-639: here mailboxes in shared memory
//                               with a barrier where MPI would wait)
//   par dot / nrm2           = weighted local dot + l1_sum  par_scalar_array.f90:303-350 (mpi_context.f90:384: allreduce)
//   cg_solve_body            LinearAlgebra/IterativeLinearSolver/cg.f90:101-238 (RES_RES, fixed number of iterations)
// out[0] assembly seconds/step (max over ranks), out[1] seconds per apply (matvec + comm), out[2] seconds of the PCG loop/step,
// out[3] |b|_2, out[4] |r|_2 after the last iteration, out[5] sum of the assembled diagonal, out[6] Jacobi set-up seconds/step
// -------------------------------------------------------------------------------------
namespace {
struct SpinBarrier {
  std::atomic<int> count{0}, gen{0};
  int n = 1;
  void wait() {
    int g = gen.load(std::memory_order_acquire);
    if (count.fetch_add(1, std::memory_order_acq_rel) == n - 1) { count.store(0, std::memory_order_relaxed); gen.fetch_add(1, std::memory_order_release); }
    else { int spins = 0; while (gen.load(std::memory_order_acquire) == g) { if (++spins > 2000) std::this_thread::yield(); } }
  }
};
}
int fo_par_run(void** hs, int np, int kind, int func, const double* fprm, const double* mat, int cg_iters, int steps, int warmup,
               double* out) {
  std::vector<Space*> S(np);
  for (int p = 0; p < np; ++p) S[p] = (Space*)hs[p];
  // mailbox (p -> q) positions: stage 1 uses the send lists, stage 2 the receive lists
  std::vector<std::vector<std::vector<double>>> box1(np), box2(np);
  std::vector<std::vector<int>> slot_of_snd(np), slot_of_rcv(np);   // for receiver p: index of p in the sender's list
  for (int p = 0; p < np; ++p) {
    box1[p].resize(S[p]->list_snd.size()); box2[p].resize(S[p]->list_rcv.size());
    for (size_t i = 0; i < S[p]->list_snd.size(); ++i) box1[p][i].resize(S[p]->snd_ptrs[i + 1] - S[p]->snd_ptrs[i]);
    for (size_t i = 0; i < S[p]->list_rcv.size(); ++i) box2[p][i].resize(S[p]->rcv_ptrs[i + 1] - S[p]->rcv_ptrs[i]);
  }
  auto find_in = [](const std::vector<int>& lst, int part) { for (size_t i = 0; i < lst.size(); ++i) if (lst[i] == part) return int(i); return -1; };
  SpinBarrier bar; bar.n = np;
  std::vector<double> red(size_t(np) * 2 * 8, 0.0);
  std::vector<double> t_asm(np, 0.0), t_app(np, 0.0), t_cg(np, 0.0), t_jac(np, 0.0), chk(8, 0.0);
  auto now = []() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  auto body = [&](int p) {
    Space& s = *S[p];
    const int n = s.n_free;
    if (s.weight.size() != size_t(n)) s.weight.assign(n, 1.0);   // serial space: every DOF is owned
    int red_flip = 0;
    auto comm = [&](double* x) {
      for (size_t i = 0; i < s.list_snd.size(); ++i)
        for (int k = s.snd_ptrs[i], o = 0; k < s.snd_ptrs[i + 1]; ++k, ++o) box1[p][i][o] = x[s.pack_idx[k] - 1];
      bar.wait();
      for (size_t i = 0; i < s.list_rcv.size(); ++i) {        // the owner adds, neighbour by neighbour in list order
        const int q = s.list_rcv[i] - 1, j = find_in(S[q]->list_snd, p + 1);
        const std::vector<double>& b = box1[q][j];
        for (int k = s.rcv_ptrs[i], o = 0; k < s.rcv_ptrs[i + 1]; ++k, ++o) x[s.unpack_idx[k] - 1] += b[o];
      }
      for (size_t i = 0; i < s.list_rcv.size(); ++i)
        for (int k = s.rcv_ptrs[i], o = 0; k < s.rcv_ptrs[i + 1]; ++k, ++o) box2[p][i][o] = x[s.unpack_idx[k] - 1];
      bar.wait();
      for (size_t i = 0; i < s.list_snd.size(); ++i) {        // the owner's total comes back
        const int q = s.list_snd[i] - 1, j = find_in(S[q]->list_rcv, p + 1);
        const std::vector<double>& b = box2[q][j];
        for (int k = s.snd_ptrs[i], o = 0; k < s.snd_ptrs[i + 1]; ++k, ++o) x[s.pack_idx[k] - 1] = b[o];
      }
    };
    auto allsum = [&](const double* loc, int cnt, double* glob) {
      double* base = &red[size_t(red_flip) * np * 8];
      for (int k = 0; k < cnt; ++k) base[size_t(p) * 8 + k] = loc[k];
      bar.wait();
      for (int k = 0; k < cnt; ++k) { double t = 0; for (int q = 0; q < np; ++q) t += base[size_t(q) * 8 + k]; glob[k] = t; }
      red_flip ^= 1;
    };
    auto wdot = [&](const double* a, const double* b) {
      double l = 0; for (int i = 0; i < n; ++i) l += s.weight[i] * a[i] * b[i];
      double g; allsum(&l, 1, &g); return g;
    };
    Func f = make_func(s.m.d, func, fprm);
    std::vector<double> b(n), x(n), r(n), z(n), pp(n), Ap(n), d(n);
    for (int it = 0; it < warmup + steps; ++it) {
      bar.wait();
      double t0 = now();
      s.assemble(kind, f, mat, false, 1);
      bar.wait();
      double t1 = now();
      std::copy(s.rhs.begin(), s.rhs.end(), b.begin());
      comm(b.data());                                        // par_sparse_assembler.f90:171-176
      bar.wait();
      double t1b = now();
      s.jacobi_setup();                                      // local diagonal (inverted), comm on the diagonal itself
      double dsum = 0;
      for (int i = 0; i < n; ++i) d[i] = 1.0 / s.dinv[i];
      comm(d.data());
      for (int i = 0; i < n; ++i) { dsum += s.weight[i] * d[i]; d[i] = 1.0 / d[i]; }
      bar.wait();
      double t2 = now(), tapp = 0;
      std::fill(x.begin(), x.end(), 0.0);
      s.spmv(x.data(), r.data()); comm(r.data());
      for (int i = 0; i < n; ++i) r[i] = b[i] - r[i];
      for (int i = 0; i < n; ++i) z[i] = d[i] * r[i];
      double r_z = wdot(r.data(), z.data());
      pp = z;
      for (int k = 0; k < cg_iters; ++k) {
        double ta = now();
        s.spmv(pp.data(), Ap.data()); comm(Ap.data());
        bar.wait();
        tapp += now() - ta;
        double App = wdot(Ap.data(), pp.data());
        double alpha = App != 0.0 ? r_z / App : 0.0;
        for (int i = 0; i < n; ++i) x[i] += alpha * pp[i];
        for (int i = 0; i < n; ++i) r[i] -= alpha * Ap[i];
        (void)wdot(r.data(), r.data());                      // RES_RES: |r|_2 every iteration
        for (int i = 0; i < n; ++i) z[i] = d[i] * r[i];
        double beta = 1.0 / r_z;
        r_z = wdot(r.data(), z.data());
        beta *= r_z;
        for (int i = 0; i < n; ++i) pp[i] = z[i] + beta * pp[i];
      }
      bar.wait();
      double t3 = now();
      if (it >= warmup) { t_asm[p] += t1 - t0; t_jac[p] += t2 - t1b; t_cg[p] += t3 - t2; t_app[p] += tapp / std::max(cg_iters, 1); }
      double c0 = wdot(b.data(), b.data()), c1 = wdot(r.data(), r.data()), c2;
      allsum(&dsum, 1, &c2);
      if (p == 0) { chk[0] = std::sqrt(c0); chk[1] = std::sqrt(c1); chk[2] = c2; }
    }
  };
  std::vector<std::thread> th;
  for (int p = 0; p < np; ++p) th.emplace_back(body, p);
  for (auto& t : th) t.join();
  out[0] = *std::max_element(t_asm.begin(), t_asm.end()) / std::max(steps, 1);
  out[1] = *std::max_element(t_app.begin(), t_app.end()) / std::max(steps, 1);
  out[2] = *std::max_element(t_cg.begin(), t_cg.end()) / std::max(steps, 1);
  out[3] = chk[0]; out[4] = chk[1]; out[5] = chk[2];
  out[6] = *std::max_element(t_jac.begin(), t_jac.end()) / std::max(steps, 1);
  return 0;
}

}
