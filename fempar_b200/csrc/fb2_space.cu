// FE space on a structured hex part, built ON THE DEVICE:
//   cell permutation (interior -> interface -> ghost), cell->DOF map (lst_dofs_gids), strong
//   Dirichlet numbering, node<->DOF lattice maps and the CSR graph.
// The reference builds all of this with serial visited-maps sweeps
//   (Integration/FESpace/sbm_serial_fe_space.i90:525-715, 2623-2875); here every (cell, n-face)
// pair finds its owner = the first cell in traversal order touching that n-face in closed form,
// per-cell creation counts are prefix-summed, and ids are written in one pass.
#include <cub/cub.cuh>
#include <algorithm>
#include <math.h>
#include <stdarg.h>
#include <string.h>
#include <map>
#include "fb2_internal.h"

// ------------------------------------------------------------------------------------------
// runtime
// ------------------------------------------------------------------------------------------
static thread_local char g_err[1024] = "";
static cudaStream_t g_stream = nullptr;
static int g_nsm = 148;
static int g_inited = 0;
static double* g_partials = nullptr;
static std::map<std::string, double> g_phase_ms;

void fb2_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
cudaStream_t fb2_stream() { return g_stream; }
static cudaStream_t g_comm_stream = nullptr;
cudaStream_t fb2_comm_stream() { return g_comm_stream; }
int fb2_num_sms() { return g_nsm; }
double* fb2_scratch_partials() { return g_partials; }
void fb2_record_phase(const char* name, double ms) { g_phase_ms[name] = ms; }

extern "C" const char* fb2_last_error_string(void) { return g_err; }
extern "C" int fb2_version(void) { return 100; }
extern "C" int fb2_init(int device) {
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    fb2_set_error("fb2_init: no CUDA device available (%s); libfempar_b200 has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    return 1;
  }
  FB2_CHECK(device >= 0 && device < ndev, "fb2_init: device index out of range");
  FB2_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  FB2_CUDA(cudaGetDeviceProperties(&prop, device));
  g_nsm = prop.multiProcessorCount;
  if (!g_inited) {
    // compute stream at the lowest priority, halo-exchange stream at the highest: CTAs of the pack/NCCL/unpack kernels are
    // scheduled ahead of pending SpMV CTAs, so the exchange overlaps the interior rows (fb2_solver.cu)
    int prio_lo = 0, prio_hi = 0;
    FB2_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    FB2_CUDA(cudaStreamCreateWithPriority(&g_stream, cudaStreamNonBlocking, prio_lo));
    FB2_CUDA(cudaStreamCreateWithPriority(&g_comm_stream, cudaStreamNonBlocking, prio_hi));
    FB2_CUDA(cudaMalloc((void**)&g_partials, sizeof(double) * FB2_MAX_PARTIALS * 5));
    g_inited = 1;
  }
  return 0;
}
extern "C" int fb2_device_synchronize(void) {
  FB2_CHECK(g_inited, "fb2_init was not called");
  FB2_CUDA(cudaStreamSynchronize(g_stream));
  FB2_CUDA(cudaGetLastError());
  return 0;
}
extern "C" int fb2_last_phase_ms(const char* phase, double* ms) {
  auto it = g_phase_ms.find(phase);
  FB2_CHECK(it != g_phase_ms.end(), "unknown phase");
  *ms = it->second;
  return 0;
}

// ------------------------------------------------------------------------------------------
// host-side reference-FE tables (own implementation; semantics of sbm_polynomial.i90:60-108,
// sbm_tensor_product_polynomial.i90:88-130, sbm_quadrature.i90:767-823)
// ------------------------------------------------------------------------------------------
static void lagrange_eq(int p, double x, double* v, double* dv) {
  double xn[8];
  for (int k = 0; k <= p; ++k) xn[k] = -1.0 + 2.0 * k / p;
  for (int k = 0; k <= p; ++k) {
    double val = 1.0, ds = 0.0;
    for (int m = 0; m <= p; ++m)
      if (m != k) val *= (x - xn[m]) / (xn[k] - xn[m]);
    for (int n = 0; n <= p; ++n) {
      if (n == k) continue;
      double t = 1.0 / (xn[k] - xn[n]);
      for (int m = 0; m <= p; ++m)
        if (m != k && m != n) t *= (x - xn[m]) / (xn[k] - xn[m]);
      ds += t;
    }
    v[k] = val;
    dv[k] = ds;
  }
}
static void gauss_legendre_rule(int n, double* x, double* w) {
  for (int i = 0; i < (n + 1) / 2; ++i) {
    double z = cos(M_PI * (i + 0.75) / (n + 0.5)), pp = 1.0;
    for (int it = 0; it < 100; ++it) {
      double p1 = 1.0, p2 = 0.0;
      for (int j = 1; j <= n; ++j) {
        double p3 = p2;
        p2 = p1;
        p1 = ((2.0 * j - 1.0) * z * p2 - (j - 1.0) * p3) / j;
      }
      pp = n * (z * p1 - p2) / (z * z - 1.0);
      double dz = p1 / pp;
      z -= dz;
      if (fabs(dz) < 1e-16) break;
    }
    double p1 = 1.0, p2 = 0.0;
    for (int j = 1; j <= n; ++j) {
      double p3 = p2;
      p2 = p1;
      p1 = ((2.0 * j - 1.0) * z * p2 - (j - 1.0) * p3) / j;
    }
    pp = n * (z * p1 - p2) / (z * z - 1.0);
    x[i] = -z;
    x[n - 1 - i] = z;
    w[i] = w[n - 1 - i] = 2.0 / ((1.0 - z * z) * pp * pp);
  }
  if (n % 2) x[n / 2] = 0.0;
}

static void tabulate(int ord, const std::vector<double>& pts, int npts, std::vector<double>& v, std::vector<double>& dv) {
  int m1 = ord + 1, nn = m1 * m1 * m1;
  v.assign((size_t)npts * nn, 0.0);
  dv.assign((size_t)npts * nn * 3, 0.0);
  double l[3][8], dl[3][8];
  for (int q = 0; q < npts; ++q) {
    for (int k = 0; k < 3; ++k) lagrange_eq(ord, pts[q * 3 + k], l[k], dl[k]);
    for (int a = 0; a < nn; ++a) {
      int a0 = a % m1, a1 = (a / m1) % m1, a2 = a / (m1 * m1);
      v[(size_t)q * nn + a] = l[0][a0] * l[1][a1] * l[2][a2];
      dv[((size_t)q * nn + a) * 3 + 0] = dl[0][a0] * l[1][a1] * l[2][a2];
      dv[((size_t)q * nn + a) * 3 + 1] = l[0][a0] * dl[1][a1] * l[2][a2];
      dv[((size_t)q * nn + a) * 3 + 2] = l[0][a0] * l[1][a1] * dl[2][a2];
    }
  }
}

static void make_ref_tables(fb2_space_s* s, int p, int ncomp) {
  RefTables& rt = s->rt;
  rt.p = p; rt.n1 = p + 1; rt.nnode = rt.n1 * rt.n1 * rt.n1; rt.ncomp = ncomp; rt.ndofs = rt.nnode * ncomp;
  rt.nq = rt.nnode;
  // hex polytope: n-faces sorted by (dim, code = free<<3|anchor)  (sbm_polytope.i90:175-227)
  std::vector<int> codes;
  for (int fr = 0; fr < 8; ++fr)
    for (int an = 0; an < 8; ++an)
      if ((fr & an) == 0) codes.push_back((fr << 3) | an);
  std::stable_sort(codes.begin(), codes.end(), [](int a, int b) {
    int da = __builtin_popcount(a >> 3), db = __builtin_popcount(b >> 3);
    return da != db ? da < db : a < b;
  });
  for (int i = 0; i < 64; ++i) rt.face_of_code[i] = -1;
  for (int f = 0; f < 27; ++f) {
    rt.code_of_face[f] = codes[f];
    rt.face_of_code[codes[f]] = f;
    int dim = __builtin_popcount(codes[f] >> 3), m = 1;
    for (int k = 0; k < dim; ++k) m *= (p - 1);
    rt.m_face[f] = m;
  }
  int nq1 = p + 1;
  double x1[8], w1[8];
  gauss_legendre_rule(nq1, x1, w1);
  std::vector<double> pts((size_t)rt.nq * 3);
  s->h_qw.assign(rt.nq, 0.0);
  for (int q = 0; q < rt.nq; ++q) {
    int q0 = q % nq1, q1 = (q / nq1) % nq1, q2 = q / (nq1 * nq1);
    pts[q * 3 + 0] = x1[q0]; pts[q * 3 + 1] = x1[q1]; pts[q * 3 + 2] = x1[q2];
    s->h_qw[q] = w1[q0] * w1[q1] * w1[q2];
  }
  s->h_b1d.assign((size_t)nq1 * rt.n1, 0.0); s->h_c1d.assign((size_t)nq1 * rt.n1, 0.0);
  for (int q = 0; q < nq1; ++q) {
    double l[8], dl[8];
    lagrange_eq(p, x1[q], l, dl);
    for (int i = 0; i < rt.n1; ++i) { s->h_b1d[q * rt.n1 + i] = l[i]; s->h_c1d[q * rt.n1 + i] = dl[i]; }
  }
  tabulate(p, pts, rt.nq, s->h_phi, s->h_dphi);
  tabulate(1, pts, rt.nq, s->h_gphi, s->h_gdphi);
}

// ------------------------------------------------------------------------------------------
// device helpers
// ------------------------------------------------------------------------------------------
struct DevTables {  // small per-space constants passed by value to kernels
  MeshDesc m;
  int p, n1, nnode, ncomp, ndofs;
  int m_face[27];
  signed char face_of_code[64];
  unsigned char code_of_face[27];
  uint32_t dir_mask;
};

static DevTables make_dev_tables(const fb2_space_s* s) {
  DevTables t;
  t.m = s->m;
  t.p = s->rt.p; t.n1 = s->rt.n1; t.nnode = s->rt.nnode; t.ncomp = s->rt.ncomp; t.ndofs = s->rt.ndofs;
  for (int f = 0; f < 27; ++f) { t.m_face[f] = s->rt.m_face[f]; t.code_of_face[f] = (unsigned char)s->rt.code_of_face[f]; }
  for (int c = 0; c < 64; ++c) t.face_of_code[c] = (signed char)s->rt.face_of_code[c];
  t.dir_mask = s->dir_mask;
  return t;
}

__device__ __forceinline__ void lex_to_ijk3(const int* n, int lex, int* ijk) {
  ijk[0] = lex % n[0];
  int r = lex / n[0];
  ijk[1] = r % n[1];
  ijk[2] = r / n[1];
}

// first cell (smallest local id) touching n-face `code` of the cell at ijk (total-box coordinates)
__device__ __forceinline__ int owner_cell(const MeshDesc& m, const int* __restrict__ perm, const int* ijk, int code) {
  int fr = code >> 3, an = code & 7;
  int lo[3], hi[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    if ((fr >> k) & 1) { lo[k] = hi[k] = ijk[k]; }
    else {
      int pos = ijk[k] + ((an >> k) & 1);
      lo[k] = max(pos - 1, 0);
      hi[k] = min(pos, m.ntot[k] - 1);
    }
  }
  if (perm == nullptr) return lo[0] + m.ntot[0] * (lo[1] + m.ntot[1] * lo[2]);
  int best = 0x7fffffff;
  for (int c2 = lo[2]; c2 <= hi[2]; ++c2)
    for (int c1 = lo[1]; c1 <= hi[1]; ++c1)
      for (int c0 = lo[0]; c0 <= hi[0]; ++c0) best = min(best, perm[c0 + m.ntot[0] * (c1 + m.ntot[1] * c2)]);
  return best;
}

// boundary id semantics of uniform_hex_mesh_generator.f90:593-630: is the n-face on a Dirichlet-flagged
// part of the domain boundary?
__device__ __forceinline__ bool vef_is_dirichlet(const DevTables& t, const int* ijk, int code) {
  int fr = code >> 3, an = code & 7, bcode = 0;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    if ((fr >> k) & 1) bcode |= (8 << k);
    else {
      int g = t.m.first[k] + ijk[k] + ((an >> k) & 1);
      if (g == 0) { /* anchor bit 0 */ }
      else if (g == t.m.N[k]) bcode |= (1 << k);
      else bcode |= (8 << k);
    }
  }
  if (bcode == 56) return false;
  int b = t.face_of_code[bcode];  // boundary id - 1
  return (t.dir_mask >> b) & 1u;
}

// ------------------------------------------------------------------------------------------
// cell permutation interior -> interface -> ghost (uniform_hex_mesh_generator.f90:316-351)
// ------------------------------------------------------------------------------------------
__global__ void k_cell_kind(MeshDesc m, int* f_int, int* f_itf, int* f_gh) {
  int lex = blockIdx.x * blockDim.x + threadIdx.x;
  if (lex >= m.n_cells) return;
  int ijk[3];
  lex_to_ijk3(m.ntot, lex, ijk);
  int kind = 0;
  for (int k = 0; k < 3; ++k)
    if ((m.L[k] && ijk[k] == 0) || (m.R[k] && ijk[k] == m.ntot[k] - 1)) { kind = 2; break; }
  if (kind == 0)
    for (int k = 0; k < 3; ++k)
      if ((m.L[k] && ijk[k] == 1) || (m.R[k] && ijk[k] == m.ntot[k] - 2)) { kind = 1; break; }
  f_int[lex] = kind == 0; f_itf[lex] = kind == 1; f_gh[lex] = kind == 2;
}
__global__ void k_cell_perm(MeshDesc m, const int* f_int, const int* s_int, const int* s_itf, const int* s_gh,
                            const int* f_itf, int* perm, int* lex_of) {
  int lex = blockIdx.x * blockDim.x + threadIdx.x;
  if (lex >= m.n_cells) return;
  int p;
  if (f_int[lex]) p = s_int[lex];
  else if (f_itf[lex]) p = m.n_local - 1 - s_itf[lex];   // interface cells are filled from the end backwards
  else p = m.n_cells - 1 - s_gh[lex];                    // and so are ghost cells
  perm[lex] = p;
  lex_of[p] = lex;
}

// ------------------------------------------------------------------------------------------
// DOF numbering
// ------------------------------------------------------------------------------------------
__global__ void k_count_created(DevTables t, const int* __restrict__ perm, const int* __restrict__ lex_of,
                                int* cnt_free, int* cnt_dir, uint32_t* own_mask, uint32_t* dirv_mask) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= t.m.n_cells) return;
  int lex = lex_of ? lex_of[c] : c;
  int ijk[3];
  lex_to_ijk3(t.m.ntot, lex, ijk);
  uint32_t om = 0, dm = 0;
  int nf = t.m_face[26] * t.ncomp, nd = 0;
  for (int f = 0; f < 26; ++f) {
    int code = t.code_of_face[f];
    bool own = owner_cell(t.m, perm, ijk, code) == c;
    bool dir = vef_is_dirichlet(t, ijk, code);
    if (own) om |= (1u << f);
    if (dir) dm |= (1u << f);
    if (own) { if (dir) nd += t.m_face[f] * t.ncomp; else nf += t.m_face[f] * t.ncomp; }
  }
  cnt_free[c] = nf; cnt_dir[c] = nd; own_mask[c] = om; dirv_mask[c] = dm;
}

__global__ void k_fill_gids(DevTables t, const int* __restrict__ perm, const int* __restrict__ lex_of,
                            const int* __restrict__ base_free, const int* __restrict__ base_dir,
                            const uint32_t* __restrict__ own_mask, const uint32_t* __restrict__ dirv_mask, int* gids) {
  i64 tid = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  i64 total = (i64)t.m.n_cells * t.ndofs;
  if (tid >= total) return;
  int c = (int)(tid / t.ndofs), l = (int)(tid % t.ndofs);
  int comp = l / t.nnode, node = l % t.nnode;
  int in[3] = {node % t.n1, (node / t.n1) % t.n1, node / (t.n1 * t.n1)};
  int code = 0, tloc = 0, stride = 1;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    if (in[k] == 0) { /* fixed, anchor 0 */ }
    else if (in[k] == t.p) code |= (1 << k);
    else { code |= (8 << k); tloc += (in[k] - 1) * stride; stride *= (t.p - 1); }
  }
  int f = t.face_of_code[code];
  int m = t.m_face[f];
  int within = comp * m + tloc;
  int lex = lex_of ? lex_of[c] : c;
  int ijk[3];
  lex_to_ijk3(t.m.ntot, lex, ijk);
  if (f == 26) { gids[tid] = base_free[c] + within + 1; return; }
  bool dir = vef_is_dirichlet(t, ijk, code);
  int oc = owner_cell(t.m, perm, ijk, code);
  int olex = lex_of ? lex_of[oc] : oc;
  int oijk[3];
  lex_to_ijk3(t.m.ntot, olex, oijk);
  int ocode = code & 0x38;
#pragma unroll
  for (int k = 0; k < 3; ++k)
    if (!((code >> (3 + k)) & 1)) ocode |= ((ijk[k] + ((code >> k) & 1) - oijk[k]) & 1) << k;
  int fo = t.face_of_code[ocode];
  uint32_t om = own_mask[oc], dm = dirv_mask[oc];
  int off = dir ? 0 : t.m_face[26] * t.ncomp;
  for (int f2 = 0; f2 < fo; ++f2)
    if (((om >> f2) & 1u) && (((dm >> f2) & 1u) == (dir ? 1u : 0u))) off += t.m_face[f2] * t.ncomp;
  gids[tid] = dir ? -(base_dir[oc] + off + within + 1) : (base_free[oc] + off + within + 1);
}

// node <-> dof lattice maps over the LOCAL cell box
__global__ void k_node_maps(DevTables t, const int* __restrict__ lex_of, const int* __restrict__ gids, int M0, int M1,
                            int M2, i64 lat_size, i64 n_free, int* node_dof, int* dof_node) {
  i64 tid = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  i64 total = (i64)t.m.n_local * t.ndofs;
  if (tid >= total) return;
  int c = (int)(tid / t.ndofs), l = (int)(tid % t.ndofs);
  int comp = l / t.nnode, node = l % t.nnode;
  int lex = lex_of ? lex_of[c] : c;
  int ijk[3];
  lex_to_ijk3(t.m.ntot, lex, ijk);
  int a0 = (ijk[0] - t.m.L[0]) * t.p + node % t.n1;
  int a1 = (ijk[1] - t.m.L[1]) * t.p + (node / t.n1) % t.n1;
  int a2 = (ijk[2] - t.m.L[2]) * t.p + node / (t.n1 * t.n1);
  i64 lat = a0 + (i64)M0 * (a1 + (i64)M1 * a2);
  int g = gids[tid];
  node_dof[comp * lat_size + lat] = g;
  if (g > 0 && g <= n_free) dof_node[g - 1] = (int)(comp * lat_size + lat);
}

template <class T>
static int exclusive_scan(const T* d_in, T* d_out, i64 n) {
  void* tmp = nullptr;
  size_t bytes = 0;
  FB2_CUDA(cub::DeviceScan::ExclusiveSum(tmp, bytes, d_in, d_out, n, fb2_stream()));
  FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
  cudaError_t e = cub::DeviceScan::ExclusiveSum(tmp, bytes, d_in, d_out, n, fb2_stream());
  cudaStreamSynchronize(fb2_stream());
  cudaFree(tmp);
  FB2_CUDA(e);
  return 0;
}

int fb2_build_node_maps(fb2_space_s* s) {
  DevTables t = make_dev_tables(s);
  if (s->d_node_dof) cudaFree(s->d_node_dof);
  if (s->d_dof_node) cudaFree(s->d_dof_node);
  FB2_TRY(fb2_dalloc(&s->d_node_dof, (size_t)s->lat_size * s->rt.ncomp));
  FB2_TRY(fb2_dalloc(&s->d_dof_node, (size_t)s->n_free));
  i64 total = (i64)s->m.n_local * s->rt.ndofs;
  k_node_maps<<<(unsigned)((total + 255) / 256), 256, 0, fb2_stream()>>>(t, s->d_lex_of, s->d_gids, s->Mlat[0], s->Mlat[1],
                                                                        s->Mlat[2], s->lat_size, s->n_free, s->d_node_dof,
                                                                        s->d_dof_node);
  FB2_CUDA(cudaGetLastError());
  return 0;
}

static int upload_ref_tables(fb2_space_s* s) {
  auto up = [&](double** d, const std::vector<double>& h) -> int {
    FB2_TRY(fb2_dalloc(d, h.size()));
    FB2_CUDA(fb2_memcpy(*d, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice));
    return 0;
  };
  FB2_TRY(up(&s->d_qw, s->h_qw)); FB2_TRY(up(&s->d_phi, s->h_phi)); FB2_TRY(up(&s->d_dphi, s->h_dphi));
  FB2_TRY(up(&s->d_gphi, s->h_gphi)); FB2_TRY(up(&s->d_gdphi, s->h_gdphi));
  return 0;
}
int fb2_space_setup_reference_fe(fb2_space_s* s, int order, int ncomp) {
  FB2_CHECK(g_inited, "fb2_init was not called (no CUDA device, no CPU fallback)");
  make_ref_tables(s, order, ncomp);
  return upload_ref_tables(s);
}

extern "C" int fb2_fe_space_create(fb2_space_t* out, const int num_cells[3], const int num_parts[3], int part_id,
                                   int order, int num_components, uint32_t dirichlet_boundary_mask) {
  FB2_CHECK(g_inited, "fb2_init was not called (no CUDA device, no CPU fallback)");
  FB2_CHECK(order >= 1 && order <= 4, "order must be 1..4");
  FB2_CHECK(num_components == 1 || num_components == 3, "num_components must be 1 or 3");
  fb2_space_s* s = new fb2_space_s();
  s->dir_mask = dirichlet_boundary_mask;
  s->part_id = part_id;
  MeshDesc& m = s->m;
  m.par = part_id > 0;
  for (int k = 0; k < 3; ++k) {
    m.N[k] = num_cells[k];
    m.P[k] = m.par ? num_parts[k] : 1;
    FB2_CHECK(m.N[k] >= 1 && m.P[k] >= 1 && m.N[k] % m.P[k] == 0, "num_cells must be a multiple of num_parts");
  }
  int pid = m.par ? part_id - 1 : 0;
  FB2_CHECK(pid < m.P[0] * m.P[1] * m.P[2], "part_id out of range");
  m.pijk[0] = pid % m.P[0]; m.pijk[1] = (pid / m.P[0]) % m.P[1]; m.pijk[2] = pid / (m.P[0] * m.P[1]);
  m.n_local = 1; m.n_cells = 1;
  i64 ncells64 = 1;
  int n_int = 1;
  for (int k = 0; k < 3; ++k) {
    m.L[k] = (m.par && m.pijk[k] > 0) ? 1 : 0;
    m.R[k] = (m.par && m.pijk[k] < m.P[k] - 1) ? 1 : 0;
    m.nloc[k] = m.N[k] / m.P[k];
    m.ntot[k] = m.nloc[k] + m.L[k] + m.R[k];
    m.first[k] = m.pijk[k] * m.nloc[k] - m.L[k];
    m.n_local *= m.nloc[k];
    ncells64 *= m.ntot[k];
    n_int = (n_int > 0 && m.nloc[k] > m.L[k] + m.R[k]) ? n_int * (m.nloc[k] - m.L[k] - m.R[k]) : 0;
  }
  FB2_CHECK(ncells64 * (i64)(order + 1) * (order + 1) * (order + 1) * num_components < ((i64)1 << 40), "mesh too large");
  FB2_CHECK(ncells64 < 0x7fffffffLL, "too many cells for int32 cell ids");
  m.n_cells = (int)ncells64;
  m.n_itfc = m.n_local - n_int;
  make_ref_tables(s, order, num_components);
  for (int k = 0; k < 3; ++k) s->Mlat[k] = order * m.nloc[k] + 1;
  s->lat_size = (i64)s->Mlat[0] * s->Mlat[1] * s->Mlat[2];
  FB2_CHECK(s->lat_size * num_components < 0x7fffffffLL, "too many DOFs for int32 DOF ids (same limit as integer(ip))");

  cudaStream_t st = fb2_stream();
  const int nc = m.n_cells;
  // 1. cell permutation
  if (m.par) {
    int *f_int, *f_itf, *f_gh, *s_int, *s_itf, *s_gh;
    FB2_TRY(fb2_dalloc(&f_int, nc)); FB2_TRY(fb2_dalloc(&f_itf, nc)); FB2_TRY(fb2_dalloc(&f_gh, nc));
    FB2_TRY(fb2_dalloc(&s_int, nc)); FB2_TRY(fb2_dalloc(&s_itf, nc)); FB2_TRY(fb2_dalloc(&s_gh, nc));
    FB2_TRY(fb2_dalloc(&s->d_perm, nc)); FB2_TRY(fb2_dalloc(&s->d_lex_of, nc));
    k_cell_kind<<<(nc + 255) / 256, 256, 0, st>>>(m, f_int, f_itf, f_gh);
    FB2_TRY(exclusive_scan(f_int, s_int, nc)); FB2_TRY(exclusive_scan(f_itf, s_itf, nc)); FB2_TRY(exclusive_scan(f_gh, s_gh, nc));
    k_cell_perm<<<(nc + 255) / 256, 256, 0, st>>>(m, f_int, s_int, s_itf, s_gh, f_itf, s->d_perm, s->d_lex_of);
    FB2_CUDA(cudaStreamSynchronize(st));
    cudaFree(f_int); cudaFree(f_itf); cudaFree(f_gh); cudaFree(s_int); cudaFree(s_itf); cudaFree(s_gh);
  }
  // 2. creation counts, prefix sums, ids
  DevTables t = make_dev_tables(s);
  int *cnt_free, *cnt_dir, *base_free, *base_dir;
  uint32_t *own_mask, *dirv_mask;
  FB2_TRY(fb2_dalloc(&cnt_free, nc + 1)); FB2_TRY(fb2_dalloc(&cnt_dir, nc + 1));
  FB2_TRY(fb2_dalloc(&base_free, nc + 1)); FB2_TRY(fb2_dalloc(&base_dir, nc + 1));
  FB2_TRY(fb2_dalloc(&own_mask, nc)); FB2_TRY(fb2_dalloc(&dirv_mask, nc));
  FB2_CUDA(cudaMemsetAsync(cnt_free + nc, 0, sizeof(int), st));
  FB2_CUDA(cudaMemsetAsync(cnt_dir + nc, 0, sizeof(int), st));
  k_count_created<<<(nc + 127) / 128, 128, 0, st>>>(t, s->d_perm, s->d_lex_of, cnt_free, cnt_dir, own_mask, dirv_mask);
  FB2_CUDA(cudaGetLastError());
  FB2_TRY(exclusive_scan(cnt_free, base_free, (i64)nc + 1));
  FB2_TRY(exclusive_scan(cnt_dir, base_dir, (i64)nc + 1));
  int h_tot[3];
  FB2_CUDA(fb2_memcpy(&h_tot[0], base_free + m.n_local, sizeof(int), cudaMemcpyDeviceToHost));
  FB2_CUDA(fb2_memcpy(&h_tot[1], base_free + nc, sizeof(int), cudaMemcpyDeviceToHost));
  FB2_CUDA(fb2_memcpy(&h_tot[2], base_dir + nc, sizeof(int), cudaMemcpyDeviceToHost));
  s->n_free = h_tot[0]; s->n_free_ghost = h_tot[1] - h_tot[0]; s->n_dirichlet = h_tot[2];
  s->n_interior = s->n_free; s->n_interface = 0;
  i64 total = (i64)nc * s->rt.ndofs;
  FB2_TRY(fb2_dalloc(&s->d_gids, (size_t)total));
  k_fill_gids<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(t, s->d_perm, s->d_lex_of, base_free, base_dir, own_mask,
                                                              dirv_mask, s->d_gids);
  FB2_CUDA(cudaGetLastError());
  FB2_CUDA(cudaStreamSynchronize(st));
  cudaFree(cnt_free); cudaFree(cnt_dir); cudaFree(base_free); cudaFree(base_dir); cudaFree(own_mask); cudaFree(dirv_mask);
  // 3. reference tables to the device
  FB2_TRY(upload_ref_tables(s));
  FB2_TRY(fb2_dalloc(&s->d_fixed, (size_t)s->n_dirichlet));
  FB2_CUDA(cudaMemsetAsync(s->d_fixed, 0, sizeof(double) * (s->n_dirichlet ? s->n_dirichlet : 1), fb2_stream()));
  // 4. par layer: interface DOFs, dof_import, interior-then-interface renumbering
  if (m.par) FB2_TRY(fb2_build_par_layer(s));
  // 5. lattice maps for the graph builder
  FB2_TRY(fb2_build_node_maps(s));
  FB2_CUDA(cudaStreamSynchronize(st));
  *out = s;
  return 0;
}

extern "C" int fb2_fe_space_destroy(fb2_space_t s) {
  if (!s) return 0;
  cudaFree(s->d_perm); cudaFree(s->d_lex_of); cudaFree(s->d_gids); cudaFree(s->d_node_dof); cudaFree(s->d_dof_node);
  cudaFree(s->d_coords); cudaFree(s->d_src_qp);
  cudaFree(s->d_fixed); cudaFree(s->d_qw); cudaFree(s->d_phi); cudaFree(s->d_dphi); cudaFree(s->d_gphi); cudaFree(s->d_gdphi);
  if (s->imp) fb2_import_destroy(s->imp);
  delete s;
  return 0;
}
extern "C" int fb2_fe_space_get_sizes(fb2_space_t s, int64_t sizes[8]) {
  sizes[0] = s->m.n_cells; sizes[1] = s->m.n_local; sizes[2] = s->rt.ndofs; sizes[3] = s->n_free;
  sizes[4] = s->n_dirichlet; sizes[5] = s->n_free_ghost; sizes[6] = s->n_interior; sizes[7] = s->n_interface;
  return 0;
}
extern "C" int fb2_fe_space_get_lst_dofs_gids(fb2_space_t s, int32_t* host_out) {
  FB2_CUDA(fb2_memcpy(host_out, s->d_gids, sizeof(int) * (size_t)s->m.n_cells * s->rt.ndofs, cudaMemcpyDeviceToHost));
  return 0;
}
extern "C" int fb2_fe_space_get_cells_gids(fb2_space_t s, int64_t* gid_out, int32_t* part_out) {
  // uniform_hex_mesh_generator.f90:468-489 (closed form; tiny, evaluated on the host from the device permutation)
  const MeshDesc& m = s->m;
  std::vector<int> perm(m.n_cells);
  if (s->d_perm) FB2_CUDA(fb2_memcpy(perm.data(), s->d_perm, sizeof(int) * m.n_cells, cudaMemcpyDeviceToHost));
  else for (int i = 0; i < m.n_cells; ++i) perm[i] = i;
  for (int lex = 0; lex < m.n_cells; ++lex) {
    int ijk[3] = {lex % m.ntot[0], (lex / m.ntot[0]) % m.ntot[1], lex / (m.ntot[0] * m.ntot[1])};
    int mp[3]; i64 g[3];
    for (int k = 0; k < 3; ++k) {
      if (m.L[k] && ijk[k] == 0) mp[k] = m.pijk[k] - 1;
      else if (m.R[k] && ijk[k] == m.ntot[k] - 1) mp[k] = m.pijk[k] + 1;
      else mp[k] = m.pijk[k];
      g[k] = m.first[k] + ijk[k];
    }
    if (gid_out) gid_out[perm[lex]] = 1 + g[0] + (i64)m.N[0] * (g[1] + (i64)m.N[1] * g[2]);
    if (part_out) part_out[perm[lex]] = 1 + mp[0] + m.P[0] * (mp[1] + m.P[1] * mp[2]);
  }
  return 0;
}
extern "C" int fb2_fe_space_get_fixed_values(fb2_space_t s, double* host_out) {
  FB2_CUDA(fb2_memcpy(host_out, s->d_fixed, sizeof(double) * (size_t)s->n_dirichlet, cudaMemcpyDeviceToHost));
  return 0;
}
extern "C" int fb2_fe_space_set_fixed_values(fb2_space_t s, const double* host_in) {
  FB2_CUDA(fb2_memcpy(s->d_fixed, host_in, sizeof(double) * (size_t)s->n_dirichlet, cudaMemcpyHostToDevice));
  return 0;
}

// ------------------------------------------------------------------------------------------
// CSR graph.  Row r <-> (component, lattice node); its columns are the free DOFs of every node in
// the union of the cells around that node, all components, ascending by DOF id -- exactly what
// coo sort_and_compress yields (base_sparse_matrix.f90:4479-4628, csr_sparse_matrix.f90:416-479).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void row_box(int p, const int* M, i64 lat, int* lo, int* hi) {
  int a[3];
  a[0] = (int)(lat % M[0]);
  i64 r = lat / M[0];
  a[1] = (int)(r % M[1]);
  a[2] = (int)(r / M[1]);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    if (a[k] % p == 0) { lo[k] = max(a[k] - p, 0); hi[k] = min(a[k] + p, M[k] - 1); }
    else { lo[k] = (a[k] / p) * p; hi[k] = lo[k] + p; }
  }
}

__global__ void k_row_count(int p, int ncomp, int M0, int M1, int M2, i64 lat_size, i64 n, int sym,
                            const int* __restrict__ dof_node, const int* __restrict__ node_dof, i64* cnt) {
  i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) { if (r == n) cnt[n] = 0; return; }
  int M[3] = {M0, M1, M2}, lo[3], hi[3];
  i64 lat = dof_node[r] % lat_size;
  row_box(p, M, lat, lo, hi);
  int c = 0;
  for (int comp = 0; comp < ncomp; ++comp)
    for (int a2 = lo[2]; a2 <= hi[2]; ++a2)
      for (int a1 = lo[1]; a1 <= hi[1]; ++a1) {
        const int* row = node_dof + comp * lat_size + (i64)M0 * (a1 + (i64)M1 * a2);
        for (int a0 = lo[0]; a0 <= hi[0]; ++a0) {
          int g = row[a0];
          c += (g > 0 && g <= n && (!sym || g >= r + 1)) ? 1 : 0;
        }
      }
  cnt[r] = c;
}

template <int SORT_N>
__global__ void k_row_fill(int p, int ncomp, int M0, int M1, int M2, i64 lat_size, i64 n, int sym,
                           const int* __restrict__ dof_node, const int* __restrict__ node_dof,
                           const i64* __restrict__ irp, int* __restrict__ ja) {
  extern __shared__ int sbuf[];
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int* sb = sbuf + w * SORT_N;
  int M[3] = {M0, M1, M2};
  for (i64 r = (i64)blockIdx.x * warps + w; r < n; r += (i64)gridDim.x * warps) {
    int lo[3], hi[3];
    i64 lat = dof_node[r] % lat_size;
    row_box(p, M, lat, lo, hi);
    int b0 = hi[0] - lo[0] + 1, b1 = hi[1] - lo[1] + 1, b2 = hi[2] - lo[2] + 1;
    int nb = b0 * b1 * b2, ncand = nb * ncomp;
    for (int e = lane; e < SORT_N; e += 32) {
      int g = 0x7fffffff;
      if (e < ncand) {
        int comp = e / nb, q = e % nb;
        int a0 = lo[0] + q % b0, a1 = lo[1] + (q / b0) % b1, a2 = lo[2] + q / (b0 * b1);
        int v = node_dof[comp * lat_size + a0 + (i64)M0 * (a1 + (i64)M1 * a2)];
        if (v > 0 && v <= n && (!sym || v >= r + 1)) g = v;
      }
      sb[e] = g;
    }
    __syncwarp();
    for (int k = 2; k <= SORT_N; k <<= 1)
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int idx = lane; idx < SORT_N; idx += 32) {
          int ixj = idx ^ j;
          if (ixj > idx) {
            int a = sb[idx], b = sb[ixj];
            bool asc = (idx & k) == 0;
            if ((a > b) == asc) { sb[idx] = b; sb[ixj] = a; }
          }
        }
        __syncwarp();
      }
    i64 start = irp[r];
    int cnt = (int)(irp[r + 1] - start);
    for (int e = lane; e < cnt; e += 32) ja[start + e] = sb[e];
    __syncwarp();
  }
}

// Same rows, sorted in registers: lane l holds candidates l*KPL .. l*KPL+KPL-1; bitonic network with in-lane exchanges for
// distances < KPL and warp shuffles above (no shared memory, no __syncwarp): rows of up to 32*KPL candidates.
template <int KPL>
__global__ void __launch_bounds__(128) k_row_fill_reg(int p, int ncomp, int M0, int M1, int M2, i64 lat_size, i64 n, int sym,
                                                      const int* __restrict__ dof_node, const int* __restrict__ node_dof,
                                                      const i64* __restrict__ irp, int* __restrict__ ja) {
  constexpr int SORT_N = 32 * KPL;
  const int warps = blockDim.x >> 5, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int M[3] = {M0, M1, M2};
  for (i64 r = (i64)blockIdx.x * warps + w; r < n; r += (i64)gridDim.x * warps) {
    int lo[3], hi[3];
    const i64 lat = dof_node[r] % lat_size;
    row_box(p, M, lat, lo, hi);
    const int b0 = hi[0] - lo[0] + 1, b1 = hi[1] - lo[1] + 1, b2 = hi[2] - lo[2] + 1;
    const int nb = b0 * b1 * b2, ncand = nb * ncomp;
    int v[KPL];
    // candidate e = lane * KPL + i  <->  (component, box position): decoded once per lane (integer divisions by run-time
    // box sizes), then advanced with carries for the lane's other KPL - 1 candidates
    int e = lane * KPL;
    int comp = e / nb, q = e - comp * nb;
    int o0 = q % b0, o1 = (q / b0) % b1, o2 = q / (b0 * b1);
#pragma unroll
    for (int i = 0; i < KPL; ++i, ++e) {
      int g = 0x7fffffff;
      if (e < ncand) {
        const int c = node_dof[comp * lat_size + (lo[0] + o0) + (i64)M0 * ((lo[1] + o1) + (i64)M1 * (lo[2] + o2))];
        if (c > 0 && c <= n && (!sym || c >= r + 1)) g = c;
      }
      v[i] = g;
      if (++o0 == b0) { o0 = 0; if (++o1 == b1) { o1 = 0; if (++o2 == b2) { o2 = 0; ++comp; } } }
    }
#pragma unroll
    for (int k = 2; k <= SORT_N; k <<= 1) {
#pragma unroll
      for (int j = k >> 1; j > 0; j >>= 1) {
        if (j >= KPL) {
          const int lj = j / KPL;
#pragma unroll
          for (int i = 0; i < KPL; ++i) {
            const int idx = lane * KPL + i;
            const int other = __shfl_xor_sync(0xffffffffu, v[i], lj);
            const bool up = (idx & k) == 0, lower = (idx & j) == 0;
            v[i] = (lower == up) ? min(v[i], other) : max(v[i], other);
          }
        } else {
#pragma unroll
          for (int i = 0; i < KPL; ++i) {
            if ((i & j) == 0) {
              const int i2 = i | j;
              const bool up = ((lane * KPL + i) & k) == 0;
              const int a = v[i], b = v[i2];
              if ((a > b) == up) { v[i] = b; v[i2] = a; }
            }
          }
        }
      }
    }
    const i64 start = irp[r];
    const int cnt = (int)(irp[r + 1] - start);
#pragma unroll
    for (int i = 0; i < KPL; ++i) {
      const int e = lane * KPL + i;
      if (e < cnt) ja[start + e] = v[i];
    }
  }
}

// Same rows again, the sort replaced by a learned permutation wherever it repeats.  A warp walks rows r, r + D, r + 2 D, ...
// (D = DOFs an interior cell creates: rows D apart sit on the same kind of n-face one cell further, so their candidate box has
// the same shape and -- away from the boundary -- the same order of DOF ids).  It keeps, per sorted slot, the lattice offset
// (relative to the box origin) of the candidate that landed there when it last sorted a row; for the next row it gathers the
// candidates straight into that order and checks: same box shape, same row length, every gathered id a free DOF (and in the
// stored triangle), ids strictly increasing.  irp already holds the exact number of admissible candidates of the row, the
// offsets are distinct positions of the box, so a strictly increasing admissible sequence of that length IS the sorted row.
// Any failed check falls back to the sort (64-bit keys: id, offset) and re-learns.  ~60 instead of ~1000 instructions per row.
template <int KPL>
__global__ void __launch_bounds__(128) k_row_fill_perm(int p, int ncomp, int M0, int M1, int M2, int lat_size, i64 n, int sym,
                                                       const int* __restrict__ dof_node, const int* __restrict__ node_dof,
                                                       const i64* __restrict__ irp, int* __restrict__ ja, int D, int J,
                                                       unsigned long long* n_sorted) {
  constexpr int SORT_N = 32 * KPL;
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31;
  const i64 gw = (i64)blockIdx.x * warps + (threadIdx.x >> 5), nw = (i64)gridDim.x * warps;
  const i64 span = (i64)D * J;
  const i64 ntask = ((n + span - 1) / span) * D;
  unsigned long long sorted_here = 0;
  for (i64 task = gw; task < ntask; task += nw) {
    const i64 blk = task / D;
    const int c = (int)(task - blk * D);
    int off[KPL];
#pragma unroll
    for (int i = 0; i < KPL; ++i) off[i] = 0;
    int pb0 = -1, pb1 = -1, pb2 = -1, pcnt = -1;
    for (int j = 0; j < J; ++j) {
      const i64 r = blk * span + (i64)j * D + c;
      if (r >= n) break;
      const unsigned lat = (unsigned)dof_node[r] % (unsigned)lat_size;
      const int a0 = (int)(lat % (unsigned)M0), a12 = (int)(lat / (unsigned)M0);
      const int a1 = a12 % M1, a2 = a12 / M1;
      int lo0, lo1, lo2, b0, b1, b2;
      {
        int hi;
        if (a0 % p == 0) { lo0 = max(a0 - p, 0); hi = min(a0 + p, M0 - 1); } else { lo0 = (a0 / p) * p; hi = lo0 + p; }
        b0 = hi - lo0 + 1;
        if (a1 % p == 0) { lo1 = max(a1 - p, 0); hi = min(a1 + p, M1 - 1); } else { lo1 = (a1 / p) * p; hi = lo1 + p; }
        b1 = hi - lo1 + 1;
        if (a2 % p == 0) { lo2 = max(a2 - p, 0); hi = min(a2 + p, M2 - 1); } else { lo2 = (a2 / p) * p; hi = lo2 + p; }
        b2 = hi - lo2 + 1;
      }
      const int base = lo0 + M0 * (lo1 + M1 * lo2);
      const i64 start = irp[r];
      const int cnt = (int)(irp[r + 1] - start);
      int v[KPL];
      bool ok = b0 == pb0 && b1 == pb1 && b2 == pb2 && cnt == pcnt;
      if (ok) {
        bool good = true;
#pragma unroll
        for (int i = 0; i < KPL; ++i) {
          const int slot = lane * KPL + i;
          v[i] = 0x7fffffff;
          if (slot < cnt) {
            const int g = node_dof[base + off[i]];
            good = good && g > 0 && g <= n && (!sym || g >= r + 1);
            v[i] = g;
          }
        }
#pragma unroll
        for (int i = 0; i + 1 < KPL; ++i) good = good && (lane * KPL + i + 1 >= cnt || v[i] < v[i + 1]);
        const int nxt = __shfl_down_sync(0xffffffffu, v[0], 1);
        good = good && (lane == 31 || lane * KPL + KPL >= cnt || v[KPL - 1] < nxt);
        ok = __all_sync(0xffffffffu, good);
      }
      if (!ok) {
        // sort (id, offset) pairs and learn the permutation
        ++sorted_here;
        const int nb = b0 * b1 * b2, ncand = nb * ncomp;
        unsigned long long key[KPL];
        int e = lane * KPL;
        int comp = e / nb, q = e - comp * nb;
        int o0 = q % b0, o1 = (q / b0) % b1, o2 = q / (b0 * b1);
#pragma unroll
        for (int i = 0; i < KPL; ++i, ++e) {
          unsigned long long k = ((unsigned long long)0x7fffffffu << 32) | (unsigned)e;
          if (e < ncand) {
            const int o = comp * lat_size + o0 + M0 * (o1 + M1 * o2);
            const int g = node_dof[base + o];
            if (g > 0 && g <= n && (!sym || g >= r + 1)) k = ((unsigned long long)(unsigned)g << 32) | (unsigned)o;
          }
          key[i] = k;
          if (++o0 == b0) { o0 = 0; if (++o1 == b1) { o1 = 0; if (++o2 == b2) { o2 = 0; ++comp; } } }
        }
#pragma unroll
        for (int k = 2; k <= SORT_N; k <<= 1) {
#pragma unroll
          for (int jj = k >> 1; jj > 0; jj >>= 1) {
            if (jj >= KPL) {
              const int lj = jj / KPL;
#pragma unroll
              for (int i = 0; i < KPL; ++i) {
                const int idx = lane * KPL + i;
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, key[i], lj);
                const bool up = (idx & k) == 0, lower = (idx & jj) == 0;
                key[i] = (lower == up) ? min(key[i], other) : max(key[i], other);
              }
            } else {
#pragma unroll
              for (int i = 0; i < KPL; ++i) {
                if ((i & jj) == 0) {
                  const int i2 = i | jj;
                  const bool up = ((lane * KPL + i) & k) == 0;
                  const unsigned long long a = key[i], b = key[i2];
                  if ((a > b) == up) { key[i] = b; key[i2] = a; }
                }
              }
            }
          }
        }
#pragma unroll
        for (int i = 0; i < KPL; ++i) { v[i] = (int)(key[i] >> 32); off[i] = (int)(unsigned)key[i]; }
        pb0 = b0; pb1 = b1; pb2 = b2; pcnt = cnt;
      }
#pragma unroll
      for (int i = 0; i < KPL; ++i) {
        const int slot = lane * KPL + i;
        if (slot < cnt) ja[start + slot] = v[i];
      }
    }
  }
  if (lane == 0 && sorted_here) atomicAdd(n_sorted, sorted_here);
}

extern "C" int fb2_csr_create(fb2_matrix_t* out, fb2_space_t s, int symmetric_storage) {
  FB2_CHECK(s != nullptr, "null space");
  fb2_matrix_s* A = new fb2_matrix_s();
  A->s = s; A->sym = symmetric_storage ? 1 : 0; A->n = s->n_free;
  cudaStream_t st = fb2_stream();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0, st);
  const i64 n = A->n;
  if (s->d_coords) {   // host-built space: no lattice, the graph comes from the cell DOF lists
    FB2_TRY(fb2_csr_graph_from_gids(A));
    cudaEventRecord(e1, st);
    FB2_CUDA(cudaStreamSynchronize(st));
    float ms_u = 0;
    cudaEventElapsedTime(&ms_u, e0, e1);
    fb2_record_phase("pattern", ms_u);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *out = A;
    return 0;
  }
  i64* cnt;
  FB2_TRY(fb2_dalloc(&cnt, (size_t)n + 1));
  FB2_TRY(fb2_dalloc(&A->d_irp, (size_t)n + 72));  // padding: the SpMV bulk-copies irp slices of up to 66 entries per tile
  FB2_CUDA(cudaMemsetAsync(A->d_irp, 0, sizeof(i64) * ((size_t)n + 72), st));
  const int p = s->rt.p, ncomp = s->rt.ncomp;
  k_row_count<<<(unsigned)((n + 1 + 127) / 128), 128, 0, st>>>(p, ncomp, s->Mlat[0], s->Mlat[1], s->Mlat[2], s->lat_size, n,
                                                              A->sym, s->d_dof_node, s->d_node_dof, cnt);
  FB2_CUDA(cudaGetLastError());
  FB2_TRY(exclusive_scan(cnt, A->d_irp, n + 1));
  FB2_CUDA(fb2_memcpy(&A->nnz, A->d_irp + n, sizeof(i64), cudaMemcpyDeviceToHost));
  cudaFree(cnt);
  FB2_TRY(fb2_dalloc(&A->d_ja, (size_t)A->nnz + 16));    // +16: TMA bulk copies read whole 16-byte groups
  FB2_TRY(fb2_dalloc(&A->d_val, (size_t)A->nnz + 16));
  FB2_CUDA(cudaMemsetAsync(A->d_ja + A->nnz, 0, 16 * sizeof(int), st));
  FB2_CUDA(cudaMemsetAsync(A->d_val, 0, sizeof(double) * ((size_t)A->nnz + 16), st));
  int maxrow = (2 * p + 1) * (2 * p + 1) * (2 * p + 1) * ncomp;
  int blocks = fb2_num_sms() * 8, threads = 128;
  if (n > 0) {
#define LAUNCH_FILL(SN)                                                                                                  \
  k_row_fill<SN><<<blocks, threads, (threads / 32) * SN * sizeof(int), st>>>(p, ncomp, s->Mlat[0], s->Mlat[1], s->Mlat[2], \
                                                                             s->lat_size, n, A->sym, s->d_dof_node,      \
                                                                             s->d_node_dof, A->d_irp, A->d_ja)
    const int fill_variant = getenv("FB2_ROW_FILL") ? atoi(getenv("FB2_ROW_FILL")) : 2;
#define LAUNCH_FILL_REG(K)                                                                                               \
  k_row_fill_reg<K><<<fb2_num_sms() * 16, 128, 0, st>>>(p, ncomp, s->Mlat[0], s->Mlat[1], s->Mlat[2], s->lat_size, n, A->sym, \
                                                        s->d_dof_node, s->d_node_dof, A->d_irp, A->d_ja)
    // 2 (default): learned permutation, needs 32-bit lattice offsets; 1: register sort of every row; 0: shared-memory sort
    const bool perm_ok = fill_variant >= 2 && maxrow <= 512 && s->lat_size * ncomp < 0x7fffffffLL;
    if (perm_ok) {
      unsigned long long* d_ns;
      FB2_TRY(fb2_dalloc(&d_ns, 1));
      FB2_CUDA(cudaMemsetAsync(d_ns, 0, sizeof(unsigned long long), st));
      int D = ncomp;
      for (int k = 0; k < 3; ++k) D *= p;                  // DOFs created by an interior cell
      if (getenv("FB2_ROW_FILL_STRIDE")) D = std::max(1, atoi(getenv("FB2_ROW_FILL_STRIDE")));
      const int J = 32;
#define LAUNCH_FILL_PERM(K)                                                                                              \
  k_row_fill_perm<K><<<fb2_num_sms() * 16, 128, 0, st>>>(p, ncomp, s->Mlat[0], s->Mlat[1], s->Mlat[2], (int)s->lat_size, n, A->sym, \
                                                         s->d_dof_node, s->d_node_dof, A->d_irp, A->d_ja, D, J, d_ns)
      if (maxrow <= 32) LAUNCH_FILL_PERM(1);
      else if (maxrow <= 128) LAUNCH_FILL_PERM(4);
      else LAUNCH_FILL_PERM(16);
#undef LAUNCH_FILL_PERM
      unsigned long long h_ns = 0;
      FB2_CUDA(cudaMemcpyAsync(&h_ns, d_ns, sizeof(h_ns), cudaMemcpyDeviceToHost, st));
      FB2_CUDA(cudaStreamSynchronize(st));
      cudaFree(d_ns);
      fb2_record_phase("pattern_rows_sorted", (double)h_ns);
    } else if (fill_variant && maxrow <= 32) LAUNCH_FILL_REG(1);
    else if (fill_variant && maxrow <= 128) LAUNCH_FILL_REG(4);
    else if (fill_variant && maxrow <= 512) LAUNCH_FILL_REG(16);
    else if (maxrow <= 32) LAUNCH_FILL(32);
    else if (maxrow <= 128) LAUNCH_FILL(128);
    else if (maxrow <= 512) LAUNCH_FILL(512);
    else if (maxrow <= 1024) LAUNCH_FILL(1024);
    else { FB2_CHECK(maxrow <= 2048, "row too long"); LAUNCH_FILL(2048); }
#undef LAUNCH_FILL
#undef LAUNCH_FILL_REG
  }
  FB2_CUDA(cudaGetLastError());
  cudaEventRecord(e1, st);
  FB2_CUDA(cudaStreamSynchronize(st));
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  fb2_record_phase("pattern", ms);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  *out = A;
  return 0;
}
extern "C" int fb2_csr_destroy(fb2_matrix_t A) {
  if (!A) return 0;
  fb2_solver_ws_free(A);
  cudaFree(A->d_irp); cudaFree(A->d_ja); cudaFree(A->d_val); cudaFree(A->d_cell_tab); cudaFree(A->d_tabs); cudaFree(A->d_tile_row); cudaFree(A->d_tile_nrows);
  cudaFree(A->d_idx16); cudaFree(A->d_dict); cudaFree(A->d_dict_ptr);
  cudaFree(A->d_tile_class); cudaFree(A->d_cls_desc); cudaFree(A->d_cls_idx); cudaFree(A->d_tile_priv);
  cudaFree(A->d_rows_plan); cudaFree(A->d_chunk_base); cudaFree(A->d_row_fast); cudaFree(A->d_slow_rows); cudaFree(A->d_slow_cells);
  if (A->owns_space) fb2_fe_space_destroy(A->s);
  delete A;
  return 0;
}
extern "C" int fb2_csr_get_sizes(fb2_matrix_t A, int64_t* num_rows, int64_t* nnz) {
  *num_rows = A->n; *nnz = A->nnz;
  return 0;
}
extern "C" int fb2_csr_get_pattern(fb2_matrix_t A, int64_t* irp_host, int32_t* ja_host) {
  if (irp_host) {
    FB2_CUDA(fb2_memcpy(irp_host, A->d_irp, sizeof(i64) * (size_t)(A->n + 1), cudaMemcpyDeviceToHost));
    for (i64 i = 0; i <= A->n; ++i) irp_host[i] += 1;  // Fortran irp(1) = 1
  }
  if (ja_host) FB2_CUDA(fb2_memcpy(ja_host, A->d_ja, sizeof(int) * (size_t)A->nnz, cudaMemcpyDeviceToHost));
  return 0;
}
extern "C" int fb2_csr_get_values(fb2_matrix_t A, double* val_host) {
  FB2_CUDA(cudaStreamSynchronize(fb2_stream()));
  FB2_CUDA(fb2_memcpy(val_host, A->d_val, sizeof(double) * (size_t)A->nnz, cudaMemcpyDeviceToHost));
  return 0;
}
extern "C" int fb2_csr_set_values(fb2_matrix_t A, const double* val_host) {
  FB2_CUDA(fb2_memcpy(A->d_val, val_host, sizeof(double) * (size_t)A->nnz, cudaMemcpyHostToDevice));
  return 0;
}
