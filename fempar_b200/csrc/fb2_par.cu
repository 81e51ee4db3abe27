// par_fe_space_t on the device: interface couplings, owner decision, dof_import_t and the interior-then-interface
// renumbering of the local DOFs.  Reference:
//   compute_raw_interface_data_by_continuity        Integration/FESpace/sbm_par_fe_space.i90:1877-2033
//   raw_interface_data_by_continuity_decide_owner                                           :2737-2851
//   generate_non_consecutive_dof_ggid                                                       :1103-1118
//   dof_import_create, fill_first_I_then_G_renumbering, renum_dofs   LinearAlgebra/Vector/dof_import.f90:102-213,365-421
//   renum_dofs_first_interior_then_interface                         sbm_par_fe_space.i90:449-474
// The reference sweeps interface vefs and the cells around them; on a structured part everything it derives per DOF
// (parts around, max part, max cell global id around the vef, local position of the DOF in that cell) has a closed form
// in the lattice coordinates of the DOF, so one thread per DOF emits its couplings, a radix sort orders them by
// (neighbour part, #parts around + 1, non-consecutive global DOF id) and one thread per coupling decides ownership.
#include <cub/cub.cuh>
#include <limits.h>
#include <algorithm>
#include "fb2_internal.h"

struct ParGeom {
  int N[3], P[3], pijk[3], nloc[3], L[3], R[3];
  int p, n1, nnode, ncomp;
  int M[3];
  i64 lat_size;
  int mypart;
};

// interface dims of a lattice node: dirs[k] = -1 / +1 when the node lies on the plane shared with the lower / upper neighbour part
__device__ __forceinline__ int node_dirs(const ParGeom& G, const int* a, int* dirs) {
  int k_itf = 0;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    dirs[k] = 0;
    if (a[k] == 0 && G.L[k]) dirs[k] = -1;
    else if (a[k] == G.p * G.nloc[k] && G.R[k]) dirs[k] = +1;
    k_itf += dirs[k] != 0;
  }
  return k_itf;
}
__device__ __forceinline__ void lat_coords(const ParGeom& G, i64 lat, int* a) {
  a[0] = (int)(lat % G.M[0]);
  i64 r = lat / G.M[0];
  a[1] = (int)(r % G.M[1]);
  a[2] = (int)(r / G.M[1]);
}

__global__ void k_par_count(ParGeom G, i64 n_free, const int* __restrict__ dof_node, int* cnt) {
  i64 g = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (g > n_free) return;
  if (g == n_free) { cnt[g] = 0; return; }
  int a[3], dirs[3];
  lat_coords(G, dof_node[g] % G.lat_size, a);
  int k = node_dirs(G, a, dirs);
  cnt[g] = (1 << k) - 1;
}

// one coupling column per (interface DOF, neighbour part)
__global__ void k_par_fill(ParGeom G, i64 n_free, const int* __restrict__ dof_node, const int* __restrict__ off,
                           unsigned* major /*part*16 + flag*/, unsigned long long* ggid, int* dof, int* maxpart) {
  i64 g = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_free) return;
  int a[3], dirs[3];
  const int node_id = dof_node[g];
  const int comp = (int)(node_id / G.lat_size);
  lat_coords(G, node_id % G.lat_size, a);
  const int k_itf = node_dirs(G, a, dirs);
  if (k_itf == 0) return;
  // max cell (global lexicographic id) around the vef that carries the DOF, and the DOF's local position in it
  i64 cell_lex = 0, mul = 1;
  int lnode = 0, lmul = 1, mp[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int gn = (G.pijk[k] * G.nloc[k]) * G.p + a[k];   // global lattice coordinate
    int cellk, lk;
    if (gn % G.p != 0) { cellk = gn / G.p; lk = gn % G.p; }
    else {
      cellk = gn / G.p; lk = 0;                             // the cell on the + side, node on its low face
      if (cellk > G.N[k] - 1) { cellk = G.N[k] - 1; lk = G.p; }
    }
    cell_lex += mul * cellk; mul *= G.N[k];
    lnode += lmul * lk; lmul *= G.n1;
    mp[k] = G.pijk[k] + (dirs[k] > 0 ? 1 : 0);
  }
  const int local_idof = comp * G.nnode + lnode;            // 0-based (reference: lpos_dof_fe_dofs - 1)
  const unsigned long long key = (unsigned long long)cell_lex + ((unsigned long long)local_idof << 44);
  const int max_part = 1 + mp[0] + G.P[0] * (mp[1] + G.P[1] * mp[2]);
  const unsigned flag = 1u << k_itf;                        // number of parts around = neighbours + 1
  int o = off[g];
  for (int m = 1; m < 8; ++m) {                             // non-empty subsets of the interface dims
    bool ok = true;
    int q[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int bit = (m >> k) & 1;
      if (bit && dirs[k] == 0) ok = false;
      q[k] = G.pijk[k] + (bit ? dirs[k] : 0);
    }
    if (!ok) continue;
    const int part = 1 + q[0] + G.P[0] * (q[1] + G.P[1] * q[2]);
    major[o] = (unsigned)part * 16u + flag;
    ggid[o] = key; dof[o] = (int)g + 1; maxpart[o] = max_part;
    ++o;
  }
}

// owner flags on the sorted couplings: -2 owner, -3 non owner, -1 uncoupled (dof_import.f90:44-46)
__global__ void k_par_owner(int ncol, int mypart, const unsigned* __restrict__ major, const int* __restrict__ maxpart, int* owner) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncol) return;
  const unsigned mk = major[c];
  const int part = (int)(mk >> 4), flag = (int)(mk & 15u);
  if (flag == 2) {
    // run of couplings with this neighbour shared by exactly two parts: split half/half, the max part owns the first half
    int lo = 0, hi = c;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (major[mid] < mk) lo = mid + 1; else hi = mid; }
    const int start = lo;
    lo = c; hi = ncol;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (major[mid] <= mk) lo = mid + 1; else hi = mid; }
    const int size = lo - start;
    const bool first_half = (c - start) < size / 2;
    const bool i_am_max = mypart > part;
    owner[c] = (first_half == i_am_max) ? -2 : -3;
  } else {
    const int mx = maxpart[c];
    owner[c] = (mypart == mx) ? -2 : (part == mx ? -3 : -1);
  }
}

// ------------------------------------------------------------------------------------------
// interior-then-interface renumbering (dof_import.f90:391-421)
// ------------------------------------------------------------------------------------------
__global__ void k_first_pos(int nseq, const int* __restrict__ seq /*1-based dofs: pack then unpack*/, int* first) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nseq) atomicMin(&first[seq[i] - 1], i);
}
__global__ void k_flag_first(int nseq, const int* __restrict__ seq, const int* __restrict__ first, int* flag) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= nseq) flag[i] = (i < nseq && first[seq[i] - 1] == i) ? 1 : 0;
}
__global__ void k_flag_interior(i64 n, const int* __restrict__ first, int* flag) {
  i64 g = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (g <= n) flag[g] = (g < n && first[g] == INT_MAX) ? 1 : 0;
}
__global__ void k_perm_interface(int nseq, const int* __restrict__ seq, const int* __restrict__ first, const int* __restrict__ rank,
                                 int n_interior, int* perm) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nseq && first[seq[i] - 1] == i) perm[seq[i] - 1] = n_interior + 1 + rank[i];
}
__global__ void k_perm_interior(i64 n, const int* __restrict__ first, const int* __restrict__ rank, int* perm) {
  i64 g = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (g < n && first[g] == INT_MAX) perm[g] = 1 + rank[g];
}
__global__ void k_apply_perm_gids(i64 total, i64 n_free, const int* __restrict__ perm, int* gids) {
  i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  int g = gids[t];
  if (g > 0 && g <= n_free) gids[t] = perm[g - 1];        // only free DOFs are renumbered (sbm_fe_cell_iterator.i90:907-933)
}
__global__ void k_apply_perm_idx(int n, const int* __restrict__ perm, const int* __restrict__ in1 /*1-based*/, int* out0 /*0-based*/,
                                 double* weight_zero /*nullable: weight[new] = 0*/) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int nw = perm[in1[i] - 1] - 1;
  out0[i] = nw;
  if (weight_zero) weight_zero[nw] = 0.0;
}
__global__ void k_fill_ones(i64 n, double* w) {
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) w[i] = 1.0;
}

template <class T>
static int scan_excl(const T* in, T* out, i64 n) {
  void* tmp = nullptr; size_t bytes = 0;
  FB2_CUDA(cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, n, fb2_stream()));
  FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
  cudaError_t e = cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, n, fb2_stream());
  cudaStreamSynchronize(fb2_stream());
  cudaFree(tmp);
  FB2_CUDA(e);
  return 0;
}
template <class K>
static int sort_pairs(K* keys_in, K* keys_out, int* vals_in, int* vals_out, int n) {
  void* tmp = nullptr; size_t bytes = 0;
  FB2_CUDA(cub::DeviceRadixSort::SortPairs(tmp, bytes, keys_in, keys_out, vals_in, vals_out, n, 0, (int)sizeof(K) * 8, fb2_stream()));
  FB2_CUDA(cudaMalloc(&tmp, bytes ? bytes : 1));
  cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, bytes, keys_in, keys_out, vals_in, vals_out, n, 0, (int)sizeof(K) * 8, fb2_stream());
  cudaStreamSynchronize(fb2_stream());
  cudaFree(tmp);
  FB2_CUDA(e);
  return 0;
}
template <class T>
__global__ void k_gather(int n, const int* __restrict__ idx, const T* __restrict__ in, T* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[idx[i]];
}
__global__ void k_fill_int(i64 n, int* v, int val) {
  i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = val;
}
__global__ void k_iota(int n, int* v) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = i;
}

#define GRID(n) (unsigned)(((n) + 255) / 256), 256, 0, st

int fb2_build_par_layer(fb2_space_s* s) {
  cudaStream_t st = fb2_stream();
  const MeshDesc& m = s->m;
  const i64 n = s->n_free;
  FB2_TRY(fb2_build_node_maps(s));            // lattice <-> DOF maps in the numbering of list_dofs (rebuilt after the renumbering)
  ParGeom G;
  for (int k = 0; k < 3; ++k) {
    G.N[k] = m.N[k]; G.P[k] = m.P[k]; G.pijk[k] = m.pijk[k]; G.nloc[k] = m.nloc[k]; G.L[k] = m.L[k]; G.R[k] = m.R[k];
    G.M[k] = s->Mlat[k];
  }
  G.p = s->rt.p; G.n1 = s->rt.n1; G.nnode = s->rt.nnode; G.ncomp = s->rt.ncomp; G.lat_size = s->lat_size; G.mypart = s->part_id;
  fb2_import_s* imp = new fb2_import_s();
  s->imp = imp;
  // 1. couplings
  int *cnt, *off;
  FB2_TRY(fb2_dalloc(&cnt, (size_t)n + 1)); FB2_TRY(fb2_dalloc(&off, (size_t)n + 1));
  k_par_count<<<GRID(n + 1)>>>(G, n, s->d_dof_node, cnt);
  FB2_TRY(scan_excl(cnt, off, n + 1));
  int ncol = 0;
  FB2_CUDA(fb2_memcpy(&ncol, off + n, sizeof(int), cudaMemcpyDeviceToHost));
  std::vector<int> h_part(ncol), h_owner(ncol), h_dof(ncol);
  if (ncol > 0) {
    unsigned *major, *major_s; unsigned long long *ggid, *ggid_s; int *dof, *maxpart, *idx0, *idx1, *idx2, *owner, *dof_s, *maxpart_s;
    FB2_TRY(fb2_dalloc(&major, ncol)); FB2_TRY(fb2_dalloc(&major_s, ncol)); FB2_TRY(fb2_dalloc(&ggid, ncol)); FB2_TRY(fb2_dalloc(&ggid_s, ncol));
    FB2_TRY(fb2_dalloc(&dof, ncol)); FB2_TRY(fb2_dalloc(&maxpart, ncol)); FB2_TRY(fb2_dalloc(&idx0, ncol)); FB2_TRY(fb2_dalloc(&idx1, ncol));
    FB2_TRY(fb2_dalloc(&idx2, ncol)); FB2_TRY(fb2_dalloc(&owner, ncol)); FB2_TRY(fb2_dalloc(&dof_s, ncol)); FB2_TRY(fb2_dalloc(&maxpart_s, ncol));
    k_par_fill<<<GRID(n)>>>(G, n, s->d_dof_node, off, major, ggid, dof, maxpart);
    k_iota<<<GRID(ncol)>>>(ncol, idx0);
    // lexicographic (major, ggid): stable LSD -- sort by ggid, then by major (sort_array_cols_by_row_section, Tools/sort.f90:118-141)
    FB2_TRY(sort_pairs(ggid, ggid_s, idx0, idx1, ncol));
    k_gather<<<GRID(ncol)>>>(ncol, idx1, major, major_s);
    FB2_TRY(sort_pairs(major_s, major, idx1, idx2, ncol));     // major now holds the sorted keys, idx2 the permutation
    k_gather<<<GRID(ncol)>>>(ncol, idx2, dof, dof_s);
    k_gather<<<GRID(ncol)>>>(ncol, idx2, maxpart, maxpart_s);
    k_par_owner<<<GRID(ncol)>>>(ncol, s->part_id, major, maxpart_s, owner);
    FB2_CUDA(cudaGetLastError());
    std::vector<unsigned> h_major(ncol);
    FB2_CUDA(fb2_memcpy(h_major.data(), major, sizeof(unsigned) * ncol, cudaMemcpyDeviceToHost));
    FB2_CUDA(fb2_memcpy(h_owner.data(), owner, sizeof(int) * ncol, cudaMemcpyDeviceToHost));
    FB2_CUDA(fb2_memcpy(h_dof.data(), dof_s, sizeof(int) * ncol, cudaMemcpyDeviceToHost));
    for (int c = 0; c < ncol; ++c) h_part[c] = (int)(h_major[c] >> 4);
    cudaFree(major); cudaFree(major_s); cudaFree(ggid); cudaFree(ggid_s); cudaFree(dof); cudaFree(maxpart); cudaFree(idx0); cudaFree(idx1);
    cudaFree(idx2); cudaFree(owner); cudaFree(dof_s); cudaFree(maxpart_s);
  }
  cudaFree(cnt); cudaFree(off);
  // 2. dof_import_create (dof_import.f90:102-213): the couplings are grouped by neighbour part already, so the lists are
  //    plain compactions; neighbours appear in increasing part id
  for (int pass = 0; pass < 2; ++pass) {
    const int want = pass == 0 ? -2 : -3;
    std::vector<int>& lst = pass == 0 ? imp->list_rcv : imp->list_snd;
    std::vector<int>& ptr = pass == 0 ? imp->rcv_ptrs : imp->snd_ptrs;
    std::vector<int>& idx = pass == 0 ? imp->h_unpack_idx : imp->h_pack_idx;
    ptr.assign(1, 0);
    for (int c = 0; c < ncol; ++c) {
      if (h_owner[c] != want) continue;
      if (lst.empty() || lst.back() != h_part[c]) { lst.push_back(h_part[c]); ptr.push_back(ptr.back()); }
      idx.push_back(h_dof[c]);
      ptr.back()++;
    }
  }
  imp->num_rcv = (int)imp->list_rcv.size(); imp->num_snd = (int)imp->list_snd.size();
  imp->n_unpack = (i64)imp->h_unpack_idx.size(); imp->n_pack = (i64)imp->h_pack_idx.size();
  // 3. interior-then-interface permutation
  const int nseq = (int)(imp->n_pack + imp->n_unpack);
  std::vector<int> h_seq(imp->h_pack_idx);
  h_seq.insert(h_seq.end(), imp->h_unpack_idx.begin(), imp->h_unpack_idx.end());
  int *seq, *first, *flag, *rank, *perm;
  FB2_TRY(fb2_dalloc(&seq, (size_t)nseq + 1)); FB2_TRY(fb2_dalloc(&first, (size_t)n)); FB2_TRY(fb2_dalloc(&perm, (size_t)n));
  const i64 nmax = n > nseq ? n : nseq;
  FB2_TRY(fb2_dalloc(&flag, (size_t)nmax + 1)); FB2_TRY(fb2_dalloc(&rank, (size_t)nmax + 1));
  if (nseq) FB2_CUDA(fb2_memcpy(seq, h_seq.data(), sizeof(int) * nseq, cudaMemcpyHostToDevice));
  k_fill_int<<<GRID(n)>>>(n, first, INT_MAX);
  if (nseq) k_first_pos<<<GRID(nseq)>>>(nseq, seq, first);
  k_flag_first<<<GRID(nseq + 1)>>>(nseq, seq, first, flag);
  FB2_TRY(scan_excl(flag, rank, (i64)nseq + 1));
  int n_itf = 0;
  FB2_CUDA(fb2_memcpy(&n_itf, rank + nseq, sizeof(int), cudaMemcpyDeviceToHost));
  s->n_interface = n_itf; s->n_interior = n - n_itf;
  if (nseq) k_perm_interface<<<GRID(nseq)>>>(nseq, seq, first, rank, (int)s->n_interior, perm);
  k_flag_interior<<<GRID(n + 1)>>>(n, first, flag);
  FB2_TRY(scan_excl(flag, rank, n + 1));
  k_perm_interior<<<GRID(n)>>>(n, first, rank, perm);
  // 4. apply: lst_dofs_gids, pack/unpack lists (kept 0-based on the device), weights
  const i64 total = (i64)m.n_cells * s->rt.ndofs;
  k_apply_perm_gids<<<GRID(total)>>>(total, n, perm, s->d_gids);
  FB2_TRY(fb2_dalloc(&imp->d_weight, (size_t)n));
  k_fill_ones<<<GRID(n)>>>(n, imp->d_weight);
  FB2_TRY(fb2_dalloc(&imp->d_pack_idx, (size_t)imp->n_pack)); FB2_TRY(fb2_dalloc(&imp->d_unpack_idx, (size_t)imp->n_unpack));
  if (imp->n_pack) k_apply_perm_idx<<<GRID(imp->n_pack)>>>((int)imp->n_pack, perm, seq, imp->d_pack_idx, imp->d_weight);
  if (imp->n_unpack) k_apply_perm_idx<<<GRID(imp->n_unpack)>>>((int)imp->n_unpack, perm, seq + imp->n_pack, imp->d_unpack_idx, nullptr);
  FB2_CUDA(cudaGetLastError());
  if (imp->n_pack) FB2_CUDA(fb2_memcpy(imp->h_pack_idx.data(), imp->d_pack_idx, sizeof(int) * imp->n_pack, cudaMemcpyDeviceToHost));
  if (imp->n_unpack) FB2_CUDA(fb2_memcpy(imp->h_unpack_idx.data(), imp->d_unpack_idx, sizeof(int) * imp->n_unpack, cudaMemcpyDeviceToHost));
  FB2_TRY(fb2_import_build_add_plan(imp));   // h_unpack_idx is still 0-based here
  for (auto& v : imp->h_pack_idx) v += 1;
  for (auto& v : imp->h_unpack_idx) v += 1;
  const i64 nbuf = imp->n_pack > imp->n_unpack ? imp->n_pack : imp->n_unpack;
  FB2_TRY(fb2_dalloc(&imp->d_snd_buf, (size_t)nbuf)); FB2_TRY(fb2_dalloc(&imp->d_rcv_buf, (size_t)nbuf));
  FB2_CUDA(cudaStreamSynchronize(st));
  cudaFree(seq); cudaFree(first); cudaFree(flag); cudaFree(rank); cudaFree(perm);
  return 0;
}

// ---- export for parity tests / the Fortran dof_import_t (Fortran conventions: ptrs and ids 1-based) ----
extern "C" int fb2_dof_import_get_sizes(fb2_space_t s, int32_t sizes[4]) {
  FB2_CHECK(s && s->imp, "the space has no par layer (part_id = 0)");
  sizes[0] = s->imp->num_rcv; sizes[1] = s->imp->num_snd; sizes[2] = (int)s->imp->n_unpack; sizes[3] = (int)s->imp->n_pack;
  return 0;
}
extern "C" int fb2_dof_import_get(fb2_space_t s, int32_t* list_rcv, int32_t* rcv_ptrs, int32_t* unpack_idx, int32_t* list_snd,
                                  int32_t* snd_ptrs, int32_t* pack_idx, double* weight) {
  FB2_CHECK(s && s->imp, "the space has no par layer (part_id = 0)");
  const fb2_import_s* imp = s->imp;
  for (int i = 0; i < imp->num_rcv; ++i) list_rcv[i] = imp->list_rcv[i];
  for (int i = 0; i <= imp->num_rcv; ++i) rcv_ptrs[i] = imp->rcv_ptrs[i] + 1;
  for (i64 i = 0; i < imp->n_unpack; ++i) unpack_idx[i] = imp->h_unpack_idx[i];
  for (int i = 0; i < imp->num_snd; ++i) list_snd[i] = imp->list_snd[i];
  for (int i = 0; i <= imp->num_snd; ++i) snd_ptrs[i] = imp->snd_ptrs[i] + 1;
  for (i64 i = 0; i < imp->n_pack; ++i) pack_idx[i] = imp->h_pack_idx[i];
  if (weight) FB2_CUDA(fb2_memcpy(weight, imp->d_weight, sizeof(double) * (size_t)s->n_free, cudaMemcpyDeviceToHost));
  return 0;
}
