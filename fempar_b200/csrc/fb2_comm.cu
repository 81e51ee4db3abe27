// Distributed layer: execution_context_t on NCCL (mpi_context.f90:384-474,522-639) and the device
// dof_import_t (dof_import.f90:49-213) used by par_scalar_array%comm() (par_scalar_array.f90:498-533).
#include <nccl.h>
#include <math.h>
#include <string.h>
#include <unistd.h>
#include <algorithm>
#include <map>
#include "fb2_internal.h"

static ncclComm_t g_comm = nullptr;
static int g_rank = 0, g_nranks = 1;

#define FB2_NCCL(call)                                                                            \
  do {                                                                                            \
    ncclResult_t r_ = (call);                                                                     \
    if (r_ != ncclSuccess) {                                                                      \
      fb2_set_error("%s:%d: %s failed: %s", __FILE__, __LINE__, #call, ncclGetErrorString(r_));  \
      return 3;                                                                                   \
    }                                                                                             \
  } while (0)

// Several sub-domains (parts) per process: part p (1-based, FEMPAR numbering) lives on rank (p-1)/parts_per_rank, e.g. the
// 8-part layout of BASELINE config 3 on 2 or 4 GPUs.  With more than one part per rank the single-part entry points do no
// communication; the *_multi entry points exchange all co-resident parts collectively.
static int g_parts_per_rank = 1;
static inline int rank_of_part(int part1) { return (part1 - 1) / g_parts_per_rank; }
int fb2_comm_parts_per_rank() { return g_parts_per_rank; }
extern "C" int fb2_comm_set_parts_per_rank(int k) {
  FB2_CHECK(k >= 1, "parts per rank must be >= 1");
  g_parts_per_rank = k;
  return 0;
}
int fb2_comm_active() { return g_comm != nullptr; }
int fb2_comm_nranks() { return g_nranks; }
int fb2_comm_rank() { return g_rank; }

extern "C" int fb2_comm_unique_id(void* id128) {
  ncclUniqueId id;
  FB2_NCCL(ncclGetUniqueId(&id));
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
  memcpy(id128, &id, 128);
  return 0;
}
// ------------------------------------------------------------------------------------------
// Scalar all-reduce over NVLink peer memory.  The two all-reduces of a PCG iteration carry 1 and 4 doubles and sit on the
// critical path (2 x ~25 us through NCCL at 8 GPUs against a 2.4 ms iteration).  Every rank owns a mailbox in its HBM, opened by all
// the others through CUDA IPC; one kernel stores the rank's contribution straight into every peer's mailbox (st.global over
// NVLink / NVSwitch, then a system-scope fence and an epoch flag), waits for the flags of the peers in its own mailbox and sums
// the contributions in rank order -- the same order on every rank, so all ranks hold bit-identical results, as after
// mpi_allreduce.  Mailboxes are double-buffered by the parity of the epoch: a rank can enter all-reduce e + 2 only after it
// received everyone's e + 1, which a peer sends after it has read e.  Used only when every rank drives its own GPU of one
// node and all peers are IPC-reachable (otherwise, and for more than 7 values, NCCL); FB2_PEER_ALLREDUCE=0 switches it off.
// ------------------------------------------------------------------------------------------
#define PEER_SLOT 8            // doubles per mailbox slot: 7 values + the epoch flag
struct PeerAllreduce {
  bool ok = false;
  double* local = nullptr;       // [2][nranks][PEER_SLOT]
  double** d_peers = nullptr;    // device array: mailbox base of every rank (own entry = local)
  std::vector<void*> opened;
  unsigned long long epoch = 0;
};
static PeerAllreduce g_peer;

// one rank's part of the exchange: executed by the first 32 threads of a block
__device__ __forceinline__ void peer_allreduce_body(double* buf, int count, double* const* peers, double* local, int me, int nranks,
                                                    unsigned long long epoch) {
  const int r = threadIdx.x;
  const size_t par = (size_t)(epoch & 1ULL) * nranks * PEER_SLOT;
  if (r < nranks) {
    double* slot = peers[r] + par + (size_t)me * PEER_SLOT;       // my slot in the mailbox of rank r
    for (int k = 0; k < count; ++k) slot[k] = buf[k];
    __threadfence_system();
    *((volatile unsigned long long*)(slot + PEER_SLOT - 1)) = epoch;
    const volatile unsigned long long* flag = (const volatile unsigned long long*)(local + par + (size_t)r * PEER_SLOT + PEER_SLOT - 1);
    while (*flag != epoch) { }
    __threadfence_system();
  }
  __syncthreads();
  if (r < count) {
    double acc = 0.0;
    for (int q = 0; q < nranks; ++q) acc += ((const volatile double*)(local + par + (size_t)q * PEER_SLOT))[r];
    buf[r] = acc;
  }
  __syncthreads();
}
__global__ void k_peer_allreduce(double* buf, int count, double* const* peers, double* local, int me, int nranks,
                                 unsigned long long epoch) {
  peer_allreduce_body(buf, count, peers, local, me, nranks, epoch);
}
// The same protocol with every rank played by one block of ONE cooperative launch (all blocks co-resident, so the waits cannot
// deadlock): lets a single-GPU box exercise the mailbox / epoch-parity logic over many consecutive all-reduces.
__global__ void k_peer_allreduce_emulated(double* bufs /*[nranks][PEER_SLOT]*/, int count, double* const* peers, int nranks,
                                          int epochs, double* sums_out /*[epochs][nranks][count]*/) {
  const int me = blockIdx.x;
  double* buf = bufs + (size_t)me * PEER_SLOT;
  for (int e = 1; e <= epochs; ++e) {
    if (threadIdx.x < count) buf[threadIdx.x] = (double)(me + 1) * 0.125 + (double)e + 1000.0 * threadIdx.x;   // this rank's contribution
    __syncthreads();
    peer_allreduce_body(buf, count, peers, peers[me], me, nranks, (unsigned long long)e);
    if (threadIdx.x < count) sums_out[((size_t)(e - 1) * nranks + me) * count + threadIdx.x] = buf[threadIdx.x];
    __syncthreads();
  }
}
extern "C" int fb2_peer_allreduce_selftest(int nranks, int count, int epochs, double* max_abs_error) {
  FB2_CHECK(nranks >= 2 && nranks <= 16 && count >= 1 && count <= PEER_SLOT - 1 && epochs >= 1 && max_abs_error, "bad argument");
  int dev = 0, coop = 0;
  FB2_CUDA(cudaGetDevice(&dev));
  FB2_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
  FB2_CHECK(coop, "cooperative launch not supported");
  cudaStream_t st = fb2_stream();
  const size_t box = (size_t)2 * nranks * PEER_SLOT;
  double *d_boxes, *d_bufs, *d_out;
  double** d_peers;
  FB2_TRY(fb2_dalloc(&d_boxes, box * nranks)); FB2_TRY(fb2_dalloc(&d_bufs, (size_t)nranks * PEER_SLOT));
  FB2_TRY(fb2_dalloc(&d_out, (size_t)epochs * nranks * count)); FB2_TRY(fb2_dalloc(&d_peers, (size_t)nranks));
  FB2_CUDA(cudaMemsetAsync(d_boxes, 0, box * nranks * sizeof(double), st));
  std::vector<double*> peers((size_t)nranks);
  for (int r = 0; r < nranks; ++r) peers[r] = d_boxes + box * r;
  FB2_CUDA(fb2_memcpy(d_peers, peers.data(), sizeof(double*) * (size_t)nranks, cudaMemcpyHostToDevice));
  double* const* peers_arg = d_peers;
  void* args[] = {(void*)&d_bufs, (void*)&count, (void*)&peers_arg, (void*)&nranks, (void*)&epochs, (void*)&d_out};
  FB2_CUDA(cudaLaunchCooperativeKernel((const void*)k_peer_allreduce_emulated, dim3(nranks), dim3(32), args, 0, st));
  std::vector<double> out((size_t)epochs * nranks * count);
  FB2_CUDA(fb2_memcpy(out.data(), d_out, sizeof(double) * out.size(), cudaMemcpyDeviceToHost));
  double err = 0.0;
  for (int e = 1; e <= epochs; ++e)
    for (int k = 0; k < count; ++k) {
      double want = 0.0;
      for (int q = 0; q < nranks; ++q) want += (double)(q + 1) * 0.125 + (double)e + 1000.0 * k;     // rank order, like the kernel
      for (int me = 0; me < nranks; ++me) err = std::max(err, fabs(out[((size_t)(e - 1) * nranks + me) * count + k] - want));
    }
  *max_abs_error = err;
  cudaFree(d_boxes); cudaFree(d_bufs); cudaFree(d_out); cudaFree(d_peers);
  return 0;
}

static void peer_allreduce_teardown() {
  for (void* p : g_peer.opened) cudaIpcCloseMemHandle(p);
  g_peer.opened.clear();
  cudaFree(g_peer.local); cudaFree(g_peer.d_peers);
  g_peer = PeerAllreduce();
}

// collective: every rank calls it right after the communicator exists
static int peer_allreduce_setup() {
  const int n = g_nranks;
  if (n < 2 || n > 16 || (getenv("FB2_PEER_ALLREDUCE") && atoi(getenv("FB2_PEER_ALLREDUCE")) == 0)) return 0;
  cudaStream_t st = fb2_stream();
  // 1. one rank per GPU?  all-gather the PCI bus ids
  int dev = 0;
  FB2_CUDA(cudaGetDevice(&dev));
  char bus[64];
  memset(bus, 0, sizeof(bus));
  FB2_CUDA(cudaDeviceGetPCIBusId(bus, 32, dev));
  char hostn[32];
  memset(hostn, 0, sizeof(hostn));
  gethostname(hostn, 31);
  memcpy(bus + 32, hostn, 32);
  char *d_all, *d_mine;
  FB2_TRY(fb2_dalloc(&d_all, (size_t)64 * n)); FB2_TRY(fb2_dalloc(&d_mine, 64));
  FB2_CUDA(cudaMemcpyAsync(d_mine, bus, 64, cudaMemcpyHostToDevice, st));
  FB2_NCCL(ncclAllGather(d_mine, d_all, 64, ncclChar, g_comm, st));
  std::vector<char> all((size_t)64 * n);
  FB2_CUDA(fb2_memcpy(all.data(), d_all, all.size(), cudaMemcpyDeviceToHost));
  int good = 1;
  for (int a = 0; a < n && good; ++a) {
    if (memcmp(&all[(size_t)64 * a + 32], hostn, 32) != 0) good = 0;          // another node
    for (int b = a + 1; b < n && good; ++b)
      if (memcmp(&all[(size_t)64 * a], &all[(size_t)64 * b], 32) == 0) good = 0;   // two ranks on one GPU: a spin wait could deadlock
  }
  // 2. mailbox + IPC handle exchange
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  const size_t box = (size_t)2 * n * PEER_SLOT;
  if (good) {
    if (cudaMalloc((void**)&g_peer.local, box * sizeof(double)) != cudaSuccess) { cudaGetLastError(); good = 0; }
    else {
      cudaMemsetAsync(g_peer.local, 0, box * sizeof(double), st);
      if (cudaIpcGetMemHandle(&mine, g_peer.local) != cudaSuccess) { cudaGetLastError(); good = 0; }
    }
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  FB2_CUDA(cudaMemcpyAsync(d_mine, &mine, 64, cudaMemcpyHostToDevice, st));
  FB2_NCCL(ncclAllGather(d_mine, d_all, 64, ncclChar, g_comm, st));
  FB2_CUDA(fb2_memcpy(all.data(), d_all, all.size(), cudaMemcpyDeviceToHost));
  std::vector<double*> peers((size_t)n, nullptr);
  if (good) {
    for (int r = 0; r < n && good; ++r) {
      if (r == g_rank) { peers[r] = g_peer.local; continue; }
      cudaIpcMemHandle_t h;
      memcpy(&h, &all[(size_t)64 * r], 64);
      void* p = nullptr;
      if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); good = 0; break; }
      g_peer.opened.push_back(p);
      peers[r] = (double*)p;
    }
  }
  // 3. everybody or nobody
  double* d_flag;
  FB2_TRY(fb2_dalloc(&d_flag, 1));
  double h_flag = good ? 1.0 : 0.0;
  FB2_CUDA(cudaMemcpyAsync(d_flag, &h_flag, sizeof(double), cudaMemcpyHostToDevice, st));
  FB2_NCCL(ncclAllReduce(d_flag, d_flag, 1, ncclDouble, ncclMin, g_comm, st));
  FB2_CUDA(fb2_memcpy(&h_flag, d_flag, sizeof(double), cudaMemcpyDeviceToHost));
  cudaFree(d_flag); cudaFree(d_all); cudaFree(d_mine);
  if (h_flag < 0.5) { peer_allreduce_teardown(); fb2_record_phase("peer_allreduce", 0.0); return 0; }
  FB2_TRY(fb2_dalloc(&g_peer.d_peers, (size_t)n));
  FB2_CUDA(fb2_memcpy(g_peer.d_peers, peers.data(), sizeof(double*) * (size_t)n, cudaMemcpyHostToDevice));
  g_peer.epoch = 0;
  g_peer.ok = true;
  fb2_record_phase("peer_allreduce", 1.0);
  return 0;
}

extern "C" int fb2_comm_init(int rank, int nranks, const void* id128) {
  FB2_CHECK(fb2_stream() != nullptr, "fb2_init must be called before fb2_comm_init");
  ncclUniqueId id;
  memcpy(&id, id128, 128);
  FB2_NCCL(ncclCommInitRank(&g_comm, nranks, id, rank));
  g_rank = rank; g_nranks = nranks;
  FB2_TRY(peer_allreduce_setup());
  return 0;
}
extern "C" int fb2_comm_finalize(void) {
  if (g_comm) {
    cudaStreamSynchronize(fb2_stream());
    peer_allreduce_teardown();
    ncclCommDestroy(g_comm); g_comm = nullptr;
  }
  g_rank = 0; g_nranks = 1;
  return 0;
}

// environment%l1_sum_scalar_rp -> mpi_allreduce(MPI_IN_PLACE, ., mpi_sum)  (mpi_context.f90:384)
int fb2_allreduce_sum(double* d_buf, int count) {
  if (!g_comm) return 0;
  if (g_peer.ok && count >= 1 && count <= PEER_SLOT - 1) {
    ++g_peer.epoch;
    k_peer_allreduce<<<1, 32, 0, fb2_stream()>>>(d_buf, count, g_peer.d_peers, g_peer.local, g_rank, g_nranks, g_peer.epoch);
    FB2_CUDA(cudaGetLastError());
    return 0;
  }
  FB2_NCCL(ncclAllReduce(d_buf, d_buf, count, ncclDouble, ncclSum, g_comm, fb2_stream()));
  return 0;
}

// environment%l1_max (mpi_context.f90 max_scalar_rp analogue)
int fb2_allreduce_max(double* d_buf, int count) {
  if (!g_comm) return 0;
  FB2_NCCL(ncclAllReduce(d_buf, d_buf, count, ncclDouble, ncclMax, g_comm, fb2_stream()));
  return 0;
}

// ------------------------------------------------------------------------------------------
// dof_import on the device
// ------------------------------------------------------------------------------------------
const double* fb2_import_weight(fb2_import_s* imp) { return imp->d_weight; }
void fb2_import_destroy(fb2_import_s* imp) {
  if (!imp) return;
  cudaFree(imp->d_pack_idx); cudaFree(imp->d_unpack_idx); cudaFree(imp->d_snd_buf); cudaFree(imp->d_rcv_buf);
  cudaFree(imp->d_weight); cudaFree(imp->d_add_dst); cudaFree(imp->d_add_ptr); cudaFree(imp->d_add_pos);
  delete imp;
}

// Gather-add plan of stage 1 of comm(): a DOF shared by more than two parts is received from several neighbours, so
// unpack_idx holds it several times.  Group the receive-buffer positions by destination DOF (positions ascending =
// neighbour list order = the order mpi_context.f90:610-632 adds them in) -> one thread per DOF, no atomics, deterministic.
// imp->h_unpack_idx must hold 0-based ids.
int fb2_import_build_add_plan(fb2_import_s* imp) {
  const int nu = (int)imp->n_unpack;
  std::vector<int> pos(nu);
  for (int i = 0; i < nu; ++i) pos[i] = i;
  std::stable_sort(pos.begin(), pos.end(), [&](int a, int b) { return imp->h_unpack_idx[a] < imp->h_unpack_idx[b]; });
  std::vector<int> dst, ptr;
  for (int i = 0; i < nu; ++i)
    if (i == 0 || imp->h_unpack_idx[pos[i]] != imp->h_unpack_idx[pos[i - 1]]) { dst.push_back(imp->h_unpack_idx[pos[i]]); ptr.push_back(i); }
  ptr.push_back(nu);
  imp->n_add_dst = (int)dst.size();
  cudaFree(imp->d_add_dst); cudaFree(imp->d_add_ptr); cudaFree(imp->d_add_pos);
  FB2_TRY(fb2_dalloc(&imp->d_add_dst, dst.size())); FB2_TRY(fb2_dalloc(&imp->d_add_ptr, ptr.size())); FB2_TRY(fb2_dalloc(&imp->d_add_pos, (size_t)nu));
  if (nu) {
    FB2_CUDA(fb2_memcpy(imp->d_add_dst, dst.data(), sizeof(int) * dst.size(), cudaMemcpyHostToDevice));
    FB2_CUDA(fb2_memcpy(imp->d_add_pos, pos.data(), sizeof(int) * (size_t)nu, cudaMemcpyHostToDevice));
  }
  FB2_CUDA(fb2_memcpy(imp->d_add_ptr, ptr.data(), sizeof(int) * ptr.size(), cudaMemcpyHostToDevice));
  return 0;
}

// pack_rp / unpack_rp of mpi_context.f90:522-639
__global__ void k_pack(const double* __restrict__ x, const int* __restrict__ idx, int lo, int hi, double* buf) {
  int i = lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (i < hi) buf[i] = x[idx[i]];
}
__global__ void k_unpack(double* __restrict__ x, const int* __restrict__ idx, int lo, int hi, const double* buf) {
  int i = lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (i < hi) x[idx[i]] = buf[i];
}
// x[dst] += buf[pos_0] + buf[pos_1] + ... added one after the other in neighbour-list order (one thread per destination DOF)
__global__ void k_unpack_add(double* __restrict__ x, int ndst, const int* __restrict__ dst, const int* __restrict__ ptr,
                             const int* __restrict__ pos, const double* __restrict__ buf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ndst) return;
  double v = x[dst[i]];
  for (int k = ptr[i]; k < ptr[i + 1]; ++k) v += buf[pos[k]];
  x[dst[i]] = v;
}

// one neighbours exchange.  forward: pack_idx -> snd neighbours, unpack into unpack_idx (add or overwrite);
// reverse: roles of the two lists swapped (stage 2 of comm()).
static int exchange_stage(fb2_import_s* imp, double* x, bool reverse, bool add, cudaStream_t s) {
  const std::vector<int>& sl = reverse ? imp->list_rcv : imp->list_snd;
  const std::vector<int>& sp = reverse ? imp->rcv_ptrs : imp->snd_ptrs;
  const std::vector<int>& rl = reverse ? imp->list_snd : imp->list_rcv;
  const std::vector<int>& rp = reverse ? imp->snd_ptrs : imp->rcv_ptrs;
  const int* sidx = reverse ? imp->d_unpack_idx : imp->d_pack_idx;
  const int* ridx = reverse ? imp->d_pack_idx : imp->d_unpack_idx;
  int ns = sp.empty() ? 0 : sp.back(), nr = rp.empty() ? 0 : rp.back();
  if (ns > 0) k_pack<<<(ns + 255) / 256, 256, 0, s>>>(x, sidx, 0, ns, imp->d_snd_buf);
  FB2_NCCL(ncclGroupStart());
  for (size_t i = 0; i < rl.size(); ++i)
    if (rp[i + 1] > rp[i]) FB2_NCCL(ncclRecv(imp->d_rcv_buf + rp[i], rp[i + 1] - rp[i], ncclDouble, rank_of_part(rl[i]), g_comm, s));
  for (size_t i = 0; i < sl.size(); ++i)
    if (sp[i + 1] > sp[i]) FB2_NCCL(ncclSend(imp->d_snd_buf + sp[i], sp[i + 1] - sp[i], ncclDouble, rank_of_part(sl[i]), g_comm, s));
  FB2_NCCL(ncclGroupEnd());
  if (add) {
    // stage 1 always unpacks into unpack_idx; additions in neighbour-list order (the reference's order)
    if (imp->n_add_dst > 0)
      k_unpack_add<<<(imp->n_add_dst + 255) / 256, 256, 0, s>>>(x, imp->n_add_dst, imp->d_add_dst, imp->d_add_ptr, imp->d_add_pos, imp->d_rcv_buf);
  } else if (nr > 0) k_unpack<<<(nr + 255) / 256, 256, 0, s>>>(x, ridx, 0, nr, imp->d_rcv_buf);   // each non-owned DOF has one owner
  FB2_CUDA(cudaGetLastError());
  return 0;
}

// par_scalar_array_comm (par_scalar_array.f90:498-533):
//  stage 1: non-owners send their partial sums, the owner receives and adds   (rcv side = owner)
//  stage 2: the owner sends the total, non-owners receive and overwrite
int fb2_import_exchange_on(fb2_import_s* imp, double* d_vec, cudaStream_t stream) {
  if (!g_comm || !imp) return 0;
  // several sub-domains per process: a single part cannot exchange on its own (its neighbours may live in this process);
  // the *_multi entry points run the exchange collectively.  Returning 0 here would leave interface values unassembled.
  FB2_CHECK(g_parts_per_rank == 1, "single-part par entry point called with several sub-domains per rank: use the *_multi entry points");
  FB2_TRY(exchange_stage(imp, d_vec, false, true, stream));
  FB2_TRY(exchange_stage(imp, d_vec, true, false, stream));
  return 0;
}
int fb2_import_exchange(fb2_import_s* imp, double* d_vec) { return fb2_import_exchange_on(imp, d_vec, fb2_stream()); }


// ------------------------------------------------------------------------------------------
// comm() of the co-resident parts of this rank, collectively: messages between two parts of the same process are
// device-to-device copies, the others one grouped ncclSend/ncclRecv.  NCCL matches the messages of a rank pair in issue
// order, so both sides order them by (source part, destination part).
// ------------------------------------------------------------------------------------------
struct Msg { int peer, src, dst; double* ptr; int len; };
static int exchange_stage_multi(int np, fb2_import_s* const* imps, const int* part_ids, double* const* xs, bool reverse,
                                cudaStream_t s) {
  std::vector<Msg> sends, recvs;
  for (int i = 0; i < np; ++i) {
    fb2_import_s* imp = imps[i];
    const std::vector<int>& sp = reverse ? imp->rcv_ptrs : imp->snd_ptrs;
    const int* sidx = reverse ? imp->d_unpack_idx : imp->d_pack_idx;
    const int ns = sp.empty() ? 0 : sp.back();
    if (ns > 0) k_pack<<<(ns + 255) / 256, 256, 0, s>>>(xs[i], sidx, 0, ns, imp->d_snd_buf);
  }
  for (int i = 0; i < np; ++i) {
    fb2_import_s* imp = imps[i];
    const std::vector<int>& sl = reverse ? imp->list_rcv : imp->list_snd;
    const std::vector<int>& sp = reverse ? imp->rcv_ptrs : imp->snd_ptrs;
    const std::vector<int>& rl = reverse ? imp->list_snd : imp->list_rcv;
    const std::vector<int>& rp = reverse ? imp->snd_ptrs : imp->rcv_ptrs;
    for (size_t a = 0; a < sl.size(); ++a) {
      const int len = sp[a + 1] - sp[a];
      if (len <= 0) continue;
      const int nb = sl[a];
      if (rank_of_part(nb) == g_rank) {
        int j = -1;
        for (int k = 0; k < np; ++k) if (part_ids[k] == nb) j = k;
        FB2_CHECK(j >= 0, "a neighbour part of this rank is missing from the multi-part call");
        fb2_import_s* dst = imps[j];
        const std::vector<int>& drl = reverse ? dst->list_snd : dst->list_rcv;
        const std::vector<int>& drp = reverse ? dst->snd_ptrs : dst->rcv_ptrs;
        int b = -1;
        for (size_t k = 0; k < drl.size(); ++k) if (drl[k] == part_ids[i]) b = (int)k;
        FB2_CHECK(b >= 0 && drp[b + 1] - drp[b] == len, "inconsistent dof_import lists between two co-resident parts");
        FB2_CUDA(cudaMemcpyAsync(dst->d_rcv_buf + drp[b], imp->d_snd_buf + sp[a], sizeof(double) * (size_t)len,
                                 cudaMemcpyDeviceToDevice, s));
      } else sends.push_back({rank_of_part(nb), part_ids[i], nb, imp->d_snd_buf + sp[a], len});
    }
    for (size_t a = 0; a < rl.size(); ++a) {
      const int len = rp[a + 1] - rp[a];
      if (len > 0 && rank_of_part(rl[a]) != g_rank) recvs.push_back({rank_of_part(rl[a]), rl[a], part_ids[i], imp->d_rcv_buf + rp[a], len});
    }
  }
  auto by_pair = [](const Msg& x, const Msg& y) { return x.src != y.src ? x.src < y.src : x.dst < y.dst; };
  std::sort(sends.begin(), sends.end(), by_pair);
  std::sort(recvs.begin(), recvs.end(), by_pair);
  if (!sends.empty() || !recvs.empty()) {
    FB2_CHECK(g_comm != nullptr, "a neighbour part lives on another rank but fb2_comm_init was not called");
    FB2_NCCL(ncclGroupStart());
    for (const Msg& m : recvs) FB2_NCCL(ncclRecv(m.ptr, m.len, ncclDouble, m.peer, g_comm, s));
    for (const Msg& m : sends) FB2_NCCL(ncclSend(m.ptr, m.len, ncclDouble, m.peer, g_comm, s));
    FB2_NCCL(ncclGroupEnd());
  }
  for (int i = 0; i < np; ++i) {
    fb2_import_s* imp = imps[i];
    if (!reverse) {
      if (imp->n_add_dst > 0)
        k_unpack_add<<<(imp->n_add_dst + 255) / 256, 256, 0, s>>>(xs[i], imp->n_add_dst, imp->d_add_dst, imp->d_add_ptr, imp->d_add_pos, imp->d_rcv_buf);
    } else {
      const int nr = imp->snd_ptrs.empty() ? 0 : imp->snd_ptrs.back();
      if (nr > 0) k_unpack<<<(nr + 255) / 256, 256, 0, s>>>(xs[i], imp->d_pack_idx, 0, nr, imp->d_rcv_buf);
    }
  }
  FB2_CUDA(cudaGetLastError());
  return 0;
}
int fb2_import_exchange_multi(int np, fb2_import_s* const* imps, const int* part_ids, double* const* xs, cudaStream_t s) {
  // without a communicator every part must be co-resident (one process holding the whole partition)
  FB2_TRY(exchange_stage_multi(np, imps, part_ids, xs, false, s));
  FB2_TRY(exchange_stage_multi(np, imps, part_ids, xs, true, s));
  return 0;
}
