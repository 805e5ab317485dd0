// mg.cu -- multi-GPU row-sharded CSR SpMM behind the C ABI (the reference is single-GPU).
// One process per GPU.  Rank g owns rows [bounds[g], bounds[g+1]) of A and of C, and the same rows of B (the C
// rows it produced as the previous layer's output).  B is replicated by an all-gather-v over NCCL (NVLink 5 /
// NVSwitch): one ncclGroup of send / recv pairs, every rank sending its rows to every peer at once, instead of
// one broadcast per owner queued behind each other.  spmmdb_sharded_csrmm_f32 cuts K into feature slabs and
// gathers slab s + 1 on a communication stream while the SpMM of slab s runs on the caller's stream: columns of
// C need only the same columns of B.  The replica is kept slab-major ([slab][k][K / slabs], library scratch), so
// every rank's rows of a slab are one contiguous message and the SpMM reads it with ldb = K / slabs.
//
// NCCL is resolved at run time (dlopen of libnccl.so.2, preferring the copy already loaded in the process,
// e.g. torch's): libspmmdb.so has no link-time dependency on it and single-GPU users never load it.
#include <dlfcn.h>
#include <nccl.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "common.cuh"

namespace spmmdb {
int csrmm_rowmajor(int m, int K, int64_t nnz, const int* rowptr, const int* colind, const float* val,
                   const float* B, int64_t ldb, float* C, int64_t ldc, float alpha, float beta, cudaStream_t st,
                   int b_bf16, const unsigned* hot);
namespace {

struct Nccl {
  void* h = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

Nccl& nccl() {
  static Nccl n;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      n.h = dlopen(nm, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);  // the copy the process already uses, if any
      if (n.h != nullptr) break;
    }
    if (n.h == nullptr) {
      const char* env = getenv("SPMMDB_NCCL_LIB");
      if (env != nullptr) n.h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
      for (const char* nm : names) {
        if (n.h != nullptr) break;
        n.h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      }
    }
    if (n.h == nullptr) return;
#define SPMMDB_SYM(field, name) n.field = reinterpret_cast<decltype(n.field)>(dlsym(n.h, name))
    SPMMDB_SYM(GetUniqueId, "ncclGetUniqueId");
    SPMMDB_SYM(CommInitRank, "ncclCommInitRank");
    SPMMDB_SYM(CommDestroy, "ncclCommDestroy");
    SPMMDB_SYM(GroupStart, "ncclGroupStart");
    SPMMDB_SYM(GroupEnd, "ncclGroupEnd");
    SPMMDB_SYM(Send, "ncclSend");
    SPMMDB_SYM(Recv, "ncclRecv");
    SPMMDB_SYM(AllGather, "ncclAllGather");
    SPMMDB_SYM(AllReduce, "ncclAllReduce");
    SPMMDB_SYM(GetErrorString, "ncclGetErrorString");
#undef SPMMDB_SYM
    n.ok = n.GetUniqueId && n.CommInitRank && n.CommDestroy && n.GroupStart && n.GroupEnd && n.Send && n.Recv &&
           n.AllGather && n.AllReduce;
  });
  return n;
}

int nccl_fail(const char* what, ncclResult_t r) {
  char msg[256];
  snprintf(msg, sizeof(msg), "%s: %s", what, nccl().GetErrorString ? nccl().GetErrorString(r) : "NCCL error");
  set_last_error(msg, cudaErrorUnknown, __FILE__, __LINE__);
  return SPMMDB_ERR_CUDA;
}
#define SPMMDB_NCCL_OK(expr)                                   \
  do {                                                         \
    ncclResult_t _r = (expr);                                  \
    if (_r != ncclSuccess) return nccl_fail(#expr, _r);        \
  } while (0)

// out[r][0..Ks) = in[r][c0..c0+Ks): one slab of this rank's rows of B into its place in the slab-major replica
__global__ void __launch_bounds__(256) pack_slab_kernel(const float* __restrict__ in, int64_t ldi, int c0, int Ks,
                                                        float* __restrict__ out, int64_t rows) {
  const int kv = Ks >> 2;
  const int64_t total = rows * kv;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t r = i / kv;
    const int j = (int)(i - r * kv);
    reinterpret_cast<float4*>(out + r * Ks)[j] = __ldcs(reinterpret_cast<const float4*>(in + r * ldi + c0) + j);
  }
}

// The same pack fused with the all-gather: every float4 read from this rank's rows of B is stored into the slab of
// EVERY rank's replica (own memory + the peers' replicas mapped through CUDA IPC): plain stores over NVLink, all peers
// written concurrently by all SMs that take part, no staging buffer and no second pass over the data.
struct PushDst { float* p[16]; };
__global__ void __launch_bounds__(256) pack_push_kernel(const float* __restrict__ in, int64_t ldi, int c0, int Ks, PushDst dst,
                                                        int ndst, int64_t rows) {
  const int kv = Ks >> 2;
  const int64_t total = rows * kv;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t r = i / kv;
    const int j = (int)(i - r * kv);
    const float4 x = __ldcs(reinterpret_cast<const float4*>(in + r * ldi + c0) + j);
#pragma unroll 8
    for (int d = 0; d < ndst; ++d) reinterpret_cast<float4*>(dst.p[d] + r * Ks)[j] = x;
  }
}

}  // namespace
}  // namespace spmmdb

using namespace spmmdb;

struct spmmdb_mg {
  ncclComm_t comm = nullptr;
  bool own_comm = false;
  int rank = 0, world = 1, device = 0;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev_ready = nullptr;             // caller's stream reached the call (B rows are final)
  cudaEvent_t ev_slab[16] = {};               // slab s gathered
  cudaEvent_t ev_done = nullptr;              // the call's last use of the replica (next call's gathers wait for it)
  float* replica = nullptr;                   // slab-major copy of B, grow-only
  size_t replica_bytes = 0;
  // Peer path: every rank's replica mapped into this process (CUDA IPC), so that a rank PUSHES its rows of a slab
  // straight into its peers' replicas with copy-engine transfers over NVLink (no SM, no staging).  Tiny NCCL
  // all-reduces order the pushes against the peers' SpMMs.
  bool peer_ok = false;
  float* peer_replica[16] = {};               // [rank] -> that rank's replica in this address space (own = replica)
  cudaStream_t push_stream[16] = {};
  cudaEvent_t push_done[16] = {};
  cudaEvent_t push_go = nullptr;
  int* flag_dev = nullptr;                    // 2 ints: operand / result of the barrier all-reduce
  char* ipc_dev = nullptr;                    // world x 64-byte IPC handles (all-gathered)
};

static int mg_init_common(spmmdb_mg* g) {
  SPMMDB_CUDA_OK(cudaGetDevice(&g->device));
  SPMMDB_CUDA_OK(cudaStreamCreateWithFlags(&g->comm_stream, cudaStreamNonBlocking));
  SPMMDB_CUDA_OK(cudaEventCreateWithFlags(&g->ev_ready, cudaEventDisableTiming));
  SPMMDB_CUDA_OK(cudaEventCreateWithFlags(&g->ev_done, cudaEventDisableTiming));
  for (auto& e : g->ev_slab) SPMMDB_CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  SPMMDB_CUDA_OK(cudaEventCreateWithFlags(&g->push_go, cudaEventDisableTiming));
  for (int p = 0; p < g->world && p < 16; ++p) {
    SPMMDB_CUDA_OK(cudaStreamCreateWithFlags(&g->push_stream[p], cudaStreamNonBlocking));
    SPMMDB_CUDA_OK(cudaEventCreateWithFlags(&g->push_done[p], cudaEventDisableTiming));
  }
  SPMMDB_CUDA_OK(cudaMalloc(&g->flag_dev, 64));
  SPMMDB_CUDA_OK(cudaMemset(g->flag_dev, 0, 64));
  SPMMDB_CUDA_OK(cudaMalloc(&g->ipc_dev, 64 * 16));
  return SPMMDB_OK;
}

static void mg_close_peers(spmmdb_mg* g) {
  for (int p = 0; p < g->world && p < 16; ++p) {
    if (p != g->rank && g->peer_replica[p] != nullptr) cudaIpcCloseMemHandle(g->peer_replica[p]);
    g->peer_replica[p] = nullptr;
  }
  g->peer_ok = false;
}

// a few bytes through NCCL on the communication stream: returns on a rank only after every rank's stream got here
static int mg_barrier(spmmdb_mg* g) {
  SPMMDB_NCCL_OK(nccl().AllReduce(g->flag_dev, g->flag_dev + 8, 1, ncclInt32, ncclSum, g->comm, g->comm_stream));
  return SPMMDB_OK;
}

// Collective, called by every rank right after it (re)allocated its replica: exchange the IPC handles of the replicas and
// map the peers' ones.  Any failure on any rank (containers without IPC, no peer access) leaves every rank on the NCCL
// send / recv path: the decision is an all-reduce.
static int mg_map_peers(spmmdb_mg* g) {
  mg_close_peers(g);
  if (g->world == 1 || g->world > 16 || SPMMDB_TUNE("MG_PEER", 2) == 0) return SPMMDB_OK;
  cudaIpcMemHandle_t mine;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handles are 64 bytes");
  int bad = cudaIpcGetMemHandle(&mine, g->replica) != cudaSuccess ? 1 : 0;
  cudaGetLastError();
  SPMMDB_CUDA_OK(cudaMemcpyAsync(g->ipc_dev + 64 * g->rank, &mine, 64, cudaMemcpyHostToDevice, g->comm_stream));
  SPMMDB_NCCL_OK(nccl().AllGather(g->ipc_dev + 64 * g->rank, g->ipc_dev, 64, ncclChar, g->comm, g->comm_stream));
  std::vector<cudaIpcMemHandle_t> all(g->world);
  SPMMDB_CUDA_OK(cudaMemcpyAsync(all.data(), g->ipc_dev, 64 * (size_t)g->world, cudaMemcpyDeviceToHost, g->comm_stream));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(g->comm_stream));
  for (int p = 0; p < g->world && !bad; ++p) {
    if (p == g->rank) { g->peer_replica[p] = g->replica; continue; }
    void* ptr = nullptr;
    if (cudaIpcOpenMemHandle(&ptr, all[p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); bad = 1; break; }
    g->peer_replica[p] = (float*)ptr;
  }
  int h_bad[2] = {bad, 0};
  SPMMDB_CUDA_OK(cudaMemcpyAsync(g->flag_dev + 2, h_bad, 4, cudaMemcpyHostToDevice, g->comm_stream));
  SPMMDB_NCCL_OK(nccl().AllReduce(g->flag_dev + 2, g->flag_dev + 3, 1, ncclInt32, ncclSum, g->comm, g->comm_stream));
  SPMMDB_CUDA_OK(cudaMemcpyAsync(h_bad + 1, g->flag_dev + 3, 4, cudaMemcpyDeviceToHost, g->comm_stream));
  SPMMDB_CUDA_OK(cudaStreamSynchronize(g->comm_stream));
  if (h_bad[1] != 0) { mg_close_peers(g); return SPMMDB_OK; }
  g->peer_ok = true;
  return SPMMDB_OK;
}

extern "C" int spmmdb_mg_unique_id(void* id128) {
  if (id128 == nullptr) return SPMMDB_ERR_INVALID_VALUE;
  if (!nccl().ok) return SPMMDB_ERR_NOT_SUPPORTED;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  SPMMDB_NCCL_OK(nccl().GetUniqueId(&id));
  memcpy(id128, &id, sizeof(id));
  return SPMMDB_OK;
}

extern "C" int spmmdb_mg_create(int rank, int world, const void* id128, spmmdb_mg** out) {
  if (out == nullptr || id128 == nullptr || world < 1 || rank < 0 || rank >= world) return SPMMDB_ERR_INVALID_VALUE;
  if (!nccl().ok) return SPMMDB_ERR_NOT_SUPPORTED;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  spmmdb_mg* g = new spmmdb_mg();
  g->rank = rank;
  g->world = world;
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  ncclResult_t r = nccl().CommInitRank(&g->comm, world, id, rank);
  if (r != ncclSuccess) { delete g; return nccl_fail("ncclCommInitRank", r); }
  g->own_comm = true;
  const int rc = mg_init_common(g);
  if (rc != SPMMDB_OK) { nccl().CommDestroy(g->comm); delete g; return rc; }
  *out = g;
  return SPMMDB_OK;
}

extern "C" int spmmdb_mg_create_from_comm(void* nccl_comm, int rank, int world, spmmdb_mg** out) {
  if (out == nullptr || nccl_comm == nullptr || world < 1 || rank < 0 || rank >= world) return SPMMDB_ERR_INVALID_VALUE;
  if (!nccl().ok) return SPMMDB_ERR_NOT_SUPPORTED;
  if (device_sm() < 0) return SPMMDB_ERR_CUDA;
  spmmdb_mg* g = new spmmdb_mg();
  g->rank = rank;
  g->world = world;
  g->comm = (ncclComm_t)nccl_comm;
  const int rc = mg_init_common(g);
  if (rc != SPMMDB_OK) { delete g; return rc; }
  *out = g;
  return SPMMDB_OK;
}

extern "C" int spmmdb_mg_destroy(spmmdb_mg* g) {
  if (g == nullptr) return SPMMDB_OK;
  cudaStreamSynchronize(g->comm_stream);
  mg_close_peers(g);
  if (g->own_comm && g->comm != nullptr) nccl().CommDestroy(g->comm);
  if (g->replica != nullptr) cudaFree(g->replica);
  for (int p = 0; p < g->world && p < 16; ++p) {
    if (g->push_stream[p]) cudaStreamDestroy(g->push_stream[p]);
    if (g->push_done[p]) cudaEventDestroy(g->push_done[p]);
  }
  if (g->push_go) cudaEventDestroy(g->push_go);
  if (g->flag_dev) cudaFree(g->flag_dev);
  if (g->ipc_dev) cudaFree(g->ipc_dev);
  cudaStreamDestroy(g->comm_stream);
  cudaEventDestroy(g->ev_ready);
  cudaEventDestroy(g->ev_done);
  for (auto& e : g->ev_slab) cudaEventDestroy(e);
  delete g;
  return SPMMDB_OK;
}

static int check_bounds(const spmmdb_mg* g, const int* bounds, int k) {
  if (bounds == nullptr || bounds[0] != 0 || bounds[g->world] != k) return SPMMDB_ERR_INVALID_VALUE;
  for (int p = 0; p < g->world; ++p)
    if (bounds[p + 1] < bounds[p]) return SPMMDB_ERR_INVALID_VALUE;
  return SPMMDB_OK;
}

// one all-gather-v: rank p's rows [bounds[p], bounds[p+1]) of a [k][row_elems] array, in place
static int allgather_v(spmmdb_mg* g, float* base, int64_t row_elems, const int* bounds, cudaStream_t st) {
  if (g->world == 1) return SPMMDB_OK;
  Nccl& n = nccl();
  const int me = g->rank;
  const int64_t mine = (int64_t)(bounds[me + 1] - bounds[me]) * row_elems;
  SPMMDB_NCCL_OK(n.GroupStart());
  for (int d = 1; d < g->world; ++d) {
    const int to = (me + d) % g->world, from = (me - d + g->world) % g->world;
    const int64_t theirs = (int64_t)(bounds[from + 1] - bounds[from]) * row_elems;
    if (mine > 0) SPMMDB_NCCL_OK(n.Send(base + (int64_t)bounds[me] * row_elems, (size_t)mine, ncclFloat, to, g->comm, st));
    if (theirs > 0)
      SPMMDB_NCCL_OK(n.Recv(base + (int64_t)bounds[from] * row_elems, (size_t)theirs, ncclFloat, from, g->comm, st));
  }
  SPMMDB_NCCL_OK(n.GroupEnd());
  return SPMMDB_OK;
}

extern "C" int spmmdb_mg_allgather_rows_f32(spmmdb_mg* g, float* B, int64_t ldb, int k, const int* bounds, void* stream) {
  if (g == nullptr || B == nullptr || ldb <= 0 || k < 0) return SPMMDB_ERR_INVALID_VALUE;
  int rc = check_bounds(g, bounds, k);
  if (rc != SPMMDB_OK) return rc;
  return allgather_v(g, B, ldb, bounds, (cudaStream_t)stream);  // rows are contiguous with their padding
}

extern "C" int spmmdb_sharded_csrmm_f32(spmmdb_mg* g, int k, int K, int64_t nnz_local, const int* rowptr,
                                        const int* colind, const float* val, const float* B_local, int64_t ldb,
                                        const int* bounds, float* C, int64_t ldc, float alpha, float beta, int slabs,
                                        void* stream) {
  if (g == nullptr || k < 0 || K < 0 || nnz_local < 0) return SPMMDB_ERR_INVALID_VALUE;
  int rc = check_bounds(g, bounds, k);
  if (rc != SPMMDB_OK) return rc;
  const int me = g->rank;
  const int lo = bounds[me], m_local = bounds[me + 1] - bounds[me];
  if (K == 0) return SPMMDB_OK;
  if (ldb < K || ldc < K) return SPMMDB_ERR_INVALID_VALUE;
  if ((m_local > 0 && (rowptr == nullptr || B_local == nullptr || C == nullptr)) ||
      (nnz_local > 0 && (colind == nullptr || val == nullptr)))
    return SPMMDB_ERR_INVALID_VALUE;
  if (K % 4 != 0 || ldb % 4 != 0 || ((uintptr_t)B_local % 16) != 0) return SPMMDB_ERR_NOT_SUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  if (slabs <= 0) slabs = SPMMDB_TUNE("MG_SLABS", 2);
  if (slabs > 16) slabs = 16;
  while (slabs > 1 && (K % slabs != 0 || (K / slabs) % 32 != 0)) --slabs;  // a slab keeps whole 128-byte lines
  const int Ks = K / slabs;
  const size_t need = sizeof(float) * (size_t)k * K;
  if (need > g->replica_bytes) {
    // (re)allocation is collective: every rank sees the same k and K in the same call
    SPMMDB_CUDA_OK(cudaStreamSynchronize(g->comm_stream));
    SPMMDB_CUDA_OK(cudaStreamSynchronize(st));
    mg_close_peers(g);
    if (g->world > 1) {  // every rank has closed its mappings of the old replicas before anybody frees one
      rc = mg_barrier(g);
      if (rc != SPMMDB_OK) return rc;
      SPMMDB_CUDA_OK(cudaStreamSynchronize(g->comm_stream));
    }
    if (g->replica != nullptr) cudaFree(g->replica);
    g->replica = nullptr;
    g->replica_bytes = 0;
    SPMMDB_CUDA_OK(cudaMalloc(&g->replica, need));
    g->replica_bytes = need;
    rc = mg_map_peers(g);
    if (rc != SPMMDB_OK) return rc;
  }
  const int nsm = device_num_sms() > 0 ? device_num_sms() : 148;
  // every slab's kernel cuts long rows like the whole call would (same bits as the one-shot product)
  const int seg_now = current_segment_override();
  ScopedSegment pin(seg_now ? seg_now : spmmdb_csrmm_segment_for(m_local, nnz_local));
  // the communication stream starts when the caller's stream reaches this call, and after the previous call's
  // SpMMs are done with the replica
  SPMMDB_CUDA_OK(cudaEventRecord(g->ev_ready, st));
  SPMMDB_CUDA_OK(cudaStreamWaitEvent(g->comm_stream, g->ev_ready, 0));
  if (g->peer_ok) {
    // every peer is done reading its replica (its previous call's SpMMs) before anybody writes into it
    rc = mg_barrier(g);
    if (rc != SPMMDB_OK) return rc;
  }
  for (int s = 0; s < slabs; ++s) {
    float* slab = g->replica + (size_t)s * k * Ks;
    const int peer_mode = g->peer_ok ? SPMMDB_TUNE("MG_PEER", 2) : 0;
    if (peer_mode == 2) {
      // fused pack + all-gather: SM stores into every rank's replica (profiles/r2i: copy-engine pushes to 7 peers reach
      // 480 GB/s inbound, NCCL send / recv 587)
      PushDst dst;
      const size_t off = (size_t)s * k * Ks + (size_t)lo * Ks;
      for (int d = 0; d < g->world; ++d) dst.p[d] = g->peer_replica[(me + d) % g->world] + off;
      if (m_local > 0) {
        const int ctas = SPMMDB_TUNE("MG_PUSH_CTAS", 0);
        pack_push_kernel<<<ctas > 0 ? ctas : nsm * 2, 256, 0, g->comm_stream>>>(B_local, ldb, s * Ks, Ks, dst, g->world, m_local);
      }
      SPMMDB_CUDA_OK(cudaGetLastError());
      rc = mg_barrier(g);  // every rank's stores into this rank's replica are complete (and visible) past this point
      if (rc != SPMMDB_OK) return rc;
      SPMMDB_CUDA_OK(cudaEventRecord(g->ev_slab[s], g->comm_stream));
      SPMMDB_CUDA_OK(cudaStreamWaitEvent(st, g->ev_slab[s], 0));
      if (m_local > 0) {
        rc = csrmm_rowmajor(m_local, Ks, nnz_local, rowptr, colind, val, slab, Ks, C + (size_t)s * Ks, ldc, alpha, beta, st,
                            0, nullptr);
        if (rc != SPMMDB_OK) return rc;
      }
      continue;
    }
    if (m_local > 0)
      pack_slab_kernel<<<nsm * 4, 256, 0, g->comm_stream>>>(B_local, ldb, s * Ks, Ks, slab + (size_t)lo * Ks, m_local);
    SPMMDB_CUDA_OK(cudaGetLastError());
    if (peer_mode == 1) {
      // push this rank's rows of the slab into every peer's replica: one copy-engine transfer per peer, all at once
      SPMMDB_CUDA_OK(cudaEventRecord(g->push_go, g->comm_stream));
      const size_t off = (size_t)s * k * Ks + (size_t)lo * Ks, bytes = sizeof(float) * (size_t)m_local * Ks;
      for (int d = 1; d < g->world; ++d) {
        const int to = (me + d) % g->world;
        SPMMDB_CUDA_OK(cudaStreamWaitEvent(g->push_stream[to], g->push_go, 0));
        if (bytes > 0)
          SPMMDB_CUDA_OK(cudaMemcpyAsync(g->peer_replica[to] + off, g->replica + off, bytes, cudaMemcpyDeviceToDevice,
                                         g->push_stream[to]));
        SPMMDB_CUDA_OK(cudaEventRecord(g->push_done[to], g->push_stream[to]));
        SPMMDB_CUDA_OK(cudaStreamWaitEvent(g->comm_stream, g->push_done[to], 0));
      }
      rc = mg_barrier(g);  // returns here only after every rank's pushes of this slab have completed
    } else {
      rc = allgather_v(g, slab, Ks, bounds, g->comm_stream);
    }
    if (rc != SPMMDB_OK) return rc;
    SPMMDB_CUDA_OK(cudaEventRecord(g->ev_slab[s], g->comm_stream));
    SPMMDB_CUDA_OK(cudaStreamWaitEvent(st, g->ev_slab[s], 0));
    if (m_local > 0) {
      rc = csrmm_rowmajor(m_local, Ks, nnz_local, rowptr, colind, val, slab, Ks, C + (size_t)s * Ks, ldc, alpha, beta, st,
                          0, nullptr);
      if (rc != SPMMDB_OK) return rc;
    }
  }
  SPMMDB_CUDA_OK(cudaEventRecord(g->ev_done, st));
  SPMMDB_CUDA_OK(cudaStreamWaitEvent(g->comm_stream, g->ev_done, 0));  // next call's packs wait for these SpMMs
  return SPMMDB_OK;
}
