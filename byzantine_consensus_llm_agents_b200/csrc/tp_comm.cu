// Tensor-parallel plumbing: one NCCL communicator per process (one process per GPU), used by the
// model steps for the two all-reduces per layer (after the row-parallel o and down projections)
// and the all-gather of vocab-parallel logits — the collectives vLLM issues for the reference when
// `tensor_parallel_size > 1` (vllm_agent.py:131,141-142).
// libnccl is the one PyTorch already loaded; it is resolved with dlopen so the library has no
// link-time dependency on it and single-GPU users never touch it.
#include <dlfcn.h>
#include <string.h>

#include "common.cuh"
#include "kernels.h"

namespace bcg {

struct NcclUniqueId { char internal[128]; };
typedef void* NcclComm;
typedef int (*GetUniqueIdFn)(NcclUniqueId*);
typedef int (*CommInitRankFn)(NcclComm*, int, NcclUniqueId, int);
typedef int (*CommDestroyFn)(NcclComm);
typedef int (*AllReduceFn)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t);
typedef int (*AllGatherFn)(const void*, void*, size_t, int, NcclComm, cudaStream_t);
typedef const char* (*GetErrorStringFn)(int);

static struct {
  void* handle = nullptr;
  GetUniqueIdFn get_unique_id = nullptr;
  CommInitRankFn comm_init_rank = nullptr;
  CommDestroyFn comm_destroy = nullptr;
  AllReduceFn all_reduce = nullptr;
  AllGatherFn all_gather = nullptr;
  GetErrorStringFn error_string = nullptr;
  NcclComm comm = nullptr;
  int rank = 0, size = 1;
} g_tp;

constexpr int kMaxTp = 8;
constexpr int kNcclFloat32 = 7;

// symmetric peer memory for the fused decode all-reduce (tp_reduce_norm.cu)
static struct {
  void* local = nullptr;
  void* peers[kMaxTp] = {};
  long long bytes = 0;
  void** d_peers = nullptr;   // device copy of peers[]
} g_sym;

constexpr int kNcclSum = 0;

static int load_nccl() {
  if (g_tp.handle) return BCG_OK;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    set_last_error("bcg_tp", "libnccl.so.2 not found (import torch first, or set LD_LIBRARY_PATH)");
    return BCG_ERR_UNSUPPORTED;
  }
  g_tp.get_unique_id = reinterpret_cast<GetUniqueIdFn>(dlsym(h, "ncclGetUniqueId"));
  g_tp.comm_init_rank = reinterpret_cast<CommInitRankFn>(dlsym(h, "ncclCommInitRank"));
  // Teardown uses ncclCommAbort: ncclCommDestroy blocks while CUDA graphs that contain NCCL
  // kernels are still alive (the engine's captured decode steps), abort never does.
  g_tp.comm_destroy = reinterpret_cast<CommDestroyFn>(dlsym(h, "ncclCommAbort"));
  g_tp.all_reduce = reinterpret_cast<AllReduceFn>(dlsym(h, "ncclAllReduce"));
  g_tp.all_gather = reinterpret_cast<AllGatherFn>(dlsym(h, "ncclAllGather"));
  g_tp.error_string = reinterpret_cast<GetErrorStringFn>(dlsym(h, "ncclGetErrorString"));
  if (!g_tp.get_unique_id || !g_tp.comm_init_rank || !g_tp.comm_destroy || !g_tp.all_reduce ||
      !g_tp.all_gather) {
    set_last_error("bcg_tp", "libnccl.so.2 lacks a required symbol");
    return BCG_ERR_UNSUPPORTED;
  }
  g_tp.handle = h;
  return BCG_OK;
}

static int nccl_check(int rc, const char* what) {
  if (rc == 0) return BCG_OK;
  set_last_error(what, g_tp.error_string ? g_tp.error_string(rc) : "NCCL error");
  return BCG_ERR_CUDA;
}

int tp_unique_id(void* out128) {
  if (!out128) return BCG_ERR_ARG;
  int rc = load_nccl();
  if (rc != BCG_OK) return rc;
  NcclUniqueId id;
  rc = nccl_check(g_tp.get_unique_id(&id), "ncclGetUniqueId");
  if (rc != BCG_OK) return rc;
  memcpy(out128, &id, sizeof(id));
  return BCG_OK;
}

int tp_init(const void* id128, int rank, int size) {
  if (!id128 || size < 1 || rank < 0 || rank >= size) {
    set_last_error("bcg_tp_init", "bad rank / size");
    return BCG_ERR_ARG;
  }
  int rc = load_nccl();
  if (rc != BCG_OK) return rc;
  if (g_tp.comm) {
    g_tp.comm_destroy(g_tp.comm);
    g_tp.comm = nullptr;
  }
  NcclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  rc = nccl_check(g_tp.comm_init_rank(&g_tp.comm, size, id, rank), "ncclCommInitRank");
  if (rc != BCG_OK) return rc;
  g_tp.rank = rank;
  g_tp.size = size;
  return BCG_OK;
}

int tp_shutdown() {
  for (int r = 0; r < kMaxTp; ++r) {
    if (g_sym.peers[r] && g_sym.peers[r] != g_sym.local) cudaIpcCloseMemHandle(g_sym.peers[r]);
    g_sym.peers[r] = nullptr;
  }
  if (g_sym.d_peers) cudaFree(g_sym.d_peers);
  if (g_sym.local) cudaFree(g_sym.local);
  g_sym.d_peers = nullptr;
  g_sym.local = nullptr;
  g_sym.bytes = 0;
  if (g_tp.comm) {
    g_tp.comm_destroy(g_tp.comm);
    g_tp.comm = nullptr;
  }
  g_tp.rank = 0;
  g_tp.size = 1;
  return BCG_OK;
}

// ---------------------------------------------------------------------------------------------
// Symmetric peer memory for the fused decode all-reduce (tp_reduce_norm.cu)
// ---------------------------------------------------------------------------------------------
int tp_symm_create(long long bytes, void* handle_out64, void** local_ptr) {
  if (bytes <= 0 || !handle_out64 || !local_ptr) return BCG_ERR_ARG;
  if (g_sym.local) {
    set_last_error("bcg_tp_symm_create", "already created");
    return BCG_ERR_ARG;
  }
  BCG_CUDA_CHECK(cudaMalloc(&g_sym.local, static_cast<size_t>(bytes)));
  BCG_CUDA_CHECK(cudaMemset(g_sym.local, 0, static_cast<size_t>(bytes)));
  cudaIpcMemHandle_t h;
  BCG_CUDA_CHECK(cudaIpcGetMemHandle(&h, g_sym.local));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle_out64, &h, 64);
  g_sym.bytes = bytes;
  *local_ptr = g_sym.local;
  return BCG_OK;
}

// handles: [size][64] bytes in rank order (this rank's own entry is ignored)
int tp_symm_open(const void* handles, int rank, int size) {
  if (!handles || !g_sym.local || size < 1 || size > kMaxTp || rank < 0 || rank >= size) {
    set_last_error("bcg_tp_symm_open", "bad arguments (tp size <= 8) or bcg_tp_symm_create not called");
    return BCG_ERR_ARG;
  }
  for (int r = 0; r < size; ++r) {
    if (r == rank) {
      g_sym.peers[r] = g_sym.local;
      continue;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const char*>(handles) + 64 * r, 64);
    BCG_CUDA_CHECK(cudaIpcOpenMemHandle(&g_sym.peers[r], h, cudaIpcMemLazyEnablePeerAccess));
  }
  BCG_CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&g_sym.d_peers), sizeof(void*) * kMaxTp));
  BCG_CUDA_CHECK(cudaMemcpy(g_sym.d_peers, g_sym.peers, sizeof(void*) * kMaxTp, cudaMemcpyHostToDevice));
  return BCG_OK;
}

void* tp_symm_local() { return g_sym.local; }
void* const* tp_symm_peers_device() { return g_sym.d_peers; }
long long tp_symm_bytes() { return g_sym.bytes; }

int tp_size() { return g_tp.comm ? g_tp.size : 1; }
int tp_rank() { return g_tp.comm ? g_tp.rank : 0; }

int tp_all_reduce_f32(float* buf, long long count, cudaStream_t stream) {
  if (!g_tp.comm) {
    set_last_error("bcg_tp", "tensor-parallel step without bcg_tp_init");
    return BCG_ERR_ARG;
  }
  return nccl_check(g_tp.all_reduce(buf, buf, static_cast<size_t>(count), kNcclFloat32, kNcclSum, g_tp.comm, stream),
                    "ncclAllReduce");
}

// recv = [size][count] floats, rank-major
int tp_all_gather_f32(const float* send, float* recv, long long count, cudaStream_t stream) {
  if (!g_tp.comm) {
    set_last_error("bcg_tp", "tensor-parallel step without bcg_tp_init");
    return BCG_ERR_ARG;
  }
  return nccl_check(g_tp.all_gather(send, recv, static_cast<size_t>(count), kNcclFloat32, g_tp.comm, stream),
                    "ncclAllGather");
}

}  // namespace bcg
