// Tensor-parallel plumbing: one NCCL communicator per process (one process per GPU), used by the
// model steps for the two all-reduces per layer (after the row-parallel o and down projections)
// and the all-gather of vocab-parallel logits — the collectives vLLM issues for the reference when
// `tensor_parallel_size > 1` (vllm_agent.py:131,141-142).
// libnccl is the one PyTorch already loaded; it is resolved with dlopen so the library has no
// link-time dependency on it and single-GPU users never touch it.
#include <dlfcn.h>

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

constexpr int kNcclFloat32 = 7;
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
  if (g_tp.comm) {
    g_tp.comm_destroy(g_tp.comm);
    g_tp.comm = nullptr;
  }
  g_tp.rank = 0;
  g_tp.size = 1;
  return BCG_OK;
}

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
