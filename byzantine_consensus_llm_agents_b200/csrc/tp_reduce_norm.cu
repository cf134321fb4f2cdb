// Fused tensor-parallel all-reduce + residual add + RMSNorm for the decode step, over NVLink /
// NVSwitch peer memory — no NCCL call, no separate reduction kernel.
//
// After a row-parallel projection (o_proj, down_proj) every rank holds split-K partial slabs of
// ITS share of the sum.  Those slabs live in a symmetric buffer that all ranks of the replica
// have mapped (cudaIpc).  Each rank then
//   1. tp_signal_kernel: bumps its epoch for the slot and stores it, with release semantics at
//      system scope, into every peer's flag array (a kernel boundary after the GEMM guarantees
//      the slabs are complete);
//   2. tp_add_rmsnorm_kernel: waits (acquire, bounded spin) until every peer's flag has reached
//      the epoch, reads ALL ranks' slabs straight from peer memory with L1-bypassing loads, adds
//      them in a fixed (rank, split) order — so every rank computes bit-identical sums — adds the
//      residual and normalises.
// A 10-row all-reduce is ~200 KB per rank: the cost is a flag round trip over NVLink plus a
// few microseconds of peer reads, instead of ~35-55 us of NCCL latency, 2 x n_layers times a step.
// Two slots (o / down) alternate, which is enough to make buffer reuse safe: a rank overwrites
// slot s only after the NEXT wait on the other slot, which every peer signals only after it has
// finished reading slot s.
//
// Symmetric buffer layout (floats): [flag area: kFlagFloats][slot 0 slabs][slot 1 slabs]
#include "common.cuh"
#include "kernels.h"

namespace bcg {

constexpr int kTpMax = 8;
constexpr long long kFlagFloats = 256;     // epochs[2] at 0..1, flags[2][8] at 64.. (ints)
constexpr int kTpNormThreads = 1024;

long long tp_symm_flag_floats() { return kFlagFloats; }

__device__ __forceinline__ void st_release_sys(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_v4(const float* p) {
  float4 r;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

__global__ void tp_signal_kernel(void* const* __restrict__ peers, int rank, int size, int slot) {
  pdl_launch_dependents();
  pdl_wait();
  int* local = reinterpret_cast<int*>(peers[rank]);
  __shared__ int epoch;
  if (threadIdx.x == 0) {
    epoch = local[slot] + 1;
    local[slot] = epoch;
    __threadfence_system();
  }
  __syncthreads();
  if (threadIdx.x < size) {
    int* flags = reinterpret_cast<int*>(peers[threadIdx.x]) + 64 + slot * kTpMax + rank;
    st_release_sys(flags, epoch);
  }
}

__global__ void __launch_bounds__(kTpNormThreads)
tp_add_rmsnorm_kernel(float* __restrict__ residual, void* const* __restrict__ peers, int rank, int size,
                      long long region_offset, int n_partials, long long partial_stride,
                      const float* __restrict__ w, __nv_bfloat16* __restrict__ xn, int slot, int d,
                      float eps) {
  __shared__ float red[kTpNormThreads / 32];
  pdl_launch_dependents();
  pdl_wait();
  const int* local = reinterpret_cast<const int*>(peers[rank]);
  if (threadIdx.x < size) {
    const int want = local[slot];   // written by this rank's tp_signal_kernel, earlier on this stream
    const int* flag = local + 64 + slot * kTpMax + threadIdx.x;
    unsigned spins = 0;
    while (ld_acquire_sys(flag) < want) {
      if (++spins > (1u << 24)) __trap();   // a peer died: fail loudly instead of hanging the GPU
    }
  }
  __syncthreads();
  const int t = blockIdx.x;
  const int nvec = d >> 2;
  float4 x[2];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int v = threadIdx.x + i * kTpNormThreads;
    if (v < nvec) {
      float4 r = *reinterpret_cast<const float4*>(residual + static_cast<long long>(t) * d + v * 4);
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      // all (rank, split) slabs as one flat list, loaded eight at a time so that a whole batch
      // of peer reads shares one NVLink round trip; summed in list order = fixed (rank, split)
      // order on every rank
      const int total = size * n_partials;
      const long long elem = static_cast<long long>(t) * d + v * 4;
      for (int i0 = 0; i0 < total; i0 += 8) {
        float4 p[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int idx = i0 + k;
          if (idx < total) {
            const int rk = idx / n_partials, sp = idx - rk * n_partials;
            p[k] = ld_peer_v4(reinterpret_cast<const float*>(peers[rk]) + region_offset + elem +
                              sp * partial_stride);
          }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (i0 + k < total) {
            acc.x += p[k].x; acc.y += p[k].y; acc.z += p[k].z; acc.w += p[k].w;
          }
      }
      r.x += acc.x; r.y += acc.y; r.z += acc.z; r.w += acc.w;
      x[i] = r;
      ss += r.x * r.x + r.y * r.y + r.z * r.z + r.w * r.w;
    }
  }
  ss = warp_sum(ss);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) red[warp] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < kTpNormThreads / 32; ++i) tot += red[i];
  const float inv = rsqrtf(tot / static_cast<float>(d) + eps);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int v = threadIdx.x + i * kTpNormThreads;
    if (v < nvec) {
      *reinterpret_cast<float4*>(residual + static_cast<long long>(t) * d + v * 4) = x[i];
      const float4 ww = *reinterpret_cast<const float4*>(w + v * 4);
      uint2 o;
      o.x = pack_bf16(x[i].x * inv * ww.x, x[i].y * inv * ww.y);
      o.y = pack_bf16(x[i].z * inv * ww.z, x[i].w * inv * ww.w);
      *reinterpret_cast<uint2*>(xn + static_cast<long long>(t) * d + v * 4) = o;
    }
  }
}

int tp_signal(int slot, cudaStream_t stream) {
  if (!tp_symm_peers_device()) {
    set_last_error("bcg_tp", "symmetric buffer not set up");
    return BCG_ERR_ARG;
  }
  BCG_CUDA_CHECK(launch_k(tp_signal_kernel, dim3(1), dim3(32), 0, stream, tp_symm_peers_device(), tp_rank(),
                          tp_size(), slot));
  return BCG_OK;
}

int tp_add_rmsnorm(float* residual, long long region_offset_floats, int n_partials, long long partial_stride,
                   const float* w, void* xn, int slot, int T, int d, float eps, cudaStream_t stream) {
  if (!tp_symm_peers_device() || !residual || !w || !xn || T <= 0 || (d % 4) || d > kTpNormThreads * 8) {
    set_last_error("bcg_tp_add_rmsnorm", "bad argument or symmetric buffer not set up");
    return BCG_ERR_ARG;
  }
  BCG_CUDA_CHECK(launch_k(tp_add_rmsnorm_kernel, dim3(T), dim3(kTpNormThreads), 0, stream, residual,
                          tp_symm_peers_device(), tp_rank(), tp_size(), region_offset_floats, n_partials,
                          partial_stride, w, reinterpret_cast<__nv_bfloat16*>(xn), slot, d, eps));
  return BCG_OK;
}

}  // namespace bcg
