// Fused tensor-parallel all-reduce + residual add + RMSNorm statistics for the decode step, over
// NVLink / NVSwitch peer memory — no NCCL call, no separate reduction or signalling kernel, no fence.
//
// After a row-parallel projection (o_proj, down_proj) every rank holds split-K partial slabs of
// ITS share of the sum in its LOCAL scratch.  One kernel per all-reduce, grid (8 segments, rows),
// the same decomposition on every rank; CTA (segment, row) of rank r
//   1. adds its local slabs in index order  ->  one fp32 partial per element (B*d floats per rank,
//      not splits*B*d: the exchange carries the reduced message only);
//   2. PUSHES that partial into inbox [r] of every peer's symmetric buffer with plain 16-byte stores
//      over NVLink, each 8-byte half carrying one value and the launch's epoch ({x, epoch, y, epoch}:
//      8-byte stores are single transactions, so a reader that sees the epoch sees the value — the
//      "LL" idea of NCCL's low-latency protocol).  Posted writes: one-way latency, no membar, no flag;
//   3. polls ITS OWN inboxes (local memory) until every peer's words carry the epoch (bounded spin)
//      and adds the size partials in rank order (its own from registers): every rank forms the same
//      sum bit for bit;
//   4. finishes like bcg_add_norm_stats: residual += sum, xn = bf16(residual * w), the segment's
//      sum of squares to sumsq[segment][row]; the projection that reads xn applies 1/RMS in its
//      epilogue (bcg_gemm_bf16_tn_fused consumer half).
// A thread only ever waits for the thread with its own index on other GPUs, so nothing needs to be
// co-resident and the kernel composes with programmatic dependent launch.  Per all-reduce a rank
// sends (size-1) * B*d*8 bytes (2.9 MB at 10 rows, d 5120, TP 8) and receives as much.  Two slots
// (o / down) alternate, which makes buffer reuse safe: a rank overwrites slot s only two all-reduces
// later, and it cannot finish the one in between before every peer has started it — which a peer
// does only after its kernel on slot s ended.  Epochs are per-CTA launch counters kept in the
// (local) header, identical on every rank because every rank makes the same sequence of launches.
//
// Symmetric buffer layout (4-byte words): header kTpHeaderWords, then [slot 2][src size][2*region].
//   header: launch counters [2][1024] (local), sampler epochs [128] (local),
//           sampler candidates [2][8][128][4]  (see sampler.cu)
#include "common.cuh"
#include "kernels.h"

namespace bcg {

constexpr int kTpMax = 8;
constexpr int kTpMaxCtas = 1024;                       // 128 rows x 8 segments
constexpr long long kTpHeaderWords = 32768;
constexpr int kTpOffEpoch = 0;                         // [2][kTpMaxCtas]
constexpr int kTpSegments = 8;
constexpr int kTpSegThreads = 256;
constexpr int kTpSegTurns = 2;                         // d <= 8 * 256 * 2 * 4 = 16384
constexpr int kTpStatsStride = 128;

long long tp_symm_flag_floats() { return kTpHeaderWords; }


__device__ __forceinline__ uint4 ld_sys_u4(const float* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}
__device__ __forceinline__ void st_sys_u4(float* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d)
               : "memory");
}

__global__ void __launch_bounds__(kTpSegThreads)
tp_add_norm_stats_kernel(float* __restrict__ residual, const float* __restrict__ partials, int n_partials,
                         long long partial_stride, const float* __restrict__ w,
                         __nv_bfloat16* __restrict__ xn, float* __restrict__ sumsq, int d,
                         void* const* __restrict__ peers, int rank, int size, int slot, long long region) {
  __shared__ float red[kTpSegThreads / 32];
  __shared__ int s_epoch;
  pdl_launch_dependents();
  const int seg = blockIdx.x, t = blockIdx.y;
  const int cta = t * kTpSegments + seg;
  int* local = reinterpret_cast<int*>(peers[rank]);
  const int nvec = d >> 2;
  const int per = (nvec + kTpSegments - 1) / kTpSegments;
  const int v_end = min(nvec, (seg + 1) * per);
  // the launch counter of this (slot, CTA) is touched by nobody else: bump it before the dependency wait
  if (threadIdx.x == 0) {
    const int e = local[kTpOffEpoch + slot * kTpMaxCtas + cta] + 1;
    local[kTpOffEpoch + slot * kTpMaxCtas + cta] = e;
    s_epoch = e;
  }
  __syncthreads();
  const uint32_t epoch = static_cast<uint32_t>(s_epoch);
  pdl_wait();
  // 1. local split-K slabs -> one partial per element; 2. push {value, epoch} pairs to every peer
  float4 mine[kTpSegTurns];
  bool have[kTpSegTurns];
  const long long slot_base = kTpHeaderWords + static_cast<long long>(slot) * size * (2 * region);
#pragma unroll
  for (int i = 0; i < kTpSegTurns; ++i) {
    const int v = seg * per + threadIdx.x + i * kTpSegThreads;
    have[i] = threadIdx.x + i * kTpSegThreads < per && v < v_end;
    mine[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (have[i]) {
      const long long off = static_cast<long long>(t) * d + v * 4;
      const float* dp = partials + off;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int s0 = 0; s0 < n_partials; s0 += 8) {
        float4 p[8];
#pragma unroll
        for (int s = 0; s < 8; ++s)   // unconditional, clamped: all eight loads in flight at once
          p[s] = *reinterpret_cast<const float4*>(dp + min(s0 + s, n_partials - 1) * partial_stride);
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          if (s0 + s < n_partials) {
            if (s0 + s == 0) {
              acc = p[s];
            } else {
              acc.x += p[s].x; acc.y += p[s].y; acc.z += p[s].z; acc.w += p[s].w;
            }
          }
        }
      }
      mine[i] = acc;
      for (int pr = 0; pr < size; ++pr) {
        if (pr == rank) continue;
        float* dst = reinterpret_cast<float*>(peers[pr]) + slot_base + rank * (2 * region) + 2 * off;
        st_sys_u4(dst, __float_as_uint(acc.x), epoch, __float_as_uint(acc.y), epoch);
        st_sys_u4(dst + 4, __float_as_uint(acc.z), epoch, __float_as_uint(acc.w), epoch);
      }
    }
  }
  // 3. poll the local inboxes; 4. rank-ordered sum, residual, bf16(r * w), segment sum of squares
  const float* inbox = reinterpret_cast<const float*>(local) + slot_base;
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kTpSegTurns; ++i) {
    if (have[i]) {
      const int v = seg * per + threadIdx.x + i * kTpSegThreads;
      const long long off = static_cast<long long>(t) * d + v * 4;
      float4 p[kTpMax];
      unsigned pending = 0;
      for (int src = 0; src < size; ++src)
        if (src != rank) pending |= 1u << src;
      unsigned spins = 0;
      while (pending) {
#pragma unroll
        for (int src = 0; src < kTpMax; ++src) {
          if (pending & (1u << src)) {
            const float* q = inbox + src * (2 * region) + 2 * off;
            const uint4 a = ld_sys_u4(q), b = ld_sys_u4(q + 4);
            if (a.y == epoch && a.w == epoch && b.y == epoch && b.w == epoch) {
              p[src] = make_float4(__uint_as_float(a.x), __uint_as_float(a.z), __uint_as_float(b.x),
                                   __uint_as_float(b.z));
              pending &= ~(1u << src);
            }
          }
        }
        if (++spins > (1u << 22)) __trap();   // a peer died: fail loudly instead of hanging the GPU
      }
      float4 tot = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int src = 0; src < kTpMax; ++src) {
        if (src < size) {
          const float4 a = (src == rank) ? mine[i] : p[src];
          if (src == 0) {
            tot = a;
          } else {
            tot.x += a.x; tot.y += a.y; tot.z += a.z; tot.w += a.w;
          }
        }
      }
      float4 x = *reinterpret_cast<const float4*>(residual + off);
      x.x += tot.x; x.y += tot.y; x.z += tot.z; x.w += tot.w;
      const float4 ww = *reinterpret_cast<const float4*>(w + v * 4);
      ss = __fadd_rn(ss, __fmaf_rn(x.w, x.w, __fmaf_rn(x.z, x.z, __fmaf_rn(x.y, x.y, __fmul_rn(x.x, x.x)))));
      *reinterpret_cast<float4*>(residual + off) = x;
      uint2 o;
      o.x = pack_bf16(x.x * ww.x, x.y * ww.y);
      o.y = pack_bf16(x.z * ww.z, x.w * ww.w);
      *reinterpret_cast<uint2*>(xn + off) = o;
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float sg = 0.f;
#pragma unroll
    for (int i = 0; i < kTpSegThreads / 32; ++i) sg += red[i];
    sumsq[seg * kTpStatsStride + t] = sg;
  }
}

int tp_add_norm_stats(float* residual, const float* partials, int n_partials, long long partial_stride,
                      const float* w, void* xn, float* sumsq, int slot, long long region_floats, int T,
                      int d, cudaStream_t stream) {
  if (!tp_symm_peers_device() || !residual || !partials || !w || !xn || !sumsq || T <= 0 || T > 128 ||
      (d % 4) || d > kTpSegments * kTpSegThreads * kTpSegTurns * 4 || n_partials < 1 || slot < 0 || slot > 1 ||
      static_cast<long long>(T) * d > region_floats || tp_size() > kTpMax ||
      (kTpHeaderWords + 4 * tp_size() * region_floats) * 4 > tp_symm_bytes()) {
    set_last_error("bcg_tp_add_norm_stats", "bad argument, more than 128 rows, or symmetric buffer too small / not set up");
    return BCG_ERR_ARG;
  }
  BCG_CUDA_CHECK(launch_k(tp_add_norm_stats_kernel, dim3(kTpSegments, T), dim3(kTpSegThreads), 0, stream,
                          residual, partials, n_partials, partial_stride, w,
                          reinterpret_cast<__nv_bfloat16*>(xn), sumsq, d, tp_symm_peers_device(), tp_rank(),
                          tp_size(), slot, region_floats));
  return BCG_OK;
}

}  // namespace bcg
