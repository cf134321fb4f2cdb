// Memory-bound glue kernels of the decoder layer:
//   embed_rmsnorm        K1+K2 : h = E[ids]; xn = rmsnorm(h) * w
//   add_rmsnorm          K2    : h += delta (bf16, or fp32 split-K partials summed in fixed order);
//                                xn = rmsnorm(h) * w
//   qknorm_rope_kvwrite  K4-K6 : per-head RMSNorm of q,k (Qwen3), rotate-half RoPE, q -> bf16,
//                                k,v -> paged bf16 KV cache
//   swiglu_reduce        K10   : silu(gate) * up over split-K partials (decode, small models)
// The residual stream is kept in fp32 (tighter against the fp32 CPU forward that is the
// oracle); every row is computed independently of the batch it is in (batch-invariant).
// Arithmetic follows transformers' Qwen3RMSNorm / apply_rotary_pos_emb in fp32
// (modeling_qwen3.py:50-64,86-148 as cited in SURVEY.md §8c).
#include <cooperative_groups.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace bcg {

constexpr int kNormThreads = 1024;
constexpr int kMaxVecPerThread = 2;  // d <= 1024 * 2 * 4 = 8192
constexpr int kMaxPartialsBatch = 8;

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = 0.f;
  // fixed order over the warps: deterministic
#pragma unroll
  for (int i = 0; i < kNormThreads / 32; ++i) t += red[i];
  __syncthreads();
  return t;
}

__device__ __forceinline__ void store_bf16x4(__nv_bfloat16* p, float a, float b, float c, float d) {
  uint2 v;
  v.x = pack_bf16(a, b);
  v.y = pack_bf16(c, d);
  *reinterpret_cast<uint2*>(p) = v;
}

__global__ void __launch_bounds__(kNormThreads)
embed_rmsnorm_kernel(const int* __restrict__ ids, const __nv_bfloat16* __restrict__ embed,
                     const float* __restrict__ w, float* __restrict__ residual,
                     __nv_bfloat16* __restrict__ xn, int d, float eps) {
  __shared__ float red[kNormThreads / 32];
  pdl_launch_dependents();
  pdl_wait();
  const int t = blockIdx.x;
  const long long row = ids[t];
  const int nvec = d >> 2;
  float4 x[kMaxVecPerThread];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxVecPerThread; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) {
      const uint2 raw = *reinterpret_cast<const uint2*>(embed + row * d + v * 4);
      x[i] = make_float4(bf16lo(raw.x), bf16hi(raw.x), bf16lo(raw.y), bf16hi(raw.y));
      ss += x[i].x * x[i].x + x[i].y * x[i].y + x[i].z * x[i].z + x[i].w * x[i].w;
    }
  }
  const float tot = block_sum(ss, red);
  const float inv = rsqrtf(tot / static_cast<float>(d) + eps);
#pragma unroll
  for (int i = 0; i < kMaxVecPerThread; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) {
      *reinterpret_cast<float4*>(residual + static_cast<long long>(t) * d + v * 4) = x[i];
      const float4 ww = *reinterpret_cast<const float4*>(w + v * 4);
      store_bf16x4(xn + static_cast<long long>(t) * d + v * 4, x[i].x * inv * ww.x,
                   x[i].y * inv * ww.y, x[i].z * inv * ww.z, x[i].w * inv * ww.w);
    }
  }
}

// ---------------------------------------------------------------------------------------
// add_rmsnorm.  Decode calls this twice per layer on <= 128 rows: one CTA per row leaves 138 SMs
// idle and makes every thread walk ~60 values through several dependent L2 round trips, so the
// kernel was pure latency (5.4 us against 0.4 us of traffic).  A row is therefore cut into
// kNormSegments segments of "virtual threads":
//     vector v (4 floats)  ->  segment s = v / per,  j = v - s * per   (per = ceil(nvec / 8))
//                              virtual thread vt = j % 256,  turn i = j / 256
// and the sum of squares is ALWAYS formed in this one order: virtual thread (turns in order)
// -> xor-shuffle tree over its 32-lane virtual warp -> the segment's 8 warps in order -> the
// 8 segments in order.  Two kernels execute that order:
//   * add_rmsnorm_cluster_kernel (T <= 128): a cluster of 8 CTAs x 256 threads per row, one
//     segment per CTA, every load of a thread in flight at once, segment sums exchanged through
//     distributed shared memory;
//   * add_rmsnorm_row_kernel (prefill-sized T): one 128-thread CTA per row, whose 4 warps play the
//     64 virtual warps in turn.
// Same order, same intrinsics -> bit-identical rows, so a row's result does not depend on how
// many other rows share the launch (batch invariance across prefill and decode shapes).
//
// src_rows == nullptr: in-place residual update over all T rows.
// src_rows != nullptr: row i reads residual/delta row src_rows[i], writes xn row i only
//                      (last-token gather in front of lm_head); residual is left untouched.
// ---------------------------------------------------------------------------------------
constexpr int kNormSegments = 8;
constexpr int kSegThreads = 256;
constexpr int kSegTurns = 2;  // d <= 8 * 256 * 2 * 4 = 16384
constexpr int kPartialsInFlight = 12;
constexpr int kNormClusterMaxRows = 128;
constexpr int kNormStatsStride = 128;     // rows per segment entry of the deferred-scale table  // decode-sized launches take the cluster kernel

__device__ __forceinline__ float sumsq4(const float4& x) {
  return __fmaf_rn(x.w, x.w, __fmaf_rn(x.z, x.z, __fmaf_rn(x.y, x.y, __fmul_rn(x.x, x.x))));
}

// residual row + delta (bf16, or fp32 split-K partials added in index order) for one vector.
// BATCH partial loads are issued before any is used.
template <int BATCH>
__device__ __forceinline__ float4 residual_plus_delta(const float* __restrict__ residual,
                                                      const void* __restrict__ delta,
                                                      int delta_is_f32, int n_partials,
                                                      long long partial_stride, long long off) {
  float4 r = *reinterpret_cast<const float4*>(residual + off);
  if (!delta) return r;
  if (!delta_is_f32) {
    const uint2 raw = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(delta) + off);
    r.x += bf16lo(raw.x); r.y += bf16hi(raw.x); r.z += bf16lo(raw.y); r.w += bf16hi(raw.y);
    return r;
  }
  const float* dp = reinterpret_cast<const float*>(delta) + off;
  if (n_partials == 1) {  // prefill: one fp32 slab, no reason to issue clamped duplicates
    const float4 p0 = *reinterpret_cast<const float4*>(dp);
    r.x += p0.x; r.y += p0.y; r.z += p0.z; r.w += p0.w;
    return r;
  }
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int s0 = 0; s0 < n_partials; s0 += BATCH) {
    float4 p[BATCH];
    // Unconditional loads (slabs past the end re-read the last one and are ignored below): with
    // a predicate per load the compiler runs out of predicate registers and interleaves loads
    // with the dependent adds, which turns one memory round trip into three.
#pragma unroll
    for (int s = 0; s < BATCH; ++s)
      p[s] = *reinterpret_cast<const float4*>(dp + min(s0 + s, n_partials - 1) * partial_stride);
#pragma unroll
    for (int s = 0; s < BATCH; ++s) {
      if (s0 + s < n_partials) {
        if (s0 + s == 0) {
          acc = p[s];
        } else {
          acc.x += p[s].x; acc.y += p[s].y; acc.z += p[s].z; acc.w += p[s].w;
        }
      }
    }
  }
  r.x += acc.x; r.y += acc.y; r.z += acc.z; r.w += acc.w;
  return r;
}

__device__ __forceinline__ void norm_store(float* __restrict__ residual, __nv_bfloat16* __restrict__ xn,
                                           bool write_residual, long long off, const float4& x,
                                           const float4& ww, float inv) {
  if (write_residual) *reinterpret_cast<float4*>(residual + off) = x;
  store_bf16x4(xn + off, x.x * inv * ww.x, x.y * inv * ww.y, x.z * inv * ww.z, x.w * inv * ww.w);
}

// Register budget: this kernel must become resident NEXT TO a GEMM CTA (6 warps x 160 registers;
// the SM sub-partitions holding two of those warps have 16,384 - 10,240 = 6,144 registers left)
// so that it — and, behind it, the next GEMM's CTAs — can start before the GEMM in front has
// drained.  Two of its warps land on each sub-partition: <= 96 registers per thread.  At 101
// registers the norm CTAs (and the GEMM CTAs queued behind them) only started once the previous
// GEMM's CTAs had exited (tools/timeline.py).
__global__ void __cluster_dims__(kNormSegments, 1, 1) __maxnreg__(96)
add_rmsnorm_cluster_kernel(float* __restrict__ residual, const void* __restrict__ delta,
                           int delta_is_f32, int n_partials, long long partial_stride,
                           const float* __restrict__ w, __nv_bfloat16* __restrict__ xn,
                           const int* __restrict__ src_rows, int d, float eps) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ float red[kSegThreads / 32];
  __shared__ float seg_ss;                 // this segment's sum of squares (read by the peers)
  __shared__ float all_ss[kNormSegments];
  pdl_launch_dependents();
  if (threadIdx.x == 0) BCG_TL(2, 0);
  pdl_wait();
  if (threadIdx.x == 0) BCG_TL(2, 1);
  const int seg = blockIdx.x;  // == rank in the cluster
  const int t = blockIdx.y;
  const long long src = src_rows ? src_rows[t] : t;
  const int nvec = d >> 2;
  const int per = (nvec + kNormSegments - 1) / kNormSegments;
  const int v_end = min(nvec, (seg + 1) * per);
  float4 x[kSegTurns], ww[kSegTurns];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kSegTurns; ++i) {
    const int v = seg * per + threadIdx.x + i * kSegThreads;
    if (threadIdx.x + i * kSegThreads < per && v < v_end) {
      ww[i] = *reinterpret_cast<const float4*>(w + v * 4);
      x[i] = residual_plus_delta<kPartialsInFlight>(residual, delta, delta_is_f32, n_partials,
                                                    partial_stride, src * d + v * 4);
      ss = __fadd_rn(ss, sumsq4(x[i]));
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float sg = 0.f;
#pragma unroll
    for (int i = 0; i < kSegThreads / 32; ++i) sg += red[i];
    seg_ss = sg;
    BCG_TL(2, 4);
  }
  cluster.sync();
  if (threadIdx.x == 0) BCG_TL(2, 5);
  if (threadIdx.x < kNormSegments) all_ss[threadIdx.x] = *cluster.map_shared_rank(&seg_ss, threadIdx.x);
  __syncthreads();
  if (threadIdx.x == 0) BCG_TL(2, 6);
  cluster.barrier_arrive();  // my reads of the peers' seg_ss are done
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < kNormSegments; ++i) tot += all_ss[i];
  const float inv = rsqrtf(tot / static_cast<float>(d) + eps);
#pragma unroll
  for (int i = 0; i < kSegTurns; ++i) {
    const int v = seg * per + threadIdx.x + i * kSegThreads;
    if (threadIdx.x + i * kSegThreads < per && v < v_end)
      norm_store(residual, xn, !src_rows, static_cast<long long>(t) * d + v * 4, x[i], ww[i], inv);
  }
  if (threadIdx.x == 0) BCG_TL(2, 7);
  cluster.barrier_wait();  // nobody leaves while a peer may still read its seg_ss
  if (threadIdx.x == 0) BCG_TL(2, 3);
}

// Deferred-scale variant (decode): the residual add and bf16(r * w) are elementwise, only 1/RMS
// needs the whole row — and the projection that consumes the activations can apply it in its
// epilogue (bcg_gemm_bf16_tn_fused, consumer half).  So this kernel stops after the segment sums:
// no cluster barrier, no DSMEM exchange on the critical path between two GEMMs.  It writes the
// residual, xn = bf16(r * w) and sumsq[segment][row] (8 segments, stride 128 rows); the segment
// sums are formed exactly as in the two kernels above.
__global__ void __maxnreg__(96)
add_norm_stats_kernel(float* __restrict__ residual, const void* __restrict__ delta, int delta_is_f32,
                      int n_partials, long long partial_stride, const float* __restrict__ w,
                      __nv_bfloat16* __restrict__ xn, float* __restrict__ sumsq, int d) {
  __shared__ float red[kSegThreads / 32];
  pdl_launch_dependents();
  if (threadIdx.x == 0) BCG_TL(2, 0);
  pdl_wait();
  if (threadIdx.x == 0) BCG_TL(2, 1);
  const int seg = blockIdx.x;
  const int t = blockIdx.y;
  const int nvec = d >> 2;
  const int per = (nvec + kNormSegments - 1) / kNormSegments;
  const int v_end = min(nvec, (seg + 1) * per);
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kSegTurns; ++i) {
    const int v = seg * per + threadIdx.x + i * kSegThreads;
    if (threadIdx.x + i * kSegThreads < per && v < v_end) {
      const long long off = static_cast<long long>(t) * d + v * 4;
      const float4 ww = *reinterpret_cast<const float4*>(w + v * 4);
      const float4 x = residual_plus_delta<kPartialsInFlight>(residual, delta, delta_is_f32, n_partials,
                                                              partial_stride, off);
      ss = __fadd_rn(ss, sumsq4(x));
      *reinterpret_cast<float4*>(residual + off) = x;
      store_bf16x4(xn + off, x.x * ww.x, x.y * ww.y, x.z * ww.z, x.w * ww.w);
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float sg = 0.f;
#pragma unroll
    for (int i = 0; i < kSegThreads / 32; ++i) sg += red[i];
    sumsq[seg * kNormStatsStride + t] = sg;
    BCG_TL(2, 3);
  }
}

// One 128-thread CTA per row (16 of them fit on an SM: the launch is bandwidth-bound, where the
// earlier 1024-thread version reached a third of the HBM rate).  Physical warp pw plays the
// virtual warps pw, pw + 4, ... of the order described above; the summed row is staged in shared
// memory between the statistics pass and the scaling pass.
constexpr int kRowThreads = 128;

__global__ void __launch_bounds__(kRowThreads)
add_rmsnorm_row_kernel(float* __restrict__ residual, const void* __restrict__ delta, int delta_is_f32,
                       int n_partials, long long partial_stride, const float* __restrict__ w,
                       __nv_bfloat16* __restrict__ xn, const int* __restrict__ src_rows, int d,
                       float eps) {
  constexpr int kVirtWarps = kNormSegments * kSegThreads / 32;  // 64
  constexpr int kPhysWarps = kRowThreads / 32;                  // 4
  extern __shared__ float4 row_stage[];                         // [d / 4]
  __shared__ float red[kVirtWarps];
  pdl_launch_dependents();
  pdl_wait();
  const int t = blockIdx.x;
  const long long src = src_rows ? src_rows[t] : t;
  const int nvec = d >> 2;
  const int per = (nvec + kNormSegments - 1) / kNormSegments;
  const int pw = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int vw = pw; vw < kVirtWarps; vw += kPhysWarps) {
    const int seg = vw >> 3;
    const int vt = (vw & 7) * 32 + lane;
    float ss = 0.f;
    if ((vw & 7) * 32 < per) {  // virtual warps past the segment's end contribute exact zeros
      const int v_end = min(nvec, (seg + 1) * per);
#pragma unroll
      for (int i = 0; i < kSegTurns; ++i) {
        const int v = seg * per + vt + i * kSegThreads;
        if (vt + i * kSegThreads < per && v < v_end) {
          const float4 x = residual_plus_delta<4>(residual, delta, delta_is_f32, n_partials,
                                                  partial_stride, src * d + v * 4);
          row_stage[v] = x;
          ss = __fadd_rn(ss, sumsq4(x));
        }
      }
      ss = warp_sum(ss);
    }
    if (lane == 0) red[vw] = ss;
  }
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int sgi = 0; sgi < kNormSegments; ++sgi) {
    float sg = 0.f;
#pragma unroll
    for (int i = 0; i < kSegThreads / 32; ++i) sg += red[sgi * (kSegThreads / 32) + i];
    tot += sg;
  }
  const float inv = rsqrtf(tot / static_cast<float>(d) + eps);
  for (int v = threadIdx.x; v < nvec; v += kRowThreads) {
    const float4 x = row_stage[v];
    const float4 ww = *reinterpret_cast<const float4*>(w + v * 4);
    norm_store(residual, xn, !src_rows, static_cast<long long>(t) * d + v * 4, x, ww, inv);
  }
}

// One warp per (token, head).  head_dim is fixed at 128: lane l owns dims {2l, 2l+1} and
// {64+2l, 64+2l+1}, i.e. both members of each rotate-half pair.
// KV cache layout (one layer): [page][2 (k,v)][n_kv_heads][page_size][128] bf16.
__global__ void __launch_bounds__(128)
qknorm_rope_kvwrite_kernel(const void* __restrict__ qkv, int qkv_is_f32, int n_partials,
                           long long partial_stride, long long ld_qkv,
                           const float* __restrict__ q_norm_w, const float* __restrict__ k_norm_w,
                           const float* __restrict__ cos_sin, const int* __restrict__ positions,
                           const int* __restrict__ slot_map, __nv_bfloat16* __restrict__ q_out,
                           __nv_bfloat16* __restrict__ kv_cache, int T, int n_heads, int n_kv_heads,
                           int page_size, float eps) {
  pdl_launch_dependents();
  if (threadIdx.x == 0) BCG_TL(3, 0);
  const int lane = threadIdx.x & 31;
  const int heads_total = n_heads + 2 * n_kv_heads;
  const long long wid = static_cast<long long>(blockIdx.x) * 4 + (threadIdx.x >> 5);
  const bool live = wid < static_cast<long long>(T) * heads_total;
  const int t = live ? static_cast<int>(wid / heads_total) : 0;
  const int h = live ? static_cast<int>(wid - static_cast<long long>(t) * heads_total) : 0;
  const bool is_q = h < n_heads;
  const bool is_k = !is_q && h < n_heads + n_kv_heads;
  // Everything that does not come from the GEMM in front is fetched BEFORE the dependency wait:
  // position, cache slot (written by the previous step's sampler, which never triggers its
  // dependents early), the RoPE table row and the norm weights.
  const float* nw = is_q ? q_norm_w : (is_k ? k_norm_w : nullptr);
  float2 wl = make_float2(1.f, 1.f), wh = make_float2(1.f, 1.f), c = make_float2(1.f, 1.f),
         sn = make_float2(0.f, 0.f);
  int slot = -1;
  if (live) {
    if (nw) {
      wl = *reinterpret_cast<const float2*>(nw + 2 * lane);
      wh = *reinterpret_cast<const float2*>(nw + 64 + 2 * lane);
    }
    if (is_q || is_k) {
      const int pos = positions[t];
      c = *reinterpret_cast<const float2*>(cos_sin + static_cast<long long>(pos) * 128 + 2 * lane);
      sn = *reinterpret_cast<const float2*>(cos_sin + static_cast<long long>(pos) * 128 + 64 + 2 * lane);
    }
    if (!is_q) slot = slot_map[t];
  }
  pdl_wait();
  if (threadIdx.x == 0) BCG_TL(3, 1);
  if (live) {
    const long long base = static_cast<long long>(t) * ld_qkv + h * 128;
    float lo0, lo1, hi0, hi1;
    if (qkv_is_f32) {
      const float* p = reinterpret_cast<const float*>(qkv) + base;
      // slabs are requested eight at a time, unconditionally (indices past the end re-read the last
      // slab and are ignored), and added in index order
      float2 a = make_float2(0.f, 0.f), b = make_float2(0.f, 0.f);
      if (n_partials == 1) {  // prefill: a single slab
        a = *reinterpret_cast<const float2*>(p + 2 * lane);
        b = *reinterpret_cast<const float2*>(p + 64 + 2 * lane);
      } else
      for (int s0 = 0; s0 < n_partials; s0 += 8) {
        float2 pa[8], pb[8];
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          const long long so = static_cast<long long>(min(s0 + s, n_partials - 1)) * partial_stride;
          pa[s] = *reinterpret_cast<const float2*>(p + so + 2 * lane);
          pb[s] = *reinterpret_cast<const float2*>(p + so + 64 + 2 * lane);
        }
#pragma unroll
        for (int s = 0; s < 8; ++s) {
          if (s0 + s < n_partials) {
            if (s0 + s == 0) {
              a = pa[s]; b = pb[s];
            } else {
              a.x += pa[s].x; a.y += pa[s].y; b.x += pb[s].x; b.y += pb[s].y;
            }
          }
        }
      }
      lo0 = a.x; lo1 = a.y; hi0 = b.x; hi1 = b.y;
    } else {
      const __nv_bfloat16* p = reinterpret_cast<const __nv_bfloat16*>(qkv) + base;
      const uint32_t a = *reinterpret_cast<const uint32_t*>(p + 2 * lane);
      const uint32_t b = *reinterpret_cast<const uint32_t*>(p + 64 + 2 * lane);
      lo0 = bf16lo(a); lo1 = bf16hi(a); hi0 = bf16lo(b); hi1 = bf16hi(b);
    }
    if (is_q || is_k) {
      if (nw) {
        const float ss = warp_sum(lo0 * lo0 + lo1 * lo1 + hi0 * hi0 + hi1 * hi1);
        const float inv = rsqrtf(ss * (1.0f / 128.0f) + eps);
        lo0 = lo0 * inv * wl.x; lo1 = lo1 * inv * wl.y;
        hi0 = hi0 * inv * wh.x; hi1 = hi1 * inv * wh.y;
      }
      const float r_lo0 = lo0 * c.x - hi0 * sn.x;
      const float r_lo1 = lo1 * c.y - hi1 * sn.y;
      const float r_hi0 = hi0 * c.x + lo0 * sn.x;
      const float r_hi1 = hi1 * c.y + lo1 * sn.y;
      lo0 = r_lo0; lo1 = r_lo1; hi0 = r_hi0; hi1 = r_hi1;
    }
    const uint32_t plo = pack_bf16(lo0, lo1), phi = pack_bf16(hi0, hi1);
    if (is_q) {
      __nv_bfloat16* o = q_out + (static_cast<long long>(t) * n_heads + h) * 128;
      *reinterpret_cast<uint32_t*>(o + 2 * lane) = plo;
      *reinterpret_cast<uint32_t*>(o + 64 + 2 * lane) = phi;
    } else {
      const int kvh = is_k ? h - n_heads : h - n_heads - n_kv_heads;
      if (slot >= 0) {
        const int page = slot / page_size, off = slot - page * page_size;
        __nv_bfloat16* o = kv_cache +
            (((static_cast<long long>(page) * 2 + (is_k ? 0 : 1)) * n_kv_heads + kvh) * page_size + off) * 128;
        *reinterpret_cast<uint32_t*>(o + 2 * lane) = plo;
        *reinterpret_cast<uint32_t*>(o + 64 + 2 * lane) = phi;
      }
    }
  }
}

// Prefill-sized launches (one fp32 slab, thousands of tokens): the kernel above starts a 4-warp CTA per
// four (token, head) pairs — 173,460 CTAs of ~1.5 us at 12.4k tokens, each lane with 16 bytes of
// loads in flight: 2.25 TB/s, a third of the HBM rate (ncu, profiles/r02_prof_prefill_glue_raw.csv).
// Here one CTA owns a TOKEN: position, cache slot, RoPE row and norm weights are fetched once per
// warp, warp w takes heads w, w+8, ... and keeps four heads' loads in flight before it touches any.
// Per head the lane mapping and every arithmetic step are those of the kernel above, so a row's
// q / k / v bits do not depend on which of the two kernels produced them (decode uses the other one).
constexpr int kRopeTokWarps = 8;
constexpr int kRopeTokBatch = 4;

__global__ void __launch_bounds__(kRopeTokWarps * 32)
qknorm_rope_kvwrite_token_kernel(const float* __restrict__ qkv, long long ld_qkv,
                                 const float* __restrict__ q_norm_w, const float* __restrict__ k_norm_w,
                                 const float* __restrict__ cos_sin, const int* __restrict__ positions,
                                 const int* __restrict__ slot_map, __nv_bfloat16* __restrict__ q_out,
                                 __nv_bfloat16* __restrict__ kv_cache, int n_heads, int n_kv_heads,
                                 int page_size, float eps) {
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int t = blockIdx.x;
  const int heads_total = n_heads + 2 * n_kv_heads;
  const int pos = positions[t];
  const int slot = slot_map[t];
  const float2 c = *reinterpret_cast<const float2*>(cos_sin + static_cast<long long>(pos) * 128 + 2 * lane);
  const float2 sn = *reinterpret_cast<const float2*>(cos_sin + static_cast<long long>(pos) * 128 + 64 + 2 * lane);
  float2 qwl = make_float2(1.f, 1.f), qwh = qwl, kwl = qwl, kwh = qwl;
  if (q_norm_w) {
    qwl = *reinterpret_cast<const float2*>(q_norm_w + 2 * lane);
    qwh = *reinterpret_cast<const float2*>(q_norm_w + 64 + 2 * lane);
  }
  if (k_norm_w) {
    kwl = *reinterpret_cast<const float2*>(k_norm_w + 2 * lane);
    kwh = *reinterpret_cast<const float2*>(k_norm_w + 64 + 2 * lane);
  }
  pdl_wait();
  const float* row = qkv + static_cast<long long>(t) * ld_qkv;
  for (int h0 = warp; h0 < heads_total; h0 += kRopeTokWarps * kRopeTokBatch) {
    float2 a[kRopeTokBatch], b[kRopeTokBatch];
#pragma unroll
    for (int i = 0; i < kRopeTokBatch; ++i) {      // clamped: all loads of the batch in flight at once
      const int h = min(h0 + i * kRopeTokWarps, heads_total - 1);
      a[i] = *reinterpret_cast<const float2*>(row + h * 128 + 2 * lane);
      b[i] = *reinterpret_cast<const float2*>(row + h * 128 + 64 + 2 * lane);
    }
#pragma unroll
    for (int i = 0; i < kRopeTokBatch; ++i) {
      const int h = h0 + i * kRopeTokWarps;
      if (h >= heads_total) break;
      const bool is_q = h < n_heads;
      const bool is_k = !is_q && h < n_heads + n_kv_heads;
      float lo0 = a[i].x, lo1 = a[i].y, hi0 = b[i].x, hi1 = b[i].y;
      if (is_q || is_k) {
        if (is_q ? (q_norm_w != nullptr) : (k_norm_w != nullptr)) {
          const float2 wl = is_q ? qwl : kwl, wh = is_q ? qwh : kwh;
          const float ss = warp_sum(lo0 * lo0 + lo1 * lo1 + hi0 * hi0 + hi1 * hi1);
          const float inv = rsqrtf(ss * (1.0f / 128.0f) + eps);
          lo0 = lo0 * inv * wl.x; lo1 = lo1 * inv * wl.y;
          hi0 = hi0 * inv * wh.x; hi1 = hi1 * inv * wh.y;
        }
        const float r_lo0 = lo0 * c.x - hi0 * sn.x;
        const float r_lo1 = lo1 * c.y - hi1 * sn.y;
        const float r_hi0 = hi0 * c.x + lo0 * sn.x;
        const float r_hi1 = hi1 * c.y + lo1 * sn.y;
        lo0 = r_lo0; lo1 = r_lo1; hi0 = r_hi0; hi1 = r_hi1;
      }
      const uint32_t plo = pack_bf16(lo0, lo1), phi = pack_bf16(hi0, hi1);
      if (is_q) {
        __nv_bfloat16* o = q_out + (static_cast<long long>(t) * n_heads + h) * 128;
        *reinterpret_cast<uint32_t*>(o + 2 * lane) = plo;
        *reinterpret_cast<uint32_t*>(o + 64 + 2 * lane) = phi;
      } else if (slot >= 0) {
        const int kvh = is_k ? h - n_heads : h - n_heads - n_kv_heads;
        const int page = slot / page_size, off = slot - page * page_size;
        __nv_bfloat16* o = kv_cache +
            (((static_cast<long long>(page) * 2 + (is_k ? 0 : 1)) * n_kv_heads + kvh) * page_size + off) * 128;
        *reinterpret_cast<uint32_t*>(o + 2 * lane) = plo;
        *reinterpret_cast<uint32_t*>(o + 64 + 2 * lane) = phi;
      }
    }
  }
}

// act[t, f] = silu(gate) * up where the GEMM's N axis is interleaved [16 gate | 16 up].
__global__ void __launch_bounds__(256)
swiglu_reduce_kernel(const float* __restrict__ partials, int n_partials, long long partial_stride,
                     __nv_bfloat16* __restrict__ act, long long total_out, int f) {
  pdl_launch_dependents();
  pdl_wait();
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < total_out) {
    const long long t = i / f;
    const int j = static_cast<int>(i - t * f);
    const long long col = static_cast<long long>(j / 16) * 32 + (j % 16);
    const float* p = partials + t * (2LL * f) + col;
    float g = 0.f, u = 0.f;
    for (int s0 = 0; s0 < n_partials; s0 += 8) {  // eight slabs in flight, added in index order
      float pg[8], pu[8];
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        const long long so = static_cast<long long>(min(s0 + s, n_partials - 1)) * partial_stride;
        pg[s] = p[so];
        pu[s] = p[so + 16];
      }
#pragma unroll
      for (int s = 0; s < 8; ++s) {
        if (s0 + s < n_partials) {
          if (s0 + s == 0) {
            g = pg[s]; u = pu[s];
          } else {
            g += pg[s]; u += pu[s];
          }
        }
      }
    }
    act[i] = __float2bfloat16_rn(g * (1.0f / (1.0f + __expf(-g))) * u);
  }
}

// ---------------------------------------------------------------------------------------

int embed_rmsnorm(const int* ids, const void* embed, const float* w, float* residual, void* xn,
                  int T, int d, float eps, cudaStream_t stream) {
  if (!ids || !embed || !w || !residual || !xn || T <= 0 || d <= 0 || (d % 4) ||
      d > kNormThreads * kMaxVecPerThread * 4) {
    set_last_error("bcg_embed_rmsnorm", "bad argument (d must be a multiple of 4 and <= 8192)");
    return BCG_ERR_ARG;
  }
  BCG_CUDA_CHECK(launch_k(embed_rmsnorm_kernel, dim3(T), dim3(kNormThreads), 0, stream, ids,
                          reinterpret_cast<const __nv_bfloat16*>(embed), w, residual,
                          reinterpret_cast<__nv_bfloat16*>(xn), d, eps));
  return BCG_OK;
}

int add_rmsnorm(float* residual, const void* delta, int delta_is_f32, int n_partials,
                long long partial_stride, const float* w, void* xn, const int* src_rows, int T,
                int d, float eps, cudaStream_t stream) {
  if (!residual || !w || !xn || T <= 0 || d <= 0 || (d % 4) ||
      d > kNormSegments * kSegThreads * kSegTurns * 4 || n_partials < 1) {
    set_last_error("bcg_add_rmsnorm", "bad argument (d must be a multiple of 4 and <= 16384)");
    return BCG_ERR_ARG;
  }
  if (T <= kNormClusterMaxRows) {
    BCG_CUDA_CHECK(launch_k(add_rmsnorm_cluster_kernel, dim3(kNormSegments, T), dim3(kSegThreads), 0,
                            stream, residual, delta, delta_is_f32, n_partials, partial_stride, w,
                            reinterpret_cast<__nv_bfloat16*>(xn), src_rows, d, eps));
  } else {
    if (const int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(add_rmsnorm_row_kernel),
                                           kNormSegments * kSegThreads * kSegTurns * 16))
      return rc;
    BCG_CUDA_CHECK(launch_k(add_rmsnorm_row_kernel, dim3(T), dim3(kRowThreads), static_cast<size_t>(d) * 4,
                            stream, residual, delta, delta_is_f32, n_partials, partial_stride, w,
                            reinterpret_cast<__nv_bfloat16*>(xn), src_rows, d, eps));
  }
  return BCG_OK;
}

int add_norm_stats(float* residual, const void* delta, int delta_is_f32, int n_partials,
                   long long partial_stride, const float* w, void* xn, float* sumsq, int T, int d,
                   cudaStream_t stream) {
  if (!residual || !w || !xn || !sumsq || T <= 0 || T > kNormStatsStride || d <= 0 || (d % 4) ||
      d > kNormSegments * kSegThreads * kSegTurns * 4 || n_partials < 1) {
    set_last_error("bcg_add_norm_stats", "bad argument (T <= 128, d a multiple of 4 and <= 16384)");
    return BCG_ERR_ARG;
  }
  BCG_CUDA_CHECK(launch_k(add_norm_stats_kernel, dim3(kNormSegments, T), dim3(kSegThreads), 0, stream,
                          residual, delta, delta_is_f32, n_partials, partial_stride, w,
                          reinterpret_cast<__nv_bfloat16*>(xn), sumsq, d));
  return BCG_OK;
}

int qknorm_rope_kvwrite(const void* qkv, int qkv_is_f32, int n_partials, long long partial_stride,
                        long long ld_qkv, const float* q_norm_w, const float* k_norm_w,
                        const float* cos_sin, const int* positions, const int* slot_map,
                        void* q_out, void* kv_cache, int T, int n_heads, int n_kv_heads,
                        int page_size, float eps, cudaStream_t stream) {
  if (!qkv || !cos_sin || !positions || !slot_map || !q_out || !kv_cache || T <= 0 ||
      n_heads <= 0 || n_kv_heads <= 0 || page_size <= 0 || n_partials < 1) {
    set_last_error("bcg_qknorm_rope_kvwrite", "bad argument");
    return BCG_ERR_ARG;
  }
  if (qkv_is_f32 && n_partials == 1 && T > 128) {   // prefill-sized: one CTA per token
    static const bool off = getenv("BCG_ROPE_TOKEN_KERNEL") && atoi(getenv("BCG_ROPE_TOKEN_KERNEL")) == 0;
    if (!off) {
      BCG_CUDA_CHECK(launch_k(qknorm_rope_kvwrite_token_kernel, dim3(T), dim3(kRopeTokWarps * 32), 0, stream,
                              reinterpret_cast<const float*>(qkv), ld_qkv, q_norm_w, k_norm_w, cos_sin, positions,
                              slot_map, reinterpret_cast<__nv_bfloat16*>(q_out),
                              reinterpret_cast<__nv_bfloat16*>(kv_cache), n_heads, n_kv_heads, page_size, eps));
      return BCG_OK;
    }
  }
  const long long warps = static_cast<long long>(T) * (n_heads + 2 * n_kv_heads);
  const unsigned grid = static_cast<unsigned>((warps + 3) / 4);
  BCG_CUDA_CHECK(launch_k(qknorm_rope_kvwrite_kernel, dim3(grid), dim3(128), 0, stream, qkv, qkv_is_f32,
                          n_partials, partial_stride, ld_qkv, q_norm_w, k_norm_w, cos_sin, positions,
                          slot_map, reinterpret_cast<__nv_bfloat16*>(q_out),
                          reinterpret_cast<__nv_bfloat16*>(kv_cache), T, n_heads, n_kv_heads, page_size,
                          eps));
  return BCG_OK;
}

int swiglu_reduce(const float* partials, int n_partials, long long partial_stride, void* act,
                  int T, int two_f, cudaStream_t stream) {
  if (!partials || !act || T <= 0 || two_f <= 0 || (two_f % 32) || n_partials < 1) {
    set_last_error("bcg_swiglu_reduce", "bad argument (2F must be a multiple of 32)");
    return BCG_ERR_ARG;
  }
  const int f = two_f / 2;
  const long long total = static_cast<long long>(T) * f;
  BCG_CUDA_CHECK(launch_k(swiglu_reduce_kernel, dim3(static_cast<unsigned>((total + 255) / 256)),
                          dim3(256), 0, stream, partials, n_partials, partial_stride,
                          reinterpret_cast<__nv_bfloat16*>(act), total, f));
  return BCG_OK;
}

}  // namespace bcg

BCG_TL_DEFINE_DUMP(bcg_tl_dump_norm)
