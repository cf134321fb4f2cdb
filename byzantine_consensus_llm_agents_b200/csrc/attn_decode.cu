// Batched decode attention over the paged bf16 KV cache:
// one query token per row, GQA (the q-heads that share a KV head share each KV load),
// HBM-bound.  Work is split along the context in FIXED 256-token chunks (16 pages), so a
// row's arithmetic — chunk boundaries and the order partial results are merged in —
// depends only on that row's own length, never on which other rows share the batch.
//
// Phase 1  decode_attn_kernel  grid (max_splits, n_heads / GH, B), 128 threads:
//          each warp streams whole pages with 128-bit non-allocating loads (two tokens per
//          load instruction, 16 loads in flight per lane), online softmax in fp32 with
//          half-warp shuffle reductions, partial (m, l, o[128]) per (row, head, chunk) to
//          scratch.
// Phase 2  decode_attn_merge_kernel  grid (n_heads, B): fixed-order merge of the chunks.
//
// KV cache layout (one layer): [page][2 (k,v)][n_kv_heads][16 tokens][128] bf16 — a
// (page, kv-head) K or V slab is 4 KB contiguous.
#include <mutex>

#include "common.cuh"
#include "kernels.h"

namespace bcg {

constexpr int kSplitTokens = 256;
constexpr int kPage = 16;

__device__ __forceinline__ float dot8(const float* q, const uint4& k) {
  float s = q[0] * bf16lo(k.x);
  s = fmaf(q[1], bf16hi(k.x), s);
  s = fmaf(q[2], bf16lo(k.y), s);
  s = fmaf(q[3], bf16hi(k.y), s);
  s = fmaf(q[4], bf16lo(k.z), s);
  s = fmaf(q[5], bf16hi(k.z), s);
  s = fmaf(q[6], bf16lo(k.w), s);
  s = fmaf(q[7], bf16hi(k.w), s);
  return s;
}

// G = q-heads per KV head, GH = q-heads handled by one CTA (GH divides G).
template <int G, int GH>
__global__ void __launch_bounds__(128)
decode_attn_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ kv_cache,
                   const int* __restrict__ block_table, int max_blocks,
                   const int* __restrict__ seq_lens, float* __restrict__ scratch_o,
                   float* __restrict__ scratch_ml, int n_heads, int n_kv_heads, int max_splits,
                   float scale_log2e) {
  __shared__ float sm_acc[4][GH][128];
  __shared__ float sm_m[4][GH];
  __shared__ float sm_l[4][GH];
  pdl_launch_dependents();
  pdl_wait();
  const int split = blockIdx.x;
  const int head0 = blockIdx.y * GH;
  const int kvh = head0 / G;
  const int b = blockIdx.z;
  const int seq_len = seq_lens[b];
  const int n_splits = (seq_len + kSplitTokens - 1) / kSplitTokens;
  if (split < n_splits) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int half = lane >> 4, dl = lane & 15;
    const int tok0 = split * kSplitTokens;
    const int tok1 = min(seq_len, tok0 + kSplitTokens);

    float qv[GH][8];
#pragma unroll
    for (int g = 0; g < GH; ++g) {
      const uint4 raw = *reinterpret_cast<const uint4*>(
          q + (static_cast<long long>(b) * n_heads + head0 + g) * 128 + 8 * dl);
      qv[g][0] = bf16lo(raw.x) * scale_log2e; qv[g][1] = bf16hi(raw.x) * scale_log2e;
      qv[g][2] = bf16lo(raw.y) * scale_log2e; qv[g][3] = bf16hi(raw.y) * scale_log2e;
      qv[g][4] = bf16lo(raw.z) * scale_log2e; qv[g][5] = bf16hi(raw.z) * scale_log2e;
      qv[g][6] = bf16lo(raw.w) * scale_log2e; qv[g][7] = bf16hi(raw.w) * scale_log2e;
    }
    float m[GH], l[GH], acc[GH][8];
#pragma unroll
    for (int g = 0; g < GH; ++g) {
      m[g] = -INFINITY;
      l[g] = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[g][i] = 0.f;
    }

    for (int pg = tok0 / kPage + warp; pg * kPage < tok1; pg += 4) {
      const int phys = block_table[static_cast<long long>(b) * max_blocks + pg];
      const __nv_bfloat16* kbase =
          kv_cache + ((static_cast<long long>(phys) * 2 * n_kv_heads) + kvh) * (kPage * 128);
      const __nv_bfloat16* vbase = kbase + static_cast<long long>(n_kv_heads) * (kPage * 128);
      uint4 kr[8], vr[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) kr[j] = ldg_nc_v4(kbase + (2 * j + half) * 128 + 8 * dl);
#pragma unroll
      for (int j = 0; j < 8; ++j) vr[j] = ldg_nc_v4(vbase + (2 * j + half) * 128 + 8 * dl);

      float s[8][GH];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const bool valid = pg * kPage + 2 * j + half < tok1;
#pragma unroll
        for (int g = 0; g < GH; ++g) {
          float d = dot8(qv[g], kr[j]);
          d += __shfl_xor_sync(0xffffffffu, d, 8);
          d += __shfl_xor_sync(0xffffffffu, d, 4);
          d += __shfl_xor_sync(0xffffffffu, d, 2);
          d += __shfl_xor_sync(0xffffffffu, d, 1);
          s[j][g] = valid ? d : -INFINITY;
        }
      }
#pragma unroll
      for (int g = 0; g < GH; ++g) {
        float mx = s[0][g];
#pragma unroll
        for (int j = 1; j < 8; ++j) mx = fmaxf(mx, s[j][g]);
        const float m_new = fmaxf(m[g], mx);
        const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
        const float corr = exp2f(m[g] - m_use);  // m == -inf -> 0
        l[g] *= corr;
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[g][i] *= corr;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float p = exp2f(s[j][g] - m_use);  // invalid -> 0
          l[g] += p;
          acc[g][0] = fmaf(p, bf16lo(vr[j].x), acc[g][0]);
          acc[g][1] = fmaf(p, bf16hi(vr[j].x), acc[g][1]);
          acc[g][2] = fmaf(p, bf16lo(vr[j].y), acc[g][2]);
          acc[g][3] = fmaf(p, bf16hi(vr[j].y), acc[g][3]);
          acc[g][4] = fmaf(p, bf16lo(vr[j].z), acc[g][4]);
          acc[g][5] = fmaf(p, bf16hi(vr[j].z), acc[g][5]);
          acc[g][6] = fmaf(p, bf16lo(vr[j].w), acc[g][6]);
          acc[g][7] = fmaf(p, bf16hi(vr[j].w), acc[g][7]);
        }
        m[g] = m_new;
      }
    }

    // merge the two half-warps (disjoint token sets), then the four warps via smem
#pragma unroll
    for (int g = 0; g < GH; ++g) {
      const float m_o = __shfl_xor_sync(0xffffffffu, m[g], 16);
      const float l_o = __shfl_xor_sync(0xffffffffu, l[g], 16);
      const float mm = fmaxf(m[g], m_o);
      const float mu = (mm == -INFINITY) ? 0.f : mm;
      // evaluate in a half-independent order so both halves hold identical values
      const float c_lo = exp2f((half ? m_o : m[g]) - mu);
      const float c_hi = exp2f((half ? m[g] : m_o) - mu);
      const float l_lo = half ? l_o : l[g], l_hi = half ? l[g] : l_o;
      l[g] = l_lo * c_lo + l_hi * c_hi;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float a_o = __shfl_xor_sync(0xffffffffu, acc[g][i], 16);
        const float a_lo = half ? a_o : acc[g][i], a_hi = half ? acc[g][i] : a_o;
        acc[g][i] = a_lo * c_lo + a_hi * c_hi;
      }
      m[g] = mm;
      if (half == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) sm_acc[warp][g][8 * dl + i] = acc[g][i];
        if (dl == 0) {
          sm_m[warp][g] = m[g];
          sm_l[warp][g] = l[g];
        }
      }
    }
    __syncthreads();
    const int tid = threadIdx.x;
#pragma unroll
    for (int g = 0; g < GH; ++g) {
      float mm = fmaxf(fmaxf(sm_m[0][g], sm_m[1][g]), fmaxf(sm_m[2][g], sm_m[3][g]));
      const float mu = (mm == -INFINITY) ? 0.f : mm;
      float o = 0.f, ll = 0.f;
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const float c = exp2f(sm_m[w][g] - mu);
        o += sm_acc[w][g][tid] * c;
        ll += sm_l[w][g] * c;
      }
      const long long idx = (static_cast<long long>(b) * n_heads + head0 + g) * max_splits + split;
      scratch_o[idx * 128 + tid] = o;
      if (tid == 0) {
        scratch_ml[idx * 2] = mm;
        scratch_ml[idx * 2 + 1] = ll;
      }
    }
  }
}

__global__ void __launch_bounds__(128)
decode_attn_merge_kernel(const float* __restrict__ scratch_o, const float* __restrict__ scratch_ml,
                         const int* __restrict__ seq_lens, __nv_bfloat16* __restrict__ out,
                         int n_heads, int max_splits) {
  pdl_launch_dependents();
  pdl_wait();
  const int head = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  const int seq_len = seq_lens[b];
  const int n_splits = (seq_len + kSplitTokens - 1) / kSplitTokens;
  const long long base = (static_cast<long long>(b) * n_heads + head) * max_splits;
  float mm = -INFINITY;
  for (int s = 0; s < n_splits; ++s) mm = fmaxf(mm, scratch_ml[(base + s) * 2]);
  const float mu = (mm == -INFINITY) ? 0.f : mm;
  float o = 0.f, ll = 0.f;
  for (int s = 0; s < n_splits; ++s) {
    const float c = exp2f(scratch_ml[(base + s) * 2] - mu);
    o += scratch_o[(base + s) * 128 + tid] * c;
    ll += scratch_ml[(base + s) * 2 + 1] * c;
  }
  const float r = ll > 0.f ? o / ll : 0.f;
  out[(static_cast<long long>(b) * n_heads + head) * 128 + tid] = __float2bfloat16_rn(r);
}

template <int G, int GH>
static int launch_decode(const void* q, const void* kv_cache, const int* block_table,
                         int max_blocks, const int* seq_lens, float* scratch_o, float* scratch_ml,
                         int B, int n_heads, int n_kv_heads, int max_splits, float scale,
                         cudaStream_t stream) {
  dim3 grid(max_splits, n_heads / GH, B);
  BCG_CUDA_CHECK(launch_k(decode_attn_kernel<G, GH>, grid, dim3(128), 0, stream,
                          reinterpret_cast<const __nv_bfloat16*>(q),
                          reinterpret_cast<const __nv_bfloat16*>(kv_cache), block_table, max_blocks,
                          seq_lens, scratch_o, scratch_ml, n_heads, n_kv_heads, max_splits,
                          scale * 1.4426950408889634f));
  return BCG_OK;
}

int decode_attn_paged(const void* q, const void* kv_cache, const int* block_table,
                      int max_blocks_per_seq, const int* seq_lens, void* out, float* scratch_o,
                      float* scratch_ml, int B, int n_heads, int n_kv_heads, int page_size,
                      int max_splits, float scale, cudaStream_t stream) {
  if (!q || !kv_cache || !block_table || !seq_lens || !out || !scratch_o || !scratch_ml || B <= 0 ||
      n_heads <= 0 || n_kv_heads <= 0 || max_splits <= 0 || (n_heads % n_kv_heads)) {
    set_last_error("bcg_decode_attn_paged", "bad argument");
    return BCG_ERR_ARG;
  }
  if (page_size != kPage) {
    set_last_error("bcg_decode_attn_paged", "page_size must be 16");
    return BCG_ERR_UNSUPPORTED;
  }
  const int G = n_heads / n_kv_heads;
  int rc;
#define BCG_DECODE_CASE(g, gh)                                                                   \
  case g:                                                                                        \
    rc = launch_decode<g, gh>(q, kv_cache, block_table, max_blocks_per_seq, seq_lens, scratch_o, \
                              scratch_ml, B, n_heads, n_kv_heads, max_splits, scale, stream);    \
    break;
  switch (G) {
    BCG_DECODE_CASE(1, 1)
    BCG_DECODE_CASE(2, 2)
    BCG_DECODE_CASE(4, 4)
    BCG_DECODE_CASE(5, 5)
    BCG_DECODE_CASE(6, 3)
    BCG_DECODE_CASE(8, 4)
    default:
      set_last_error("bcg_decode_attn_paged", "unsupported GQA group size (1,2,4,5,6,8)");
      return BCG_ERR_UNSUPPORTED;
  }
#undef BCG_DECODE_CASE
  if (rc != BCG_OK) return rc;
  BCG_CUDA_CHECK(launch_k(decode_attn_merge_kernel, dim3(n_heads, B), dim3(128), 0, stream,
                          static_cast<const float*>(scratch_o), static_cast<const float*>(scratch_ml),
                          seq_lens, reinterpret_cast<__nv_bfloat16*>(out), n_heads, max_splits));
  return BCG_OK;
}

// ---------------------------------------------------------------------------------------
// Allocation-free variant: bcg_decode_attn_paged_static
//
// bcg_decode_attn_paged needs two scratch buffers (the per-chunk partial outputs and their
// softmax statistics) that the caller must allocate.  Some hosts cannot do that at the point
// where they call attention: a serving process that handed its whole device-memory budget to
// weights and KV pages at start-up has nothing left for a late cudaMalloc (and a failed
// allocation in the middle of a game round would abort the round), and a latency-critical
// decode loop must stay away from cudaMalloc / cudaFree because both synchronise the whole
// device.  For those callers the scratch lives inside the library, in the fixed-size device
// array below; the caller still passes q, the KV cache, the block table, seq_lens and out.
//
//   * Limit: B * n_heads * max_splits * 130 floats <= kStaticScratchFloats (4 Mi floats =
//     16 MiB).  Larger problems are refused with BCG_STATUS_TOO_LARGE; use
//     bcg_decode_attn_paged with caller-owned scratch instead.
//   * Cost: every GPU that loads the library keeps these 16 MiB until the process exits.
//   * The first call on each device looks the array's device address up
//     (cudaGetSymbolAddress with that device current) and caches it; make that first call
//     before any stream capture begins.  Later calls only launch kernels.
//   * All calls on a device share the one array, and it keeps whatever the previous call left
//     in it.  Nothing has to be cleared: phase 1 writes every (row, head, chunk) entry that
//     phase 2 reads in the same call, and entries of chunks beyond a row's length are never
//     read.  For the same reason two calls that overlap on the GPU (different streams) race on
//     the array — issue them on one stream, or order them with events.
// ---------------------------------------------------------------------------------------
constexpr long long kStaticScratchFloats = 4ll * 1024 * 1024;
constexpr int kMaxDevices = 64;

// NOTE: in host code this name is only a host-side handle for the symbol, never a device
// address; the device address is obtained once per device below.
static __device__ __align__(256) float g_decode_scratch[kStaticScratchFloats];

static std::mutex g_scratch_mu;
static float* g_scratch_addr[kMaxDevices] = {};

static int static_scratch_on_current_device(float** out) {
  int dev = -1;
  BCG_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_last_error("bcg_decode_attn_paged_static", "device ordinal out of range");
    return BCG_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    BCG_CUDA_CHECK(cudaGetSymbolAddress(&p, g_decode_scratch));
    g_scratch_addr[dev] = static_cast<float*>(p);
  }
  *out = g_scratch_addr[dev];
  return BCG_OK;
}

int decode_attn_paged_static(const void* q, const void* kv_cache, const int* block_table,
                             int max_blocks_per_seq, const int* seq_lens, void* out, int B,
                             int n_heads, int n_kv_heads, int page_size, int max_splits, float scale,
                             cudaStream_t stream) {
  if (B <= 0 || n_heads <= 0 || max_splits <= 0) {
    set_last_error("bcg_decode_attn_paged_static", "bad argument");
    return BCG_ERR_ARG;
  }
  const long long n_o = static_cast<long long>(B) * n_heads * max_splits * 128;
  const long long n_ml = static_cast<long long>(B) * n_heads * max_splits * 2;
  if (n_o + n_ml > kStaticScratchFloats) {
    set_last_error("bcg_decode_attn_paged_static",
                   "B * n_heads * max_splits * 130 floats exceeds the built-in 16 MiB scratch; "
                   "use bcg_decode_attn_paged with caller-owned scratch");
    return BCG_ERR_TOO_LARGE;
  }
  float* scratch = nullptr;
  const int rc = static_scratch_on_current_device(&scratch);
  if (rc != BCG_OK) return rc;
  return decode_attn_paged(q, kv_cache, block_table, max_blocks_per_seq, seq_lens, out, scratch,
                           scratch + n_o, B, n_heads, n_kv_heads, page_size, max_splits, scale,
                           stream);
}

}  // namespace bcg
