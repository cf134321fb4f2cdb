// Batched decode attention over the paged bf16 KV cache:
// one query token per row, GQA (the q-heads that share a KV head share each KV load),
// HBM-bound.  Work is split along the context in FIXED 256-token chunks (16 pages), so a
// row's arithmetic — chunk boundaries and the order partial results are merged in —
// depends only on that row's own length, never on which other rows share the batch.
//
// Phase 1  decode_attn_kernel  grid (n_kv_heads, B, max_splits), 96 threads, 4 CTAs per SM:
//          a producer warp streams the chunk's pages through a 6-stage TMA ring, two consumer
//          warps run QK^T and PV on m16n8k16 bf16 MMAs with an fp32 online softmax, partial
//          (m, l, o[128]) per (row, head, chunk) go to scratch.
// Phase 2  (same kernel) the chunk CTA that finishes last — found with one atomic counter per
//          (row, KV head) that it resets to zero — merges all chunks in chunk order.
//
// KV cache layout (one layer): [page][2 (k,v)][n_kv_heads][16 tokens][128] bf16 — a
// (page, kv-head) K or V slab is 4 KB contiguous.
#include <mutex>

#include "common.cuh"
#include "kernels.h"

namespace bcg {

constexpr int kSplitTokens = 256;
constexpr int kPage = 16;

// Rows 8..15 of every product are zero padding: their A registers are 0, their accumulator
// inputs are 0 and their outputs land in write-only dummies, so they cost no live registers.
__device__ __forceinline__ void mma_top8(float& c0, float& c1, uint32_t a0, uint32_t a2, uint32_t b0,
                                         uint32_t b1) {
  float d2, d3;
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %8, %5, %8}, "
      "{%6, %7}, {%0, %1, %9, %9};"
      : "+f"(c0), "+f"(c1), "=f"(d2), "=f"(d3)
      : "r"(a0), "r"(a2), "r"(b0), "r"(b1), "r"(0u), "f"(0.f));
}

constexpr int kAttnStages = 6;                       // ring of {K page, V page} = 8 KB stages
constexpr int kAttnStageBytes = 2 * kPage * 128 * 2;  // 8192
constexpr int kAttnConsumers = 2;                     // consumer warps; + 1 TMA producer warp
constexpr int kAttnThreads = (kAttnConsumers + 1) * 32;
// Page p of a chunk lives in stage p % kAttnStages and is consumed by warp p % kAttnConsumers.
// kAttnStages is a multiple of kAttnConsumers, so every use of a stage is consumed by the SAME
// warp, in order — which is what makes waiting on a one-bit phase parity safe when a stage is
// reused more than twice within a chunk.
static_assert(kAttnStages % kAttnConsumers == 0, "a ring stage must always be consumed by the same warp");

__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(addr));
  return r;
}

// G = q-heads per KV head (1..8).  One CTA = one (row, KV head, 256-token chunk = 16 pages).
//
// Data movement: the last warp is a TMA producer.  The layer's KV cache is described to TMA as one
// 2-D bf16 tensor [token-rows, 128]; a page's K (or V) slab is 16 consecutive rows and arrives
// as two 16 x 64 boxes in the 128B-swizzled layout.  The producer looks up the chunk's block-
// table entries once and issues every page of the chunk as fast as the 6-stage ring allows, so
// up to 48 KB per CTA (192 KB per SM at 4 CTAs) is in flight without holding a single register
// — the kernel is bandwidth-bound, not latency-bound.
//
// Math: both products run on bf16 tensor-core MMAs (m16n8k16, fp32 accumulate) with the G query
// heads as MMA rows 0..G-1 (rows 8..15 are zero padding: free in an HBM-bound kernel).  Fragments
// are read from shared memory with conflict-free 128-bit loads and NO transposes, because the MMA
// k-index is permuted to match memory:
//   S = Q K^T : lane (g,t) reads the 16-byte chunks c = 2t + (j&1) + 8(j>>1), j = 0..3, of token
//               g's K row and uses them as the B fragments of k-steps 2j, 2j+1; Q's A fragments
//               are loaded in the same dim order (a dot product does not care).
//   O = P V   : lane (g,t) reads chunks g and 8+g of V rows 2t, 2t+1, 8+2t, 9+2t — the tokens
//               whose probabilities it already holds in its S accumulators — and interleaves
//               them pairwise (PRMT) into B fragments; output column g of n-tile j is head dim
//               64(j>>3) + 8g + (j&7).
// FUSED variant (bcg_decode_attn_paged_qkv): the qk-norm / RoPE / KV-write kernel disappears.
// Input is the qkv projection's raw split-K slabs.  Every chunk CTA rebuilds the queries of its
// G heads itself (64 consumer threads, one round trip of slab loads, per-head RMSNorm by warp
// reduction, rotate-half partner = lane ^ 16) and parks them in shared memory as bf16; the CTA of
// the row's LAST chunk also builds the new token's K and V row (producer warp, the arithmetic of
// qknorm_rope_kvwrite_kernel to the bit), writes it to the cache for later steps and patches it
// into the page it has just streamed in — so that page, too, can be requested before the
// dependency wait, and the wait is on the qkv GEMM instead of on a kernel behind it.
struct RopeFusion {
  const float* qkv;        // [n_partials][T, ld_qkv] fp32 split-K slabs: q heads | k heads | v heads
  int n_partials;
  long long partial_stride;
  long long ld_qkv;
  const float* q_norm_w;   // [128] or nullptr (no QK-norm)
  const float* k_norm_w;
  const float* cos_sin;    // [max_pos][128]: cos[0..63] | sin[0..63]
  const int* positions;    // [B]
  const int* slot_map;     // [B], negative = do not write the cache
  __nv_bfloat16* kv_cache;
  float eps;
};

template <int G, bool FUSED>
__global__ void __launch_bounds__(kAttnThreads, 4)
decode_attn_kernel(const __grid_constant__ CUtensorMap kv_map, const __nv_bfloat16* __restrict__ q,
                   const int* __restrict__ block_table, int max_blocks,
                   const int* __restrict__ seq_lens, float* __restrict__ scratch_o,
                   float* __restrict__ scratch_ml, int* __restrict__ counters,
                   __nv_bfloat16* __restrict__ out, int n_heads, int n_kv_heads, int max_splits,
                   float scale_log2e, const RopeFusion rf) {
  extern __shared__ uint8_t attn_smem_raw[];
  uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(attn_smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  __shared__ uint64_t full_bar[kAttnStages];
  __shared__ uint64_t empty_bar[kAttnStages];
  // per-warp partial outputs reuse the ring once every page has been consumed (keeps the CTA at
  // 51 KB of shared memory -> 4 CTAs per SM)
  float (*sm_acc)[G][128] = reinterpret_cast<float (*)[G][128]>(ring);
  __shared__ float sm_m[kAttnConsumers][G];
  __shared__ float sm_l[kAttnConsumers][G];
  __shared__ int sm_last;
  __shared__ uint64_t patch_bar;                      // FUSED: the newest token's K/V row is in place
  __shared__ __align__(16) __nv_bfloat16 sm_q[G][128];  // FUSED: the CTA's queries
  pdl_launch_dependents();
  if (threadIdx.x == 0) BCG_TL(4, 0);

  // chunk index slowest: CTAs are dispatched in linear order, so every row's early (full) chunks
  // become resident first and the late ones — short, or past the row's end — come last
  const int split = blockIdx.z;
  const int kvh = blockIdx.x;
  const int head0 = kvh * G;
  const int b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&kv_map);
    for (int i = 0; i < kAttnStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(&patch_bar, 1);
    fence_barrier_init();
  }
  __syncthreads();
  // ---- Before griddepcontrol.wait (i.e. possibly while the kernels in front of this one are
  // still running) the CTA touches only data that no kernel of the running step writes: the
  // row's length and block table (final before the step's first kernel starts: their writers —
  // host copies and the sampler — never trigger dependents early) and KV pages other than the
  // one that receives the row's newest token (written by qknorm_rope_kvwrite just in front of
  // us).  So the producer fills the ring with old pages right away; only the queries and that
  // last page wait for the predecessor.
  const int g = lane >> 2, t = lane & 3;
  const int seq_len = seq_lens[b];
  const int n_splits = (seq_len + kSplitTokens - 1) / kSplitTokens;
  // A row longer than 256 * max_splits tokens would never see its last chunk arrive: the output
  // would stay unwritten and the row's chunk counter non-zero (fatal for the shared counters of
  // bcg_decode_attn_paged_static).  Fail the launch instead of computing a silent wrong answer.
  if (n_splits > max_splits) __trap();
  const bool live = split < n_splits;
  // A CTA whose chunk lies past the row's end leaves at once, BEFORE the dependency wait: parked at
  // the wait it would hold one of the SM's four CTA slots while live chunks of other rows queue
  // behind it with their KV prefetch unissued (10 rows x 8 KV heads x 8 chunks = 640 CTAs on 592
  // slots, 80 of them idle: 5.61 vs 5.41 ms per step at 1,700 tokens of context).  It reads
  // nothing the kernels in front write and writes nothing at all.
  if (!live) return;  // whole CTA
  const int tok0 = split * kSplitTokens;
  const int tok1 = min(seq_len, tok0 + kSplitTokens);
  const int n_pages = live ? (tok1 - tok0 + kPage - 1) / kPage : 0;
  auto issue_page = [&](int p, int phys) {
    const int stage = p % kAttnStages;
    uint8_t* dst = ring + stage * kAttnStageBytes;
    const int row_k = ((phys * 2) * n_kv_heads + kvh) * kPage;
    const int row_v = ((phys * 2 + 1) * n_kv_heads + kvh) * kPage;
    mbar_arrive_expect_tx(&full_bar[stage], kAttnStageBytes);
    tma_load_2d(dst, &kv_map, &full_bar[stage], 0, row_k, kEvictFirst);
    tma_load_2d(dst + 2048, &kv_map, &full_bar[stage], 64, row_k, kEvictFirst);
    tma_load_2d(dst + 4096, &kv_map, &full_bar[stage], 0, row_v, kEvictFirst);
    tma_load_2d(dst + 6144, &kv_map, &full_bar[stage], 64, row_v, kEvictFirst);
  };
  int my_phys = 0, n_pre = 0;
  if (live && warp == kAttnConsumers) {
    const int my_pg = split * (kSplitTokens / kPage) + lane;
    if (lane < kSplitTokens / kPage && my_pg < max_blocks)
      my_phys = block_table[static_cast<long long>(b) * max_blocks + my_pg];
    // unfused: the row's last page waits for qknorm_rope_kvwrite; fused: it is patched below
    n_pre = min(kAttnStages, n_pages - ((!FUSED && split == n_splits - 1) ? 1 : 0));
    for (int p = 0; p < n_pre; ++p) {
      const int phys = __shfl_sync(0xffffffffu, my_phys, p);
      if (lane == 0) issue_page(p, phys);
    }
  }
  // FUSED: what the query / new-token math needs besides the GEMM's slabs is stable too
  constexpr int kQSlots = (G * 32 + 63) / 64;  // float4 slots of the G x 128 queries per consumer thread
  const bool last_chunk = live && split == n_splits - 1;
  float4 qw[kQSlots], qcos, qsin;              // consumers: q-norm weights per slot, RoPE row
  float2 kwl, kwh, kc, ks;                     // producer: k-norm weights and RoPE row (rope-kernel layout)
  int new_slot = -1;
  if constexpr (FUSED) {
    if (live) {
      const int pos = rf.positions[b];
      const float* cs = rf.cos_sin + static_cast<long long>(pos) * 128;
      if (warp != kAttnConsumers) {
        const int d4 = threadIdx.x & 31;       // slot f = tid + 64 k  ->  head (tid >> 5) + 2 k, float4 d4
        qcos = *reinterpret_cast<const float4*>(cs + (d4 & 15) * 4);
        qsin = *reinterpret_cast<const float4*>(cs + 64 + (d4 & 15) * 4);
#pragma unroll
        for (int k = 0; k < kQSlots; ++k)
          qw[k] = rf.q_norm_w ? *reinterpret_cast<const float4*>(rf.q_norm_w + d4 * 4)
                              : make_float4(1.f, 1.f, 1.f, 1.f);
      } else if (last_chunk) {
        new_slot = rf.slot_map[b];
        kc = *reinterpret_cast<const float2*>(cs + 2 * lane);
        ks = *reinterpret_cast<const float2*>(cs + 64 + 2 * lane);
        kwl = rf.k_norm_w ? *reinterpret_cast<const float2*>(rf.k_norm_w + 2 * lane) : make_float2(1.f, 1.f);
        kwh = rf.k_norm_w ? *reinterpret_cast<const float2*>(rf.k_norm_w + 64 + 2 * lane) : make_float2(1.f, 1.f);
      }
    }
  }
  pdl_wait();
  if (threadIdx.x == 0) BCG_TL(4, 1);
  uint32_t qa[8][2];
  if constexpr (FUSED) {
    if (warp != kAttnConsumers) {
      // ---- queries: sum the slabs (index order), per-head RMSNorm, rotate-half RoPE, bf16 -> sm_q
      const int d4 = threadIdx.x & 31;
      const float* qrow = rf.qkv + static_cast<long long>(b) * rf.ld_qkv + static_cast<long long>(head0) * 128;
      float4 x[kQSlots];
#pragma unroll
      for (int k = 0; k < kQSlots; ++k) x[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int s0 = 0; s0 < rf.n_partials; s0 += 6) {
        float4 pp[kQSlots][6];
#pragma unroll
        for (int k = 0; k < kQSlots; ++k) {
          const int hd = min((threadIdx.x >> 5) + 2 * k, G - 1);  // clamp: surplus slots re-read, unused
#pragma unroll
          for (int sp = 0; sp < 6; ++sp)
            pp[k][sp] = *reinterpret_cast<const float4*>(
                qrow + static_cast<long long>(min(s0 + sp, rf.n_partials - 1)) * rf.partial_stride + hd * 128 + d4 * 4);
        }
#pragma unroll
        for (int k = 0; k < kQSlots; ++k) {
#pragma unroll
          for (int sp = 0; sp < 6; ++sp) {
            if (s0 + sp < rf.n_partials) {
              if (s0 + sp == 0) {
                x[k] = pp[k][sp];
              } else {
                x[k].x += pp[k][sp].x; x[k].y += pp[k][sp].y; x[k].z += pp[k][sp].z; x[k].w += pp[k][sp].w;
              }
            }
          }
        }
      }
#pragma unroll
      for (int k = 0; k < kQSlots; ++k) {
        const int hd = (threadIdx.x >> 5) + 2 * k;  // warp-uniform
        if (hd < G) {
          float4 v = x[k];
          if (rf.q_norm_w) {
            const float ss = warp_sum(v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w);
            const float inv = rsqrtf(ss * (1.0f / 128.0f) + rf.eps);
            v.x = v.x * inv * qw[k].x; v.y = v.y * inv * qw[k].y;
            v.z = v.z * inv * qw[k].z; v.w = v.w * inv * qw[k].w;
          }
          // rotate-half partner (dim +-64) sits 16 lanes away
          float4 o4;
          o4.x = __shfl_xor_sync(0xffffffffu, v.x, 16);
          o4.y = __shfl_xor_sync(0xffffffffu, v.y, 16);
          o4.z = __shfl_xor_sync(0xffffffffu, v.z, 16);
          o4.w = __shfl_xor_sync(0xffffffffu, v.w, 16);
          float4 r;
          if (d4 < 16) {  // low half: lo*cos - hi*sin
            r.x = v.x * qcos.x - o4.x * qsin.x; r.y = v.y * qcos.y - o4.y * qsin.y;
            r.z = v.z * qcos.z - o4.z * qsin.z; r.w = v.w * qcos.w - o4.w * qsin.w;
          } else {        // high half: hi*cos + lo*sin
            r.x = v.x * qcos.x + o4.x * qsin.x; r.y = v.y * qcos.y + o4.y * qsin.y;
            r.z = v.z * qcos.z + o4.z * qsin.z; r.w = v.w * qcos.w + o4.w * qsin.w;
          }
          uint2 pk;
          pk.x = pack_bf16(r.x, r.y);
          pk.y = pack_bf16(r.z, r.w);
          *reinterpret_cast<uint2*>(&sm_q[hd][d4 * 4]) = pk;
        }
      }
      asm volatile("bar.sync 1, %0;" ::"n"(kAttnConsumers * 32) : "memory");
    }
  }
  if (warp != kAttnConsumers) {
    // Q as A fragments: [k-step][a0 (k 2t,2t+1), a2 (k 8+2t,9+2t)]; rows 8..15 (a1, a3) are zero
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint4 raw = make_uint4(0u, 0u, 0u, 0u);
      if (g < G) {
        const int c = 2 * t + (j & 1) + 8 * (j >> 1);
        if constexpr (FUSED) {
          raw = *reinterpret_cast<const uint4*>(&sm_q[g][8 * c]);
        } else {
          raw = *reinterpret_cast<const uint4*>(q + (static_cast<long long>(b) * n_heads + head0 + g) * 128 + 8 * c);
        }
      }
      qa[2 * j][0] = raw.x; qa[2 * j][1] = raw.y;
      qa[2 * j + 1][0] = raw.z; qa[2 * j + 1][1] = raw.w;
    }
  }

  if (warp == kAttnConsumers) {
    // ------------------------------ TMA producer ------------------------------
    uint32_t k_lo = 0, k_hi = 0, v_lo = 0, v_hi = 0;
    if constexpr (FUSED) {
      if (last_chunk) {
        // the new token's K (norm + RoPE) and V row of this KV head — qknorm_rope_kvwrite_kernel's
        // arithmetic and lane layout (dims 2l, 2l+1, 64+2l, 64+2l+1) to the bit
        const float* krow = rf.qkv + static_cast<long long>(b) * rf.ld_qkv +
                            static_cast<long long>(n_heads + kvh) * 128;
        const float* vrow = krow + static_cast<long long>(n_kv_heads) * 128;
        float2 ka = make_float2(0.f, 0.f), kb = ka, va = ka, vb = ka;
        for (int s0 = 0; s0 < rf.n_partials; s0 += 8) {
          float2 pka[8], pkb[8], pva[8], pvb[8];
#pragma unroll
          for (int sp = 0; sp < 8; ++sp) {
            const long long so = static_cast<long long>(min(s0 + sp, rf.n_partials - 1)) * rf.partial_stride;
            pka[sp] = *reinterpret_cast<const float2*>(krow + so + 2 * lane);
            pkb[sp] = *reinterpret_cast<const float2*>(krow + so + 64 + 2 * lane);
            pva[sp] = *reinterpret_cast<const float2*>(vrow + so + 2 * lane);
            pvb[sp] = *reinterpret_cast<const float2*>(vrow + so + 64 + 2 * lane);
          }
#pragma unroll
          for (int sp = 0; sp < 8; ++sp) {
            if (s0 + sp < rf.n_partials) {
              if (s0 + sp == 0) {
                ka = pka[sp]; kb = pkb[sp]; va = pva[sp]; vb = pvb[sp];
              } else {
                ka.x += pka[sp].x; ka.y += pka[sp].y; kb.x += pkb[sp].x; kb.y += pkb[sp].y;
                va.x += pva[sp].x; va.y += pva[sp].y; vb.x += pvb[sp].x; vb.y += pvb[sp].y;
              }
            }
          }
        }
        float lo0 = ka.x, lo1 = ka.y, hi0 = kb.x, hi1 = kb.y;
        if (rf.k_norm_w) {
          const float ss = warp_sum(lo0 * lo0 + lo1 * lo1 + hi0 * hi0 + hi1 * hi1);
          const float inv = rsqrtf(ss * (1.0f / 128.0f) + rf.eps);
          lo0 = lo0 * inv * kwl.x; lo1 = lo1 * inv * kwl.y;
          hi0 = hi0 * inv * kwh.x; hi1 = hi1 * inv * kwh.y;
        }
        const float r_lo0 = lo0 * kc.x - hi0 * ks.x;
        const float r_lo1 = lo1 * kc.y - hi1 * ks.y;
        const float r_hi0 = hi0 * kc.x + lo0 * ks.x;
        const float r_hi1 = hi1 * kc.y + lo1 * ks.y;
        k_lo = pack_bf16(r_lo0, r_lo1); k_hi = pack_bf16(r_hi0, r_hi1);
        v_lo = pack_bf16(va.x, va.y); v_hi = pack_bf16(vb.x, vb.y);
        if (new_slot >= 0) {  // for the steps to come
          const int page = new_slot / kPage, off = new_slot - page * kPage;
          __nv_bfloat16* kd = rf.kv_cache + ((static_cast<long long>(page) * 2 * n_kv_heads + kvh) * kPage + off) * 128;
          __nv_bfloat16* vd = kd + static_cast<long long>(n_kv_heads) * kPage * 128;
          *reinterpret_cast<uint32_t*>(kd + 2 * lane) = k_lo;
          *reinterpret_cast<uint32_t*>(kd + 64 + 2 * lane) = k_hi;
          *reinterpret_cast<uint32_t*>(vd + 2 * lane) = v_lo;
          *reinterpret_cast<uint32_t*>(vd + 64 + 2 * lane) = v_hi;
        }
      }
    }
    for (int p = n_pre; p < n_pages; ++p) {
      const int phys = __shfl_sync(0xffffffffu, my_phys, p);
      if (lane == 0) {
        if (p >= kAttnStages) mbar_wait(&empty_bar[p % kAttnStages], ((p / kAttnStages) - 1) & 1);
        issue_page(p, phys);
      }
    }
    if constexpr (FUSED) {
      if (last_chunk) {
        // the row's last page has landed (whatever the cache held in the new token's slot): put
        // the fresh K/V row into its swizzled place, then let that page's consumer go
        const int p_last = n_pages - 1;
        const int stage = p_last % kAttnStages;
        mbar_wait(&full_bar[stage], (p_last / kAttnStages) & 1);
        const int r = (seq_len - 1) & (kPage - 1);
        uint8_t* sk = ring + stage * kAttnStageBytes;
        const int o_lo = r * 128 + (((lane >> 2) ^ (r & 7)) << 4) + (lane & 3) * 4;
        *reinterpret_cast<uint32_t*>(sk + o_lo) = k_lo;
        *reinterpret_cast<uint32_t*>(sk + 2048 + o_lo) = k_hi;
        *reinterpret_cast<uint32_t*>(sk + 4096 + o_lo) = v_lo;
        *reinterpret_cast<uint32_t*>(sk + 6144 + o_lo) = v_hi;
        __syncwarp();
        if (lane == 0) mbar_arrive(&patch_bar);
      }
    }
  } else {
    // ------------------------------ consumers ---------------------------------
    float o[16][2];
#pragma unroll
    for (int j = 0; j < 16; ++j) o[j][0] = o[j][1] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;  // row g (replicated over the quad; l is a quad-partial)
    const int vt[4] = {2 * t, 2 * t + 1, 8 + 2 * t, 9 + 2 * t};
    const uint32_t ring_u32 = smem_u32(ring);

    for (int p = warp; p < n_pages; p += kAttnConsumers) {
      const int stage = p % kAttnStages;
      mbar_wait(&full_bar[stage], (p / kAttnStages) & 1);
      if constexpr (FUSED) {
        if (last_chunk && p == n_pages - 1) mbar_wait(&patch_bar, 0);
      }
      const uint32_t sk = ring_u32 + stage * kAttnStageBytes;
      const uint32_t sv = sk + 4096;
      // ---- S = Q K^T
      float s[2][2];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        s[nt][0] = s[nt][1] = 0.f;
        const int row = nt * 8 + g;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int cc = 2 * t + (j & 1);  // chunk inside the 64-column box j>>1
          const uint4 kf = lds_v4(sk + (j >> 1) * 2048 + row * 128 + ((cc ^ (row & 7)) << 4));
          mma_top8(s[nt][0], s[nt][1], qa[2 * j][0], qa[2 * j][1], kf.x, kf.y);
          mma_top8(s[nt][0], s[nt][1], qa[2 * j + 1][0], qa[2 * j + 1][1], kf.z, kf.w);
        }
      }
      // ---- online softmax of row g over this page's 16 tokens (this lane: tokens vt[0..3])
      const int base_tok = tok0 + p * kPage;
      // only a row's last page can hold tokens past its length (warp-uniform test)
      const bool ragged = base_tok + kPage > tok1;
      float sv4[4];
      bool ok[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        ok[i] = !ragged || base_tok + vt[i] < tok1;
        sv4[i] = ok[i] ? s[i >> 1][i & 1] * scale_log2e : -INFINITY;
      }
      float mx = fmaxf(fmaxf(sv4[0], sv4[1]), fmaxf(sv4[2], sv4[3]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float m_new = fmaxf(m_run, mx);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float corr = exp2f(m_run - m_use);
      m_run = m_new;
      float pr[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) pr[i] = exp2f(sv4[i] - m_use);
      l_run = l_run * corr + (pr[0] + pr[1]) + (pr[2] + pr[3]);
      const uint32_t pa0 = pack_bf16(pr[0], pr[1]);
      const uint32_t pa2 = pack_bf16(pr[2], pr[3]);
      // ---- O = O * corr + P V   (corr is exactly 1 whenever no row's running maximum moved:
      // skipping the multiply then changes nothing, and most pages after the first few qualify)
      const bool rescale = __any_sync(0xffffffffu, corr != 1.f);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint4 vf[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          vf[i] = lds_v4(sv + h * 2048 + vt[i] * 128 + ((g ^ (vt[i] & 7)) << 4));
        }
        if (ragged) {
          // never multiply stale cache contents (possibly non-finite) by p = 0
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (!ok[i]) vf[i] = make_uint4(0u, 0u, 0u, 0u);
        }
        if (rescale) {
#pragma unroll
          for (int jj = 0; jj < 8; ++jj) {
            o[h * 8 + jj][0] *= corr;
            o[h * 8 + jj][1] *= corr;
          }
        }
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
          const int j = h * 8 + jj;
          const int w = jj >> 1;
          const uint32_t sel = (jj & 1) ? 0x7632u : 0x5410u;
          const uint32_t* v0 = reinterpret_cast<const uint32_t*>(&vf[0]);
          const uint32_t* v1 = reinterpret_cast<const uint32_t*>(&vf[1]);
          const uint32_t* v2 = reinterpret_cast<const uint32_t*>(&vf[2]);
          const uint32_t* v3 = reinterpret_cast<const uint32_t*>(&vf[3]);
          mma_top8(o[j][0], o[j][1], pa0, pa2, __byte_perm(v0[w], v1[w], sel),
                   __byte_perm(v2[w], v3[w], sel));
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[stage]);
    }
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 1);
    l_run += __shfl_xor_sync(0xffffffffu, l_run, 2);
    asm volatile("bar.sync 1, %0;" ::"n"(kAttnConsumers * 32) : "memory");  // consumers are done with the ring
    if (g < G) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int dim = 64 * (j >> 3) + (j & 7);
        sm_acc[warp][g][dim + 8 * (2 * t)] = o[j][0];
        sm_acc[warp][g][dim + 8 * (2 * t + 1)] = o[j][1];
      }
      if (t == 0) {
        sm_m[warp][g] = m_run;
        sm_l[warp][g] = l_run;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) BCG_TL(4, 2);
  // ---- the CTA's two warps -> one (m, l, o[128]) per head for this chunk.  A thread owns
  // 4 consecutive dims of one head ("quad"): G * 32 quads over 96 threads.
  const int tid = threadIdx.x;
  constexpr int kQuads = G * 32;
  const bool single = (n_splits == 1);  // short row: this chunk IS the result, no scratch round trip
  for (int qd = tid; qd < kQuads; qd += kAttnThreads) {
    const int gg = qd >> 5, d = (qd & 31) * 4;
    float mm = -INFINITY;
#pragma unroll
    for (int w = 0; w < kAttnConsumers; ++w) mm = fmaxf(mm, sm_m[w][gg]);
    const float mu = (mm == -INFINITY) ? 0.f : mm;
    float4 oo = make_float4(0.f, 0.f, 0.f, 0.f);
    float ll = 0.f;
#pragma unroll
    for (int w = 0; w < kAttnConsumers; ++w) {
      const float c = exp2f(sm_m[w][gg] - mu);  // a warp without pages has m = -inf -> weight 0
      const float4 a = *reinterpret_cast<const float4*>(&sm_acc[w][gg][d]);
      oo.x += a.x * c; oo.y += a.y * c; oo.z += a.z * c; oo.w += a.w * c;
      ll += sm_l[w][gg] * c;
    }
    if (single) {
      // identical to merging one chunk: its weight is exp2(m - m) = 1
      __nv_bfloat16* dst = out + (static_cast<long long>(b) * n_heads + head0 + gg) * 128 + d;
      const bool nz = ll > 0.f;
      uint2 pk;
      pk.x = pack_bf16(nz ? oo.x / ll : 0.f, nz ? oo.y / ll : 0.f);
      pk.y = pack_bf16(nz ? oo.z / ll : 0.f, nz ? oo.w / ll : 0.f);
      *reinterpret_cast<uint2*>(dst) = pk;
    } else {
      const long long idx = (static_cast<long long>(b) * n_heads + head0 + gg) * max_splits + split;
      *reinterpret_cast<float4*>(&scratch_o[idx * 128 + d]) = oo;
      if (d == 0) *reinterpret_cast<float2*>(&scratch_ml[idx * 2]) = make_float2(mm, ll);
    }
  }
  if (single) {
    if (threadIdx.x == 0) BCG_TL(4, 3);
    return;
  }
  // ---- the chunk that finishes last merges all chunks of this (row, KV head), always in chunk
  // order, so the result does not depend on which CTA happens to be last.  Release/acquire runs
  // through one thread: the CTA barrier orders every thread's partial-result stores before
  // thread 0's fence + counter increment, and thread 0's fence after the increment orders the
  // other chunks' stores before the loads that follow the second barrier (which read L2).
  __syncthreads();
  if (tid == 0) {
    fence_acq_rel_gpu();
    const int prev = atomicAdd(&counters[b * n_kv_heads + kvh], 1);
    fence_acq_rel_gpu();
    sm_last = (prev == n_splits - 1);
  }
  __syncthreads();
  if (!sm_last) {
    if (threadIdx.x == 0) BCG_TL(4, 3);
    return;
  }
  const long long base0 = (static_cast<long long>(b) * n_heads + head0) * max_splits;
  if (n_splits <= 8) {
    // every partial of a quad is requested before any is used: one L2 round trip per quad
    for (int qd = tid; qd < kQuads; qd += kAttnThreads) {
      const int gg = qd >> 5, d = (qd & 31) * 4;
      const long long base = base0 + static_cast<long long>(gg) * max_splits;
      float2 ml[8];
      float4 ov[8];
      // unconditional loads (chunks past the end re-read the last one and are ignored below), so
      // that no load waits behind a predicate or an add
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int kk = min(k, n_splits - 1);
        ml[k] = __ldcg(reinterpret_cast<const float2*>(&scratch_ml[(base + kk) * 2]));
        ov[k] = __ldcg(reinterpret_cast<const float4*>(&scratch_o[(base + kk) * 128 + d]));
      }
      float mm = -INFINITY;
#pragma unroll
      for (int k = 0; k < 8; ++k) mm = fmaxf(mm, ml[k].x);  // duplicates of the last chunk do not change a max
      const float mu = (mm == -INFINITY) ? 0.f : mm;
      float4 oo = make_float4(0.f, 0.f, 0.f, 0.f);
      float ll = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (k < n_splits) {
          const float c = exp2f(ml[k].x - mu);
          oo.x += ov[k].x * c; oo.y += ov[k].y * c; oo.z += ov[k].z * c; oo.w += ov[k].w * c;
          ll += ml[k].y * c;
        }
      }
      const bool nz = ll > 0.f;
      uint2 pk;
      pk.x = pack_bf16(nz ? oo.x / ll : 0.f, nz ? oo.y / ll : 0.f);
      pk.y = pack_bf16(nz ? oo.z / ll : 0.f, nz ? oo.w / ll : 0.f);
      *reinterpret_cast<uint2*>(out + (static_cast<long long>(b) * n_heads + head0 + gg) * 128 + d) = pk;
    }
  } else {
    // long rows (> 2048 tokens): same arithmetic, chunks fetched eight at a time
    for (int qd = tid; qd < kQuads; qd += kAttnThreads) {
      const int gg = qd >> 5, d = (qd & 31) * 4;
      const long long base = base0 + static_cast<long long>(gg) * max_splits;
      float mm = -INFINITY;
      for (int s0 = 0; s0 < n_splits; s0 += 8) {
        float mv[8];
#pragma unroll
        for (int k = 0; k < 8; ++k)
          mv[k] = (s0 + k < n_splits) ? __ldcg(&scratch_ml[(base + s0 + k) * 2]) : -INFINITY;
#pragma unroll
        for (int k = 0; k < 8; ++k) mm = fmaxf(mm, mv[k]);
      }
      const float mu = (mm == -INFINITY) ? 0.f : mm;
      float4 oo = make_float4(0.f, 0.f, 0.f, 0.f);
      float ll = 0.f;
      for (int s0 = 0; s0 < n_splits; s0 += 8) {
        float2 ml[8];
        float4 ov[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const bool in = s0 + k < n_splits;
          ml[k] = in ? __ldcg(reinterpret_cast<const float2*>(&scratch_ml[(base + s0 + k) * 2]))
                     : make_float2(-INFINITY, 0.f);
          ov[k] = in ? __ldcg(reinterpret_cast<const float4*>(&scratch_o[(base + s0 + k) * 128 + d]))
                     : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (s0 + k < n_splits) {
            const float c = exp2f(ml[k].x - mu);
            oo.x += ov[k].x * c; oo.y += ov[k].y * c; oo.z += ov[k].z * c; oo.w += ov[k].w * c;
            ll += ml[k].y * c;
          }
        }
      }
      const bool nz = ll > 0.f;
      uint2 pk;
      pk.x = pack_bf16(nz ? oo.x / ll : 0.f, nz ? oo.y / ll : 0.f);
      pk.y = pack_bf16(nz ? oo.z / ll : 0.f, nz ? oo.w / ll : 0.f);
      *reinterpret_cast<uint2*>(out + (static_cast<long long>(b) * n_heads + head0 + gg) * 128 + d) = pk;
    }
  }
  if (tid == 0) counters[b * n_kv_heads + kvh] = 0;  // leave the counters zero for the next call
  if (tid == 0) BCG_TL(4, 5);
}

template <int G, bool FUSED>
static int launch_decode(const void* q, const void* kv_cache, const int* block_table,
                         int max_blocks, const int* seq_lens, float* scratch_o, float* scratch_ml,
                         int* counters, void* out, int B, int n_heads, int n_kv_heads,
                         int max_splits, float scale, const RopeFusion& rf, cudaStream_t stream) {
  constexpr int SMEM = kAttnStages * kAttnStageBytes + 1024;
  if (const int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(decode_attn_kernel<G, FUSED>), SMEM))
    return rc;
  // The layer's cache as a [token-rows, 128] bf16 tensor; the row count only bounds TMA's
  // out-of-range check, the block table never points past the real allocation.
  CUtensorMap kv_map;
  const int rc = make_tmap_bf16_sw128(&kv_map, kv_cache, 1ll << 30, 128, 128, kPage);
  if (rc != BCG_OK) return rc;
  dim3 grid(n_kv_heads, B, max_splits);
  BCG_CUDA_CHECK(launch_k(decode_attn_kernel<G, FUSED>, grid, dim3(kAttnThreads), SMEM, stream, kv_map,
                          reinterpret_cast<const __nv_bfloat16*>(q), block_table, max_blocks, seq_lens,
                          scratch_o, scratch_ml, counters, reinterpret_cast<__nv_bfloat16*>(out),
                          n_heads, n_kv_heads, max_splits, scale * 1.4426950408889634f, rf));
  return BCG_OK;
}

template <bool FUSED>
static int dispatch_decode(const void* q, const void* kv_cache, const int* block_table, int max_blocks,
                           const int* seq_lens, float* scratch_o, float* scratch_ml, int* counters,
                           void* out, int B, int n_heads, int n_kv_heads, int max_splits, float scale,
                           const RopeFusion& rf, cudaStream_t stream, const char* who) {
  switch (n_heads / n_kv_heads) {
#define BCG_DECODE_CASE(g)                                                                          \
  case g:                                                                                           \
    return launch_decode<g, FUSED>(q, kv_cache, block_table, max_blocks, seq_lens, scratch_o,      \
                                   scratch_ml, counters, out, B, n_heads, n_kv_heads, max_splits,   \
                                   scale, rf, stream);
    BCG_DECODE_CASE(1)
    BCG_DECODE_CASE(2)
    BCG_DECODE_CASE(3)
    BCG_DECODE_CASE(4)
    BCG_DECODE_CASE(5)
    BCG_DECODE_CASE(6)
    BCG_DECODE_CASE(7)
    BCG_DECODE_CASE(8)
#undef BCG_DECODE_CASE
    default:
      set_last_error(who, "unsupported GQA group size (1..8)");
      return BCG_ERR_UNSUPPORTED;
  }
}

// `counters`: B * n_kv_heads ints, zero on entry, zero again on exit.
int decode_attn_paged(const void* q, const void* kv_cache, const int* block_table,
                      int max_blocks_per_seq, const int* seq_lens, void* out, float* scratch_o,
                      float* scratch_ml, int* counters, int B, int n_heads, int n_kv_heads,
                      int page_size, int max_splits, float scale, cudaStream_t stream) {
  if (!q || !kv_cache || !block_table || !seq_lens || !out || !scratch_o || !scratch_ml || !counters ||
      B <= 0 ||
      n_heads <= 0 || n_kv_heads <= 0 || max_splits <= 0 || (n_heads % n_kv_heads)) {
    set_last_error("bcg_decode_attn_paged", "bad argument");
    return BCG_ERR_ARG;
  }
  if (page_size != kPage) {
    set_last_error("bcg_decode_attn_paged", "page_size must be 16");
    return BCG_ERR_UNSUPPORTED;
  }
  if ((reinterpret_cast<uintptr_t>(scratch_o) & 15) || (reinterpret_cast<uintptr_t>(scratch_ml) & 7) ||
      (reinterpret_cast<uintptr_t>(out) & 7)) {
    set_last_error("bcg_decode_attn_paged", "scratch_o must be 16-byte, scratch_ml and out 8-byte aligned");
    return BCG_ERR_ARG;
  }
  return dispatch_decode<false>(q, kv_cache, block_table, max_blocks_per_seq, seq_lens, scratch_o,
                                scratch_ml, counters, out, B, n_heads, n_kv_heads, max_splits, scale,
                                RopeFusion{}, stream, "bcg_decode_attn_paged");
}

// Decode attention straight from the qkv projection's split-K slabs (no qknorm_rope_kvwrite launch):
// see the FUSED notes above decode_attn_kernel.  kv_cache is read AND written (the new token's row).
int decode_attn_paged_qkv(const float* qkv, int n_partials, long long partial_stride, long long ld_qkv,
                          const float* q_norm_w, const float* k_norm_w, const float* cos_sin,
                          const int* positions, const int* slot_map, float eps, void* kv_cache,
                          const int* block_table, int max_blocks_per_seq, const int* seq_lens, void* out,
                          float* scratch_o, float* scratch_ml, int* counters, int B, int n_heads,
                          int n_kv_heads, int page_size, int max_splits, float scale,
                          cudaStream_t stream) {
  if (!qkv || n_partials < 1 || !cos_sin || !positions || !slot_map || !kv_cache || !block_table ||
      !seq_lens || !out || !scratch_o || !scratch_ml || !counters || B <= 0 || n_heads <= 0 ||
      n_kv_heads <= 0 || max_splits <= 0 || (n_heads % n_kv_heads) || (ld_qkv % 4) ||
      (partial_stride % 4) || (reinterpret_cast<uintptr_t>(qkv) & 15)) {
    set_last_error("bcg_decode_attn_paged_qkv", "bad argument");
    return BCG_ERR_ARG;
  }
  if (page_size != kPage) {
    set_last_error("bcg_decode_attn_paged_qkv", "page_size must be 16");
    return BCG_ERR_UNSUPPORTED;
  }
  if ((reinterpret_cast<uintptr_t>(scratch_o) & 15) || (reinterpret_cast<uintptr_t>(scratch_ml) & 7) ||
      (reinterpret_cast<uintptr_t>(out) & 7)) {
    set_last_error("bcg_decode_attn_paged_qkv", "scratch_o must be 16-byte, scratch_ml and out 8-byte aligned");
    return BCG_ERR_ARG;
  }
  RopeFusion rf;
  rf.qkv = qkv;
  rf.n_partials = n_partials;
  rf.partial_stride = partial_stride;
  rf.ld_qkv = ld_qkv;
  rf.q_norm_w = q_norm_w;
  rf.k_norm_w = k_norm_w;
  rf.cos_sin = cos_sin;
  rf.positions = positions;
  rf.slot_map = slot_map;
  rf.kv_cache = reinterpret_cast<__nv_bfloat16*>(kv_cache);
  rf.eps = eps;
  return dispatch_decode<true>(nullptr, kv_cache, block_table, max_blocks_per_seq, seq_lens, scratch_o,
                               scratch_ml, counters, out, B, n_heads, n_kv_heads, max_splits, scale, rf,
                               stream, "bcg_decode_attn_paged_qkv");
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
//     16 MiB) and B * n_kv_heads <= 4096.  Larger problems are refused with BCG_STATUS_TOO_LARGE; use
//     bcg_decode_attn_paged with caller-owned scratch instead.
//   * Cost: every GPU that loads the library keeps these 16 MiB until the process exits.
//   * The first call on each device looks the array's device address up
//     (cudaGetSymbolAddress with that device current) and caches it; make that first call
//     before any stream capture begins.  Later calls only launch kernels.
//   * All calls on a device share the one array, and it keeps whatever the previous call left
//     in it.  The partial results need no clearing: every (row, head, chunk) entry is written
//     before it is read in the same call, and entries of chunks beyond a row's length are never
//     read.  The per-(row, KV head) chunk counters DO have to start at zero; they live in a fixed
//     region of their own (so partial results of another shape never land on them), start out
//     zero like all static device memory, and each call puts every counter it used back to zero.
//     For the same reason two calls that overlap on the GPU (different streams) race on the
//     array — issue them on one stream, or order them with events.
// ---------------------------------------------------------------------------------------
constexpr long long kStaticScratchFloats = 4ll * 1024 * 1024;
constexpr int kStaticCounters = 4096;
constexpr int kMaxDevices = 64;

// NOTE: in host code this name is only a host-side handle for the symbol, never a device
// address; the device address is obtained once per device below.
static __device__ __align__(256) float g_decode_scratch[kStaticScratchFloats + kStaticCounters];

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
  if (n_o + n_ml > kStaticScratchFloats || static_cast<long long>(B) * n_kv_heads > kStaticCounters) {
    set_last_error("bcg_decode_attn_paged_static",
                   "B * n_heads * max_splits * 130 floats exceeds the built-in 16 MiB scratch; "
                   "use bcg_decode_attn_paged with caller-owned scratch");
    return BCG_ERR_TOO_LARGE;
  }
  float* scratch = nullptr;
  const int rc = static_scratch_on_current_device(&scratch);
  if (rc != BCG_OK) return rc;
  return decode_attn_paged(q, kv_cache, block_table, max_blocks_per_seq, seq_lens, out, scratch,
                           scratch + n_o, reinterpret_cast<int*>(scratch + kStaticScratchFloats), B,
                           n_heads, n_kv_heads, page_size, max_splits, scale, stream);
}

}  // namespace bcg

BCG_TL_DEFINE_DUMP(bcg_tl_dump_attn)
