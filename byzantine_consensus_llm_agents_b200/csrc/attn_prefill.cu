// Fused causal prefill attention over the paged bf16 KV cache with a cached-prefix offset
// (stands in for the FlashInfer `fmhaSm100a...PagedKv...Context` kernels
// vLLM dispatches on sm_100).  New tokens of every sequence are concatenated ("ragged") in q;
// K/V for *all* positions — the reused prefix pages and the pages just written by
// qknorm_rope_kvwrite — are read through the block table, so prefix-cache hits and cold
// prompts take the same code path.
//
// Blackwell-native: both contractions run on tcgen05 tensor cores with TMEM accumulators, every
// operand tile arrives by TMA.
//
//   grid (n_heads, n_tiles), 192 threads.  One CTA = 128 query rows of one head; KV streamed in
//   BN-token blocks (BN = 64: 112 KB of shared memory and 256 TMEM columns per CTA, so TWO CTAs share
//   an SM and each hides the other's softmax / pipeline bubbles; BN = 128: one CTA per SM).
//     warp 0      TMA producer: the Q tile once (two 64-column halves of a [rows, n_heads*128]
//                 tensor map), then per block the BN/16 K pages and V pages of this KV head
//                 (16-row x 64-column boxes of the cache seen as a [token-rows, 128] tensor — this
//                 GPU's own allocation only), into two-stage K and V rings, 128B-swizzled.
//     warp 1      MMA issuer: S_j = Q K_j^T (8 x tcgen05.mma m128 nBN k16, both operands K-major)
//                 into one of two S accumulators in TMEM; O += P_j V_j (BN/16 x m128 n128 k16, A = P
//                 K-major from shared memory, B = V in its natural [token][head_dim] layout =
//                 MN-major) into the ONE output accumulator, which stays in TMEM for the whole tile.
//                 Issue order S0 S1 PV0 S2 PV1 ...: the next S is queued while softmax runs.
//     warps 2-5   softmax: thread = query row = TMEM lane.  Per block: one tcgen05.ld of the whole
//                 S row into registers (the S buffer is released at once), row max, p = exp2(s*scale
//                 - m), row sum, bf16 P written to shared memory in the swizzled K-major layout the
//                 MMA reads.  The running maximum m is only raised when the block's maximum exceeds
//                 it by more than 8 (a factor 256, harmless in fp32 / bf16): then — rarely after the
//                 first blocks — the warp rescales its 32 rows of O in TMEM (tcgen05.ld, multiply,
//                 tcgen05.st) before the block's PV is issued.  The result is exact: numerator and
//                 denominator carry the same reference maximum.
//   TMEM columns: S0 | S1 (BN fp32 columns each) | O (128).
//
// tile_info[tile] = {sequence index, first q row (global), position of that row, rows (<= 128)}
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace bcg {

constexpr int kPfPage = 16;
constexpr int kPfBM = 128;               // query rows per CTA
constexpr int kPfThreads = 192;
constexpr int kPfQHalf = 128 * 64 * 2;   // [128 rows][64 bf16], 128B-swizzled: 16 KB
constexpr int kPfQTile = 2 * kPfQHalf;
constexpr int kPfStages = 2;
constexpr float kPfRescaleGap = 8.0f;    // log2 units

template <int BN>
struct PfCfg {
  static constexpr int kKvHalf = BN * 64 * 2;            // [BN tokens][64 bf16]
  static constexpr int kKvTile = 2 * kKvHalf;            // [BN tokens][128]
  static constexpr int kPTile = 128 * BN * 2;            // [128 rows][BN tokens] as BN/64 halves of 16 KB
  static constexpr int kOffQ = 0;
  static constexpr int kOffK = kPfQTile;
  static constexpr int kOffV = kOffK + kPfStages * kKvTile;
  static constexpr int kOffP = kOffV + kPfStages * kKvTile;
  static constexpr int kOffBar = kOffP + kPTile;
  static constexpr int kSlack = BN == 64 ? 768 : 1024;   // room to align the base to 1024 bytes
  static constexpr int kSmem = kOffBar + 256 + kSlack;   // BN=64: 113 KB (two CTAs per SM); BN=128: 193.25 KB
  static constexpr uint32_t kTmemCols = BN == 64 ? 256 : 512;   // 2*BN + 128, rounded to a power of two
  static constexpr int kCtasPerSm = BN == 64 ? 2 : 1;
};

// tcgen05.ld / st: 32 lanes x 32 consecutive fp32 columns; thread t of the warp owns lane (base + t)
__device__ __forceinline__ void tmem_ld_x32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
        "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]),
        "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]),
        "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_st_x32(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
        "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
        "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() {
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// The softmax warps are bound by instruction issue (one slot per sub-partition and clock), so the
// exponent path uses the cheapest encodings the ISA has: packed two-lane fp32 FMA / ADD
// (fma.rn.f32x2, add.rn.f32x2 — one issue slot for two elements) and the bare ex2.approx.ftz
// (exp2f() wraps it in a denormal-range rescue of three more instructions; the arguments here are
// <= 8 and flush-to-zero at the bottom is exactly what a softmax tail wants).
__device__ __forceinline__ uint64_t pack_f32x2(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ void unpack_f32x2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma_f32x2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t add_f32x2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// MN-major, 128B-swizzled B operand: V as it lies in the cache, [token][head_dim], staged as two
// 64-column halves of [BN tokens][64] (128-byte rows).  In the canonical MN-major layout
// ((8,8,m),(8,k)):((1,8,LBO),(64,SBO)) (elements; cute/atom/mma_traits_sm100.hpp) that is: 64
// head_dim elements contiguous, the next 64 (m) LBO = one half away, consecutive tokens (K) 128
// bytes apart, groups of 8 tokens SBO = 1024 bytes apart.
__device__ __forceinline__ uint64_t make_mnmajor_sw128_desc(uint32_t smem_addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(lbo_bytes >> 4) << 16;
  d |= static_cast<uint64_t>(1024 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}

template <int BN>
__global__ void __launch_bounds__(kPfThreads, PfCfg<BN>::kCtasPerSm)
prefill_attn_kernel(const __grid_constant__ CUtensorMap q_map, const __grid_constant__ CUtensorMap kv_map,
                    const int* __restrict__ block_table, int max_blocks,
                    const int* __restrict__ tile_info, int n_tiles, __nv_bfloat16* __restrict__ out,
                    int n_heads, int n_kv_heads, float scale_log2e) {
  using C = PfCfg<BN>;
  constexpr int kPages = BN / kPfPage;
  // kind::f16 instruction descriptors, D = f32, A = B = bf16, M = 128; PV: bit 16 = B is MN-major
  constexpr uint32_t kIdescQK = make_idesc_bf16_m128(BN);
  constexpr uint32_t kIdescPV = make_idesc_bf16_m128(128) | (1u << 16);
  extern __shared__ __align__(1024) uint8_t pf_raw[];
  // The swizzled tiles need a 1024-byte aligned base.  Dynamic shared memory of a kernel without
  // static shared memory starts 1024-aligned in practice; the slack covers a lesser alignment.
  const uint32_t pad = (1024u - (smem_u32(pf_raw) & 1023u)) & 1023u;
  if (pad > static_cast<uint32_t>(C::kSlack)) __trap();
  uint8_t* smem = pf_raw + pad;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kOffBar);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;    // [2]
  uint64_t* k_empty = bars + 3;   // [2]
  uint64_t* v_full = bars + 5;    // [2]
  uint64_t* v_empty = bars + 7;   // [2]
  uint64_t* s_full = bars + 9;    // [2]
  uint64_t* s_free = bars + 11;   // [2]
  uint64_t* p_full = bars + 13;
  uint64_t* pv_done = bars + 14;  // completes once per block, when that block's PV MMAs have finished
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();

  const int head = blockIdx.x;
  const int tile = n_tiles - 1 - static_cast<int>(blockIdx.y);   // later (longer) tiles first
  const int kvh = head / (n_heads / n_kv_heads);
  const int b = tile_info[tile * 4 + 0];
  const int row0 = tile_info[tile * 4 + 1];
  const int pos0 = tile_info[tile * 4 + 2];
  const int n_rows = tile_info[tile * 4 + 3];
  const int kv_total = pos0 + n_rows;
  const int nb = (kv_total + BN - 1) / BN;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&q_map);
    tma_prefetch_desc(&kv_map);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&s_free[i], 128);
    }
    mbar_init(p_full, 128);
    mbar_init(pv_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, C::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_o = tmem_base + 2 * BN;

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      pdl_wait();   // q and this step's KV rows come from the kernels in front of this one
      mbar_arrive_expect_tx(q_full, kPfQTile);
      tma_load_2d(smem + C::kOffQ, &q_map, q_full, head * 128, row0, kEvictFirst);
      tma_load_2d(smem + C::kOffQ + kPfQHalf, &q_map, q_full, head * 128 + 64, row0, kEvictFirst);
      const int* bt = block_table + static_cast<long long>(b) * max_blocks;
      const int last_page = (kv_total - 1) / kPfPage;   // pages past it are never read (clamped)
      for (int j = 0; j < nb; ++j) {
        const int st = j & 1;
        const uint32_t par = (j >> 1) & 1;
        int phys[kPages];
#pragma unroll
        for (int pg = 0; pg < kPages; ++pg) phys[pg] = bt[min(j * kPages + pg, last_page)];
        uint8_t* dk = smem + C::kOffK + st * C::kKvTile;
        mbar_wait(&k_empty[st], par ^ 1);
        mbar_arrive_expect_tx(&k_full[st], C::kKvTile);
#pragma unroll
        for (int pg = 0; pg < kPages; ++pg) {
          const int row_k = ((phys[pg] * 2) * n_kv_heads + kvh) * kPfPage;
          tma_load_2d(dk + pg * 2048, &kv_map, &k_full[st], 0, row_k, kEvictNormal);
          tma_load_2d(dk + C::kKvHalf + pg * 2048, &kv_map, &k_full[st], 64, row_k, kEvictNormal);
        }
        uint8_t* dv = smem + C::kOffV + st * C::kKvTile;
        mbar_wait(&v_empty[st], par ^ 1);
        mbar_arrive_expect_tx(&v_full[st], C::kKvTile);
#pragma unroll
        for (int pg = 0; pg < kPages; ++pg) {
          const int row_v = ((phys[pg] * 2 + 1) * n_kv_heads + kvh) * kPfPage;
          tma_load_2d(dv + pg * 2048, &kv_map, &v_full[st], 0, row_v, kEvictNormal);
          tma_load_2d(dv + C::kKvHalf + pg * 2048, &kv_map, &v_full[st], 64, row_v, kEvictNormal);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    const uint32_t sq = smem_u32(smem + C::kOffQ);
    const uint32_t sp = smem_u32(smem + C::kOffP);
    auto issue_s = [&](int j) {
      const int st = j & 1;
      const uint32_t par = (j >> 1) & 1;
      mbar_wait(&k_full[st], par);
      mbar_wait(&s_free[st], par ^ 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sk = smem_u32(smem + C::kOffK + st * C::kKvTile);
        const uint32_t d = tmem_base + static_cast<uint32_t>(st * BN);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {   // head_dim 128 = 8 x k16; halves of 64, 32 B per step
          umma_bf16(d, make_kmajor_sw128_desc(sq + (ks >> 2) * kPfQHalf + (ks & 3) * 32),
                    make_kmajor_sw128_desc(sk + (ks >> 2) * C::kKvHalf + (ks & 3) * 32), kIdescQK,
                    ks > 0 ? 1u : 0u);
        }
        umma_commit(&k_empty[st]);
        umma_commit(&s_full[st]);
      }
      __syncwarp();
    };
    auto issue_pv = [&](int j) {
      const int st = j & 1;
      mbar_wait(p_full, j & 1);          // P_j written (and O rescaled where that was needed)
      mbar_wait(&v_full[st], (j >> 1) & 1);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t sv = smem_u32(smem + C::kOffV + st * C::kKvTile);
#pragma unroll
        for (int kk = 0; kk < BN / 16; ++kk) {
          // P: K-major, 64-token halves of 16 KB, 32 B per k16 step; V: 16 token rows of 128 B per step
          umma_bf16(tmem_o, make_kmajor_sw128_desc(sp + (kk >> 2) * kPfQHalf + (kk & 3) * 32),
                    make_mnmajor_sw128_desc(sv + kk * 16 * 128, C::kKvHalf), kIdescPV,
                    (j > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(&v_empty[st]);
        umma_commit(pv_done);
      }
      __syncwarp();
    };
    mbar_wait(q_full, 0);
    issue_s(0);
    for (int j = 0; j < nb; ++j) {
      if (j + 1 < nb) issue_s(j + 1);
      issue_pv(j);
    }
  } else {
    // ------------------------------ softmax + output --------------------------
    const int q4 = warp & 3;              // TMEM sub-partition of this warp
    const int r = q4 * 32 + lane;         // query row of the tile = TMEM lane
    const uint32_t lane_addr = static_cast<uint32_t>(q4 * 32) << 16;
    const int limit = pos0 + r;           // last KV position this row may see
    uint8_t* p_row = smem + C::kOffP + r * 128;
    const int sw = r & 7;
    float m_run = -INFINITY, l_run = 0.f;

    for (int j = 0; j < nb; ++j) {
      const int st = j & 1;
      const int col0 = j * BN;
      const bool need_mask = col0 + BN - 1 > limit;
      mbar_wait(&s_full[st], (j >> 1) & 1);
      tc_fence_after();
      uint32_t s[BN];
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) tmem_ld_x32(tmem_base + lane_addr + st * BN + c * 32, s + c * 32);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&s_free[st]);     // S_j is in registers: the MMA warp may overwrite this accumulator
      float mx = -INFINITY;
      if (need_mask) {
#pragma unroll
        for (int i = 0; i < BN; ++i) {
          if (col0 + i > limit) s[i] = 0xff800000u;   // -inf
          mx = fmaxf(mx, __uint_as_float(s[i]));
        }
      } else {
#pragma unroll
        for (int i = 0; i < BN; ++i) mx = fmaxf(mx, __uint_as_float(s[i]));
      }
      const float m_blk = mx * scale_log2e;
      const bool grow = m_blk > m_run + kPfRescaleGap;   // always true for block 0 (m_run = -inf)
      const float m_new = grow ? m_blk : m_run;
      // PV_{j-1} must be complete before P is overwritten and before O may be rescaled
      mbar_wait(pv_done, (j & 1) ^ 1);
      if (j > 0 && __any_sync(0xffffffffu, grow)) {
        tc_fence_after();
        const float corr = grow ? ex2_approx(m_run - m_new) : 1.0f;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t o[32];
          tmem_ld_x32(tmem_o + lane_addr + c * 32, o);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
          tmem_st_x32(tmem_o + lane_addr + c * 32, o);
        }
        tmem_st_wait();
        l_run *= corr;
      }
      m_run = m_new;
      const uint64_t scale2 = pack_f32x2(__float_as_uint(scale_log2e), __float_as_uint(scale_log2e));
      const uint64_t negm2 = pack_f32x2(__float_as_uint(-m_run), __float_as_uint(-m_run));
      uint64_t l2 = pack_f32x2(0u, 0u);
#pragma unroll
      for (int c = 0; c < BN / 8; ++c) {   // 8 tokens = one 16-byte chunk of the K-major P tile
        uint32_t pk[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float x0, x1;
          unpack_f32x2(fma_f32x2(pack_f32x2(s[c * 8 + 2 * i], s[c * 8 + 2 * i + 1]), scale2, negm2), x0, x1);
          const float p0 = ex2_approx(x0), p1 = ex2_approx(x1);
          l2 = add_f32x2(l2, pack_f32x2(__float_as_uint(p0), __float_as_uint(p1)));
          pk[i] = pack_bf16(p0, p1);
        }
        *reinterpret_cast<uint4*>(p_row + (c >> 3) * kPfQHalf + (((c & 7) ^ sw) << 4)) =
            make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
      float l_lo, l_hi;
      unpack_f32x2(l2, l_lo, l_hi);
      l_run += l_lo + l_hi;
      tc_fence_before();            // orders the O rescale (tcgen05.st) before the MMA warp's next PV
      fence_proxy_async();          // P stores -> visible to the tensor core's async proxy
      mbar_arrive(p_full);
    }
    // the last block's PV
    mbar_wait(pv_done, (nb - 1) & 1);
    tc_fence_after();
    const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
    __nv_bfloat16* dst = out + (static_cast<long long>(row0) + r) * (static_cast<long long>(n_heads) * 128) +
                         head * 128;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      tmem_ld_x32(tmem_o + lane_addr + c * 32, o);
      tmem_ld_wait();
      if (r < n_rows) {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 w;
          w.x = pack_bf16(__uint_as_float(o[8 * g + 0]) * inv, __uint_as_float(o[8 * g + 1]) * inv);
          w.y = pack_bf16(__uint_as_float(o[8 * g + 2]) * inv, __uint_as_float(o[8 * g + 3]) * inv);
          w.z = pack_bf16(__uint_as_float(o[8 * g + 4]) * inv, __uint_as_float(o[8 * g + 5]) * inv);
          w.w = pack_bf16(__uint_as_float(o[8 * g + 6]) * inv, __uint_as_float(o[8 * g + 7]) * inv);
          reinterpret_cast<uint4*>(dst)[c * 4 + g] = w;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, C::kTmemCols);
}

int prefill_attn_paged(const void* q, const void* kv_cache, const int* block_table,
                       int max_blocks_per_seq, const int* tile_info, int n_tiles, void* out,
                       int n_q_rows, int n_heads, int n_kv_heads, int page_size, float scale,
                       cudaStream_t stream) {
  if (!q || !kv_cache || !block_table || !tile_info || !out || n_tiles <= 0 || n_q_rows <= 0 ||
      n_heads <= 0 || n_kv_heads <= 0 || (n_heads % n_kv_heads) || max_blocks_per_seq <= 0) {
    set_last_error("bcg_prefill_attn_paged", "bad argument");
    return BCG_ERR_ARG;
  }
  if (page_size != kPfPage) {
    set_last_error("bcg_prefill_attn_paged", "page_size must be 16");
    return BCG_ERR_UNSUPPORTED;
  }
  if ((reinterpret_cast<uintptr_t>(q) & 15) || (reinterpret_cast<uintptr_t>(out) & 15) ||
      (reinterpret_cast<uintptr_t>(kv_cache) & 15)) {
    set_last_error("bcg_prefill_attn_paged", "q, out and kv_cache must be 16-byte aligned");
    return BCG_ERR_ARG;
  }
  // q as [n_q_rows, n_heads*128]: 128-row x 64-column boxes, rows past the end read as zero.
  // The cache as a [token-rows, 128] tensor: 16-row x 64-column boxes (one page of one KV head);
  // the row count only bounds TMA's range check, the block table never points past the allocation.
  CUtensorMap q_map, kv_map;
  int rc = make_tmap_bf16_sw128(&q_map, q, n_q_rows, static_cast<long long>(n_heads) * 128,
                                static_cast<long long>(n_heads) * 128, kPfBM);
  if (rc != BCG_OK) return rc;
  rc = make_tmap_bf16_sw128(&kv_map, kv_cache, 1ll << 30, 128, 128, kPfPage);
  if (rc != BCG_OK) return rc;
  static const int bn = getenv("BCG_PREFILL_ATTN_BN") ? atoi(getenv("BCG_PREFILL_ATTN_BN")) : 64;   // tuning knob
  if (bn == 128) {
    if ((rc = ensure_dynamic_smem(reinterpret_cast<const void*>(prefill_attn_kernel<128>), PfCfg<128>::kSmem))) return rc;
    BCG_CUDA_CHECK(launch_k(prefill_attn_kernel<128>, dim3(n_heads, n_tiles), dim3(kPfThreads),
                            PfCfg<128>::kSmem, stream, q_map, kv_map, block_table, max_blocks_per_seq, tile_info,
                            n_tiles, reinterpret_cast<__nv_bfloat16*>(out), n_heads, n_kv_heads,
                            scale * 1.4426950408889634f));
  } else {
    if ((rc = ensure_dynamic_smem(reinterpret_cast<const void*>(prefill_attn_kernel<64>), PfCfg<64>::kSmem))) return rc;
    BCG_CUDA_CHECK(launch_k(prefill_attn_kernel<64>, dim3(n_heads, n_tiles), dim3(kPfThreads),
                            PfCfg<64>::kSmem, stream, q_map, kv_map, block_table, max_blocks_per_seq, tile_info,
                            n_tiles, reinterpret_cast<__nv_bfloat16*>(out), n_heads, n_kv_heads,
                            scale * 1.4426950408889634f));
  }
  return BCG_OK;
}

}  // namespace bcg
