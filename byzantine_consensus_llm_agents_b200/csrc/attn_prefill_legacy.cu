// DEVELOPMENT FALLBACK (mma.sync), reachable only with BCG_PREFILL_ATTN_LEGACY=1 while the tcgen05
// kernel in attn_prefill.cu is being validated on hardware; to be deleted afterwards.
// Fused causal prefill attention over the paged bf16 KV cache with a cached-prefix offset
// (SURVEY.md §2.1 row K7).  New tokens of every sequence are concatenated ("ragged") in q;
// K/V for *all* positions — the reused prefix pages and the pages just written by
// qknorm_rope_kvwrite — are read through the block table, so prefix-cache hits and cold
// prompts take the same code path.
//
// grid (n_tiles, n_heads), 128 threads.  One CTA = 64 query rows of one head; KV streamed in
// 64-token blocks (4 pages) through a double-buffered cp.async ring into XOR-swizzled smem;
// S = QK^T and O += PV on bf16 tensor-core MMAs (m16n8k16, fp32 accumulate) with an fp32
// online softmax in registers.  Attention is ~2.6 % of prefill FLOPs at the reference's
// prompt lengths; the projections carry the tcgen05 path.
//
// tile_info[tile] = {sequence index, first q row (global), position of that row, rows (<=64)}
#include "common.cuh"
#include "kernels.h"

namespace bcg {

constexpr int kLgPage = 16;
constexpr int kLgBM = 64;
constexpr int kLgBN = 64;

__device__ __forceinline__ void cp_async16(void* smem, const void* g) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void ldmatrix_x4(uint32_t* r, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t* r, uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, "
      "{%8, %9}, {%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// byte offset of 16-byte chunk `c` (0..15) of row `r` in a [rows][128] bf16 tile
__device__ __forceinline__ uint32_t swz(int r, int c) {
  return static_cast<uint32_t>((r * 16 + (c ^ (r & 7))) * 16);
}

__global__ void __launch_bounds__(128)
prefill_attn_legacy_kernel(const __nv_bfloat16* __restrict__ q, const __nv_bfloat16* __restrict__ kv_cache,
                    const int* __restrict__ block_table, int max_blocks,
                    const int* __restrict__ tile_info, __nv_bfloat16* __restrict__ out, int n_heads,
                    int n_kv_heads, float scale_log2e) {
  extern __shared__ __align__(128) uint8_t pf_smem[];
  uint8_t* sQ = pf_smem;                       // 16 KB
  uint8_t* sK = pf_smem + 16384;               // 2 x 16 KB
  uint8_t* sV = pf_smem + 16384 + 2 * 16384;   // 2 x 16 KB
  pdl_launch_dependents();
  pdl_wait();

  const int tile = blockIdx.x, head = blockIdx.y;
  const int kvh = head / (n_heads / n_kv_heads);
  const int b = tile_info[tile * 4 + 0];
  const int row0 = tile_info[tile * 4 + 1];
  const int pos0 = tile_info[tile * 4 + 2];
  const int n_rows = tile_info[tile * 4 + 3];
  const int kv_total = pos0 + n_rows;
  const int n_blocks = (kv_total + kLgBN - 1) / kLgBN;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int* bt = block_table + static_cast<long long>(b) * max_blocks;
  const long long q_ld = static_cast<long long>(n_heads) * 128;

  for (int i = tid; i < kLgBM * 16; i += 128) {
    const int r = i >> 4, c = i & 15;
    const int rr = min(r, n_rows - 1);
    cp_async16(sQ + swz(r, c), q + (row0 + rr) * q_ld + head * 128 + c * 8);
  }
  auto load_kv = [&](int it) {
    uint8_t* dk = sK + (it & 1) * 16384;
    uint8_t* dv = sV + (it & 1) * 16384;
    for (int i = tid; i < kLgBN * 16; i += 128) {
      const int r = i >> 4, c = i & 15;
      const int tok = min(it * kLgBN + r, kv_total - 1);
      const int phys = bt[tok / kLgPage];
      const __nv_bfloat16* kp = kv_cache +
          ((static_cast<long long>(phys) * 2 * n_kv_heads + kvh) * kLgPage + (tok % kLgPage)) * 128 + c * 8;
      cp_async16(dk + swz(r, c), kp);
      cp_async16(dv + swz(r, c), kp + static_cast<long long>(n_kv_heads) * kLgPage * 128);
    }
  };
  load_kv(0);
  cp_async_commit();

  uint32_t qf[8][4];
  float o[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};
  const int g = lane >> 2, t4 = lane & 3;
  const int qrow[2] = {warp * 16 + g, warp * 16 + g + 8};

  for (int it = 0; it < n_blocks; ++it) {
    cp_async_wait_all();
    __syncthreads();
    if (it == 0) {
      const int r = warp * 16 + (lane & 7) + 8 * ((lane >> 3) & 1);
#pragma unroll
      for (int ks = 0; ks < 8; ++ks) ldmatrix_x4(qf[ks], smem_u32(sQ + swz(r, 2 * ks + (lane >> 4))));
    }
    if (it + 1 < n_blocks) {
      load_kv(it + 1);
      cp_async_commit();
    }
    const uint8_t* ck = sK + (it & 1) * 16384;
    const uint8_t* cv = sV + (it & 1) * 16384;

    float s[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        uint32_t kb[4];
        ldmatrix_x4(kb, smem_u32(ck + swz(nt * 8 + (lane & 7), 4 * kk + (lane >> 3))));
        mma_bf16_16816(s[nt], qf[2 * kk], kb[0], kb[1]);
        mma_bf16_16816(s[nt], qf[2 * kk + 1], kb[2], kb[3]);
      }
    }
    const bool need_mask = it * kLgBN + kLgBN - 1 > pos0 + warp * 16;
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float v = s[nt][e] * scale_log2e;
        if (need_mask) {
          const int col = it * kLgBN + nt * 8 + 2 * t4 + (e & 1);
          if (col > pos0 + qrow[e >> 1]) v = -INFINITY;
        }
        s[nt][e] = v;
        mx[e >> 1] = fmaxf(mx[e >> 1], v);
      }
    }
    float corr[2], m_use[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      const float m_new = fmaxf(m_run[r], mx[r]);
      m_use[r] = (m_new == -INFINITY) ? 0.f : m_new;
      corr[r] = exp2f(m_run[r] - m_use[r]);
      m_run[r] = m_new;
      l_run[r] *= corr[r];
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      o[i][0] *= corr[0]; o[i][1] *= corr[0];
      o[i][2] *= corr[1]; o[i][3] *= corr[1];
    }
    uint32_t pf[4][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const float p0 = exp2f(s[nt][0] - m_use[0]), p1 = exp2f(s[nt][1] - m_use[0]);
      const float p2 = exp2f(s[nt][2] - m_use[1]), p3 = exp2f(s[nt][3] - m_use[1]);
      l_run[0] += p0 + p1;
      l_run[1] += p2 + p3;
      pf[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16(p0, p1);
      pf[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
    }
#pragma unroll
    for (int kt = 0; kt < 4; ++kt) {
      const int r = kt * 16 + (lane & 7) + 8 * ((lane >> 3) & 1);
#pragma unroll
      for (int dt = 0; dt < 8; ++dt) {
        uint32_t vb[4];
        ldmatrix_x4_trans(vb, smem_u32(cv + swz(r, 2 * dt + (lane >> 4))));
        mma_bf16_16816(o[2 * dt], pf[kt], vb[0], vb[1]);
        mma_bf16_16816(o[2 * dt + 1], pf[kt], vb[2], vb[3]);
      }
    }
  }

#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    if (qrow[r] < n_rows) {
      const float inv = l_run[r] > 0.f ? 1.0f / l_run[r] : 0.f;
      __nv_bfloat16* dst = out + (row0 + qrow[r]) * q_ld + head * 128 + 2 * t4;
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
        *reinterpret_cast<uint32_t*>(dst + nt * 8) =
            pack_bf16(o[nt][2 * r] * inv, o[nt][2 * r + 1] * inv);
      }
    }
  }
}

int prefill_attn_paged_legacy(const void* q, const void* kv_cache, const int* block_table,
                       int max_blocks_per_seq, const int* tile_info, int n_tiles, void* out,
                       int n_heads, int n_kv_heads, int page_size, float scale,
                       cudaStream_t stream) {
  if (!q || !kv_cache || !block_table || !tile_info || !out || n_tiles <= 0 || n_heads <= 0 ||
      n_kv_heads <= 0 || (n_heads % n_kv_heads)) {
    set_last_error("bcg_prefill_attn_paged", "bad argument");
    return BCG_ERR_ARG;
  }
  if (page_size != kLgPage) {
    set_last_error("bcg_prefill_attn_paged", "page_size must be 16");
    return BCG_ERR_UNSUPPORTED;
  }
  constexpr int SMEM = 16384 * 5;
  if (const int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(prefill_attn_legacy_kernel), SMEM)) return rc;
  prefill_attn_legacy_kernel<<<dim3(n_tiles, n_heads), 128, SMEM, stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(q), reinterpret_cast<const __nv_bfloat16*>(kv_cache),
      block_table, max_blocks_per_seq, tile_info, reinterpret_cast<__nv_bfloat16*>(out), n_heads,
      n_kv_heads, scale * 1.4426950408889634f);
  BCG_CUDA_CHECK(cudaGetLastError());
  return BCG_OK;
}

}  // namespace bcg
