// Prefill projection GEMM on CTA PAIRS:  Y[M,N] = X[M,K] · W[N,K]^T  with `tcgen05.mma.cta_group::2`.
//
// The single-CTA kernel (gemm_tcgen05.cu) feeds each 128x256x64 MMA step from 48 KB of shared memory
// (A 16 KB + B 32 KB) — 96 B per clock at the tensor pipe's full rate, three quarters of an SM's
// shared-memory bandwidth — and measured 1,180-1,350 TFLOP/s where cuBLAS reaches 1,450-1,550.
// Here two CTAs of a cluster (one TPC) share a 256x256 output tile: each holds 128 rows of A and HALF
// of the B tile (128 of its 256 weight rows), the leader CTA issues one M=256 MMA that reads both
// halves, and every CTA's accumulator (its 128 rows x 256 columns) stays in its own TMEM.  Per CTA and
// k-block that is 32 KB of operands for the same tensor work: 64 B per clock.
// Measured on a layer's four projections at 12,390 tokens (tools/prefill_gemm_time.py): 1,285 TFLOP/s
// single-CTA -> 1,317 (6 stages, raster groups of 8 m-tiles) -> 1,357 (groups of 12) -> 1,367 (7 stages);
// cuBLAS on the same shapes averages ~1,480.
//
//   cluster (2,1,1), 192 threads per CTA, persistent over 256x256 tiles.
//   warp 0   TMA producer in BOTH CTAs: its A rows and its half of B (tensor maps over this GPU's
//            buffers; `.cta_group::2` loads whose completion bytes land on the LEADER's full barrier —
//            no multicast), ring of 7 x 32 KB stages; slots are freed by the leader's MMA commits,
//            multicast to both CTAs' empty barriers.
//   warp 1   leader CTA: MMA issuer (4 x tcgen05.mma.cta_group::2 m256 n256 k16 per k-block), two
//            accumulator stages of 256 TMEM columns; both CTAs: TMEM alloc / dealloc (cta_group::2).
//   warps 2-5  epilogue in both CTAs (tcgen05.ld -> bf16 / fp32 / SwiGLU -> global), releasing the
//            accumulator stage on the leader's barrier.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace bcg {

constexpr int k2BlockM = 128;       // rows of A per CTA (256 per pair)
constexpr int k2BlockN = 256;       // output columns per tile; each CTA stages 128 of the B rows
constexpr int k2BlockK = 64;
constexpr int k2Threads = 192;
constexpr int k2ABytes = k2BlockM * k2BlockK * 2;          // 16 KB
constexpr int k2BBytes = (k2BlockN / 2) * k2BlockK * 2;    // 16 KB
constexpr int k2StageBytes = k2ABytes + k2BBytes;
constexpr int smem_2sm(int stages) { return stages * k2StageBytes + 1024 + 256; }
constexpr uint32_t k2TmemCols = 512;
// kind::f16, D = f32, A = B = bf16, both K-major, M = 256 (pair), N = 256
constexpr uint32_t k2Idesc = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(k2BlockN >> 3) << 17) |
                             (static_cast<uint32_t>(256 >> 4) << 24);
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // shared::cluster address of the same offset in the pair's even CTA

struct Gemm2Args {
  int M, N, K;
  int tiles_m, tiles_n;   // 256-row / 256-column tiles
  int k_blocks;
  int group_m;
  int epi;
  void* out;
  long long ld_out;
  uint64_t hint_a, hint_b;   // L2 eviction priority of the activation / weight loads
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish_2sm() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// TMA load whose completion bytes are credited to the barrier at the same offset in the LEADER CTA
__device__ __forceinline__ void tma_load_2d_2sm(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1,
                                                uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once all MMAs issued so far have completed) on the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(static_cast<uint16_t>(3))
      : "memory");
}
// arrive on the barrier at the same offset in the leader CTA (rank 0 of the pair)
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
  asm volatile(
      "{\n"
      ".reg .b32 ra;\n"
      "mapa.shared::cluster.u32 ra, %0, 0;\n"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n"
      "}\n" ::"r"(smem_u32(bar))
      : "memory");
}

__device__ __forceinline__ float silu_mul2(float g, float u) { return g * (1.0f / (1.0f + __expf(-g))) * u; }

template <int k2Stages>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(k2Threads, 1)
gemm2sm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                       const Gemm2Args args) {
  extern __shared__ uint8_t smem_raw2[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw2) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + k2Stages * k2StageBytes);
  uint64_t* empty_bar = full_bar + k2Stages;
  uint64_t* tmem_full = empty_bar + k2Stages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  pdl_launch_dependents();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < k2Stages; ++s) {
      mbar_init(&full_bar[s], 1);     // used in the leader: its own arrive.expect_tx; bytes come from both CTAs
      mbar_init(&empty_bar[s], 1);    // one multicast commit per use
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 256); // used in the leader: the epilogue threads of both CTAs
    }
    fence_barrier_init();
  }
  cluster_sync_all();                 // the peer's barriers exist before anything signals them
  if (warp == 1) {
    tmem_alloc_2sm(tmem_slot, k2TmemCols);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int n_clusters = gridDim.x >> 1;
  const int cluster_id = blockIdx.x >> 1;
  const int total = args.tiles_m * args.tiles_n;
  const int group_tiles = args.group_m * args.tiles_n;
  auto decode_item = [&](int item, int& tm, int& tn) {
    const int g = item / group_tiles;
    const int r = item - g * group_tiles;
    const int m0 = g * args.group_m;
    const int gm = min(args.group_m, args.tiles_m - m0);
    tn = r / gm;
    tm = m0 + (r - tn * gm);
  };

  if (warp == 0) {
    // ------------------------------ TMA producer (both CTAs) ------------------------------
    if (lane == 0) {
      pdl_wait();
      int stage = 0;
      uint32_t phase = 0;
      for (int item = cluster_id; item < total; item += n_clusters) {
        int tm, tn;
        decode_item(item, tm, tn);
        const int row_a = tm * 256 + static_cast<int>(rank) * k2BlockM;
        const int row_b = tn * k2BlockN + static_cast<int>(rank) * (k2BlockN / 2);
        for (int kb = 0; kb < args.k_blocks; ++kb) {
          uint8_t* sa = smem + stage * k2StageBytes;
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * k2StageBytes);
          tma_load_2d_2sm(sa, &tmap_a, &full_bar[stage], kb * k2BlockK, row_a, args.hint_a);
          tma_load_2d_2sm(sa + k2ABytes, &tmap_b, &full_bar[stage], kb * k2BlockK, row_b, args.hint_b);
          if (++stage == k2Stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer (leader CTA) --------------------------------
    if (leader) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int item = cluster_id; item < total; item += n_clusters) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * k2BlockN);
        for (int kb = 0; kb < args.k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          if (lane == 0) {
            const uint32_t sa = smem_u32(smem + stage * k2StageBytes);
            const uint64_t da = make_kmajor_sw128_desc(sa);
            const uint64_t db = make_kmajor_sw128_desc(sa + k2ABytes);
#pragma unroll
            for (int k = 0; k < k2BlockK / 16; ++k)
              umma_bf16_2sm(tmem_d, da + 2 * k, db + 2 * k, k2Idesc, (kb > 0 || k > 0) ? 1u : 0u);
            umma_commit_2sm(&empty_bar[stage]);
            if (kb == args.k_blocks - 1) umma_commit_2sm(&tmem_full[acc]);
          }
          __syncwarp();
          if (++stage == k2Stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------ epilogue (both CTAs) -----------------------------------
    pdl_wait();
    const int q = warp & 3;
    const int row = q * 32 + lane;
    int acc = 0;
    uint32_t acc_phase = 0;
    const int M = args.M, N = args.N;
    const long long ld = args.ld_out;
    for (int item = cluster_id; item < total; item += n_clusters) {
      int tm, tn;
      decode_item(item, tm, tn);
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(acc * k2BlockN);
      const int m = tm * 256 + static_cast<int>(rank) * k2BlockM + row;
      const int n0 = tn * k2BlockN;
      if (args.epi == BCG_EPI_SWIGLU) {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(args.out);
#pragma unroll 1
        for (int c = 0; c < k2BlockN; c += 32) {
          uint32_t g[16], u[16];
          tmem_ld_x16(taddr + c, g);
          tmem_ld_x16(taddr + c + 16, u);
          tmem_ld_wait();
          uint32_t p[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            p[j] = pack_bf16(silu_mul2(__uint_as_float(g[2 * j]), __uint_as_float(u[2 * j])),
                             silu_mul2(__uint_as_float(g[2 * j + 1]), __uint_as_float(u[2 * j + 1])));
          if (m < M && n0 + c < N) {
            uint4* dst = reinterpret_cast<uint4*>(o + m * ld + (n0 + c) / 2);
            dst[0] = make_uint4(p[0], p[1], p[2], p[3]);
            dst[1] = make_uint4(p[4], p[5], p[6], p[7]);
          }
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < k2BlockN; c += 32) {
          uint32_t v0[16], v1[16];
          tmem_ld_x16(taddr + c, v0);
          tmem_ld_x16(taddr + c + 16, v1);
          tmem_ld_wait();
          if (args.epi == BCG_EPI_F32) {
            float* o = reinterpret_cast<float*>(args.out);
            if (m < M && n0 + c < N) {
              float4* dst = reinterpret_cast<float4*>(o + m * ld + n0 + c);
#pragma unroll
              for (int j = 0; j < 4; ++j)
                dst[j] = make_float4(__uint_as_float(v0[4 * j]), __uint_as_float(v0[4 * j + 1]),
                                     __uint_as_float(v0[4 * j + 2]), __uint_as_float(v0[4 * j + 3]));
#pragma unroll
              for (int j = 0; j < 4; ++j)
                dst[4 + j] = make_float4(__uint_as_float(v1[4 * j]), __uint_as_float(v1[4 * j + 1]),
                                         __uint_as_float(v1[4 * j + 2]), __uint_as_float(v1[4 * j + 3]));
            }
          } else {
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(args.out);
            uint32_t p[16];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              p[j] = pack_bf16(__uint_as_float(v0[2 * j]), __uint_as_float(v0[2 * j + 1]));
              p[8 + j] = pack_bf16(__uint_as_float(v1[2 * j]), __uint_as_float(v1[2 * j + 1]));
            }
            if (m < M && n0 + c < N) {
              uint4* dst = reinterpret_cast<uint4*>(o + m * ld + n0 + c);
              dst[0] = make_uint4(p[0], p[1], p[2], p[3]);
              dst[1] = make_uint4(p[4], p[5], p[6], p[7]);
              dst[2] = make_uint4(p[8], p[9], p[10], p[11]);
              dst[3] = make_uint4(p[12], p[13], p[14], p[15]);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive_leader(&tmem_empty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  }

  // The leader's MMAs read the peer's shared memory and both CTAs hold TMEM of the pair: nobody leaves
  // before everybody is done.
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_2sm(tmem_base, k2TmemCols);
}

// Returns BCG_ERR_UNSUPPORTED when the shape is not for this kernel (the caller then uses the
// single-CTA kernel).
int gemm2sm_bf16_tn(const void* X, const void* W, void* out, int M, int N, int K, long long ldx, long long ldw,
                    long long ld_out, int epi, cudaStream_t stream) {
  if (M < 512 || (N % k2BlockN) != 0) return BCG_ERR_UNSUPPORTED;
  static const bool off = getenv("BCG_GEMM_2SM") && atoi(getenv("BCG_GEMM_2SM")) == 0;   // tuning knob
  if (off) return BCG_ERR_UNSUPPORTED;
  Gemm2Args a = {};
  a.M = M;
  a.N = N;
  a.K = K;
  a.tiles_m = (M + 255) / 256;
  a.tiles_n = N / k2BlockN;
  a.k_blocks = (K + k2BlockK - 1) / k2BlockK;
  static const int group = getenv("BCG_GEMM_2SM_GROUP") ? atoi(getenv("BCG_GEMM_2SM_GROUP")) : 12;   // tuning knobs
  static const int stages = getenv("BCG_GEMM_2SM_STAGES") ? atoi(getenv("BCG_GEMM_2SM_STAGES")) : 7;
  a.group_m = a.tiles_m < group ? a.tiles_m : group;
  // In this raster a group's activation panel (group_m x 256 rows) is what every wave of the group
  // re-reads, a weight tile is shared by the group's m-tiles within ONE wave and then dead: keep the
  // panel (evict-last), let the weights pass (normal).  Measured over a 14B layer at 12,390 tokens:
  // 1,355 TFLOP/s vs 1,314-1,329 with the hints the other way round (weights evict-first: 1,245).
  static const int ha = getenv("BCG_GEMM_2SM_HINT_A") ? atoi(getenv("BCG_GEMM_2SM_HINT_A")) : 2;
  static const int hb = getenv("BCG_GEMM_2SM_HINT_B") ? atoi(getenv("BCG_GEMM_2SM_HINT_B")) : 0;
  const uint64_t hints[3] = {kEvictNormal, kEvictFirst, kEvictLast};
  a.hint_a = hints[ha < 0 || ha > 2 ? 2 : ha];
  a.hint_b = hints[hb < 0 || hb > 2 ? 0 : hb];
  a.epi = epi;
  a.out = out;
  a.ld_out = ld_out;
  CUtensorMap tma, tmb;
  int rc;
  if ((rc = make_tmap_bf16_sw128(&tma, X, M, K, ldx, k2BlockM)) != BCG_OK) return rc;
  if ((rc = make_tmap_bf16_sw128(&tmb, W, N, K, ldw, k2BlockN / 2)) != BCG_OK) return rc;

  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int clusters = sms / 2;
  const int total = a.tiles_m * a.tiles_n;
  if (clusters > total) clusters = total;
  if (stages != 6) {
    if ((rc = ensure_dynamic_smem(reinterpret_cast<const void*>(gemm2sm_tcgen05_kernel<7>), smem_2sm(7)))) return rc;
    BCG_CUDA_CHECK(launch_k(gemm2sm_tcgen05_kernel<7>, dim3(2 * clusters), dim3(k2Threads), smem_2sm(7), stream, tma,
                            tmb, a));
  } else {
    if ((rc = ensure_dynamic_smem(reinterpret_cast<const void*>(gemm2sm_tcgen05_kernel<6>), smem_2sm(6)))) return rc;
    BCG_CUDA_CHECK(launch_k(gemm2sm_tcgen05_kernel<6>, dim3(2 * clusters), dim3(k2Threads), smem_2sm(6), stream, tma,
                            tmb, a));
  }
  return BCG_OK;
}

}  // namespace bcg
