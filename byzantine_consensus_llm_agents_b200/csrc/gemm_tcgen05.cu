// Dense projection GEMM for the agent-step engine:  Y[M,N] = X[M,K] · W[N,K]^T
// (bf16 in, fp32 accumulate in TMEM).  Replaces the cuBLAS F.linear calls the
// reference reaches through vLLM.
//
// One persistent, warp-specialised kernel serves both regimes:
//   * prefill (M large, tensor-bound): A-operand = activations (128-row tiles on the
//     TMEM lanes), B-operand = weights (BN = 128/256 columns).
//   * decode (M <= 128, HBM-bound): "swap-AB" — A-operand = weights (128 weight rows
//     on the TMEM lanes), B-operand = the few activation rows (BN = 16..128), with
//     optional split-K so every SM streams weights.  Split-K partials are written as
//     fp32 [split][M][N] and summed in a FIXED order by the consumer kernel
//     (add_rmsnorm / qknorm_rope / swiglu_reduce): deterministic, batch-invariant.
//
// Roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM alloc + MMA issuer,
// warps 2..5 = epilogue (TMEM -> registers -> global).  smem ring of STAGES
// {A 128x64, B BNx64} bf16 tiles in the 128B-swizzled K-major layout shared by TMA and
// the UMMA descriptors; two TMEM accumulators so the epilogue of tile i overlaps the
// mainloop of tile i+1.
#include <cuda.h>

#include "common.cuh"
#include "kernels.h"

namespace bcg {

constexpr int kBlockA = 128;
constexpr int kBlockK = 64;
constexpr int kUmmaK = 16;
constexpr int kGemmThreads = 192;

constexpr int kSumsqStride = 128;  // rows per tile entry of the sum-of-squares tables (M <= 128)
// Decode batches run TWO CTAs per SM with a short ring each (6 x 18 KB at 16 rows, 5 x 20 KB at 32,
// 4 x 24 KB at 64, 3 x 32 KB at 128) instead of one CTA with a 10-stage ring: 296 work slots make every decode GEMM
// of the 14B model a single, perfectly balanced wave (272-280 items), more bytes are in flight per
// SM (216 KB), and the next kernel's CTAs move into a slot as soon as ONE of the two CTAs on an SM
// has finished.  Step time 5.76 -> 5.33 ms (81 -> 88 % of the measured HBM peak).
constexpr int kTwoCtaRingBytes = 112 * 1024;  // a CTA with a ring this small shares its SM with a second one
constexpr bool two_cta(int bn, int stages) { return stages * (128 * 64 * 2 + bn * 64 * 2) <= kTwoCtaRingBytes; }
constexpr int kPreWaitStages = 6;  // weight slots requested before griddepcontrol.wait

struct GemmArgs {
  int M, N, K;              // logical problem
  int tiles_a, tiles_b;     // tiles along the A-operand rows / B-operand rows
  int k_blocks;             // ceil(K / 64)
  int splits;               // split-K factor (>=1)
  int group_a;              // raster group size along A tiles
  int swap;                 // 1: A-operand = W (n on TMEM lanes), B-operand = X
  int epi;                  // BCG_EPI_*
  void* out;
  long long ld_out;         // elements
  uint64_t hint_a, hint_b;
  int pdl_wait_b;           // 1: launched with programmatic stream serialization (see the producer)
  // ---- decode-only fusions of the residual add + RMSNorm that sits between two projections
  // consumer side: the activations are r*w (NOT yet divided by the row's RMS); every output row m is
  // multiplied by rsqrt(sum_t scale_sumsq[t][m] / scale_d + scale_eps) in the epilogue
  const float* scale_sumsq;  // [scale_tiles][kSumsqStride]
  int scale_tiles;
  int scale_d;
  float scale_eps;
  // producer side (split-K fp32 partials): the item that completes a 128-column tile adds the
  // partials (index order) to the residual, writes it back, writes bf16(r*w) and the tile's
  // per-row sum of squares
  float* red_residual;       // [M][N] fp32, N == ld_out
  const float* red_w;        // [N]
  __nv_bfloat16* red_xn;     // [M][N]
  float* red_sumsq;          // [tiles_a][kSumsqStride]
  int* red_counters;         // [tiles_a], zero on entry, zero again on exit
};

__device__ __forceinline__ float silu_mul(float g, float u) {
  return g * (1.0f / (1.0f + __expf(-g))) * u;
}

template <int BN, int STAGES>
__global__ void __launch_bounds__(kGemmThreads, two_cta(BN, STAGES) ? 2 : 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_a,
                    const __grid_constant__ CUtensorMap tmap_b, const GemmArgs args) {
  constexpr int A_BYTES = kBlockA * kBlockK * 2;
  constexpr int B_BYTES = BN * kBlockK * 2;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t TMEM_COLS = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128
                                 : (2 * BN <= 256) ? 256 : 512;
  constexpr uint32_t IDESC = make_idesc_bf16_m128(BN);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  __shared__ float row_scale[kSumsqStride];
  __shared__ int red_last;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_launch_dependents();
  if (threadIdx.x == 0) BCG_TL(1, 0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_b);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 128);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_items = args.tiles_a * args.tiles_b * args.splits;
  const int group_tiles = args.group_a * args.tiles_b;

  auto decode_item = [&](int item, int& ta, int& tb, int& split, int& kb0, int& kb1) {
    split = item % args.splits;
    const int t = item / args.splits;
    const int g = t / group_tiles;
    const int r = t - g * group_tiles;
    const int a0 = g * args.group_a;
    const int ga = min(args.group_a, args.tiles_a - a0);
    tb = r / ga;
    ta = a0 + (r - tb * ga);
    kb0 = static_cast<int>((static_cast<long long>(split) * args.k_blocks) / args.splits);
    kb1 = static_cast<int>((static_cast<long long>(split + 1) * args.k_blocks) / args.splits);
  };

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      // Under programmatic dependent launch this CTA may start while the previous kernel is still
      // running.  Weights never depend on it, activations do: so the weight halves of the first
      // kPreWaitStages ring slots are requested right away and only then does the producer wait
      // for the previous kernel and add the activation halves.  (Measured on the 14B decode step
      // with the two-CTA configuration: 1 slot 5.61 ms, 2: 5.57, 3: 5.47, 4: 5.45, all 6: 5.40.
      // With the earlier one-CTA 10-stage ring 6 of 10 was best; pulling further tiles into L2 with
      // cp.async.bulk.prefetch made the step monotonically slower, 6.16 -> 6.93 ms at 64 tiles.)
      int n_pre = 0;
      if (args.pdl_wait_b) {
        if (args.swap) {
          constexpr int pre = STAGES < kPreWaitStages ? STAGES : kPreWaitStages;
          int idx = 0;
          for (int item = blockIdx.x; item < total_items && idx < pre; item += gridDim.x) {
            int ta, tb, split, kb0, kb1;
            decode_item(item, ta, tb, split, kb0, kb1);
            for (int kb = kb0; kb < kb1 && idx < pre; ++kb, ++idx) {
              mbar_arrive_expect_tx(&full_bar[idx], STAGE_BYTES);
              tma_load_2d(smem + idx * STAGE_BYTES, &tmap_a, &full_bar[idx], kb * kBlockK,
                          ta * kBlockA, args.hint_a);
              n_pre = idx + 1;
            }
          }
        }
        pdl_wait();
      }
      BCG_TL(1, 1);
      int stage = 0;
      uint32_t phase = 0;
      int idx = 0;
      for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
        int ta, tb, split, kb0, kb1;
        decode_item(item, ta, tb, split, kb0, kb1);
        for (int kb = kb0; kb < kb1; ++kb, ++idx) {
          uint8_t* sa = smem + stage * STAGE_BYTES;
          if (idx >= n_pre) {
            mbar_wait(&empty_bar[stage], phase ^ 1);
            mbar_arrive_expect_tx(&full_bar[stage], STAGE_BYTES);
            tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * kBlockK, ta * kBlockA, args.hint_a);
          }
          tma_load_2d(sa + A_BYTES, &tmap_b, &full_bar[stage], kb * kBlockK, tb * BN,
                      args.hint_b);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
      BCG_TL(1, 2);
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer --------------------------------
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
      int ta, tb, split, kb0, kb1;
      decode_item(item, ta, tb, split, kb0, kb1);
      mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc * BN);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
          const uint64_t da = make_kmajor_sw128_desc(sa);
          const uint64_t db = make_kmajor_sw128_desc(sa + A_BYTES);
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
            // advance 16 bf16 = 32 B inside the 128 B swizzle atom: +2 in the (addr>>4) field
            umma_bf16(tmem_d, da + 2 * k, db + 2 * k, IDESC, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (kb == kb1 - 1) umma_commit(&tmem_full[acc]);
        }
        __syncwarp();
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  } else {
    // ------------------------------ epilogue ----------------------------------
    if (args.pdl_wait_b) pdl_wait();  // `out` may still be read by the previous kernel's consumers
    const int q = warp & 3;  // TMEM sub-partition this warp may read
    const int row = q * 32 + lane;
    const int etid = threadIdx.x - 64;  // 0..127 over the four epilogue warps
    if (args.scale_sumsq) {
      // 1/RMS of every activation row from the per-tile sums the producing GEMM left behind,
      // tiles added in index order (deterministic, the same for every CTA)
      if (etid < args.M) {
        float tot = 0.f;
        for (int t0 = 0; t0 < args.scale_tiles; t0 += 8) {
          float v[8];
#pragma unroll
          for (int t = 0; t < 8; ++t)
            v[t] = (t0 + t < args.scale_tiles) ? __ldcg(args.scale_sumsq + (t0 + t) * kSumsqStride + etid) : 0.f;
#pragma unroll
          for (int t = 0; t < 8; ++t)
            if (t0 + t < args.scale_tiles) tot += v[t];
        }
        row_scale[etid] = rsqrtf(tot / static_cast<float>(args.scale_d) + args.scale_eps);
      }
      asm volatile("bar.sync 2, 128;" ::: "memory");
    }
    int acc = 0;
    uint32_t acc_phase = 0;
    const int M = args.M, N = args.N;
    const long long ld = args.ld_out;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
      int ta, tb, split, kb0, kb1;
      decode_item(item, ta, tb, split, kb0, kb1);
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                             static_cast<uint32_t>(acc * BN);
      if (args.swap) {
        // lanes = weight rows (n), columns = activation rows (m)
        const int n = ta * kBlockA + row;
        const int m0 = tb * BN;
#pragma unroll 1
        for (int c = 0; c < BN; c += 16) {
          uint32_t v[16];
          tmem_ld_x16(taddr + c, v);
          tmem_ld_wait();
          if (args.scale_sumsq) {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              v[j] = __float_as_uint(__uint_as_float(v[j]) * row_scale[(m0 + c + j) & (kSumsqStride - 1)]);
          }
          if (args.epi == BCG_EPI_F32) {
            float* o = reinterpret_cast<float*>(args.out) + static_cast<long long>(split) * M * ld;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int m = m0 + c + j;
              if (m < M && n < N) o[m * ld + n] = __uint_as_float(v[j]);
            }
          } else if (args.epi == BCG_EPI_BF16) {
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(args.out);
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int m = m0 + c + j;
              if (m < M && n < N) o[m * ld + n] = __float2bfloat16_rn(__uint_as_float(v[j]));
            }
          } else {  // SWIGLU: rows interleaved [16 gate | 16 up]
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(args.out);
            const int nb = (ta * kBlockA + q * 32) / 2 + lane;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float g = __uint_as_float(v[j]);
              const float u = __shfl_down_sync(0xffffffffu, g, 16);
              const int m = m0 + c + j;
              if (lane < 16 && m < M && n < N) o[m * ld + nb] = __float2bfloat16_rn(silu_mul(g, u));
            }
          }
        }
      } else {
        // lanes = activation rows (m), columns = weight rows (n)
        const int m = ta * kBlockA + row;
        const int n0 = tb * BN;
        if (args.epi == BCG_EPI_SWIGLU) {
          __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(args.out);
#pragma unroll 1
          for (int c = 0; c < BN; c += 32) {
            uint32_t g[16], u[16];
            tmem_ld_x16(taddr + c, g);
            tmem_ld_x16(taddr + c + 16, u);
            tmem_ld_wait();
            uint32_t p[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              p[j] = pack_bf16(silu_mul(__uint_as_float(g[2 * j]), __uint_as_float(u[2 * j])),
                               silu_mul(__uint_as_float(g[2 * j + 1]), __uint_as_float(u[2 * j + 1])));
            }
            if (m < M && n0 + c < N) {
              uint4* dst = reinterpret_cast<uint4*>(o + m * ld + (n0 + c) / 2);
              dst[0] = make_uint4(p[0], p[1], p[2], p[3]);
              dst[1] = make_uint4(p[4], p[5], p[6], p[7]);
            }
          }
        } else {
#pragma unroll 1
          for (int c = 0; c < BN; c += 32) {
            uint32_t v0[16], v1[16];
            tmem_ld_x16(taddr + c, v0);
            tmem_ld_x16(taddr + c + 16, v1);
            tmem_ld_wait();
            if (args.epi == BCG_EPI_F32) {
              float* o = reinterpret_cast<float*>(args.out) +
                         static_cast<long long>(split) * M * ld;
              if (m < M) {
                float4* dst = reinterpret_cast<float4*>(o + m * ld + n0 + c);
                if (n0 + c < N) {
#pragma unroll
                  for (int j = 0; j < 4; ++j)
                    dst[j] = make_float4(__uint_as_float(v0[4 * j]), __uint_as_float(v0[4 * j + 1]),
                                         __uint_as_float(v0[4 * j + 2]), __uint_as_float(v0[4 * j + 3]));
                }
                if (n0 + c + 16 < N) {
#pragma unroll
                  for (int j = 0; j < 4; ++j)
                    dst[4 + j] = make_float4(__uint_as_float(v1[4 * j]), __uint_as_float(v1[4 * j + 1]),
                                             __uint_as_float(v1[4 * j + 2]), __uint_as_float(v1[4 * j + 3]));
                }
              }
            } else {
              __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(args.out);
              uint32_t p[16];
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                p[j] = pack_bf16(__uint_as_float(v0[2 * j]), __uint_as_float(v0[2 * j + 1]));
                p[8 + j] = pack_bf16(__uint_as_float(v1[2 * j]), __uint_as_float(v1[2 * j + 1]));
              }
              if (m < M) {
                uint4* dst = reinterpret_cast<uint4*>(o + m * ld + n0 + c);
                if (n0 + c < N) {
                  dst[0] = make_uint4(p[0], p[1], p[2], p[3]);
                  dst[1] = make_uint4(p[4], p[5], p[6], p[7]);
                }
                if (n0 + c + 16 < N) {
                  dst[2] = make_uint4(p[8], p[9], p[10], p[11]);
                  dst[3] = make_uint4(p[12], p[13], p[14], p[15]);
                }
              }
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
      if (args.red_counters) {
        // Which item completes this tile?  The four epilogue warps have stored their slab; one
        // thread publishes that (barrier -> fence -> counter) and learns whether all `splits`
        // slabs of the tile are now in L2.  Nobody waits for anybody: the last arrival reduces.
        asm volatile("bar.sync 2, 128;" ::: "memory");
        if (etid == 0) {
          fence_acq_rel_gpu();
          const int prev = atomicAdd(&args.red_counters[ta], 1);
          fence_acq_rel_gpu();
          red_last = (prev == args.splits - 1);
          if (red_last) args.red_counters[ta] = 0;  // every arrival of this launch has happened
        }
        asm volatile("bar.sync 2, 128;" ::: "memory");
        if (red_last) {
          // warp e owns rows e, e+4, ...; a lane owns 4 consecutive columns of the tile.  Loads
          // bypass L1 (other SMs wrote the data); one row's worth is in flight at a time.
          const int e = etid >> 5;
          const int n = ta * kBlockA + lane * 4;
          const float4 wv = *reinterpret_cast<const float4*>(args.red_w + n);
          const float* P = reinterpret_cast<const float*>(args.out);
          const long long slab = static_cast<long long>(M) * ld;
          for (int m = e; m < M; m += 4) {
            float4 p[16];
            // unconditional (clamped) loads so that all of them are in flight before the first add
            float4 x = __ldcg(reinterpret_cast<const float4*>(args.red_residual + m * ld + n));
#pragma unroll
            for (int sp = 0; sp < 16; ++sp)
              p[sp] = __ldcg(reinterpret_cast<const float4*>(P + min(sp, args.splits - 1) * slab + m * ld + n));
            float4 a = p[0];
#pragma unroll
            for (int sp = 1; sp < 16; ++sp) {
              if (sp < args.splits) {
                a.x += p[sp].x; a.y += p[sp].y; a.z += p[sp].z; a.w += p[sp].w;
              }
            }
            x.x += a.x; x.y += a.y; x.z += a.z; x.w += a.w;
            *reinterpret_cast<float4*>(args.red_residual + m * ld + n) = x;
            uint2 pk;
            pk.x = pack_bf16(x.x * wv.x, x.y * wv.y);
            pk.y = pack_bf16(x.z * wv.z, x.w * wv.w);
            *reinterpret_cast<uint2*>(args.red_xn + m * ld + n) = pk;
            float ss = __fmaf_rn(x.w, x.w, __fmaf_rn(x.z, x.z, __fmaf_rn(x.y, x.y, __fmul_rn(x.x, x.x))));
            ss = warp_sum(ss);
            if (lane == 0) args.red_sumsq[ta * kSumsqStride + m] = ss;
          }
        }
      }
    }
    if (threadIdx.x == 64) BCG_TL(1, 3);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) !=
          cudaSuccess ||
      qres != cudaDriverEntryPointSuccess) {
    return nullptr;
  }
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

// rows x K bf16, row pitch ld elements; box = 64 (K) x box_rows, 128B swizzle, OOB -> 0.
int make_tmap_bf16_sw128(CUtensorMap* m, const void* ptr, long long rows, long long K, long long ld,
                         int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) {
    set_last_error("cuTensorMapEncodeTiled", "driver entry point unavailable");
    return BCG_ERR_CUDA;
  }
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kBlockK), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled", "encode failed (alignment / stride?)");
    return BCG_ERR_CUDA;
  }
  return BCG_OK;
}

static int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

template <int BN, int STAGES>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& a, int pdl,
                       cudaStream_t stream) {
  constexpr int STAGE_BYTES = kBlockA * kBlockK * 2 + BN * kBlockK * 2;
  constexpr int SMEM = STAGES * STAGE_BYTES + 1024 + 256;
  if (const int rc = ensure_dynamic_smem(reinterpret_cast<const void*>(gemm_tcgen05_kernel<BN, STAGES>), SMEM))
    return rc;
  const int total = a.tiles_a * a.tiles_b * a.splits;
  cudaLaunchConfig_t cfg = {};
  const int slots = two_cta(BN, STAGES) ? 2 * num_sms() : num_sms();
  cfg.gridDim = dim3(static_cast<unsigned>(total < slots ? total : slots));
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = SMEM;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  BCG_CUDA_CHECK(cudaLaunchKernelEx(&cfg, gemm_tcgen05_kernel<BN, STAGES>, ta, tb, a));
  return BCG_OK;
}

// Split-K factor for the decode path.  Depends on (N, K) only — never on the number of rows — so
// a row's arithmetic is the same whatever batch it is in.  Work items = tiles x splits are dealt
// round-robin to the 296 work slots of the decode configuration (two CTAs per SM); the factor is
// chosen to fill whole waves (e.g. 40 tiles x 7 = 280 items = 0.95 wave, every CTA one item,
// instead of 40 x 8 = 320 = 1.08 waves -> 54 %), discounted by the fp32 partial traffic it adds
// (4*s*16/K of the weight bytes at 16 rows).  Batches of 33-128 rows (wider tiles, shorter rings,
// still two CTAs per SM) use the same factors.
int gemm_splits_for(int M, int N, int K) {
  if (M > 128) return 1;
  const int tiles = (N + kBlockA - 1) / kBlockA;
  const int kb = (K + kBlockK - 1) / kBlockK;
  const int sms = 2 * 148;  // work slots of the decode configuration (two CTAs per SM)
  int best = 1;
  double best_score = -1.0;
  for (int s = 1; s <= 16; ++s) {
    if (s > 1 && kb / s < 4) break;  // keep >= 4 k-blocks (32 KB of weights) per item
    const int items = tiles * s;
    const int rounds = (items + sms - 1) / sms;
    const double eff = static_cast<double>(items) / (static_cast<double>(rounds) * sms);
    const double extra = (s > 1) ? 4.0 * s * 16.0 / K : 0.0;
    const double score = eff / (1.0 + extra);
    if (score > best_score + 0.02) {  // need a real gain to split further
      best_score = score;
      best = s;
    }
  }
  return best;
}

int gemm_bf16_tn(const void* X, const void* W, void* out, int M, int N, int K, long long ldx,
                 long long ldw, long long ld_out, int epi, int splits, int flags,
                 cudaStream_t stream) {
  return gemm_bf16_tn_fused(X, W, out, M, N, K, ldx, ldw, ld_out, epi, splits, flags, nullptr, stream);
}

int gemm_bf16_tn_fused(const void* X, const void* W, void* out, int M, int N, int K, long long ldx,
                       long long ldw, long long ld_out, int epi, int splits, int flags,
                       const bcg_gemm_fusion* fu, cudaStream_t stream) {
  if (!X || !W || !out || M <= 0 || N <= 0 || K <= 0) {
    set_last_error("bcg_gemm_bf16_tn", "null pointer or non-positive dimension");
    return BCG_ERR_ARG;
  }
  if ((ldx % 8) || (ldw % 8) || (reinterpret_cast<uintptr_t>(X) & 15) ||
      (reinterpret_cast<uintptr_t>(W) & 15) || (reinterpret_cast<uintptr_t>(out) & 15)) {
    set_last_error("bcg_gemm_bf16_tn", "X/W/out must be 16-byte aligned, ldx/ldw multiples of 8");
    return BCG_ERR_ARG;
  }
  if (splits < 1) splits = 1;
  if (splits > 1 && epi != BCG_EPI_F32) {
    set_last_error("bcg_gemm_bf16_tn", "split-K needs the fp32 partial epilogue");
    return BCG_ERR_ARG;
  }
  if (epi == BCG_EPI_SWIGLU && (N % 32)) {
    set_last_error("bcg_gemm_bf16_tn", "SwiGLU epilogue needs N % 32 == 0");
    return BCG_ERR_ARG;
  }
  const bool swap = (M <= 128);
  if (!swap && ((epi == BCG_EPI_SWIGLU ? (ld_out % 8) : (epi == BCG_EPI_F32 ? (ld_out % 4) : (ld_out % 8))))) {
    set_last_error("bcg_gemm_bf16_tn", "ld_out must keep 16-byte row alignment");
    return BCG_ERR_ARG;
  }
  if (!swap && splits == 1 && !(fu && (fu->scale_sumsq || fu->red_counters))) {
    // prefill-sized: 256x256 tiles on CTA pairs when the shape allows it
    const int rc2 = gemm2sm_bf16_tn(X, W, out, M, N, K, ldx, ldw, ld_out, epi, stream);
    if (rc2 != BCG_ERR_UNSUPPORTED) return rc2;
  }
  GemmArgs a = {};
  a.M = M;
  a.N = N;
  a.K = K;
  a.k_blocks = (K + kBlockK - 1) / kBlockK;
  if (splits > a.k_blocks) splits = a.k_blocks;
  a.splits = splits;
  a.swap = swap ? 1 : 0;
  a.epi = epi;
  a.out = out;
  a.ld_out = ld_out;
  a.pdl_wait_b = ((flags | launch_flags()) & BCG_GEMM_PDL) ? 1 : 0;
  if (fu && (fu->scale_sumsq || fu->red_counters)) {
    if (!swap || (fu->red_counters && (N % kBlockA))) {
      set_last_error("bcg_gemm_bf16_tn_fused", "norm fusion needs M <= 128 (and N % 128 == 0 on the producer side)");
      return BCG_ERR_ARG;
    }
    if (fu->scale_sumsq) {
      if (fu->scale_tiles <= 0 || fu->scale_d <= 0) {
        set_last_error("bcg_gemm_bf16_tn_fused", "scale_tiles and scale_d must be positive");
        return BCG_ERR_ARG;
      }
      a.scale_sumsq = fu->scale_sumsq;
      a.scale_tiles = fu->scale_tiles;
      a.scale_d = fu->scale_d;
      a.scale_eps = fu->scale_eps;
    }
    if (fu->red_counters) {
      if (epi != BCG_EPI_F32 || splits > 16 || ld_out != N || !fu->red_residual || !fu->red_w ||
          !fu->red_xn || !fu->red_sumsq || (reinterpret_cast<uintptr_t>(fu->red_residual) & 15) ||
          (reinterpret_cast<uintptr_t>(fu->red_w) & 15) || (reinterpret_cast<uintptr_t>(fu->red_xn) & 7)) {
        set_last_error("bcg_gemm_bf16_tn_fused",
                       "producer fusion needs BCG_EPI_F32, splits <= 16, ld_out == N and aligned "
                       "residual / weight / xn / sumsq buffers");
        return BCG_ERR_ARG;
      }
      a.red_residual = fu->red_residual;
      a.red_w = fu->red_w;
      a.red_xn = reinterpret_cast<__nv_bfloat16*>(fu->red_xn);
      a.red_sumsq = fu->red_sumsq;
      a.red_counters = fu->red_counters;
    }
  }
  int bn;
  if (swap) {
    bn = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : 128;
    a.tiles_a = (N + kBlockA - 1) / kBlockA;
    a.tiles_b = (M + bn - 1) / bn;
    a.group_a = a.tiles_a;
    a.hint_a = kEvictFirst;  // weights are streamed exactly once per step
    a.hint_b = kEvictLast;
  } else {
    bn = (N % 256 == 0) ? 256 : 128;
    if (epi == BCG_EPI_F32 && bn == 256 && N < 2048) bn = 128;
    a.tiles_a = (M + kBlockA - 1) / kBlockA;
    a.tiles_b = (N + bn - 1) / bn;
    a.group_a = 16;
    if (a.group_a > a.tiles_a) a.group_a = a.tiles_a;
    a.hint_a = kEvictNormal;
    a.hint_b = kEvictLast;  // weight tiles are re-read by every m-group
  }
  CUtensorMap tma, tmb;
  int rc;
  if (swap) {
    if ((rc = make_tmap_bf16_sw128(&tma, W, N, K, ldw, kBlockA)) != BCG_OK) return rc;
    if ((rc = make_tmap_bf16_sw128(&tmb, X, M, K, ldx, bn)) != BCG_OK) return rc;
  } else {
    if ((rc = make_tmap_bf16_sw128(&tma, X, M, K, ldx, kBlockA)) != BCG_OK) return rc;
    if ((rc = make_tmap_bf16_sw128(&tmb, W, N, K, ldw, bn)) != BCG_OK) return rc;
  }
  const int pdl = a.pdl_wait_b;
  switch (bn) {
    case 16: return launch_gemm<16, 6>(tma, tmb, a, pdl, stream);
    case 32: return launch_gemm<32, 5>(tma, tmb, a, pdl, stream);
    case 64: return launch_gemm<64, 4>(tma, tmb, a, pdl, stream);   // decode, 33-64 rows: two CTAs per SM
    case 128: return swap ? launch_gemm<128, 3>(tma, tmb, a, pdl, stream)  // decode, 65-128 rows: two CTAs per SM
                          : launch_gemm<128, 6>(tma, tmb, a, pdl, stream);
    default: return launch_gemm<256, 4>(tma, tmb, a, pdl, stream);
  }
}

}  // namespace bcg

BCG_TL_DEFINE_DUMP(bcg_tl_dump_gemm)
