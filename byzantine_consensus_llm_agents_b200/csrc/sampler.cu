// Grammar-constrained sampler fused with the per-row decode
// bookkeeping, so the decode loop never returns to the host between tokens.
//
// One thread-block CLUSTER (8 CTAs x 512 threads) per row: each CTA scans 1/8 of the
// vocabulary with 128-bit loads, applies the token bitmask of the row's DFA state,
// computes the race score, reduces (score, index) with warp shuffles, and the eight CTA
// winners meet in CTA 0's shared memory through DSMEM.  (north_star words this as a
// "warp-per-row" kernel; at V = 151,936 and ~10 rows a single warp per row would leave
// the chip idle for ~100 us per token, so the row is spread over a cluster instead —
// same fusion, same arithmetic.)
//
// Arithmetic contract (restated bit-for-bit in oracle/sampler.py):
//   greedy (T < 1e-5):  argmax_i logit_i over allowed i, lowest index on ties
//   T >= 1e-5:          argmax_i ( logit_i / T  -  log(q_i) ),  q_i = -log(u_i)
//                       u_i = (k_i + 0.5) * 2^-23, k_i = top 23 bits of Philox4x32-10
//                       (key = seed[b], counter = (i>>2, step, 0, 0), lane i&3)
//   log() is rlog() below: IEEE +,*,/ only, each rounded to nearest (no FMA contraction),
//   so numpy float32 reproduces it exactly.
// This is vLLM's sampler (softmax(logits/T) / Exp(1) -> argmax, sampler.py:223-297,
// topk_topp_sampler.py:395-416) in the log domain; the RNG is the engine's own (H5).
#include <cooperative_groups.h>

#include "common.cuh"
#include "kernels.h"

namespace cg = cooperative_groups;

namespace bcg {

constexpr int kSampCluster = 8;
constexpr int kSampThreads = 512;

__device__ __forceinline__ float rlog(float x) {
  const uint32_t bits = __float_as_uint(x);
  int e = static_cast<int>(bits >> 23) - 127;
  float m = __uint_as_float((bits & 0x007fffffu) | 0x3f800000u);
  if (m > 1.41421356f) {
    m = __fmul_rn(m, 0.5f);
    e += 1;
  }
  const float f = __fadd_rn(m, -1.0f);
  const float s = __fdiv_rn(f, __fadd_rn(2.0f, f));
  const float z = __fmul_rn(s, s);
  float p = __fadd_rn(0.22222222f, __fmul_rn(z, 0.18181818f));
  p = __fadd_rn(0.2857143f, __fmul_rn(z, p));
  p = __fadd_rn(0.4f, __fmul_rn(z, p));
  p = __fadd_rn(0.6666667f, __fmul_rn(z, p));
  const float r = __fadd_rn(__fmul_rn(2.0f, s), __fmul_rn(s, __fmul_rn(z, p)));
  return __fadd_rn(__fmul_rn(static_cast<float>(e), 0.6931472f), r);
}

__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                               uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ float race_score(float logit, float temp, uint32_t rnd) {
  const float u = __fmul_rn(__fadd_rn(static_cast<float>(rnd >> 9), 0.5f), 1.1920929e-07f);
  const float qexp = -rlog(u);
  return __fadd_rn(__fdiv_rn(logit, temp), -rlog(qexp));
}

struct Best {
  float s;
  int i;
};
__device__ __forceinline__ Best better(Best a, Best b) {
  return (b.s > a.s || (b.s == a.s && b.i < a.i)) ? b : a;
}

struct SamplerDev {
  bcg_sampler_args a;
  void* const* peers;   // symmetric buffers of the tensor-parallel replica (a.tp_size > 1), else null
};

// Vocab-parallel sampling (tensor parallelism): every rank scans only ITS shard of the vocabulary
// and the ranks exchange one (score, token id) candidate per row instead of gathering [B, V] logits.
// A candidate travels as two 8-byte words {score bits, epoch} {token id, epoch} written with single
// 64-bit stores into every peer's symmetric buffer; the reader spins until both words carry the
// row's current epoch (the per-row launch counter every rank keeps in lock step), so no separate
// flag or fence is needed.  Two parity slots per (source rank, row).
constexpr int kCandOffEpoch = 2 * 1024 + 2 * 8 * 1024;      // after tp_reduce_norm.cu's counters and flags
constexpr int kCandOff = kCandOffEpoch + 128;               // [2][8][128][4] words
__device__ __forceinline__ void st_sys_v2(int* p, int a, int b) {
  asm volatile("st.relaxed.sys.global.v2.s32 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ int2 ld_sys_v2(const int* p) {
  int2 r;
  asm volatile("ld.relaxed.sys.global.v2.s32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory");
  return r;
}

__global__ void __cluster_dims__(kSampCluster, 1, 1) __launch_bounds__(kSampThreads)
mask_sample_kernel(const SamplerDev p) {
  const bcg_sampler_args& a = p.a;
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ Best warp_best[kSampThreads / 32];
  __shared__ Best cta_best[kSampCluster];  // only CTA 0's copy is used
  // No early griddepcontrol.launch_dependents here: this kernel writes seq_lens / slot_map, which
  // the next step's decode attention reads before its own dependency wait (attn_decode.cu).
  pdl_wait();

  const int row = blockIdx.y;
  const unsigned rank = cluster.block_rank();
  const bool row_done = a.finished[row] != 0;

  Best best = {-INFINITY, 0x7fffffff};
  if (!row_done) {
    const int gid = a.grammar_id[row];
    const int32_t* mask = nullptr;
    if (gid >= 0) mask = a.mask_tables[gid] + static_cast<long long>(a.state[row]) * a.mask_words;
    const float temp = a.temperature[row];
    const bool greedy = temp < 1e-5f;
    const uint32_t k0 = a.seed[2 * row], k1 = a.seed[2 * row + 1];
    const uint32_t step = static_cast<uint32_t>(a.n_out[row]);
    const float* lg = a.logits + static_cast<long long>(row) * (a.shard_v > 0 ? a.shard_v : a.ld_logits);
    // slice in units of 128 elements so float4 chunks never straddle; a group of 4 tokens starts at
    // a multiple of 4 of the GLOBAL id (vocab_base % 4 == 0), so its mask bits sit in one word
    const int per = ((a.V + kSampCluster * 128 - 1) / (kSampCluster * 128)) * 128;
    const int i0 = static_cast<int>(rank) * per;
    const int i1 = min(a.V, i0 + per);
    for (int i = i0 + 4 * threadIdx.x; i < i1; i += 4 * kSampThreads) {
      const float4 x = *reinterpret_cast<const float4*>(
          a.shard_v > 0 ? lg + static_cast<long long>(i / a.shard_v) * a.shard_stride + i % a.shard_v : lg + i);
      const int gi = a.vocab_base + i;   // token id (this rank's shard starts at vocab_base)
      uint32_t bits = 0xFu;
      if (mask) bits = (static_cast<uint32_t>(mask[gi >> 5]) >> (gi & 31)) & 0xFu;
      if (bits == 0) continue;
      float s0 = x.x, s1 = x.y, s2 = x.z, s3 = x.w;
      if (!greedy) {
        const uint4 r = philox4x32_10(static_cast<uint32_t>(gi >> 2), step, 0u, 0u, k0, k1);
        s0 = race_score(x.x, temp, r.x);
        s1 = race_score(x.y, temp, r.y);
        s2 = race_score(x.z, temp, r.z);
        s3 = race_score(x.w, temp, r.w);
      }
      if (bits & 1u) best = better(best, Best{s0, gi});
      if (bits & 2u) best = better(best, Best{s1, gi + 1});
      if (bits & 4u) best = better(best, Best{s2, gi + 2});
      if (bits & 8u) best = better(best, Best{s3, gi + 3});
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    Best other;
    other.s = __shfl_xor_sync(0xffffffffu, best.s, o);
    other.i = __shfl_xor_sync(0xffffffffu, best.i, o);
    best = better(best, other);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) warp_best[warp] = best;
  __syncthreads();
  if (threadIdx.x == 0) {
    Best b = warp_best[0];
    for (int w = 1; w < kSampThreads / 32; ++w) b = better(b, warp_best[w]);
    Best* dst = cluster.map_shared_rank(cta_best, 0);
    dst[rank] = b;
  }
  cluster.sync();

  if (rank == 0 && threadIdx.x == 0 && !row_done) {
    Best b = cta_best[0];
    for (int r = 1; r < kSampCluster; ++r) b = better(b, cta_best[r]);
    if (a.tp_size > 1) {
      // b is the best of this rank's vocabulary shard: trade it for every peer's and keep the winner
      // (same total order on every rank: higher score, then lower token id)
      int* local = reinterpret_cast<int*>(p.peers[a.tp_rank]);
      const int ep = local[kCandOffEpoch + row] + 1;
      local[kCandOffEpoch + row] = ep;
      const int par = ep & 1;
      for (int pr = 0; pr < a.tp_size; ++pr) {
        if (pr == a.tp_rank) continue;
        int* dst = reinterpret_cast<int*>(p.peers[pr]) + kCandOff + ((par * 8 + a.tp_rank) * 128 + row) * 4;
        st_sys_v2(dst, __float_as_int(b.s), ep);
        st_sys_v2(dst + 2, b.i, ep);
      }
      for (int src = 0; src < a.tp_size; ++src) {
        if (src == a.tp_rank) continue;
        const int* in = local + kCandOff + ((par * 8 + src) * 128 + row) * 4;
        int2 w0, w1;
        unsigned spins = 0;
        do {
          w0 = ld_sys_v2(in);
          w1 = ld_sys_v2(in + 2);
          if (++spins > (1u << 24)) __trap();   // a peer died: fail loudly instead of hanging
        } while (w0.y != ep || w1.y != ep);
        b = better(b, Best{__int_as_float(w0.x), w1.x});
      }
    }
    const int step = a.n_out[row];
    int fin = 0;
    int chosen = b.i;
    if (chosen == 0x7fffffff) {  // nothing allowed: grammar dead end
      chosen = a.eos_id0 >= 0 ? a.eos_id0 : 0;
      fin = 3;
    }
    int tok = chosen;
    if (a.forced) {
      const int f = a.forced[static_cast<long long>(row) * a.max_steps + step];
      if (f >= 0) tok = f;
    }
    if (a.chosen) a.chosen[static_cast<long long>(row) * a.max_steps + step] = chosen;
    a.out_tokens[static_cast<long long>(row) * a.max_steps + step] = tok;
    if (tok == a.eos_id0 || tok == a.eos_id1) {
      if (fin == 0) fin = 1;
    } else {
      const int gid = a.grammar_id[row];
      if (gid >= 0) {
        const int16_t* tr = a.trans_tables[gid];
        int st = a.state[row];
        const int b0 = a.tok_offsets[tok], b1 = a.tok_offsets[tok + 1];
        if (b1 == b0) st = -1;  // byte-less (special) token inside a grammar
        for (int k = b0; k < b1 && st >= 0; ++k) st = tr[st * 256 + a.tok_bytes[k]];
        if (st < 0) {
          fin = 3;
          st = 0;
        }
        a.state[row] = st;
      }
    }
    const int n = step + 1;
    a.n_out[row] = n;
    if (fin == 0 && n >= a.max_tokens[row]) fin = 2;
    if (fin == 0 && n >= a.max_steps) fin = 2;
    a.next_ids[row] = tok;
    if (fin == 0) {
      const int pos = a.seq_lens[row];
      a.positions[row] = pos;
      a.seq_lens[row] = pos + 1;
      a.slot_map[row] =
          a.block_table[static_cast<long long>(row) * a.max_blocks_per_seq + pos / a.page_size] *
              a.page_size + pos % a.page_size;
    } else {
      a.finished[row] = fin;
      a.slot_map[row] = -1;
    }
  }
}

// n_active[0] = number of rows still running (one small block; lets the host poll one int).
__global__ void count_active_kernel(const int* __restrict__ finished, int B, int* n_active) {
  // No early griddepcontrol.launch_dependents here: this kernel writes seq_lens / slot_map, which
  // the next step's decode attention reads before its own dependency wait (attn_decode.cu).
  pdl_wait();
  int c = 0;
  for (int i = threadIdx.x; i < B; i += 32) c += (finished[i] == 0);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (threadIdx.x == 0) *n_active = c;
}

int mask_sample(const bcg_sampler_args* a, cudaStream_t stream) {
  if (!a || !a->logits || a->B <= 0 || a->V <= 0 || (a->V % 4) || (a->ld_logits % 4) ||
      !a->grammar_id || !a->temperature || !a->seed || !a->max_tokens || !a->state || !a->n_out ||
      !a->finished || !a->out_tokens || !a->next_ids || !a->positions || !a->seq_lens ||
      !a->slot_map || !a->block_table || !a->tok_offsets || !a->tok_bytes || a->page_size <= 0 ||
      a->mask_words < (a->vocab_base + a->V + 31) / 32 || (a->shard_v > 0 && (a->shard_v % 4)) ||
      a->vocab_base < 0 || (a->vocab_base % 4)) {
    set_last_error("bcg_mask_sample", "bad argument (null pointer, V % 4, vocab_base % 4, or mask_words too small)");
    return BCG_ERR_ARG;
  }
  SamplerDev p;
  p.a = *a;
  p.peers = nullptr;
  if (a->tp_size > 1) {
    if (a->tp_size > 8 || a->tp_rank < 0 || a->tp_rank >= a->tp_size || a->B > 128 || !tp_symm_peers_device() ||
        tp_size() != a->tp_size) {
      set_last_error("bcg_mask_sample", "vocab-parallel sampling needs bcg_tp_init + the symmetric buffer, "
                                        "tp_size <= 8 and at most 128 rows");
      return BCG_ERR_ARG;
    }
    p.peers = tp_symm_peers_device();
  }
  BCG_CUDA_CHECK(launch_k(mask_sample_kernel, dim3(kSampCluster, a->B), dim3(kSampThreads), 0, stream, p));
  if (a->n_active) {
    BCG_CUDA_CHECK(launch_k(count_active_kernel, dim3(1), dim3(32), 0, stream,
                            static_cast<const int*>(a->finished), a->B, a->n_active));
  }
  return BCG_OK;
}

}  // namespace bcg
