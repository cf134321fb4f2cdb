/* bcg_b200.h — C ABI of the B200-native agent-step engine kernels.
 *
 * The reference (leorugli/byzantine-consensus-llm-agents) has no FFI of its own: its hot
 * path is `VLLMAgent.generate_json / batch_generate_json`
 * (byzantine_consensus_game/vllm_agent.py:294-455), which hands everything to
 * `vllm.LLM.generate` (vllm_agent.py:195,331,430,502).  The Python class in
 * byzantine_consensus_llm_agents_b200/vllm_agent.py replaces that module one-for-one; the
 * functions below are what that class binds (ctypes) in place of the third-party engine's
 * CUDA ops.  Each entry cites the engine op it stands in for.
 *
 * Conventions: plain pointers and sizes only; every pointer is DEVICE memory owned by the
 * caller unless the comment says host; `stream` is a cudaStream_t passed as void*; no entry
 * point allocates or frees device memory; return value 0 = ok, otherwise a BCG_ERR_* code
 * and bcg_last_error() holds a message.  Re-entrant per stream.  sm_100a only.
 */
#ifndef BCG_B200_H_
#define BCG_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BCG_ABI_VERSION 2

/* status codes */
#define BCG_STATUS_OK 0
#define BCG_STATUS_ARG 1
#define BCG_STATUS_CUDA 2
#define BCG_STATUS_UNSUPPORTED 3
#define BCG_STATUS_TOO_LARGE 4

/* GEMM epilogues */
#define BCG_EPI_BF16 0   /* out bf16 [M, ld_out]                                         */
#define BCG_EPI_F32 1    /* out fp32 [splits][M, ld_out] (split-K partials, or logits)   */
#define BCG_EPI_SWIGLU 2 /* W rows interleaved [16 gate | 16 up]; out bf16 [M, N/2]      */
/* GEMM flags */
#define BCG_GEMM_PDL 1   /* programmatic dependent launch: weights prefetch before the wait */

int bcg_abi_version(void);
/* Last error message of the calling thread's most recent failing call (host string). */
const char* bcg_last_error(void);

/* Y[M,N] = X[M,K] · W[N,K]^T, bf16 inputs, fp32 accumulation on tcgen05 tensor cores with
 * TMEM accumulators and TMA-fed operands.  Stands in for the cuBLAS `F.linear` calls of
 * qkv_proj / o_proj / gate_up_proj (+ `_C.silu_and_mul`) / down_proj / lm_head
 * (vLLM model_executor/models/qwen3.py:145-162,216-236).
 * M <= 128 selects the weight-streaming (swap-AB, optional split-K) decode path.
 * ldx, ldw, ld_out are row pitches in elements.  splits > 1 requires BCG_EPI_F32 and an
 * `out` buffer of splits*M*ld_out floats; the consumer sums the partials in index order. */
int bcg_gemm_bf16_tn(const void* X, const void* W, void* out, int M, int N, int K, long long ldx,
                     long long ldw, long long ld_out, int epilogue, int splits, int flags,
                     void* stream);
/* Decode-only fusion of the `fused_add_rms_norm` that vLLM runs between two projections
 * (vLLM model_executor/layers/layernorm.py RMSNorm.forward_cuda with a residual;
 * models/qwen3.py:216-236 decoder layer) into the projections on either side of it, so that the
 * norm costs no kernel of its own.  M <= 128, N % 128 == 0, ld_out == N for the producer side.
 *   producer (red_counters != NULL; BCG_EPI_F32, splits <= 16): the work item that completes a
 *     128-column tile adds the split-K partials (index order) to red_residual, writes it back,
 *     writes red_xn = bf16(residual * red_w) — NOT yet divided by the row's RMS — and the tile's
 *     per-row sum of squares to red_sumsq[tile][128].  red_counters: N/128 ints, zero on entry,
 *     zero again on exit.
 *   consumer (scale_sumsq != NULL): X is such a red_xn; every output row m is multiplied by
 *     rsqrt(sum_t scale_sumsq[t][m] / scale_d + scale_eps) in the epilogue (tiles added in order).
 * Either half may be unused (NULL). */
typedef struct bcg_gemm_fusion {
  const float* scale_sumsq;
  int scale_tiles;
  int scale_d;
  float scale_eps;
  float* red_residual;
  const float* red_w;
  void* red_xn;
  float* red_sumsq;
  int* red_counters;
} bcg_gemm_fusion;
int bcg_gemm_bf16_tn_fused(const void* X, const void* W, void* out, int M, int N, int K,
                           long long ldx, long long ldw, long long ld_out, int epilogue, int splits,
                           int flags, const bcg_gemm_fusion* fusion, void* stream);
/* Split-K factor the engine uses for a decode GEMM of this shape (host helper, no GPU work). */
int bcg_gemm_splits_for(int M, int N, int K);

/* residual[t,:] = embed[ids[t],:] (fp32); xn[t,:] = rmsnorm(residual[t,:]) * w  (bf16).
 * Stands in for `F.embedding` + `_C.rms_norm`. */
int bcg_embed_rmsnorm(const int* ids, const void* embed, const float* w, float* residual, void* xn,
                      int T, int d, float eps, void* stream);

/* residual += delta; xn = rmsnorm(residual) * w.  delta is bf16 [T,d] (delta_is_f32 = 0) or
 * n_partials fp32 slabs [T,d] spaced partial_stride floats apart, summed in index order;
 * delta may be NULL.  If src_rows != NULL, output row i is computed from row src_rows[i] and
 * residual is not written (last-token gather in front of lm_head).
 * Stands in for `_C.fused_add_rms_norm`. */
int bcg_add_rmsnorm(float* residual, const void* delta, int delta_is_f32, int n_partials,
                    long long partial_stride, const float* w, void* xn, const int* src_rows, int T,
                    int d, float eps, void* stream);

/* Decode-only variant of bcg_add_rmsnorm with the 1/RMS factor deferred to the consuming GEMM:
 * residual += delta (same operand rules), xn = bf16(residual * w) — NOT divided by the RMS — and
 * sumsq[s * 128 + t] = sum of squares of segment s (of 8) of row t.  Pass `sumsq` as
 * bcg_gemm_fusion.scale_sumsq with scale_tiles = 8, scale_d = d to the projection that reads xn.
 * T <= 128.  Takes the row-wide reduction (a cluster barrier and a DSMEM exchange in
 * bcg_add_rmsnorm) off the critical path between two GEMMs.  Stands in for the same vllm op,
 * `_C.fused_add_rms_norm`, with its last multiply moved into the next projection. */
int bcg_add_norm_stats(float* residual, const void* delta, int delta_is_f32, int n_partials,
                       long long partial_stride, const float* w, void* xn, float* sumsq, int T, int d,
                       void* stream);

/* Per-head q/k RMSNorm (weights may be NULL: Mistral), rotate-half RoPE from a
 * [max_pos][128] fp32 table (cos[0..63] | sin[0..63]), q -> bf16 [T, n_heads*128], k/v ->
 * paged cache at slot_map[t] (slot = page*page_size + offset; negative = skip).
 * qkv is bf16 [T, ld_qkv] or fp32 split-K partials.  head_dim is 128.
 * KV cache layout (one layer): [page][2][n_kv_heads][page_size][128] bf16.
 * Stands in for `_C.rms_norm` on heads + `_C.rotary_embedding` + `reshape_and_cache_flash`
 * (SURVEY.md K4-K6). */
int bcg_qknorm_rope_kvwrite(const void* qkv, int qkv_is_f32, int n_partials,
                            long long partial_stride, long long ld_qkv, const float* q_norm_w,
                            const float* k_norm_w, const float* cos_sin, const int* positions,
                            const int* slot_map, void* q_out, void* kv_cache, int T, int n_heads,
                            int n_kv_heads, int page_size, float eps, void* stream);

/* act[t, j] = silu(gate) * up summed over fp32 split-K partials of the interleaved gate_up
 * GEMM (only needed when that GEMM ran with splits > 1).  Stands in for `_C.silu_and_mul`. */
int bcg_swiglu_reduce(const float* partials, int n_partials, long long partial_stride, void* act,
                      int T, int two_f, void* stream);

/* One query token per row over the paged KV cache (GQA, head_dim 128, page_size 16).
 * seq_lens[b] counts the current token.  scratch_o: B*n_heads*max_splits*128 floats;
 * scratch_ml: B*n_heads*max_splits*2 floats; chunk_counters: B*n_kv_heads int32 that must be
 * ZERO before the first call and that every call leaves zero again (keep them in a buffer of
 * their own: partial results written by a call of another shape must never land on them);
 * max_splits >= ceil(max seq_len / 256) — a longer row makes the kernel trap (the launch fails with
 * a CUDA error) rather than leave its output unwritten.
 * Stands in for the FlashInfer paged decode kernel vLLM dispatches on sm_100. */
int bcg_decode_attn_paged(const void* q, const void* kv_cache, const int* block_table,
                          int max_blocks_per_seq, const int* seq_lens, void* out, float* scratch_o,
                          float* scratch_ml, int* chunk_counters, int B, int n_heads, int n_kv_heads,
                          int page_size, int max_splits, float scale, void* stream);

/* bcg_qknorm_rope_kvwrite + bcg_decode_attn_paged in ONE launch (decode only): takes the qkv
 * projection's fp32 split-K slabs [n_partials][B, ld_qkv] (q heads | k heads | v heads) and the
 * RoPE / norm / slot arguments of bcg_qknorm_rope_kvwrite.  Every chunk CTA rebuilds its queries
 * itself; the CTA of a row's last chunk also builds the new token's K/V row, writes it to
 * kv_cache (slot_map[b], negative = skip) and patches it into the page it streamed in.  Stands in
 * for the same vLLM ops as the two entry points it replaces (`_C.rotary_embedding`,
 * `reshape_and_cache_flash`, FlashInfer paged decode).  Other arguments as
 * for bcg_decode_attn_paged. */
int bcg_decode_attn_paged_qkv(const float* qkv, int n_partials, long long partial_stride,
                              long long ld_qkv, const float* q_norm_w, const float* k_norm_w,
                              const float* cos_sin, const int* positions, const int* slot_map,
                              float eps, void* kv_cache, const int* block_table,
                              int max_blocks_per_seq, const int* seq_lens, void* out, float* scratch_o,
                              float* scratch_ml, int* chunk_counters, int B, int n_heads,
                              int n_kv_heads, int page_size, int max_splits, float scale,
                              void* stream);

/* Same computation as bcg_decode_attn_paged, for callers that cannot allocate device memory.
 * No counterpart in the reference (the reference never manages device memory itself); it exists
 * for hosts of this library that have no memory left to give, or no safe moment to ask for it:
 * a server whose whole device budget already went to weights and KV pages, or a decode loop that
 * must avoid cudaMalloc/cudaFree because they synchronise the device.  The scratch is a 16 MiB
 * array inside the library (held by every GPU that loads it, until exit), so the problem must
 * satisfy B * n_heads * max_splits * 130 <= 4 * 1024 * 1024 floats and B * n_kv_heads <= 4096;
 * larger calls return BCG_STATUS_TOO_LARGE, and every seq_lens[b] must be <= 256 * max_splits (a longer
 * row traps, as above).  The first call on a device resolves the array's address — make it
 * before any stream capture.  Calls share the array: keep them on one stream (overlapping calls
 * race); the caller never has to clear anything (the chunk counters inside start at zero and every
 * call returns them to zero).  Allocates nothing, frees nothing. */
int bcg_decode_attn_paged_static(const void* q, const void* kv_cache, const int* block_table,
                                 int max_blocks_per_seq, const int* seq_lens, void* out, int B,
                                 int n_heads, int n_kv_heads, int page_size, int max_splits,
                                 float scale, void* stream);

/* Causal attention of ragged new-token rows against cached prefix + new tokens, both read
 * through the block table.  q and out are bf16 [n_q_rows, n_heads*128]; tile_info: int32
 * [n_tiles][4] = {sequence, first q row, position of that row in its sequence, rows (<= 128)}.
 * Both contractions run on tcgen05 tensor cores (accumulators in TMEM); q tiles and KV pages are
 * fetched by TMA through tensor maps over the caller's q buffer and KV cache.  Every KV page a tile
 * reads (pages 0 .. ceil((position + rows) / 16) - 1 of its sequence's block-table row) must hold
 * finite values, including the unwritten tail of the last page (zero-initialise the cache).
 * Stands in for the FlashInfer paged prefill kernel. */
int bcg_prefill_attn_paged(const void* q, const void* kv_cache, const int* block_table,
                           int max_blocks_per_seq, const int* tile_info, int n_tiles, void* out,
                           int n_q_rows, int n_heads, int n_kv_heads, int page_size, float scale,
                           void* stream);

/* Grammar-constrained sampler + per-row decode bookkeeping, one thread-block cluster per row.
 * Applies the JSON-schema token bitmask of the row's DFA state, then greedy argmax
 * (temperature < 1e-5) or temperature sampling as an exponential race
 * argmax_i(logit_i / T - log(q_i)), q_i = -log(u_i), u_i from Philox4x32-10(seed[b]; i, step),
 * with a reproducible fp32 log (see oracle/sampler.py); advances the DFA over the token's
 * bytes; appends the token; updates next_ids / positions / seq_lens / slot_map so the next
 * decode step needs no host round-trip.
 * Stands in for xgrammar `apply_token_bitmask_inplace` + vLLM `Sampler` + the scheduler's
 * accept_token/stop bookkeeping (vllm v1/sample/sampler.py:91,223-297,
 * v1/sample/ops/topk_topp_sampler.py:395-416, v1/structured_output/backend_xgrammar.py:148-192). */
typedef struct bcg_sampler_args {
  const float* logits;     /* [B, ld_logits]; or vocab-sharded, see shard_v */
  long long ld_logits;
  int B;
  int V;                   /* multiple of 4 */
  int mask_words;          /* int32 words per mask row = ceil(V/32) */
  int max_steps;           /* row pitch of out_tokens / chosen / forced */
  int eos_id0, eos_id1;    /* stop ids (-1 = unused) */
  int page_size;
  int max_blocks_per_seq;
  /* grammar tables */
  const int* grammar_id;              /* [B], -1 = unconstrained */
  const int32_t* const* mask_tables;  /* [n_grammars] -> int32 [n_states][mask_words], bit=1 allowed, LSB first */
  const int16_t* const* trans_tables; /* [n_grammars] -> int16 [n_states][256], -1 = reject */
  const uint8_t* tok_bytes;           /* concatenated token byte strings */
  const int* tok_offsets;             /* [V+1] */
  /* per-row sampling state */
  const float* temperature;  /* [B] */
  const uint32_t* seed;      /* [B][2] */
  const int* max_tokens;     /* [B] */
  const int* forced;         /* [B, max_steps] teacher-forced ids (<0 = none) or NULL */
  int* state;                /* [B] DFA state, in/out */
  int* n_out;                /* [B] tokens emitted so far, in/out */
  int* finished;             /* [B] in/out: 0 running, 1 eos, 2 length, 3 grammar dead end */
  int* out_tokens;           /* [B, max_steps] accepted tokens */
  int* chosen;               /* [B, max_steps] the sampler's own pick (== out_tokens unless forced) or NULL */
  /* decode bookkeeping */
  int* next_ids;             /* [B] */
  int* positions;            /* [B] */
  int* seq_lens;             /* [B] */
  int* slot_map;             /* [B] */
  const int* block_table;    /* [B, max_blocks_per_seq] */
  int* n_active;             /* [1] number of unfinished rows after this call, or NULL */
  /* vocab-parallel logits as NCCL all-gather leaves them: shard_v > 0 means token i of row b is at
   * logits[(i / shard_v) * shard_stride + b * shard_v + i % shard_v]; 0 = plain [B, ld_logits] */
  int shard_v;
  long long shard_stride;
  /* vocab-parallel sampling without moving logits (tensor parallelism): tp_size > 1
   * means `logits` holds only this rank's shard — V tokens starting at token id vocab_base (a
   * multiple of 4; mask rows still span the whole vocabulary) — and the ranks trade one
   * (score, token id) candidate per row through the symmetric buffer (bcg_tp_symm_create / _open)
   * inside the kernel; every rank ends with the same token and does the same bookkeeping.
   * B <= 128.  0 or 1 = single GPU. */
  int tp_rank, tp_size;
  int vocab_base;
} bcg_sampler_args;

/* Launches the sampler described above on `stream` (stands in for xgrammar's mask apply + vLLM's
 * Sampler + accept_token bookkeeping).  `args` is a HOST struct of device pointers. */
int bcg_mask_sample(const bcg_sampler_args* args, void* stream);

/* Tensor parallelism (Qwen3-32B TP=8, Mistral-22B TP=2 in the reference's presets): one process per
 * GPU, one NCCL communicator per process.  Rank 0 obtains an id with bcg_tp_unique_id, ships the
 * 128 bytes to the other ranks by any means, everyone calls bcg_tp_init (collective).  The model
 * steps then all-reduce the outputs of the row-parallel o / down projections (2 per layer); the
 * vocab-parallel logits stay sharded and the sampler exchanges one candidate per row, standing in
 * for vLLM's TP collectives
 * (vllm_agent.py:131,141-142).  libnccl.so.2 is resolved at run time. */
int bcg_tp_unique_id(void* out_128_bytes /* host */);
int bcg_tp_init(const void* id_128_bytes /* host */, int rank, int size);
int bcg_tp_shutdown(void);
/* The symmetric buffer every rank of the replica maps (cudaIpc over NVLink).  It carries the
 * sampler's per-row candidate exchange (bcg_sampler_args.tp_size > 1) and, when
 * bcg_workspace.tp_region_floats > 0, the decode step's two per-layer all-reduces, fused into the
 * norm-statistics kernel that follows each row-parallel projection (bcg_tp_add_norm_stats: {value,
 * epoch} pushes over NVLink, no flag, no fence, no NCCL launch); with tp_region_floats == 0 the decode step all-reduces with
 * NCCL.  create: cudaMalloc's `bytes` (the only allocation this library ever makes; freed by
 * bcg_tp_shutdown) and returns the 64-byte IPC handle; open: takes all ranks' handles in rank
 * order.  bytes = bcg_tp_symm_bytes_for(region, size) = 4 * (32768 + 4 * size * region) with
 * region = tp_region_floats >= B_max * d_model (0 if only the sampler uses the buffer): two slots,
 * one inbox per source rank, every value stored beside the launch's epoch. */
long long bcg_tp_symm_bytes_for(long long region_floats, int size);
int bcg_tp_symm_create(long long bytes, void* handle_out_64_bytes /* host */);
int bcg_tp_symm_open(const void* handles_size_x_64_bytes /* host */, int rank, int size);
/* Tensor-parallel counterpart of bcg_add_norm_stats (decode, T <= 128): `partials` are THIS rank's
 * n_partials fp32 split-K slabs [T, d] of a row-parallel projection (o_proj / down_proj); the kernel
 * adds them, exchanges the reduced [T, d] message with the other ranks through the symmetric buffer
 * (slot 0 or 1, alternate between consecutive calls), adds the ranks' contributions in rank order
 * (bit-identical on every rank), then residual += sum, xn = bf16(residual * w) and
 * sumsq[s * 128 + t] as bcg_add_norm_stats.  Collective: every rank of the replica must make the same
 * sequence of calls.  Stands in for vLLM's tensor_model_parallel_all_reduce after RowParallelLinear
 * + `_C.fused_add_rms_norm`. */
int bcg_tp_add_norm_stats(float* residual, const float* partials, int n_partials, long long partial_stride,
                          const float* w, void* xn, float* sumsq, int slot, long long region_floats, int T,
                          int d, void* stream);

/* Whole-model steps: the layer stack launched from C++ on one stream so a decode step is one
 * host call (captured into a CUDA graph by the engine).  See DESIGN.md §4. */
typedef struct bcg_model {
  /* With tensor parallelism n_heads, n_kv_heads, d_ff and vocab are THIS RANK's shard sizes
   * (column-parallel qkv / gate_up, row-parallel o / down, vocab-parallel lm_head). */
  int n_layers, d_model, n_heads, n_kv_heads, head_dim, d_ff, vocab;
  float rms_eps;
  int page_size;
  const void* embed;             /* bf16 [vocab, d_model] */
  const void* lm_head;           /* bf16 [vocab, d_model] */
  const float* final_norm;       /* [d_model] */
  const float* const* ln1;       /* host array [n_layers] of device [d_model] */
  const float* const* ln2;
  const void* const* w_qkv;      /* bf16 [(n_heads+2*n_kv_heads)*128, d_model] */
  const void* const* w_o;        /* bf16 [d_model, n_heads*128] */
  const void* const* w_gate_up;  /* bf16 [2*d_ff, d_model], rows interleaved 16 gate / 16 up */
  const void* const* w_down;     /* bf16 [d_model, d_ff] */
  const float* const* q_norm;    /* [128] per layer or NULL array (Mistral) */
  const float* const* k_norm;
  void* const* kv_cache;         /* per-layer cache base */
  const float* cos_sin;          /* [max_pos][128] */
  int tp_size;                   /* 1, or the size given to bcg_tp_init */
} bcg_model;

typedef struct bcg_workspace {
  float* residual;    /* [T_max, d_model] */
  void* xn;           /* bf16 [T_max, d_model] */
  float* qkv;         /* fp32 [T_max, qkv_dim]   (prefill)  */
  void* q;            /* bf16 [T_max, n_heads*128] */
  void* attn;         /* bf16 [T_max, n_heads*128] */
  void* act;          /* bf16 [T_max, d_ff] */
  float* delta;       /* fp32 [T_max, d_model]   (prefill)  */
  float* partials;    /* fp32 split-K scratch, partials_floats long (decode) */
  long long partials_floats;
  float* attn_scratch_o;
  float* attn_scratch_ml;
  int* attn_counters;  /* int32 [B_max * n_kv_heads], zero-initialised once */
  int max_splits;
  float* logits;      /* fp32 [B_max, vocab] (with TP: this rank's vocabulary shard, vocab = shard size) */
  float* logits_local; /* unused (kept for layout compatibility) */
  long long tp_region_floats; /* > 0: fused peer-memory all-reduce in decode; floats per (slot, source
                                 rank) region of the symmetric buffer, >= B_max * d_model */
  void* xn_last;      /* bf16 [B_max, d_model] */
  int T_max, B_max;
  /* decode norm fusion (bcg_gemm_bf16_tn_fused); both NULL = separate add_rmsnorm kernels */
  float* row_sumsq;   /* fp32 [2][max(d_model/128, 8)][128]: row sums of squares per tile / segment, one table per norm */
  int* tile_counters; /* int32 [d_model/128], zero-initialised once */
} bcg_workspace;

/* T new tokens (ragged over sequences) -> KV cache filled, logits of each sequence's last
 * token in ws->logits[0..n_seqs).  last_rows: device int32 [n_seqs].  Stands in for one prefill
 * pass of vLLM's Qwen3Model.forward + LogitsProcessor (models/qwen3.py:216-236, 328-333). */
int bcg_model_prefill(const bcg_model* m, const bcg_workspace* ws, const int* ids,
                      const int* positions, const int* slot_map, const int* block_table,
                      int max_blocks_per_seq, const int* tile_info, int n_tiles,
                      const int* last_rows, int n_seqs, int T, void* stream);
/* One decode step for B rows: ids/positions/slot_map/seq_lens are the arrays bcg_mask_sample
 * maintains; logits land in ws->logits[0..B).  Stands in for one decode step of the same vLLM
 * forward.  flags: BCG_GEMM_PDL, plus how the residual add + RMSNorm between two
 * projections is executed on a single GPU (both variants need ws->row_sumsq / ws->tile_counters):
 *   default                   bcg_add_rmsnorm kernels (complete norm, then the projection)
 *   BCG_STEP_DEFERRED_NORM    bcg_add_norm_stats + 1/RMS applied in the next projection's epilogue
 *   BCG_STEP_FUSED_NORM       the producing projection finishes the residual itself
 *                             (bcg_gemm_bf16_tn_fused producer half; measured slower) */
#define BCG_STEP_FUSED_NORM 0x20000
#define BCG_STEP_DEFERRED_NORM 0x40000
#define BCG_STEP_FUSED_ROPE 0x80000 /* bcg_decode_attn_paged_qkv instead of rope kernel + attention */
int bcg_model_decode(const bcg_model* m, const bcg_workspace* ws, const int* ids,
                     const int* positions, const int* slot_map, const int* seq_lens,
                     const int* block_table, int max_blocks_per_seq, int B, int flags,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BCG_B200_H_ */
