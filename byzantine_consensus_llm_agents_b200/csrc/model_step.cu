// Whole-model prefill / decode steps: the decoder-layer stack of Qwen3 / Mistral
// (transformers modeling_qwen3.py:305-334 as the arithmetic reference; vLLM
// models/qwen3.py:216-236 as the op sequence the reference exercises) launched from C++ on
// one stream, so that one host call = one step and the engine can capture a decode step
// (plus the sampler) into a single CUDA graph.
#include "common.cuh"
#include "kernels.h"

namespace bcg {

#define BCG_TRY(expr)            \
  do {                           \
    int _rc = (expr);            \
    if (_rc != BCG_OK) return _rc; \
  } while (0)

static int check_model(const bcg_model* m, const bcg_workspace* ws) {
  if (!m || !ws || m->head_dim != 128 || m->n_layers <= 0 || !m->embed || !m->lm_head ||
      !m->final_norm || !m->ln1 || !m->ln2 || !m->w_qkv || !m->w_o || !m->w_gate_up ||
      !m->w_down || !m->kv_cache || !m->cos_sin || !ws->residual || !ws->xn || !ws->q ||
      !ws->attn || !ws->act || !ws->logits || !ws->xn_last) {
    set_last_error("bcg_model", "incomplete model / workspace description (head_dim must be 128)");
    return BCG_ERR_ARG;
  }
  return BCG_OK;
}

int model_prefill(const bcg_model* m, const bcg_workspace* ws, const int* ids, const int* positions,
                  const int* slot_map, const int* block_table, int max_blocks, const int* tile_info,
                  int n_tiles, const int* last_rows, int n_seqs, int T, cudaStream_t st) {
  BCG_TRY(check_model(m, ws));
  if (T <= 0 || T > ws->T_max || n_seqs <= 0 || n_seqs > ws->B_max || !ws->qkv || !ws->delta) {
    set_last_error("bcg_model_prefill", "T / n_seqs exceed the workspace");
    return BCG_ERR_TOO_LARGE;
  }
  const int d = m->d_model, hq = m->n_heads, hkv = m->n_kv_heads;
  const int qkv_dim = (hq + 2 * hkv) * 128;
  const float scale = 0.08838834764831845f;  // 1/sqrt(128)
  const int tp = m->tp_size > 1 ? m->tp_size : 1;
  if (tp > 1 && tp_size() != tp) {
    set_last_error("bcg_model_prefill", "tp_size does not match bcg_tp_init");
    return BCG_ERR_ARG;
  }
  for (int l = 0; l < m->n_layers; ++l) {
    if (l == 0) {
      BCG_TRY(embed_rmsnorm(ids, m->embed, m->ln1[0], ws->residual, ws->xn, T, d, m->rms_eps, st));
    } else {
      BCG_TRY(add_rmsnorm(ws->residual, ws->delta, 1, 1, 0, m->ln1[l], ws->xn, nullptr, T, d,
                          m->rms_eps, st));
    }
    // projections that feed fp32 consumers keep their fp32 accumulators (no bf16 round trip)
    BCG_TRY(gemm_bf16_tn(ws->xn, m->w_qkv[l], ws->qkv, T, qkv_dim, d, d, d, qkv_dim, BCG_EPI_F32, 1,
                         0, st));
    BCG_TRY(qknorm_rope_kvwrite(ws->qkv, 1, 1, 0, qkv_dim, m->q_norm ? m->q_norm[l] : nullptr,
                                m->k_norm ? m->k_norm[l] : nullptr, m->cos_sin, positions, slot_map,
                                ws->q, m->kv_cache[l], T, hq, hkv, m->page_size, m->rms_eps, st));
    BCG_TRY(prefill_attn_paged(ws->q, m->kv_cache[l], block_table, max_blocks, tile_info, n_tiles,
                               ws->attn, T, hq, hkv, m->page_size, scale, st));
    BCG_TRY(gemm_bf16_tn(ws->attn, m->w_o[l], ws->delta, T, d, hq * 128, hq * 128, hq * 128, d,
                         BCG_EPI_F32, 1, 0, st));
    if (tp > 1) BCG_TRY(tp_all_reduce_f32(ws->delta, static_cast<long long>(T) * d, st));
    BCG_TRY(add_rmsnorm(ws->residual, ws->delta, 1, 1, 0, m->ln2[l], ws->xn, nullptr, T, d,
                        m->rms_eps, st));
    BCG_TRY(gemm_bf16_tn(ws->xn, m->w_gate_up[l], ws->act, T, 2 * m->d_ff, d, d, d, m->d_ff,
                         BCG_EPI_SWIGLU, 1, 0, st));
    BCG_TRY(gemm_bf16_tn(ws->act, m->w_down[l], ws->delta, T, d, m->d_ff, m->d_ff, m->d_ff, d,
                         BCG_EPI_F32, 1, 0, st));
    if (tp > 1) BCG_TRY(tp_all_reduce_f32(ws->delta, static_cast<long long>(T) * d, st));
  }
  BCG_TRY(add_rmsnorm(ws->residual, ws->delta, 1, 1, 0, m->final_norm, ws->xn_last, last_rows,
                      n_seqs, d, m->rms_eps, st));
  // With tensor parallelism m->vocab is this rank's shard and the logits stay sharded: the sampler
  // exchanges one candidate per row instead (bcg_sampler_args.tp_size).
  BCG_TRY(gemm_bf16_tn(ws->xn_last, m->lm_head, ws->logits, n_seqs, m->vocab, d, d, d, m->vocab,
                       BCG_EPI_F32, 1, 0, st));
  return BCG_OK;
}

int model_decode(const bcg_model* m, const bcg_workspace* ws, const int* ids, const int* positions,
                 const int* slot_map, const int* seq_lens, const int* block_table, int max_blocks,
                 int B, int flags, cudaStream_t st) {
  BCG_TRY(check_model(m, ws));
  if (B <= 0 || B > ws->B_max || B > 128 || !ws->partials || !ws->attn_scratch_o ||
      !ws->attn_scratch_ml || !ws->attn_counters) {
    set_last_error("bcg_model_decode", "B exceeds the workspace (or 128), or decode scratch missing");
    return BCG_ERR_TOO_LARGE;
  }
  const int d = m->d_model, hq = m->n_heads, hkv = m->n_kv_heads, ff = m->d_ff;
  const int qkv_dim = (hq + 2 * hkv) * 128;
  const float scale = 0.08838834764831845f;
  const int s_qkv = gemm_splits_for(B, qkv_dim, d);
  const int s_o = gemm_splits_for(B, d, hq * 128);
  const int s_gu = gemm_splits_for(B, 2 * ff, d);
  const int s_dn = gemm_splits_for(B, d, ff);
  long long need = static_cast<long long>(s_qkv) * B * qkv_dim;
  const long long n_o = static_cast<long long>(s_o) * B * d;
  const long long n_gu = s_gu > 1 ? static_cast<long long>(s_gu) * B * 2 * ff : 0;
  const long long n_dn = static_cast<long long>(s_dn) * B * d;
  if (n_o > need) need = n_o;
  if (n_gu > need) need = n_gu;
  if (n_dn > need) need = n_dn;
  if (need > ws->partials_floats) {
    set_last_error("bcg_model_decode", "split-K scratch too small");
    return BCG_ERR_TOO_LARGE;
  }
  const int tp = m->tp_size > 1 ? m->tp_size : 1;
  if (tp > 1 && tp_size() != tp) {
    set_last_error("bcg_model_decode", "tp_size does not match bcg_tp_init");
    return BCG_ERR_ARG;
  }
  // Tensor parallelism, fused path (tp_reduce_norm.cu) when the symmetric buffer has data regions:
  // the all-reduce after each row-parallel projection happens inside the norm-statistics kernel
  // that follows it (local slab sum -> push over NVLink -> flags -> rank-ordered sum), and 1/RMS is
  // applied by the next projection's epilogue.  Otherwise NCCL all-reduces + complete norm kernels.
  const bool fused = tp > 1 && ws->tp_region_floats > 0 && tp_symm_local() != nullptr && ws->row_sumsq;
  if (fused && static_cast<long long>(B) * d > ws->tp_region_floats) {
    set_last_error("bcg_model_decode", "symmetric TP buffer too small");
    return BCG_ERR_TOO_LARGE;
  }
  const int gf = flags & BCG_GEMM_PDL;
  // profiling aids (tools/step_ablation.py): leave one kernel family out of the step to read its
  // in-pipeline cost off the step time.  Results are garbage with any of these set.
  const bool skip_attn = flags & 0x100, skip_norm = flags & 0x200, skip_rope = flags & 0x400;
  const bool skip_qkv = flags & 0x1000, skip_o = flags & 0x2000, skip_gu = flags & 0x4000,
             skip_dn = flags & 0x8000, skip_lm = flags & 0x10000;
  struct FlagScope {
    int saved;
    explicit FlagScope(int f) : saved(launch_flags()) { launch_flags() = f; }
    ~FlagScope() { launch_flags() = saved; }
  } scope(gf);
  float* P = ws->partials;
  // qk-norm / RoPE / KV-write folded into decode attention (bcg_decode_attn_paged_qkv)
  const bool fuse_rope = (flags & BCG_STEP_FUSED_ROPE) && !skip_rope;
  // Norm fusion (single GPU, opt-in with BCG_STEP_FUSED_NORM): the o / down projections finish the
  // residual add and leave bf16(residual * w) plus per-tile sums of squares; the projection that
  // follows applies 1/RMS in its epilogue.  No add_rmsnorm kernel runs after layer 0's embedding
  // norm.  Measured SLOWER than the separate 8-CTA-cluster norm kernel (14B, 10 rows: 6.07 vs
  // 5.93 ms per step): the fence + counter + reload chain at the end of the o / down GEMMs runs
  // while HBM is saturated and costs 7-10 us, the stand-alone norm 4-5 us on an idle memory system.
  const bool nf = tp == 1 && ws->row_sumsq && ws->tile_counters && (d % 128) == 0 && s_o <= 16 &&
                  s_dn <= 16 && (flags & BCG_STEP_FUSED_NORM) && !skip_norm;
  // Deferred norm (single GPU, opt-in with BCG_STEP_DEFERRED_NORM): bcg_add_norm_stats finishes the
  // residual and writes bf16(residual * w) + 8 segment sums per row; the projection that follows
  // applies 1/RMS in its epilogue, so the row-wide reduction leaves the critical path.
  const bool dn = !nf && tp == 1 && ws->row_sumsq && B <= 128 && (flags & BCG_STEP_DEFERRED_NORM) && !skip_norm;
  const bool scaled = nf || dn || fused;   // the consuming projection applies 1/RMS in its epilogue
  const int n_tiles_d = (dn || fused) ? 8 : d / 128;
  const int table_floats = (d / 128 > 8 ? d / 128 : 8) * 128;
  float* sumsq_attn = ws->row_sumsq;                      // written after o, read by gate_up
  float* sumsq_mlp = ws->row_sumsq ? ws->row_sumsq + table_floats : nullptr;  // after down -> qkv / lm_head
  auto consumer = [&](const float* table) {
    bcg_gemm_fusion f = {};
    f.scale_sumsq = table;
    f.scale_tiles = n_tiles_d;
    f.scale_d = d;
    f.scale_eps = m->rms_eps;
    return f;
  };
  auto producer = [&](float* table, const void* norm_w, void* xn_out) {
    bcg_gemm_fusion f = {};
    f.red_residual = ws->residual;
    f.red_w = static_cast<const float*>(norm_w);
    f.red_xn = xn_out;
    f.red_sumsq = table;
    f.red_counters = ws->tile_counters;
    return f;
  };
  for (int l = 0; l < m->n_layers; ++l) {
    if (l == 0) {
      BCG_TRY(embed_rmsnorm(ids, m->embed, m->ln1[0], ws->residual, ws->xn, B, d, m->rms_eps, st));
    } else if (nf) {
      // ws->xn and sumsq_mlp were produced by the previous layer's down projection
    } else if (dn) {
      BCG_TRY(add_norm_stats(ws->residual, P, 1, s_dn, static_cast<long long>(B) * d, m->ln1[l], ws->xn,
                             sumsq_mlp, B, d, st));
    } else if (fused) {
      BCG_TRY(tp_add_norm_stats(ws->residual, P, s_dn, static_cast<long long>(B) * d, m->ln1[l], ws->xn,
                                sumsq_mlp, 1, ws->tp_region_floats, B, d, st));
    } else if (!skip_norm) {
      BCG_TRY(add_rmsnorm(ws->residual, P, 1, s_dn, static_cast<long long>(B) * d, m->ln1[l], ws->xn,
                          nullptr, B, d, m->rms_eps, st));
    }
    if (!skip_qkv) {
      const bcg_gemm_fusion f = consumer(sumsq_mlp);
      BCG_TRY(gemm_bf16_tn_fused(ws->xn, m->w_qkv[l], P, B, qkv_dim, d, d, d, qkv_dim, BCG_EPI_F32, s_qkv,
                                 gf, (scaled && l > 0) ? &f : nullptr, st));
    }
    if (fuse_rope) {
      if (!skip_attn)
      BCG_TRY(decode_attn_paged_qkv(P, s_qkv, static_cast<long long>(B) * qkv_dim, qkv_dim,
                                    m->q_norm ? static_cast<const float*>(m->q_norm[l]) : nullptr,
                                    m->k_norm ? static_cast<const float*>(m->k_norm[l]) : nullptr,
                                    static_cast<const float*>(m->cos_sin), positions, slot_map, m->rms_eps,
                                    m->kv_cache[l], block_table, max_blocks, seq_lens, ws->attn,
                                    ws->attn_scratch_o, ws->attn_scratch_ml, ws->attn_counters, B, hq, hkv,
                                    m->page_size, ws->max_splits, scale, st));
    } else {
    if (!skip_rope)
    BCG_TRY(qknorm_rope_kvwrite(P, 1, s_qkv, static_cast<long long>(B) * qkv_dim, qkv_dim,
                                m->q_norm ? m->q_norm[l] : nullptr,
                                m->k_norm ? m->k_norm[l] : nullptr, m->cos_sin, positions, slot_map,
                                ws->q, m->kv_cache[l], B, hq, hkv, m->page_size, m->rms_eps, st));
    if (!skip_attn)
    BCG_TRY(decode_attn_paged(ws->q, m->kv_cache[l], block_table, max_blocks, seq_lens, ws->attn,
                              ws->attn_scratch_o, ws->attn_scratch_ml, ws->attn_counters, B, hq, hkv,
                              m->page_size, ws->max_splits, scale, st));
    }
    if (fused) {
      BCG_TRY(gemm_bf16_tn(ws->attn, m->w_o[l], P, B, d, hq * 128, hq * 128, hq * 128, d, BCG_EPI_F32, s_o,
                           gf, st));
      BCG_TRY(tp_add_norm_stats(ws->residual, P, s_o, static_cast<long long>(B) * d, m->ln2[l], ws->xn,
                                sumsq_attn, 0, ws->tp_region_floats, B, d, st));
    } else if (nf) {
      const bcg_gemm_fusion f = producer(sumsq_attn, m->ln2[l], ws->xn);
      if (!skip_o)
      BCG_TRY(gemm_bf16_tn_fused(ws->attn, m->w_o[l], P, B, d, hq * 128, hq * 128, hq * 128, d, BCG_EPI_F32,
                                 s_o, gf, &f, st));
    } else {
    if (!skip_o)
    BCG_TRY(gemm_bf16_tn(ws->attn, m->w_o[l], P, B, d, hq * 128, hq * 128, hq * 128, d, BCG_EPI_F32,
                         s_o, gf, st));
    // row-parallel projection: sum the split-K slabs over the ranks in place (slab s of every rank
    // adds up to slab s); the norm kernel then adds the slabs as usual
    if (tp > 1) BCG_TRY(tp_all_reduce_f32(P, static_cast<long long>(s_o) * B * d, st));
    if (dn) {
      BCG_TRY(add_norm_stats(ws->residual, P, 1, s_o, static_cast<long long>(B) * d, m->ln2[l], ws->xn,
                             sumsq_attn, B, d, st));
    } else if (!skip_norm)
    BCG_TRY(add_rmsnorm(ws->residual, P, 1, s_o, static_cast<long long>(B) * d, m->ln2[l], ws->xn,
                        nullptr, B, d, m->rms_eps, st));
    }
    const bcg_gemm_fusion f_gu = consumer(sumsq_attn);
    if (skip_gu) {
    } else if (s_gu == 1) {
      BCG_TRY(gemm_bf16_tn_fused(ws->xn, m->w_gate_up[l], ws->act, B, 2 * ff, d, d, d, ff, BCG_EPI_SWIGLU, 1,
                                 gf, scaled ? &f_gu : nullptr, st));
    } else {
      BCG_TRY(gemm_bf16_tn_fused(ws->xn, m->w_gate_up[l], P, B, 2 * ff, d, d, d, 2 * ff, BCG_EPI_F32, s_gu,
                                 gf, scaled ? &f_gu : nullptr, st));
      BCG_TRY(swiglu_reduce(P, s_gu, static_cast<long long>(B) * 2 * ff, ws->act, B, 2 * ff, st));
    }
    if (fused) {
      BCG_TRY(gemm_bf16_tn(ws->act, m->w_down[l], P, B, d, ff, ff, ff, d, BCG_EPI_F32, s_dn, gf, st));
    } else if (nf) {
      const bool last = (l + 1 == m->n_layers);
      const bcg_gemm_fusion f = producer(sumsq_mlp, last ? m->final_norm : m->ln1[l + 1],
                                         last ? ws->xn_last : ws->xn);
      if (!skip_dn)
      BCG_TRY(gemm_bf16_tn_fused(ws->act, m->w_down[l], P, B, d, ff, ff, ff, d, BCG_EPI_F32, s_dn, gf, &f, st));
    } else {
      if (!skip_dn)
      BCG_TRY(gemm_bf16_tn(ws->act, m->w_down[l], P, B, d, ff, ff, ff, d, BCG_EPI_F32, s_dn, gf, st));
      if (tp > 1) BCG_TRY(tp_all_reduce_f32(P, static_cast<long long>(s_dn) * B * d, st));
    }
  }
  if (nf) {
    // ws->xn_last and sumsq_mlp come from the last layer's down projection
  } else if (dn) {
    BCG_TRY(add_norm_stats(ws->residual, P, 1, s_dn, static_cast<long long>(B) * d, m->final_norm,
                           ws->xn_last, sumsq_mlp, B, d, st));
  } else if (fused) {
    BCG_TRY(tp_add_norm_stats(ws->residual, P, s_dn, static_cast<long long>(B) * d, m->final_norm, ws->xn_last,
                              sumsq_mlp, 1, ws->tp_region_floats, B, d, st));
  } else {
    BCG_TRY(add_rmsnorm(ws->residual, P, 1, s_dn, static_cast<long long>(B) * d, m->final_norm,
                        ws->xn_last, nullptr, B, d, m->rms_eps, st));
  }
  if (!skip_lm) {
    // (tensor parallelism: m->vocab is this rank's shard; the logits stay sharded for the sampler)
    const bcg_gemm_fusion f = consumer(sumsq_mlp);
    BCG_TRY(gemm_bf16_tn_fused(ws->xn_last, m->lm_head, ws->logits, B, m->vocab, d, d, d, m->vocab,
                               BCG_EPI_F32, 1, gf, scaled ? &f : nullptr, st));
  }
  return BCG_OK;
}

}  // namespace bcg
