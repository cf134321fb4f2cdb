// extern "C" surface declared in include/bcg_b200.h.
#include <string.h>

#include <mutex>
#include <string>
#include <unordered_set>

#include "common.cuh"
#include "kernels.h"

namespace bcg {

static thread_local std::string g_last_error;

int& launch_flags() {
  static thread_local int flags = 0;
  return flags;
}

// Function attributes belong to the (kernel, device) pair, not to the kernel: a process that drives
// several GPUs (bcg_decode_attn_paged_static documents that use) must set them once on EACH device.
static bool first_use_on_current_device(std::unordered_set<unsigned long long>& done, const void* kernel) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
  // kernel addresses are at least 4-byte aligned in practice; fold the ordinal into the high bits
  const unsigned long long key = reinterpret_cast<unsigned long long>(kernel) ^
                                 (static_cast<unsigned long long>(dev + 1) << 56);
  return done.insert(key).second;
}

void prefer_max_smem_carveout(const void* kernel) {
  static std::mutex mu;
  static std::unordered_set<unsigned long long> done;
  std::lock_guard<std::mutex> lock(mu);
  if (first_use_on_current_device(done, kernel) &&
      cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                           cudaSharedmemCarveoutMaxShared) != cudaSuccess)
    (void)cudaGetLastError();  // a hint; failing to set it is harmless
}

int ensure_dynamic_smem(const void* kernel, int bytes) {
  static std::mutex mu;
  static std::unordered_set<unsigned long long> done;
  std::lock_guard<std::mutex> lock(mu);
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) dev = 0;
  const unsigned long long key = reinterpret_cast<unsigned long long>(kernel) ^
                                 (static_cast<unsigned long long>(dev + 1) << 56);
  if (done.count(key)) return BCG_OK;
  BCG_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  done.insert(key);
  return BCG_OK;
}

void set_last_error(const char* what, const char* detail) {
  g_last_error = std::string(what ? what : "?") + ": " + (detail ? detail : "");
}

int model_prefill(const bcg_model* m, const bcg_workspace* ws, const int* ids, const int* positions,
                  const int* slot_map, const int* block_table, int max_blocks, const int* tile_info,
                  int n_tiles, const int* last_rows, int n_seqs, int T, cudaStream_t st);
int model_decode(const bcg_model* m, const bcg_workspace* ws, const int* ids, const int* positions,
                 const int* slot_map, const int* seq_lens, const int* block_table, int max_blocks,
                 int B, int flags, cudaStream_t st);

}  // namespace bcg

using namespace bcg;

static inline cudaStream_t S(void* s) { return reinterpret_cast<cudaStream_t>(s); }

extern "C" {

int bcg_abi_version(void) { return BCG_ABI_VERSION; }

const char* bcg_last_error(void) { return g_last_error.c_str(); }

int bcg_gemm_bf16_tn(const void* X, const void* W, void* out, int M, int N, int K, long long ldx,
                     long long ldw, long long ld_out, int epilogue, int splits, int flags,
                     void* stream) {
  return gemm_bf16_tn(X, W, out, M, N, K, ldx, ldw, ld_out, epilogue, splits, flags, S(stream));
}

int bcg_gemm_bf16_tn_fused(const void* X, const void* W, void* out, int M, int N, int K,
                           long long ldx, long long ldw, long long ld_out, int epilogue, int splits,
                           int flags, const bcg_gemm_fusion* fusion, void* stream) {
  return gemm_bf16_tn_fused(X, W, out, M, N, K, ldx, ldw, ld_out, epilogue, splits, flags, fusion,
                            S(stream));
}

int bcg_gemm_splits_for(int M, int N, int K) { return gemm_splits_for(M, N, K); }

int bcg_embed_rmsnorm(const int* ids, const void* embed, const float* w, float* residual, void* xn,
                      int T, int d, float eps, void* stream) {
  return embed_rmsnorm(ids, embed, w, residual, xn, T, d, eps, S(stream));
}

int bcg_add_rmsnorm(float* residual, const void* delta, int delta_is_f32, int n_partials,
                    long long partial_stride, const float* w, void* xn, const int* src_rows, int T,
                    int d, float eps, void* stream) {
  return add_rmsnorm(residual, delta, delta_is_f32, n_partials, partial_stride, w, xn, src_rows, T,
                     d, eps, S(stream));
}

int bcg_add_norm_stats(float* residual, const void* delta, int delta_is_f32, int n_partials,
                       long long partial_stride, const float* w, void* xn, float* sumsq, int T, int d,
                       void* stream) {
  return add_norm_stats(residual, delta, delta_is_f32, n_partials, partial_stride, w, xn, sumsq, T, d,
                        S(stream));
}

int bcg_qknorm_rope_kvwrite(const void* qkv, int qkv_is_f32, int n_partials,
                            long long partial_stride, long long ld_qkv, const float* q_norm_w,
                            const float* k_norm_w, const float* cos_sin, const int* positions,
                            const int* slot_map, void* q_out, void* kv_cache, int T, int n_heads,
                            int n_kv_heads, int page_size, float eps, void* stream) {
  return qknorm_rope_kvwrite(qkv, qkv_is_f32, n_partials, partial_stride, ld_qkv, q_norm_w,
                             k_norm_w, cos_sin, positions, slot_map, q_out, kv_cache, T, n_heads,
                             n_kv_heads, page_size, eps, S(stream));
}

int bcg_swiglu_reduce(const float* partials, int n_partials, long long partial_stride, void* act,
                      int T, int two_f, void* stream) {
  return swiglu_reduce(partials, n_partials, partial_stride, act, T, two_f, S(stream));
}

int bcg_decode_attn_paged(const void* q, const void* kv_cache, const int* block_table,
                          int max_blocks_per_seq, const int* seq_lens, void* out, float* scratch_o,
                          float* scratch_ml, int* chunk_counters, int B, int n_heads, int n_kv_heads,
                          int page_size, int max_splits, float scale, void* stream) {
  return decode_attn_paged(q, kv_cache, block_table, max_blocks_per_seq, seq_lens, out, scratch_o,
                           scratch_ml, chunk_counters, B, n_heads, n_kv_heads, page_size, max_splits,
                           scale, S(stream));
}

int bcg_decode_attn_paged_qkv(const float* qkv, int n_partials, long long partial_stride,
                              long long ld_qkv, const float* q_norm_w, const float* k_norm_w,
                              const float* cos_sin, const int* positions, const int* slot_map,
                              float eps, void* kv_cache, const int* block_table,
                              int max_blocks_per_seq, const int* seq_lens, void* out, float* scratch_o,
                              float* scratch_ml, int* chunk_counters, int B, int n_heads,
                              int n_kv_heads, int page_size, int max_splits, float scale,
                              void* stream) {
  return decode_attn_paged_qkv(qkv, n_partials, partial_stride, ld_qkv, q_norm_w, k_norm_w, cos_sin,
                               positions, slot_map, eps, kv_cache, block_table, max_blocks_per_seq,
                               seq_lens, out, scratch_o, scratch_ml, chunk_counters, B, n_heads,
                               n_kv_heads, page_size, max_splits, scale, S(stream));
}

int bcg_decode_attn_paged_static(const void* q, const void* kv_cache, const int* block_table,
                                 int max_blocks_per_seq, const int* seq_lens, void* out, int B,
                                 int n_heads, int n_kv_heads, int page_size, int max_splits,
                                 float scale, void* stream) {
  return decode_attn_paged_static(q, kv_cache, block_table, max_blocks_per_seq, seq_lens, out, B,
                                  n_heads, n_kv_heads, page_size, max_splits, scale, S(stream));
}

int bcg_prefill_attn_paged(const void* q, const void* kv_cache, const int* block_table,
                           int max_blocks_per_seq, const int* tile_info, int n_tiles, void* out,
                           int n_q_rows, int n_heads, int n_kv_heads, int page_size, float scale,
                           void* stream) {
  return prefill_attn_paged(q, kv_cache, block_table, max_blocks_per_seq, tile_info, n_tiles, out,
                            n_q_rows, n_heads, n_kv_heads, page_size, scale, S(stream));
}

int bcg_mask_sample(const bcg_sampler_args* args, void* stream) {
  return mask_sample(args, S(stream));
}

int bcg_tp_unique_id(void* out_128_bytes) { return tp_unique_id(out_128_bytes); }
int bcg_tp_init(const void* id_128_bytes, int rank, int size) { return tp_init(id_128_bytes, rank, size); }
int bcg_tp_shutdown(void) { return tp_shutdown(); }
long long bcg_tp_symm_bytes_for(long long region_floats, int size) {
  if (region_floats < 0 || size < 1) return 0;
  return 4 * (tp_symm_flag_floats() + 4 * static_cast<long long>(size) * region_floats);
}
int bcg_tp_add_norm_stats(float* residual, const float* partials, int n_partials, long long partial_stride,
                          const float* w, void* xn, float* sumsq, int slot, long long region_floats, int T,
                          int d, void* stream) {
  return tp_add_norm_stats(residual, partials, n_partials, partial_stride, w, xn, sumsq, slot, region_floats,
                           T, d, S(stream));
}
int bcg_tp_symm_create(long long bytes, void* handle_out_64_bytes) {
  void* local = nullptr;
  return tp_symm_create(bytes, handle_out_64_bytes, &local);
}
int bcg_tp_symm_open(const void* handles, int rank, int size) { return tp_symm_open(handles, rank, size); }

int bcg_model_prefill(const bcg_model* m, const bcg_workspace* ws, const int* ids,
                      const int* positions, const int* slot_map, const int* block_table,
                      int max_blocks_per_seq, const int* tile_info, int n_tiles,
                      const int* last_rows, int n_seqs, int T, void* stream) {
  return model_prefill(m, ws, ids, positions, slot_map, block_table, max_blocks_per_seq, tile_info,
                       n_tiles, last_rows, n_seqs, T, S(stream));
}

int bcg_model_decode(const bcg_model* m, const bcg_workspace* ws, const int* ids,
                     const int* positions, const int* slot_map, const int* seq_lens,
                     const int* block_table, int max_blocks_per_seq, int B, int flags,
                     void* stream) {
  return model_decode(m, ws, ids, positions, slot_map, seq_lens, block_table, max_blocks_per_seq, B,
                      flags, S(stream));
}

}  // extern "C"
