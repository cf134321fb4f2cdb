// Internal C++ declarations shared by the .cu files; the public C ABI is include/bcg_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cuda.h>

#include "../../include/bcg_b200.h"

namespace bcg {

int gemm_bf16_tn(const void* X, const void* W, void* out, int M, int N, int K, long long ldx,
                 long long ldw, long long ld_out, int epi, int splits, int flags,
                 cudaStream_t stream);
int gemm_bf16_tn_fused(const void* X, const void* W, void* out, int M, int N, int K, long long ldx,
                       long long ldw, long long ld_out, int epi, int splits, int flags,
                       const bcg_gemm_fusion* fusion, cudaStream_t stream);
int gemm_splits_for(int M, int N, int K);
// Prefill shapes on CTA pairs (tcgen05.mma.cta_group::2); BCG_ERR_UNSUPPORTED = not a shape for it
int gemm2sm_bf16_tn(const void* X, const void* W, void* out, int M, int N, int K, long long ldx, long long ldw,
                    long long ld_out, int epi, cudaStream_t stream);
// TMA descriptor of a row-major bf16 matrix (rows x K, pitch ld elements): boxes of box_rows x 64
// elements land in shared memory 128B-swizzled, out-of-bounds elements read as zero.
int make_tmap_bf16_sw128(CUtensorMap* m, const void* ptr, long long rows, long long K, long long ld,
                         int box_rows);

int embed_rmsnorm(const int* ids, const void* embed, const float* w, float* residual, void* xn,
                  int T, int d, float eps, cudaStream_t stream);
int add_norm_stats(float* residual, const void* delta, int delta_is_f32, int n_partials,
                   long long partial_stride, const float* w, void* xn, float* sumsq, int T, int d,
                   cudaStream_t stream);
int add_rmsnorm(float* residual, const void* delta, int delta_is_f32, int n_partials,
                long long partial_stride, const float* w, void* xn, const int* src_rows, int T,
                int d, float eps, cudaStream_t stream);
int qknorm_rope_kvwrite(const void* qkv, int qkv_is_f32, int n_partials, long long partial_stride,
                        long long ld_qkv, const float* q_norm_w, const float* k_norm_w,
                        const float* cos_sin, const int* positions, const int* slot_map,
                        void* q_out, void* kv_cache, int T, int n_heads, int n_kv_heads,
                        int page_size, float eps, cudaStream_t stream);
int swiglu_reduce(const float* partials, int n_partials, long long partial_stride, void* act,
                  int T, int two_f, cudaStream_t stream);

int decode_attn_paged(const void* q, const void* kv_cache, const int* block_table,
                      int max_blocks_per_seq, const int* seq_lens, void* out, float* scratch_o,
                      float* scratch_ml, int* counters, int B, int n_heads, int n_kv_heads,
                      int page_size, int max_splits, float scale, cudaStream_t stream);
int decode_attn_paged_qkv(const float* qkv, int n_partials, long long partial_stride, long long ld_qkv,
                          const float* q_norm_w, const float* k_norm_w, const float* cos_sin,
                          const int* positions, const int* slot_map, float eps, void* kv_cache,
                          const int* block_table, int max_blocks_per_seq, const int* seq_lens, void* out,
                          float* scratch_o, float* scratch_ml, int* counters, int B, int n_heads,
                          int n_kv_heads, int page_size, int max_splits, float scale,
                          cudaStream_t stream);
int decode_attn_paged_static(const void* q, const void* kv_cache, const int* block_table,
                             int max_blocks_per_seq, const int* seq_lens, void* out, int B,
                             int n_heads, int n_kv_heads, int page_size, int max_splits, float scale,
                             cudaStream_t stream);
int prefill_attn_paged(const void* q, const void* kv_cache, const int* block_table,
                       int max_blocks_per_seq, const int* tile_info, int n_tiles, void* out,
                       int n_q_rows, int n_heads, int n_kv_heads, int page_size, float scale,
                       cudaStream_t stream);

int mask_sample(const bcg_sampler_args* a, cudaStream_t stream);

// tensor parallelism (tp_comm.cu)
int tp_unique_id(void* out128);
int tp_init(const void* id128, int rank, int size);
int tp_shutdown();
int tp_size();
int tp_rank();
int tp_all_reduce_f32(float* buf, long long count, cudaStream_t stream);
int tp_all_gather_f32(const float* send, float* recv, long long count, cudaStream_t stream);
int tp_symm_create(long long bytes, void* handle_out64, void** local_ptr);
int tp_symm_open(const void* handles, int rank, int size);
void* tp_symm_local();
void* const* tp_symm_peers_device();
long long tp_symm_bytes();
// fused cross-GPU reduction + residual add + RMSNorm statistics over peer memory (tp_reduce_norm.cu)
int tp_add_norm_stats(float* residual, const float* partials, int n_partials, long long partial_stride,
                      const float* w, void* xn, float* sumsq, int slot, long long region_floats, int T,
                      int d, cudaStream_t stream);
long long tp_symm_flag_floats();

}  // namespace bcg
