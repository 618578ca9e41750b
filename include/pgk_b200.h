/*
 * pgk_b200.h -- C ABI of libpgk_b200.so: the B200 (sm_100a) decode hot path behind
 * PyGPUkit's operator API.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; every pointer is a DEVICE pointer unless the
 *     parameter name starts with `host_`.  Arguments are borrowed, never retained.
 *   - every op is asynchronous on `stream` (a cudaStream_t passed as void*; NULL = default
 *     stream) and returns 0 on success, non-zero on failure; pgk_last_error() then holds a
 *     message of the form "<fn> failed: <cudaGetErrorString>" (the reference's
 *     std::runtime_error text, native/ops/common/error.cuh:28-37).
 *   - bf16 / f16 tensors are 16-bit storage, row-major contiguous; "u8" tensors carry
 *     FP8-E4M3 codes or packed NVF4 (E2M1) nibbles.
 *   - each entry point cites the reference interface it replaces (paths relative to the
 *     PyGPUkit tree).
 */
#ifndef PGK_B200_H
#define PGK_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ runtime
 * replaces native/core/{memory,stream,event,cuda_graph,device}.cpp as bound by
 * native/bindings/core_bindings.cpp:29-509 */
const char* pgk_last_error(void);
int pgk_version(void);
int pgk_device_count(int* count);
int pgk_device_set(int device);
int pgk_device_get(int* device);
typedef struct {
    char name[128];
    int sm_count;
    int cc_major, cc_minor;
    size_t total_mem;
    size_t l2_bytes;
    int max_smem_per_block;
} pgk_device_props_t;
int pgk_device_props(int device, pgk_device_props_t* out);
int pgk_device_sync(void);
int pgk_mem_info(size_t* free_bytes, size_t* total_bytes);

int pgk_malloc(void** ptr, size_t bytes);            /* memory.cpp:29 cuMemAlloc */
int pgk_free(void* ptr);
int pgk_host_alloc(void** host_ptr, size_t bytes);   /* pinned host memory */
int pgk_host_free(void* host_ptr);
int pgk_memset(void* ptr, int value, size_t bytes, void* stream);
int pgk_memcpy_h2d(void* dst, const void* host_src, size_t bytes, void* stream);
int pgk_memcpy_d2h(void* host_dst, const void* src, size_t bytes, void* stream);
int pgk_memcpy_d2d(void* dst, const void* src, size_t bytes, void* stream);

int pgk_stream_create(void** stream);                /* native/core/stream.cpp */
int pgk_stream_destroy(void* stream);
int pgk_stream_sync(void* stream);
int pgk_event_create(void** event);                  /* native/core/event.cpp */
int pgk_event_destroy(void* event);
int pgk_event_record(void* event, void* stream);
int pgk_event_sync(void* event);
int pgk_event_elapsed_ms(void* start, void* stop, float* ms);

/* CUDA-graph capture of a decode step: native/core/cuda_graph.cpp, used by
 * llm/decode/m1_graph.py:248-457 */
int pgk_graph_begin(void* stream);
int pgk_graph_end(void* stream, void** graph_exec);
int pgk_graph_launch(void* graph_exec, void* stream);
int pgk_graph_destroy(void* graph_exec);

/* programmatic dependent launch on/off for all subsequent launches (default off) */
int pgk_set_pdl(int enabled);
/* number of kernels this library has launched since load (or since reset) */
int64_t pgk_launch_count(void);
void pgk_launch_count_reset(void);
/* write >L2-size scratch so the next kernel starts with a cold L2 (bench hygiene) */
int pgk_flush_l2(void* stream);

/* ------------------------------------------------------------------ W8A16 FP8 GEMV
 * C[M,N] = A[M,K] . dequant(B[N,K])^T,  B = FP8-E4M3 codes, S = bf16 [ceil(N/128), ceil(K/128)].
 * replaces gemv_fp8_bf16 / gemv_fp8_bf16_batched: ops/matmul/gemv.py:73-195,
 * native/bindings/gemv/fp8xfp8_bf16.cpp:27-58, launch_gemv_fp8_opt(_batched)
 * fp8_opt_kernels.cu:15-69.  Weights are read ONCE for all M rows. */
int pgk_gemv_fp8_bf16(const void* A, const void* B_nk, const void* S, void* C, int M, int K,
                      int N, void* stream);
/* fused variants used by the llm layers (same math, reference rounding points kept):
 *   residual != NULL : C = bf16( bf16(acc) + residual )        (add_inplace, elementwise_kernels.cuh:381-390)
 *   bias     != NULL : C = bf16( bf16(acc) + bias[n] )         (bias_add,   elementwise_kernels.cuh:57-68) */
int pgk_gemv_fp8_bf16_ex(const void* A, const void* B_nk, const void* S, void* C, int M, int K,
                         int N, const void* bias, const void* residual, void* stream);
/* C[M,I] = bf16( bf16(silu(bf16(A.Wg^T))) * bf16(A.Wu^T) ): gate/up GEMV + silu + mul
 * (models/causal.py:489-514, activation_kernels.cuh:164-171, elementwise_kernels.cuh:429-438) */
int pgk_gemv_fp8_swiglu_bf16(const void* A, const void* Wg_nk, const void* Sg, const void* Wu_nk,
                             const void* Su, void* C, int M, int K, int N, void* stream);

/* Decode batch 5..64 on the tensor cores (tcgen05.mma kind::f16, BF16 x BF16 -> F32 in TMEM): same operands
 * and epilogues as pgk_gemv_fp8_bf16_ex; the weights are read ONCE for all M rows, converted exactly to BF16
 * on the way into shared memory, the 128x128 block scale is folded into the activation slab (rounded to BF16
 * where the reference rounds weight*scale, w8a16_gemm.cu:104-339).  Needs 1 <= M <= 64 (rows up to the MMA
 * width 16/32/64 are zero-filled), K % 128 == 0, N % 128 == 0.  pgk_gemv_fp8_bf16(_ex) dispatch here by
 * themselves for M >= 5 (measured crossover against the CUDA-core slab GEMV; env PGK_TC_MIN_M) when the shape
 * allows (replaces gemv_fp8_bf16_batched, fp8_opt.cuh:220-295, and w8a16_gemm_sm120's mma.sync path). */
int pgk_gemm_fp8_bf16_tc(const void* A, const void* B_nk, const void* S, void* C, int M, int K, int N,
                         const void* bias, const void* residual, void* stream);

/* The same contract with NO CUDA-core work on the weight stream: raw E4M3 tiles by TMA (cp.async.bulk.tensor, 2-D tensor
 * map over the [N,K] matrix, 128-byte swizzle) straight into tcgen05.mma.kind::f8f6f4; the activations are split once
 * (pre-kernel) into two exact E4M3 terms a' = t0 + t1/16 after a per-row power-of-two shift; the 128x128 block scale is
 * applied per k block when the partial products are promoted from TMEM to FP32 registers.  1 <= M <= 64, K % 128 == 0,
 * N % 128 == 0, needs cuTensorMapEncodeTiled (driver >= 12.0). */
int pgk_gemm_fp8_bf16_f8mma(const void* A, const void* B_nk, const void* S, void* C, int M, int K, int N,
                            const void* bias, const void* residual, void* stream);

/* W8A16 GEMM with the reference's [K,N] weight layout: C[M,N] = A[M,K] . dequant(B[K,N]),
 * S = bf16 [ceil(K/128), ceil(N/128)].  replaces pygpukit_w8a16_gemm_sm120
 * (w8a16_gemm.cu:417-456, ops/matmul/w8a16.py:38-119). */
int pgk_w8a16_gemm(const void* A, const void* B_kn, const void* S, void* C, int M, int K, int N,
                   void* stream);

/* ------------------------------------------------------------------ NVF4 (W4A16)
 * C[M,N] = alpha * A[M,K] . dequant(B), B_data u8 [K/2, N] (low nibble = even k),
 * B_scale u8 UE4M3 [ceil(K/32), N].  replaces pygpukit_gemv_nvf4_bf16 (nvf4.cu:161-178,
 * ops/matmul/nvf4.py:132-194). */
int pgk_gemv_nvf4_bf16(const void* A, const void* B_data, const void* B_scale, void* C, int M,
                       int K, int N, float alpha, void* stream);
int pgk_gemv_nvf4_bf16_ex(const void* A, const void* B_data, const void* B_scale, void* C, int M,
                          int K, int N, float alpha, const void* bias, const void* residual,
                          void* stream);
/* Decode batch 5..64 on the tensor cores: E2M1 x UE4M3 has <= 6 significant bits, so the converter warps write EXACT BF16
 * weight tiles (F2FP.E2M1 unpack, HMUL2 by the scale) and one tcgen05.mma.kind::f16 per 16 k computes exact products with
 * F32 accumulation in TMEM -- the FP32-accumulate class of the GEMV, weights read once for all M rows.  Needs
 * 1 <= M <= 64, K % 128 == 0, N % 128 == 0; pgk_gemv_nvf4_bf16(_ex) dispatch here for M >= 5.  Replaces the per-row use of
 * gemv_nvf4_bf16_kernel (nvf4_kernels.cu:79-233) and the CUTLASS NVF4 GEMM of gemm/w4a16_bf16/sm120/nvf4_cutlass.cu. */
int pgk_gemm_nvf4_bf16_tc(const void* A, const void* B_data, const void* B_scale, void* C, int M, int K, int N,
                          float alpha, const void* bias, const void* residual, void* stream);
/* out[M,I] = bf16(bf16(silu(g)) * u) from a [M, 2I] tensor whose columns are gate/up interleaved in groups of 8 (the
 * fused NVF4 MLP layout below): the epilogue of the batched (tensor-core) gate/up projection */
int pgk_swiglu_interleaved8_bf16(const void* gate_up, void* out, int M, int I, void* stream);
/* gate/up GEMV + SiLU*up on ONE fused weight tensor: Wgu_data u8 [K/2, 2I], Wgu_scale
 * [K/32, 2I] whose columns are gate/up interleaved in groups of 8
 * ([g0..g7, u0..u7, g8..g15, u8..]); N2 = 2I, C = [M, I].  The layout is produced by
 * pygpukit_b200.llm.layers.LinearNVF4.fuse_gate_up (net-new: the reference has no NVF4 MLP). */
int pgk_gemv_nvf4_swiglu_bf16(const void* A, const void* Wgu_data, const void* Wgu_scale, void* C,
                              int M, int K, int N2, void* stream);
/* BF16 [K,N] row-major -> NVF4 data/scale, bit-exact with quantize_bf16_to_nvf4_kernel
 * (nvf4_kernels.cu:239-324; ops/matmul/nvf4.py:32-77). */
int pgk_quantize_bf16_to_nvf4(const void* in_kn, void* out_data, void* out_scale, int K, int N,
                              void* stream);

/* ------------------------------------------------------------------ BF16 GEMV
 * C[M,N] = A[M,K] . B[N,K]^T (lm_head, router).  replaces pygpukit_gemv_bf16_opt
 * (bf16_opt.cu:36-47, ops/matmul/gemv.py:15-70). */
int pgk_gemv_bf16(const void* A, const void* B_nk, void* C, int M, int K, int N, void* stream);
int pgk_gemv_bf16_ex(const void* A, const void* B_nk, void* C, int M, int K, int N,
                     const void* bias, const void* residual, void* stream);
/* lm_head + greedy sampling: logits (optional, may be NULL) and argmax token (int32,
 * first maximum wins as np.argmax: llm/sampling.py:59-61). */
int pgk_gemv_bf16_argmax(const void* A, const void* B_nk, void* logits, int32_t* token, int K,
                         int N, void* stream);

/* ------------------------------------------------------------------ fused RMSNorm -> GEMV
 * y = GEMV( bf16(rmsnorm(A, gamma, eps)) ): the pre-norm of a decoder block folded into the staging of
 * the activation rows of the GEMV that consumes it (one launch instead of two).  Every CTA recomputes
 * the row statistics in the summation order of pgk_rmsnorm (= rmsnorm_bf16_kernel,
 * norm_kernels.cuh:535-585), so results are bit-identical to pgk_rmsnorm followed by the plain GEMV
 * (decode/m1.py:74-82 + layers/linear.py:241-251).  Needs K >= 512 and the fast-path alignment. */
int pgk_rmsnorm_gemv_fp8_bf16(const void* A, const void* gamma, float eps, const void* B_nk, const void* S,
                              void* C, int M, int K, int N, const void* bias, const void* residual, void* stream);
int pgk_rmsnorm_gemv_fp8_swiglu_bf16(const void* A, const void* gamma, float eps, const void* Wg_nk,
                                     const void* Sg, const void* Wu_nk, const void* Su, void* C, int M, int K,
                                     int N, void* stream);
int pgk_rmsnorm_gemv_nvf4_bf16(const void* A, const void* gamma, float eps, const void* B_data,
                               const void* B_scale, void* C, int M, int K, int N, float alpha, const void* bias,
                               const void* residual, void* stream);
int pgk_rmsnorm_gemv_nvf4_swiglu_bf16(const void* A, const void* gamma, float eps, const void* Wgu_data,
                                      const void* Wgu_scale, void* C, int M, int K, int N2, void* stream);
int pgk_rmsnorm_gemv_bf16(const void* A, const void* gamma, float eps, const void* B_nk, void* C, int M, int K,
                          int N, const void* bias, const void* residual, void* stream);
int pgk_rmsnorm_gemv_bf16_argmax(const void* A, const void* gamma, float eps, const void* B_nk, void* logits,
                                 int32_t* token, int K, int N, void* stream);

/* ------------------------------------------------------------------ norm / rope / glue
 * dtype codes: 0 = f32, 1 = f16, 2 = bf16 */
int pgk_rmsnorm(const void* x, const void* gamma, void* y, int rows, int features, float eps,
                int dtype, void* stream);                    /* norm/rmsnorm.inl:9-57 */
int pgk_layernorm(const void* x, const void* gamma, const void* beta, void* y, int rows,
                  int features, float eps, int dtype, void* stream); /* norm/layernorm.inl */
/* y = rmsnorm(x + residual) (sum NOT written back): fused_kernels.cuh:36-220 */
int pgk_rmsnorm_residual(const void* x, const void* residual, const void* gamma, void* y, int rows,
                         int features, float eps, int dtype, void* stream);
/* NeoX half-split rotation in place; cos/sin [seq, head_dim] (table_dtype 0 = f32 table,
 * otherwise same dtype as q/k): rope/rope_inplace.inl:8-149 */
int pgk_rope_inplace(void* q, void* k, const void* cos, const void* sin, int seq, int n_heads_q,
                     int n_heads_k, int head_dim, int dtype, int table_dtype, void* stream);
int pgk_silu(const void* x, void* y, size_t n, int dtype, void* stream);
int pgk_gelu(const void* x, void* y, size_t n, int dtype, void* stream);
int pgk_swiglu(const void* gate, const void* up, void* y, size_t n, int dtype, void* stream);
/* y = bf16(bf16(silu(gate)) * up): pgk_silu followed by pgk_mul_inplace in one launch (two roundings) */
int pgk_silu_mul(const void* gate, const void* up, void* y, size_t n, int dtype, void* stream);
int pgk_add(const void* a, const void* b, void* c, size_t n, int dtype, void* stream);
int pgk_mul(const void* a, const void* b, void* c, size_t n, int dtype, void* stream);
int pgk_add_inplace(void* a, const void* b, size_t n, int dtype, void* stream);
int pgk_mul_inplace(void* a, const void* b, size_t n, int dtype, void* stream);
int pgk_bias_add_inplace(void* x, const void* bias, int rows, int features, int dtype,
                         void* stream);
int pgk_cast(const void* src, void* dst, size_t n, int src_dtype, int dst_dtype, void* stream);
int pgk_embedding_lookup(const void* table, void* out, int hidden, int token, int elem_bytes,
                         void* stream);                      /* embedding_kernels.cuh:27-60 */
int pgk_embedding_lookup_ptr(const void* table, void* out, int hidden, const int32_t* token_buf,
                             int n_tokens, int elem_bytes, void* stream);
/* GPT-2 input embedding: out[t] = wte[token_buf[t]] + wpe[pos + t] (one rounding; models/causal.py:137-146 adds
 * the two rows on the host); pos from *pos_buf when not NULL.  dtype 0 f32 / 1 f16 / 2 bf16. */
int pgk_embedding_lookup_pos(const void* wte, const void* wpe, void* out, int hidden, const int32_t* token_buf,
                             int n_tokens, int pos, const int32_t* pos_buf, int dtype, void* stream);
int pgk_argmax(const void* logits, int32_t* out, int n, int dtype, void* stream);
/* transpose [d0,d1,d2] -> [d1,d0,d2] (transpose_3d_021) */
int pgk_transpose_021(const void* in, void* out, int d0, int d1, int d2, int elem_bytes,
                      void* stream);

/* KV cache: new_kv [seq, n_kv, hd] -> cache [num_heads_out, max_seq, hd] at position pos..pos+seq-1.
 * num_heads_out == n_kv writes the un-expanded cache; num_heads_out = G*n_kv replicates each KV
 * head G times (the reference's GQA-expanded layout, kv_cache_kernels.cuh:176-421).  pos_buf != NULL
 * reads the position from device memory (the `_ptr` variant). */
int pgk_kv_cache_update(const void* new_kv, void* cache, int seq, int n_kv, int num_heads_out,
                        int head_dim, int max_seq, int pos, const int32_t* pos_buf, int elem_bytes,
                        void* stream);

/* fused decode bracket (q_len == 1): per head QK-RMSNorm (optional, gamma NULL = off), RoPE from f32
 * tables at position *pos_buf (or pos), K/V written into the un-expanded cache
 * [n_kv, max_seq, hd].  Same rounding points as rmsnorm -> rope_inplace_f32table ->
 * kv_cache_update_gqa (layers/attention.py:326-351). qkv = [.., (Hq + 2*Hkv) * hd] bf16, M rows
 * are M consecutive positions of one sequence.  cos_table == NULL: no rotation (learned-position
 * models, GPT-2) -- q stays as projected and K/V rows are copied into the cache. */
int pgk_qk_norm_rope_kv_write(void* qkv, const void* q_gamma, const void* k_gamma, float eps,
                              const float* cos_table, const float* sin_table, void* k_cache,
                              void* v_cache, int M, int n_heads_q, int n_kv, int head_dim,
                              int max_seq, int pos, const int32_t* pos_buf, void* stream);

/* ------------------------------------------------------------------ attention
 * softmax(Q K^T * scale) V over the first context_len cache rows, causal offset
 * kv_len - q_len for q_len > 1.  Q [Hq, q_len, hd] (q_stride_h/q_stride_t in elements allow
 * reading straight out of a qkv row), K/V cache [n_kv, max_seq, hd] (n_kv == Hq is the
 * reference's expanded layout), out [Hq, q_len, hd].  Split-KV flash-decoding with an
 * in-kernel last-block combine.  replaces sdpa_causal_fixed_cache(_ptr)
 * (sdpa_causal.inl:736-769, flash_decoding.cuh:75-377).  ctx_buf != NULL reads context_len
 * from device memory. */
int pgk_sdpa_decode(const void* Q, const void* K, const void* V, void* out, int n_heads_q,
                    int n_kv, int q_len, int head_dim, int max_seq, int context_len,
                    const int32_t* ctx_buf, float scale, long q_stride_h, long q_stride_t,
                    long o_stride_h, long o_stride_t, int dtype, void* stream);
/* The whole decode attention bracket of ONE new token in one launch (bf16, head_dim 128): per-head
 * QK-RMSNorm (gamma NULL = off) -> RoPE at *pos_buf (or pos) from f32 tables -> K/V row written into the
 * un-expanded cache -> split-KV attention over the first *ctx_buf (or context_len) positions, new row
 * included.  qkv = [(Hq + 2 Hkv) * 128] bf16 (left untouched), out = [Hq * 128] bf16.  Same rounding points
 * as rmsnorm -> rope_inplace_f32table -> kv_cache_update_gqa -> sdpa_causal_fixed_cache_ptr
 * (layers/attention.py:326-368).  K/V tiles are requested before griddepcontrol.wait (they do not depend on
 * the qkv projection), so under PDL the HBM round trip overlaps the previous kernel. */
int pgk_rope_sdpa_decode(const void* qkv, const void* q_gamma, const void* k_gamma, float eps,
                         const float* cos_table, const float* sin_table, void* k_cache, void* v_cache,
                         void* out, int n_heads_q, int n_kv, int head_dim, int max_seq, int pos,
                         const int32_t* pos_buf, int context_len, const int32_t* ctx_buf, float scale,
                         void* stream);
size_t pgk_sdpa_workspace_bytes(int n_heads_q, int q_len, int head_dim, int max_context);

/* ------------------------------------------------------------------ MoE (config 4)
 * native/bindings/moe.cpp:43-223, native/ops/moe/{topk,permute,moe}_kernels.cuh; bf16 data */
int pgk_moe_topk(const void* logits, void* values, int32_t* indices, int num_tokens,
                 int num_experts, int k, int dtype, void* stream);      /* topk_kernels.cuh:69-186 */
int pgk_moe_softmax_topk(void* values, int num_tokens, int k, int dtype, void* stream); /* :192-256 */
/* counts [E], offsets [E+1], permute [T*k], reverse [T*k] (permute_kernels.cuh:21-88,245-256) */
int pgk_moe_compute_permutation(const int32_t* expert_indices, int32_t* expert_counts,
                                int32_t* expert_offsets, int32_t* permute_indices,
                                int32_t* reverse_perm, int num_tokens, int num_experts, int k,
                                void* stream);
int pgk_moe_expand_expert_offsets(const int32_t* expert_offsets, int32_t* row_expert_ids,
                                  int num_experts, int m_total, void* stream); /* moe_kernels.cuh:236-259 */
int pgk_moe_gather(const void* hidden, const int32_t* permute_indices, void* gathered,
                   int num_tokens, int hidden_size, int k, void* stream);      /* permute_kernels.cuh:95-182 */
int pgk_moe_scatter(const void* expert_out, const void* router_weights, const int32_t* reverse_perm,
                    void* output, int num_tokens, int hidden_size, int k, void* stream); /* :218-242 */
/* row m of A[M,K] uses expert row_expert_ids[m] of B[E,N,K] (FP8) / S[E,N/128,K/128]:
 * grouped_gemm_fp8_bf16 (ops/matmul/grouped.py:36-131, grouped_gemm.cu:77-201) */
int pgk_moe_grouped_gemv_fp8(const void* A, const void* B_enk, const void* S, void* C,
                             const int32_t* row_expert_ids, int M, int K, int N, void* stream);

/* expert-parallel form: this rank holds experts [expert_offset, expert_offset + num_local_experts) in B / S;
 * rows routed elsewhere are written as zeros (their owner contributes them through the all-reduce) */
int pgk_moe_grouped_gemv_fp8_ep(const void* A, const void* B_enk, const void* S, void* C,
                                const int32_t* row_expert_ids, int M, int K, int N, int expert_offset,
                                int num_local_experts, void* stream);

/* NVF4 experts (net-new; the reference's MoE is FP8-only): B_data u8 [E_local, K/2, N], B_scale u8
 * [E_local, K/32, N] (per-expert reference NVF4 layout); rows whose expert id falls outside
 * [expert_offset, expert_offset + num_local_experts) are written as zeros. */
int pgk_moe_grouped_gemv_nvf4(const void* A, const void* B_data, const void* B_scale, void* C,
                              const int32_t* row_expert_ids, int M, int K, int N, int expert_offset,
                              int num_local_experts, void* stream);
/* _ex: a_stride = elements between consecutive rows of A; 0 broadcasts one activation row to all M rows
 * (decode: every routed expert reads the same token, so no gather launch is needed). */
int pgk_moe_grouped_gemv_fp8_ex(const void* A, int a_stride, const void* B_enk, const void* S, void* C,
                                const int32_t* row_expert_ids, int M, int K, int N, int expert_offset,
                                int num_local_experts, void* stream);
int pgk_moe_grouped_gemv_nvf4_ex(const void* A, int a_stride, const void* B_data, const void* B_scale, void* C,
                                 const int32_t* row_expert_ids, int M, int K, int N, int expert_offset,
                                 int num_local_experts, void* stream);

/* ------------------------------------------------------------------ token sampling on the device
 * sample_token (llm/sampling.py:11-63: temperature -> softmax -> top-k -> top-p -> argmax / draw in index order) for one
 * row of bf16 logits; replaces the D2H of the logits + NumPy that generate() does per token (models/causal.py:62-71) and
 * the reference's sample_token_gpu (ops/sampling.py:11-41 -> native/ops/sampling/sampling.cu:336-362), whose top-k kernel
 * is timing-dependent and whose top-p kernel does not truncate the distribution (see sampling.cu).  The uniform number is
 * the caller's: random_val in [0, 1), or random_buf (device, one float) when not NULL -- update it and replay a graph.
 * temperature == 0: the first maximum.  Ties at a top-k / top-p cut keep the highest indices (a stable argsort's choice).
 * vocab_size <= 1048576, logits 16-byte aligned; token_out is an int32 device buffer. */
int pgk_sample_token_bf16(const void* logits, int vocab_size, float temperature, int top_k, float top_p,
                          float random_val, const float* random_buf, int32_t* token_out, void* stream);

/* Decode (num_tokens == 1) forms of MoELayer.__call__ (llm/layers/moe.py:153-318).  With one token the k
 * rows can stay in slot order, so permutation / expand / gather drop out; each entry reproduces the generic
 * kernels' arithmetic exactly (bf16 only).
 *   route_t1:          topk_with_indices + softmax_topk; row_expert_ids[j] = indices[j]
 *   gateup_fp8_t1:     act[j, n] = bf16(bf16(silu(gate_j[n])) * up_j[n])  (gate/up GEMVs + silu + mul)
 *   down_combine_fp8_t1: out[n] = bf16(sum_j w_j * bf16(down_j[n])) (+ residual[n], second rounding)
 *   combine_t1:        the scatter + residual add alone, rows in slot order */
int pgk_moe_route_t1(const void* logits, void* weights, int32_t* indices, int32_t* row_expert_ids, int num_experts,
                     int k, void* stream);
int pgk_moe_gateup_fp8_t1(const void* x, const void* Bg_enk, const void* Sg, const void* Bu_enk, const void* Su,
                          void* act, const int32_t* row_expert_ids, int k, int K, int N, int expert_offset,
                          int num_local_experts, void* stream);
/* NVF4 experts: gate/up GEMVs + silu + multiply of the k routed experts in one launch (x broadcast to the k rows) */
int pgk_moe_gateup_nvf4_t1(const void* x, const void* Bg_data, const void* Bg_scale, const void* Bu_data,
                           const void* Bu_scale, void* act, const int32_t* row_expert_ids, int k, int K, int N,
                           int expert_offset, int num_local_experts, void* stream);
int pgk_moe_down_combine_fp8_t1(const void* act, const void* Bd_enk, const void* Sd, const void* router_weights,
                                const int32_t* row_expert_ids, const void* residual, void* out, int k, int K, int N,
                                int expert_offset, int num_local_experts, void* stream);
int pgk_moe_combine_t1(const void* expert_out, const void* router_weights, const void* residual, void* out,
                       int hidden_size, int k, void* stream);

/* ------------------------------------------------------------------ synthetic init (net-new)
 * The benchmarks use random-init models (no checkpoint I/O): BF16 masters N(0, std^2) are drawn on
 * the device and quantised there.  FP8 block quantiser = the definition of LinearFP8's storage
 * (llm/layers/linear.py:101-179): codes [N,K] + bf16 scale_inv [ceil(N/128), ceil(K/128)]. */
int pgk_fill_normal_bf16(void* out, size_t n, uint32_t seed, float std, void* stream);
int pgk_quantize_bf16_to_fp8_block(const void* w_nk, void* codes, void* scale, int N, int K, void* stream);
/* *pos_buf += delta; *ctx_buf += delta (either may be NULL): device-side counters of the
 * graph-replayed decode step (position_buf of llm/decode/m1_graph.py:329-331,505-506). */
int pgk_decode_advance(int32_t* pos_buf, int32_t* ctx_buf, int delta, void* stream);

/* Multi-sequence decode (continuous batching; replaces native/ops/attention/paged_attention.cuh:46-283 and
 * native/ops/batch/continuous_batching.cuh:70-245 on the decode path): n_seq independent sequences in ONE launch of the
 * fused bracket above.  Row s of qkv [n_seq, (Hq + 2 Hkv) * 128] / out [n_seq, Hq * 128] belongs to sequence s, whose
 * cache region starts cache_stride elements after the previous one (k_cache / v_cache = [n_seq, n_kv, max_seq, 128]) and
 * whose position / context length are pos_buf[s] / ctx_buf[s] (device arrays, so one CUDA graph serves every step). */
int pgk_rope_sdpa_decode_batch(const void* qkv, const void* q_gamma, const void* k_gamma, float eps,
                               const float* cos_table, const float* sin_table, void* k_cache, void* v_cache, void* out,
                               int n_heads_q, int n_kv, int head_dim, int max_seq, int n_seq, const int32_t* pos_buf,
                               const int32_t* ctx_buf, long cache_stride, float scale, void* stream);
/* greedy tokens of a decode batch: out[r] (and out2[r] when not NULL) = argmax of row r of logits [rows, n], first
 * maximum wins (llm/sampling.py:59-61) */
int pgk_argmax_rows(const void* logits, int32_t* out, int32_t* out2, int rows, int n, int dtype, void* stream);
int pgk_decode_advance_n(int32_t* pos_buf, int32_t* ctx_buf, int delta, int n, void* stream);

/* ------------------------------------------------------------------ one-launch decode step ("step program")
 * net-new orchestration, same math.  Replaces the per-layer launch sequence of DecodeM1.step
 * (src/pygpukit/llm/decode/m1.py:40-121: ~25 launches per layer upstream, 5 per layer with the fused per-op
 * kernels above) by ONE persistent kernel per token: one CTA per SM, a producer warp streams every weight row
 * and K/V tile of the token through a shared-memory ring with cp.async.bulk (TMA) + mbarriers and runs ahead of
 * the consumer warps across op boundaries, so the HBM stream never stops while the grid synchronises (device-side
 * barrier) and re-stages the activation vector between ops.  The host describes the token's op sequence once;
 * the GEMV ops use the same per-row instruction sequences as pgk_gemv_fp8_bf16_ex / pgk_rmsnorm_gemv_* /
 * pgk_gemv_bf16_argmax, so their outputs are bit-identical to those entry points.
 *
 * kinds: 1 = W8A16 FP8 GEMV (W [N,K] e4m3, S bf16 [N/128,K/128]), 2 = BF16 GEMV (W [N,K]), 3 = NVF4 GEMV over
 * ROW-PACKED weights (pgk_nvf4_pack_rows: per output row K/2 nibble bytes then K/32 UE4M3 scale bytes),
 * 4 = QK-norm + RoPE + KV-cache write + split-KV attention (+ combine) of ONE new token (= pgk_rope_sdpa_decode).
 * flags: 1 = SWIGLU (out[n] = bf16(bf16(silu(bf16(W0 row n . x))) * bf16(W1 row n . x)), pgk_gemv_fp8_swiglu_bf16),
 * 2 = ACT_EMBED (act is an embedding table [V,K]; the row is token_buf[0]), 4 = RES_EMBED (same for residual),
 * 8 = ARGMAX (greedy token over the bf16-rounded outputs, lowest index wins ties; written to next_token AND
 * token_buf, position/context advanced by one when the program was created with advance = 1). */
typedef struct {
    int32_t kind, flags, K, N;
    const void* act;        /* bf16 [K] input vector (kind 4: the qkv row [(hq + 2 n_kv) * 128]) */
    const void* gamma;      /* bf16 [K]: RMSNorm fused into the staging (pgk_rmsnorm_gemv_*), or NULL */
    float eps, alpha;       /* alpha: NVF4 output scale */
    const void* W0; const void* S0; const void* W1; const void* S1;
    const void* residual;   /* bf16 [N] or NULL: out = bf16(acc) + residual with one rounding (hadd) */
    void* out;              /* bf16 [N] (kind 4: [hq * 128]); may be NULL with ARGMAX */
    const void* q_gamma; const void* k_gamma;         /* kind 4: QK-norm weights [128] or NULL */
    const float* cos_table; const float* sin_table;   /* kind 4: f32 [max_seq, 128] */
    void* k_cache; void* v_cache;                     /* kind 4: bf16 [n_kv, max_seq, 128] */
    int32_t hq, n_kv, max_seq, reserved;
    float scale;            /* kind 4: softmax scale (<= 0: 1/sqrt(128)) */
    float attn_eps;         /* kind 4: QK-norm eps */
} pgk_step_op_t;
/* host_ops: HOST array (copied).  token_buf / pos_buf / ctx_buf / next_token: int32 device words of the decode state
 * (position_buf of llm/decode/m1_graph.py:329-331).  Fails with a message when an op does not fit the kernel. */
int pgk_step_create(const pgk_step_op_t* host_ops, int n_ops, int32_t* token_buf, int32_t* pos_buf, int32_t* ctx_buf,
                    int32_t* next_token, int vocab, int advance, void** handle);
/* flag 16 = XCHG (tensor parallel, needs pgk_p2p_open): the op is a row-parallel GEMV whose rows are partial sums; the
 * kernel pushes them to every rank as 8-byte {payload, exchange number} words over NVLink peer memory (no fence, no
 * separate flag: one hop of latency), sums every rank's partials in rank order and writes out = bf16(sum) (+ residual)
 * -- the all-reduce of the row-parallel layers without a launch.  With a sharded lm_head (ARGMAX) the ranks exchange
 * their (value, global index) pairs the same way; `vocab` is then the global size and this sets the shard's offset. */
int pgk_step_set_vocab_offset(void* handle, int vocab_offset);
int pgk_step_run(void* handle, void* stream);           /* ONE kernel launch; CUDA-graph capturable */
int pgk_step_error(void* handle, int* host_code);       /* synchronises; 0 = every wait of the last runs completed */
int pgk_step_info(void* handle, int* ring_slots, int* act_bytes, int* smem_bytes, int* ctas);
/* profiling aid: (n_ops + 1) x 8 globaltimer (ns) stamps taken by CTA 0 during the last run (see csrc/step.cu) */
int pgk_step_trace(void* handle, unsigned long long* host_out, int count);
int pgk_step_destroy(void* handle);
/* reference NVF4 layout (data u8 [K/2,ld], scale u8 [K/32,ld], ops/matmul/nvf4.py:14-26) -> row-packed [N, K/2 + K/32];
 * output row n takes source column (n / group) * stride + n % group + col0 (plain matrix: ld = group = N, col0 = 0;
 * the gate half of a gate/up tensor interleaved in groups of 8 columns: ld = 2N, group = 8, stride = 16, col0 = 0 / 8) */
int pgk_nvf4_pack_rows(const void* data_kn, const void* scale_kn, void* packed, int K, int N, int ld, int col0, int group,
                       int stride, void* stream);

/* ------------------------------------------------------------------ tensor parallel (NCCL)
 * net-new (the reference is single-GPU): one communicator per process. unique id is 128 bytes. */
int pgk_nccl_unique_id(void* host_id128);
int pgk_nccl_init(const void* host_id128, int rank, int world);
int pgk_nccl_allreduce_bf16(void* buf, size_t count, void* stream);
int pgk_nccl_allgather(const void* send, void* recv, size_t bytes_per_rank, void* stream);
int pgk_nccl_destroy(void);

/* ------------------------------------------------------------------ tensor parallel: peer-memory all-reduce
 * net-new.  One-shot all-reduce fused with the residual add over NVLink peer memory (cudaIpc-mapped
 * buffers, flag words published with st.release.sys, sums taken in rank order so every rank holds the same
 * bits): hidden <- bf16(hidden + bf16(sum_r partial_r)).  Replaces ncclAllReduce + add_inplace for the
 * row-parallel o_proj / down_proj of the decode step (NCCL costs ~10 us per 16 KB message; this is one
 * kernel of a few us and is CUDA-graph replayable).  Setup: every rank pgk_p2p_create()s its buffer, the
 * handles are all-gathered over the host channel, every rank pgk_p2p_open()s them. */
size_t pgk_p2p_handle_bytes(void);
int pgk_p2p_create(int rank, int world, void* host_handle_out);
int pgk_p2p_open(const void* host_handles_all);
int pgk_p2p_ready(void);
/* largest message in bytes that pgk_p2p_allreduce_add_bf16 accepts (callers fall back to NCCL above it) */
size_t pgk_p2p_max_bytes(void);
int pgk_p2p_allreduce_add_bf16(const void* partial, const void* residual, void* out, size_t n, void* stream);
/* LL ("low latency") forms for the decode step: every 8-byte word carries {payload, exchange number}, so there is no
 * system fence and no separate flag -- one NVLink hop.  One decode row (n <= 8192); same result as the call above. */
int pgk_p2p_ll_allreduce_add_bf16(const void* partial, const void* residual, void* out, size_t n, void* stream);
/* greedy token of a vocabulary-sharded lm_head: token_buf[0] = local argmax index in, global token out (16 bytes per
 * rank over NVLink instead of an all-gather of the logits) */
int pgk_p2p_ll_argmax_bf16(const void* logits_local, int32_t* token_buf, int vocab_offset, void* stream);
/* Row-parallel linear of ONE token + the all-reduce after it (+ residual) in ONE launch: the thread that finishes output
 * element n publishes its bf16 partial to every rank as an LL word over peer memory and sums the ranks' words in rank order
 * -- pgk_gemv_*_bf16 followed by pgk_p2p_ll_allreduce_add_bf16, bit for bit, without the second launch and with the NVLink
 * hop hidden behind the rows still being computed.  `out` may alias `residual` (the hidden state updated in place).
 * Needs pgk_p2p_open, N <= 8192 and the streaming kernels' shape rules (they fail loudly otherwise). */
int pgk_gemv_fp8_allreduce_bf16(const void* A, const void* B_nk, const void* S, const void* residual, void* out, int K,
                                int N, void* stream);
int pgk_gemv_nvf4_allreduce_bf16(const void* A, const void* B_data, const void* B_scale, const void* residual, void* out,
                                 int K, int N, float alpha, void* stream);

/* Allocation-free variant of pgk_gemv_nvf4_bf16.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * It calls no cudaMalloc / cudaFree / allocating library routine: the split-K scratch is a file-scope device array of
 * 1028 KB (1 MB of partial sums + 1024 arrival counters) that every GPU loading the library holds until exit and that
 * all calls share -- calls that overlap on the GPU race on it.  C[M, N] = alpha * A[M, K] . dequant(B) (rows served one at
 * a time by the M = 1 kernel), same storage as pgk_gemv_nvf4_bf16.  Limits, rejected with an error otherwise:
 * K % 32 == 0, 32 <= K <= 101888, N % 16 == 0, 16 <= N <= 524288, 16-byte aligned pointers.  Results agree with
 * pgk_gemv_nvf4_bf16 within its f32-accumulate bar (the K split may differ, so the summation order may). */
int pgk_gemv_nvf4_bf16_noalloc(const void* A, const void* B_data, const void* B_scale, void* C, int M, int K, int N,
                               float alpha, void* stream);
int pgk_p2p_error(int* host_flag);
int pgk_p2p_destroy(void);

#ifdef __cplusplus
}
#endif
#endif /* PGK_B200_H */
