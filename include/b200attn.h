/*
 * b200attn.h — C ABI of libb200attn.so: the sm_100a paged-decode attention hot path.
 *
 * This is the drop-in boundary for the AR-stage hot path of ht-vllm-omni.
 * The reference has no native code of its own for this path: it calls vLLM's compiled ops
 * from `AttentionImpl.do_kv_cache_update` / `AttentionImpl.forward`.  Each entry point below
 * names the reference-side call it replaces.  `$VLLM` = vllm 0.22.0 site-packages (the
 * un-vendored dependency the reference delegates the arithmetic to; reference pins
 * vllm==0.18.0 in docs/getting_started/quickstart.md:22).
 *
 * Conventions
 *   - plain pointers and sizes only; every data pointer is a DEVICE pointer unless it says host;
 *   - strides are in ELEMENTS of the pointed-to dtype;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - stateless and capture-safe: no allocation, no host sync, no device->host copy inside;
 *   - return 0 on success, a negative B200ATTN_E* code otherwise; b200attn_last_error() then
 *     returns a thread-local, human-readable host string.  Nothing throws.
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *     B200ATTN_ECUDA.
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ATTN_ABI_VERSION 8

#if defined(__GNUC__)
#define B200ATTN_API __attribute__((visibility("default")))
#else
#define B200ATTN_API
#endif

/* element types */
#define B200ATTN_BF16 0
#define B200ATTN_FP16 1
#define B200ATTN_FP8_E4M3 2 /* KV cache only; 1 byte per element */

/* error codes */
#define B200ATTN_OK 0
#define B200ATTN_EINVAL (-1)  /* bad argument / unsupported shape */
#define B200ATTN_ECUDA (-2)   /* CUDA runtime / driver error (incl. no device) */
#define B200ATTN_ENOSPACE (-3) /* plan or workspace buffer too small */

/* b200attn_decode flags */
/* b200attn_rope_kv_write flags */
#define B200ATTN_ROPE_NEOX 1           /* partners are the halves of the rotary span (else even / odd elements) */
#define B200ATTN_ROPE_ROUND_EACH_OP 2  /* torch-native rounding of every elementary op (default: fp32, one rounding) */
#define B200ATTN_ROPE_MROPE_INTERLEAVED 4 /* M-RoPE sections interleaved [THWTHW...TT] instead of chunked [T..H..W..] */

#define B200ATTN_KV_DYN_IEEE_DIV 1 /* b200attn_kv_write_per_token_head: IEEE divisions instead of the compiled reference's div.full.f32 */

#define B200ATTN_DECODE_FUSED_MERGE 1 /* merge split-KV partials inside the decode kernel (no b200attn_merge call) */

/* geometry constants the host side needs */
#define B200ATTN_TILE_TOKENS 64 /* scheduling granule of the split-KV decode kernel */
#define B200ATTN_MAX_GROUP 16   /* query heads sharing one KV head handled per CTA */
#define B200ATTN_PEER_CTAS 32   /* CTAs of the peer-exchange kernels */

B200ATTN_API int b200attn_abi_version(void);
B200ATTN_API const char* b200attn_last_error(void);

/* Number of SMs of the current device (0 if none); decode grids are sized from it. */
B200ATTN_API int b200attn_sm_count(void);

/*
 * Kernel 1 — KV write + in-register FP8 quantisation ("reshape_and_cache").
 * Replaces: torch.ops._C_cache_ops.reshape_and_cache_flash as called from
 *   $VLLM/v1/attention/backends/flashinfer.py:1828-1855 (FlashInferImpl.do_kv_cache_update) and
 *   $VLLM/v1/attention/backends/triton_attn.py:720-763 (spec: ops/triton_reshape_and_cache_flash.py:17-124),
 *   reached from vllm_omni/worker/gpu_ar_model_runner.py:228-298 via unified_kv_cache_update.
 *
 * For every token t with slot = slot_mapping[t] >= 0:
 *   cache[slot / block_size][slot % block_size][h][:] = cvt(key[t][h][:] / k_scale[h*scale_stride])
 * fp8: fp32 IEEE division then round-to-nearest-even, saturate-to-finite e4m3; 16-bit: bit copy.
 * k_cache / v_cache are the base pointers of the K and V planes; both use the same
 * (stride_block, stride_slot, stride_head) and a unit stride along D.
 */
B200ATTN_API int b200attn_kv_write(const void* key, const void* value,
                      int64_t key_stride_t, int64_t key_stride_h,
                      int64_t value_stride_t, int64_t value_stride_h,
                      void* k_cache, void* v_cache,
                      int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                      const int64_t* slot_mapping, int32_t num_tokens,
                      int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                      int32_t in_dtype, int32_t kv_dtype,
                      const float* k_scale, const float* v_scale, int32_t scale_stride,
                      void* stream);

/*
 * Kernel 1 with per-(token, head) dynamic FP8 scales (kv_cache_dtype "fp8_per_token_head").
 * Replaces: triton_reshape_and_cache_flash_per_token_head_quant
 *   ($VLLM/v1/attention/ops/triton_reshape_and_cache_flash.py:138-300, called from TritonAttentionImpl.do_kv_cache_update,
 *   $VLLM/v1/attention/backends/triton_attn.py:733-747).
 * For every token t with slot >= 0 and every kv head h, for K and for V:
 *   scale = max(max_d |x[t][h][d]| / 448, 1e-6);  cache byte = e4m3(clamp(x * (1 / scale), -448, 448));
 *   scale_cache[slot / block_size][slot % block_size][h] = scale    (float32, strides in float32 elements).
 * Both divisions are the compiled reference's div.full.f32 (bit-identical bytes and scales to the Triton kernel on the GPU);
 * B200ATTN_KV_DYN_IEEE_DIV selects IEEE division (what numpy and the Triton interpreter compute).  head_dim: power of two.
 */
B200ATTN_API int b200attn_kv_write_per_token_head(const void* key, const void* value,
                      int64_t key_stride_t, int64_t key_stride_h,
                      int64_t value_stride_t, int64_t value_stride_h,
                      void* k_cache, void* v_cache,
                      int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                      float* k_scale_cache, float* v_scale_cache,
                      int64_t scale_stride_block, int64_t scale_stride_slot, int64_t scale_stride_head,
                      const int64_t* slot_mapping, int32_t num_tokens,
                      int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                      int32_t in_dtype, int32_t flags, void* stream);

/*
 * Kernel 1b -- fused rotary embedding + KV write (the step before the path).
 * Replaces, in one launch: RotaryEmbedding.forward(positions, q, k)
 *   ($VLLM/model_executor/layers/rotary_embedding/base.py:130-180, common.py:144-183; torch.ops._C.rotary_embedding)
 *   followed by reshape_and_cache_flash, i.e. the body of AttentionImpl.do_rope_and_kv_cache_update
 *   ($VLLM/v1/attention/backend.py:815-840; call order $VLLM/model_executor/models/qwen3.py:153-160).
 *
 * query [T][Hq][D] and key [T][Hkv][D] are rotated IN PLACE over their first rotary_dim elements (neox: halves of the span
 * are partners, B200ATTN_ROPE_NEOX; otherwise even/odd elements):  o1 = rnd(x1*cos - x2*sin), o2 = rnd(x2*cos + x1*sin),
 * computed in fp32 and rounded once to the activation dtype -- bit-identical to the compiled op a GPU run of the reference
 * executes.  With B200ATTN_ROPE_ROUND_EACH_OP every elementary operation is rounded instead, bit-identical to
 * RotaryEmbedding.forward_native (torch bf16/fp16 tensor arithmetic).  cos_sin_cache is vLLM's
 * [max_position][rotary_dim] table (cos half, then sin half) in the activation dtype.  The rotated key and the value are
 * then written to the cache with b200attn_kv_write's semantics (slot < 0: no cache write; the rotation still happens).
 * rotary_dim must be a multiple of 16.
 */
B200ATTN_API int b200attn_rope_kv_write(void* query, int64_t q_stride_t, int64_t q_stride_h, int32_t num_q_heads,
                           void* key, const void* value,
                           int64_t key_stride_t, int64_t key_stride_h,
                           int64_t value_stride_t, int64_t value_stride_h,
                           const int64_t* positions, const void* cos_sin_cache, int32_t rotary_dim, int32_t rope_flags,
                           void* k_cache, void* v_cache,
                           int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                           const int64_t* slot_mapping, int32_t num_tokens,
                           int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                           int32_t in_dtype, int32_t kv_dtype,
                           const float* k_scale, const float* v_scale, int32_t scale_stride,
                           void* stream);

/*
 * Kernel 1c -- q/k RMSNorm + (multimodal) rotary embedding + KV write, one launch: b200attn_rope_kv_write with the two steps
 * the Qwen3 stages put around it.
 * Replaces: q_norm / k_norm ($VLLM/model_executor/models/qwen3.py:153-158; RMSNorm.forward,
 *   $VLLM/model_executor/layers/layernorm.py, csrc/layernorm_kernels.cu), the rotary embedding -- RotaryEmbedding.forward or,
 *   for the Qwen3-Omni thinker's 3-axis positions (vllm_omni/model_executor/models/qwen3_omni/qwen3_omni_moe_thinker.py:
 *   1633-1728), MRotaryEmbedding.forward ($VLLM/model_executor/layers/rotary_embedding/mrope.py:263-360) -- and
 *   reshape_and_cache_flash; the sequence vLLM itself fuses as fused_qk_norm_rope + unified_kv_cache_update.
 *
 *   q_norm_weight / k_norm_weight  nullable, [head_dim] in the activation dtype: every q (k) head becomes
 *                          rnd(rnd(x * rsqrt(mean(x^2) + rms_eps)) * w) before the rotation (fp32 statistics, the two roundings
 *                          of the reference's kernel); elements beyond rotary_dim are normalised too.  In place.
 *   positions_axis_stride  0: positions is int64 [T] (plain RoPE, mrope_* ignored);
 *                          else positions is int64 [3][T] (temporal, height, width), axes this many elements apart, and rotary
 *                          frequency f reads cos / sin from the cache row of the position of ITS axis: chunked sections
 *                          [0, t) -> T, [t, t + h) -> H, rest -> W, or interleaved (B200ATTN_ROPE_MROPE_INTERLEAVED): f % 3 == 1
 *                          and f < 3 h -> H, f % 3 == 2 and f < 3 w -> W, else T.  mrope_t + mrope_h + mrope_w = rotary_dim / 2.
 * Everything else as b200attn_rope_kv_write.
 */
B200ATTN_API int b200attn_norm_rope_kv_write(void* query, int64_t q_stride_t, int64_t q_stride_h, int32_t num_q_heads,
                           void* key, const void* value,
                           int64_t key_stride_t, int64_t key_stride_h,
                           int64_t value_stride_t, int64_t value_stride_h,
                           const int64_t* positions, int64_t positions_axis_stride,
                           int32_t mrope_t, int32_t mrope_h, int32_t mrope_w,
                           const void* cos_sin_cache, int32_t rotary_dim, int32_t rope_flags,
                           const void* q_norm_weight, const void* k_norm_weight, float rms_eps,
                           void* k_cache, void* v_cache,
                           int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                           const int64_t* slot_mapping, int32_t num_tokens,
                           int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                           int32_t in_dtype, int32_t kv_dtype,
                           const float* k_scale, const float* v_scale, int32_t scale_stride,
                           void* stream);

/*
 * Split-KV work plan (once per decode step, shared by all layers).
 * Replaces: the per-step planning in AttentionMetadataBuilder.build
 *   ($VLLM/v1/attention/backends/flashinfer.py:886-1248 `plan`, triton_attn.py:216-268).
 *
 * Cuts the flat list of (sequence, kv head, 64-token tile) work items into equal contiguous
 * ranges, one per persistent CTA, entirely on the device (no host sync, capture-safe), and copies into
 * each CTA's 256-byte plan entry everything its first TMA load needs (sequence length, block-table row,
 * the first 32 block ids), so a decode CTA starts streaming after ONE dependent load.
 * `block_table` / `block_table_stride` / `row_map` (nullable) / `block_size` must be the ones later given to
 * b200attn_decode (the step's metadata: the same for every layer).
 * `plan` must hold b200attn_plan_bytes(num_seqs) bytes.  `num_ctas` is the decode grid: use
 * b200attn_default_ctas(kv_dtype) (2 CTAs/SM); tests pass small values
 * to force many split points.  The SAME num_ctas must be given to decode and merge (the decode kernel traps on a
 * plan built for another grid or batch).
 */
B200ATTN_API size_t b200attn_plan_bytes(int32_t max_num_seqs);
B200ATTN_API int b200attn_default_ctas(int32_t kv_dtype);
B200ATTN_API int b200attn_decode_plan(const int32_t* seq_lens, int32_t num_seqs,
                         int32_t num_q_heads, int32_t num_kv_heads,
                         int32_t num_ctas,
                         const int32_t* block_table, int64_t block_table_stride,
                         const int32_t* row_map /* nullable, [num_seqs] */, int32_t block_size,
                         void* plan, size_t plan_bytes, void* stream);

/* Bytes of split-KV workspace (partial o, m, l) the decode + merge kernels need. */
B200ATTN_API size_t b200attn_workspace_bytes(int32_t head_dim, int32_t num_ctas);

/*
 * Kernel 2 — split-KV paged flash-decoding.
 * Replaces: the decode branch of AttentionImpl.forward, i.e.
 *   flashinfer.decode.trtllm_batch_decode_with_kv_cache ($VLLM/v1/attention/backends/flashinfer.py:1751-1822)
 *   and kernel_unified_attention ($VLLM/v1/attention/ops/triton_unified_attention.py:178-643).
 *
 * out[b][hq][:] = softmax(sm_scale * k_scale * q[b][hq] . K[b][hq/G]^T) . V[b][hq/G] * v_scale over the
 * first seq_lens[b] tokens addressed through block_table[row(b)][*], row(b) = row_map ? row_map[b] : b
 * (row_map lets several decode rows -- the query tokens of one request in a spec-decode / chunked-prefill
 * batch, each with its own causal seq_len -- share one block-table row).  Sequences whose whole range falls
 * inside one CTA are written to `out` directly; the others leave (o, m, l) partials in `workspace`.  With
 * B200ATTN_DECODE_FUSED_MERGE the last CTA to deliver a partial of a pair reduces them in-kernel (arrival
 * counters live in the plan buffer); without it b200attn_merge must follow.
 * q/out dtype = q_dtype; kv_dtype fp8 accepts either q dtype, 16-bit KV must equal q_dtype.  head_dim 128 runs the
 * tensor-core kernels (K/V planes read through TMA); head_dim 64 and 256 run a CUDA-core split-KV kernel that ignores the
 * plan and folds its own partials in `workspace` (decode_simt.cu).  block_size a multiple of 16; base pointers and byte
 * strides 16-byte aligned.
 * fresh_tokens: upper bound on how many TRAILING positions of any row the kernel launched immediately before
 * this one on `stream` may have written (1 for a decode step's KV write, max_query_len in general); K/V tiles
 * below seq_len - fresh_tokens are prefetched before the programmatic-dependent-launch wait.  Pass -1 when
 * unknown (e.g. the preceding kernel filled the whole cache): nothing is read early.
 */
B200ATTN_API int b200attn_decode(const void* q, int64_t q_stride_t, int64_t q_stride_h,
                    void* out, int64_t out_stride_t, int64_t out_stride_h,
                    const void* k_cache, const void* v_cache,
                    int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                    int64_t num_blocks,
                    const int32_t* block_table, int64_t block_table_stride,
                    const int32_t* seq_lens, int32_t num_seqs,
                    int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                    int32_t block_size, int32_t q_dtype, int32_t kv_dtype,
                    float sm_scale, const float* k_scale, const float* v_scale,
                    int32_t scale_stride,
                    const void* plan, int32_t num_ctas,
                    void* workspace, size_t workspace_bytes,
                    float* lse_out /* nullable, [num_seqs][num_q_heads] natural-log LSE */,
                    const int32_t* row_map /* nullable, [num_seqs] */, int32_t fresh_tokens, int32_t flags,
                    void* stream);

/*
 * Kernels 1 + 2 in one call for a PURE-DECODE step (one query token per row): the new token of row b -- key[b], value[b],
 * destined for slot_mapping[b] (< 0: not written) -- is quantised and stored to the cache with b200attn_kv_write's
 * arithmetic (bit-identical bytes), and attention then runs over seq_lens[b] positions that include it.
 * Replaces: do_kv_cache_update + forward of a backend that declares `forward_includes_kv_cache_update = True`
 *   ($VLLM/v1/attention/backend.py:66; the runner then hands forward() the UNPADDED slot_mapping,
 *   vllm_omni/worker/gpu_ar_model_runner.py:231-244), i.e. reshape_and_cache_flash followed by the decode branch.
 * With an FP8 cache, <= 8 query heads per kv head and no row_map, the write happens INSIDE the decode kernel: the CTA that
 * streams the tile holding a row's last position stores the row's new token right before its TMA reads the tile back
 * (generic-proxy store, fence.proxy.async, async-proxy load) -- one launch per layer instead of two and no kernel
 * boundary between the write and the read.  Every other shape (16-bit cache, > 8 heads per kv head, row_map != NULL, or
 * option "fused_write" = 0) runs the KV-write kernel and then the decode kernel from this one call: same bytes, same output.
 * key / value dtype = q_dtype.  All other arguments as b200attn_decode.
 */
B200ATTN_API int b200attn_decode_kv_write(const void* q, int64_t q_stride_t, int64_t q_stride_h,
                    void* out, int64_t out_stride_t, int64_t out_stride_h,
                    const void* key, const void* value,
                    int64_t key_stride_t, int64_t key_stride_h, int64_t value_stride_t, int64_t value_stride_h,
                    const int64_t* slot_mapping,
                    void* k_cache, void* v_cache,
                    int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                    int64_t num_blocks,
                    const int32_t* block_table, int64_t block_table_stride,
                    const int32_t* seq_lens, int32_t num_seqs,
                    int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                    int32_t block_size, int32_t q_dtype, int32_t kv_dtype,
                    float sm_scale, const float* k_scale, const float* v_scale,
                    int32_t scale_stride,
                    const void* plan, int32_t num_ctas,
                    void* workspace, size_t workspace_bytes,
                    float* lse_out /* nullable */,
                    const int32_t* row_map /* nullable */, int32_t fresh_tokens, int32_t flags,
                    void* stream);

/*
 * Kernel 2 over a cache with per-(token, head) dynamic FP8 scales (kv_cache_dtype "fp8_per_token_head").
 * Replaces: kernel_unified_attention with KV_QUANT_MODE = FP8_PER_TOKEN_HEAD
 *   ($VLLM/v1/attention/ops/triton_unified_attention.py:472-545: S = (Q.K8^T) * (scale * k_scale[token]),
 *   acc += (P * v_scale[token]) . V8).
 * k_scale_cache / v_scale_cache: float32 views [block][slot][head] of the side caches b200attn_kv_write_per_token_head
 * fills, with slot stride 1 (the layout this library's backend allocates: 16 consecutive scales are one 64-byte copy),
 * block / head strides in float32 elements.  fp8 cache only; every other argument as b200attn_decode (no static scales).
 * The merge of split-KV partials is in-kernel (B200ATTN_DECODE_FUSED_MERGE) or by b200attn_merge with v_scale = NULL.
 */
B200ATTN_API int b200attn_decode_per_token_head(const void* q, int64_t q_stride_t, int64_t q_stride_h,
                    void* out, int64_t out_stride_t, int64_t out_stride_h,
                    const void* k_cache, const void* v_cache,
                    int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                    int64_t num_blocks,
                    const float* k_scale_cache, const float* v_scale_cache,
                    int64_t scale_stride_block, int64_t scale_stride_slot, int64_t scale_stride_head,
                    const int32_t* block_table, int64_t block_table_stride,
                    const int32_t* seq_lens, int32_t num_seqs,
                    int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                    int32_t block_size, int32_t q_dtype, float sm_scale,
                    const void* plan, int32_t num_ctas,
                    void* workspace, size_t workspace_bytes,
                    float* lse_out /* nullable */, const int32_t* row_map /* nullable */,
                    int32_t fresh_tokens, int32_t flags, void* stream);

/*
 * Plan + split-KV paged decode WITHOUT caller-provided scratch and without any device allocation.
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 *
 * Same result as b200attn_decode_plan followed by b200attn_decode(..., B200ATTN_DECODE_FUSED_MERGE) with
 * row_map = NULL and num_ctas = min(b200attn_default_ctas(kv_dtype), B200ATTN_STATIC_MAX_CTAS); the caller still passes the
 * inputs and `out` (and `lse_out`, nullable).  The plan buffer and the split-KV workspace live in ONE file-scope
 * `static __device__` array inside the library (csrc/decode_static.cu; < 6 MiB, held by every GPU that loads the library
 * until the process exits), whose device address is looked up once per device with cudaGetSymbolAddress.  The call performs
 * no cudaMalloc / cudaFree and calls no allocating library routine.
 * Limits: head_dim 128 only (B200ATTN_EINVAL otherwise); num_seqs <= B200ATTN_STATIC_MAX_SEQS (B200ATTN_ENOSPACE beyond, with
 * a message naming the limit).  All other arguments, dtypes and alignment rules as b200attn_decode.
 * Every call shares the array: nothing in it needs to survive between calls (the plan entries and merge counters are
 * rewritten / zeroed by each call's plan kernel), and two calls that OVERLAP on the GPU -- different streams, or graph
 * branches without an edge between them -- race on it.  Calls on one stream are ordered and safe.
 */
#define B200ATTN_STATIC_MAX_SEQS 1024   /* decode rows per call */
#define B200ATTN_STATIC_MAX_CTAS 320    /* decode grid the static workspace is sized for (2 CTAs/SM on 148 SMs = 296) */
B200ATTN_API int b200attn_decode_static(const void* q, int64_t q_stride_t, int64_t q_stride_h,
                    void* out, int64_t out_stride_t, int64_t out_stride_h,
                    const void* k_cache, const void* v_cache,
                    int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                    int64_t num_blocks,
                    const int32_t* block_table, int64_t block_table_stride,
                    const int32_t* seq_lens, int32_t num_seqs,
                    int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                    int32_t block_size, int32_t q_dtype, int32_t kv_dtype,
                    float sm_scale, const float* k_scale, const float* v_scale,
                    int32_t scale_stride,
                    float* lse_out /* nullable, [num_seqs][num_q_heads] natural-log LSE */,
                    int32_t fresh_tokens, void* stream);

/*
 * Kernel 3 — merge of split-KV partials (online-softmax reduction with warp shuffles).
 * Replaces: reduce_segments ($VLLM/v1/attention/ops/triton_unified_attention.py:645-733) and the
 *   in-kernel reduction of the TRT-LLM-gen "MultiCtasKv" cubins.
 * Must follow a b200attn_decode launched WITHOUT B200ATTN_DECODE_FUSED_MERGE on the same stream with the same
 * plan / num_ctas / workspace / out.
 */
B200ATTN_API int b200attn_merge(const void* plan, int32_t num_ctas, const void* workspace,
                   void* out, int64_t out_stride_t, int64_t out_stride_h,
                   int32_t num_seqs, int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim,
                   int32_t out_dtype, const float* v_scale, int32_t scale_stride,
                   float* lse_out, void* stream);

/*
 * Cross-rank / cross-shard merge by log-sum-exp (sequence-split "DCP" exchange):
 *   out = sum_p o_p * exp(lse_p - lse) / sum_p exp(lse_p - lse).
 * Replaces: _C.merge_attn_states ($VLLM/v1/attention/ops/merge_attn_states.py:7-101) and
 *   cp_lse_ag_out_rs ($VLLM/v1/attention/backends/flashinfer.py:1714-1738), generalised to P parts.
 * parts_o: [P][n][Hq][D] (dtype), parts_lse: [P][n][Hq] fp32 (natural log).  -inf LSE = empty part.
 */
B200ATTN_API int b200attn_merge_lse(const void* parts_o, const float* parts_lse, int32_t num_parts,
                       int64_t part_stride_o, int64_t part_stride_lse,
                       void* out, float* lse_out, int32_t n, int32_t num_q_heads,
                       int32_t head_dim, int32_t dtype, void* stream);

/*
 * The same combine over PEER memory (NVLink / NVSwitch): part p is read from parts_o_ptrs[p] + o_offset (elements of
 * dtype, laid out [n][Hq][D]) and parts_lse_ptrs[p] + lse_offset (fp32 [n][Hq]) -- DEVICE arrays of num_parts pointers,
 * e.g. torch symmetric-memory `buffer_ptrs_dev`, so the ranks of a sequence-split group exchange their partials with
 * plain peer loads inside the merge kernel instead of two NCCL all-gathers.  The caller orders the peers' writes before
 * this launch (a symmetric-memory barrier) and alternates two offsets so a buffer is not rewritten while a peer reads it.
 * Replaces: cp_lse_ag_out_rs ($VLLM/v1/attention/backends/flashinfer.py:1714-1738).
 */
B200ATTN_API int b200attn_merge_lse_peers(const void* const* parts_o_ptrs, const float* const* parts_lse_ptrs,
                             int32_t num_parts, int64_t o_offset, int64_t lse_offset,
                             void* out, float* lse_out, int32_t n, int32_t num_q_heads,
                             int32_t head_dim, int32_t dtype, void* stream);

/*
 * Sequence-split exchange by PUSHES over NVLink peer memory (tp > Hkv: the ranks of a replica group share one KV head, each
 * holds 1/R of the context).  Two launches per layer around the decode; no collective, no barrier, no memory fence:
 *
 *   b200attn_peer_gather_heads   every rank pushes its heads_local query heads to every peer and unpacks what the peers pushed:
 *                                afterwards the LOCAL buffer `gathered` [n][R * heads_local][head_dim] holds the group's query
 *                                heads (rank j's at heads [j * heads_local, (j + 1) * heads_local)).  Replaces the all-gather of
 *                                query heads of the DCP flow ($VLLM/v1/attention/backends/flashinfer.py:1714-1738).
 *   b200attn_peer_combine        every rank pushes, for each peer, the slice of its partial (o_part [n][R * heads_local][D],
 *                                lse_part [n][R * heads_local], natural log, -inf = empty) that covers THAT peer's heads, and
 *                                merges its own slice with the slices it receives by log-sum-exp into out [n][heads_local][D].
 *                                Replaces cp_lse_ag_out_rs / merge_attn_states
 *                                ($VLLM/v1/attention/ops/merge_attn_states.py:7-101).
 *
 * Transport: "LL" words -- each 8-byte store carries 4 payload bytes and the 4-byte launch count of the exchange; the receiver
 * polls its LOCAL receive buffer until the count matches, so the data is its own flag.  `ll_ptrs` is a DEVICE array of
 * num_ranks per-rank base pointers of a symmetric, ZERO-INITIALISED allocation (e.g. torch symmetric memory
 * `buffer_ptrs_dev`) of b200attn_peer_ll_bytes(...) bytes per rank; each op needs its own allocation and its own LOCAL
 * zero-initialised `epoch` words (2 x int32).  Nothing is ever reset, so CUDA-graph replay is safe.  Every rank of the group
 * must issue the same sequence of calls with the same n.  A peer that never arrives traps the kernel after ~4 s instead of
 * hanging the device.
 */
B200ATTN_API size_t b200attn_peer_ll_bytes(int32_t rows_capacity, int32_t heads_local, int32_t head_dim, int32_t num_ranks,
                              int32_t with_lse /* 0: gather_heads, 1: combine */);
B200ATTN_API int b200attn_peer_gather_heads(const void* local, int64_t stride_t, int64_t stride_h, int32_t n,
                               int32_t rows_capacity, int32_t heads_local, int32_t head_dim, int32_t dtype,
                               void* gathered, void* const* ll_ptrs, int32_t* epoch,
                               int32_t rank, int32_t num_ranks, void* stream);
B200ATTN_API int b200attn_peer_combine(const void* o_part, const float* lse_part, int32_t n, int32_t rows_capacity,
                          int32_t heads_local, int32_t head_dim, int32_t dtype,
                          void* const* ll_ptrs, int32_t* epoch, int32_t rank, int32_t num_ranks,
                          void* out, int64_t out_stride_t, int64_t out_stride_h, void* stream);

/*
 * Prefill / mixed-batch rows (q_len > 1): varlen causal flash-attention forward over the paged FP8 (e4m3) cache.
 * Replaces the prefill branch of AttentionImpl.forward: kernel_unified_attention with q_len > 1
 * ($VLLM/v1/attention/ops/triton_unified_attention.py:286-303, 440-545) and FlashInfer's
 * trtllm_batch_context_with_kv_cache ($VLLM/v1/attention/backends/flashinfer.py:1500-1689).  A query token at position
 * t of a request's chunk of q_len tokens attends to cache positions [0, seq_len - q_len + t + 1).
 *
 *   b200attn_prefill_tile_tokens  query tokens per 128-row tile (128 / query heads per kv head; 0 for unsupported heads)
 *   b200attn_prefill_plan         tile_prefix[num_reqs + 1] (int32, device) = exclusive scan of ceil(q_len / tile_tokens) over
 *                                 the requests with q_len >= min_query_len (the others get no tile: in a mixed batch the
 *                                 decode kernel serves them, so the routing needs no host-side split and no batch
 *                                 reordering); once per scheduler step (shared by all layers), capture-safe, no host sync
 *   b200attn_prefill              one launch: grid = (ceil(num_tokens / tile_tokens) + num_reqs) * kv heads CTAs (host-side
 *                                 bound, surplus CTAs exit), tcgen05 (kind::f8f6f4) S = Q.K^T and O += P.V with TMEM accumulators.
 *
 * seq_lens / query_start_loc / block_table / tile_prefix are the rows of the requests being prefilled (pointers pre-offset by
 * the caller when decode rows come first); q and out are the full [tokens, heads, head_dim] tensors and `first_token` is
 * subtracted from every query_start_loc entry before indexing them (pass query_start_loc[first request] when q / out are
 * pre-offset, else 0).  `num_tokens` bounds the grid: the number of query tokens covered by these requests.
 * fp8_e4m3 cache, bf16 / fp16 q and out, head_dim 128, per-tensor (scale_stride 0) or per-head (1) scales.
 */
B200ATTN_API int b200attn_prefill_tile_tokens(int32_t num_q_heads, int32_t num_kv_heads);
B200ATTN_API int b200attn_prefill_plan(const int32_t* query_start_loc, int32_t num_reqs, int32_t num_q_heads,
                                       int32_t num_kv_heads, int32_t min_query_len, int32_t* tile_prefix, void* stream);
B200ATTN_API int b200attn_prefill(const void* q, int64_t q_stride_t, int64_t q_stride_h,
                                  void* out, int64_t out_stride_t, int64_t out_stride_h,
                                  const void* k_cache, const void* v_cache,
                                  int64_t stride_block, int64_t stride_slot, int64_t stride_head, int64_t num_blocks,
                                  const int32_t* block_table, int64_t block_table_stride,
                                  const int32_t* seq_lens, const int32_t* query_start_loc,
                                  int32_t num_reqs, int32_t num_tokens, int32_t first_token,
                                  int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                                  int32_t q_dtype, int32_t kv_dtype, float sm_scale,
                                  const float* k_scale, const float* v_scale, int32_t scale_stride,
                                  const int32_t* tile_prefix, void* stream);

/*
 * KV hand-off gather (the step after the path).
 * Replaces: the per-layer index -> flatten -> truncate loop of OmniKVTransferManager._extract_kv_cache
 *   (vllm_omni/distributed/omni_connectors/kv_transfer_manager.py:224-301; layouts utils/kv_utils.py:13-85).
 * For every layer l and plane in {K, V}:  out[l][plane][i][h][:] = plane_l[block_ids[i / block_size]][i % block_size][h][:]
 * for i < n_tokens (the caller passes n_tokens = min(seq_len, n_ids * block_size) and only in-range ids, as the reference
 * filters them; an out-of-range id is zero-filled, never dereferenced).  Raw bytes: row_bytes = head_dim * element size.
 * k_planes / v_planes are DEVICE arrays of num_layers plane base pointers; all layers share the three BYTE strides.
 * `out` holds num_layers * 2 * n_tokens * num_kv_heads * row_bytes bytes; one pinned D2H of it replaces 2 x num_layers copies.
 */
B200ATTN_API int b200attn_kv_gather(const void* const* k_planes, const void* const* v_planes, int32_t num_layers,
                       int64_t stride_block_bytes, int64_t stride_slot_bytes, int64_t stride_head_bytes,
                       int64_t num_blocks, const int32_t* block_ids, int32_t n_tokens, int32_t num_kv_heads,
                       int32_t block_size, int32_t row_bytes, void* out, void* stream);

/*
 * Runtime knobs (host, process-wide): "pdl" (programmatic dependent launch, default 1), "tc" (tcgen05 consumer for
 * fp8 KV with group size <= 8, default 1), "cluster" (thread-block cluster size of that kernel: partials of a split
 * pair whose CTAs share a cluster are merged through distributed shared memory; default 2, 1 = off), "fused_write"
 * (b200attn_decode_kv_write stores the new token inside the decode kernel when the shape allows, default 1), "big_box"
 * (tcgen05 decode: one TMA box per min(block_size, 64) tokens instead of one per 16, default 1).
 * Returns the previous value or B200ATTN_EINVAL for an unknown name.
 */
B200ATTN_API int b200attn_set_option(const char* name, int32_t value);

/*
 * Debug only (tests/tools/timeline*.py): a DEVICE buffer of [num_ctas][8] uint64 that the next decode launches fill with
 * %globaltimer stamps (CTA entry, set-up done, first TMA, first tile, last tile, exit, publish / merge time).  While a
 * buffer is set, b200attn_decode launches the instrumented instantiation of the kernel; NULL (the default) restores the
 * production one, which carries no instrumentation at all.
 */
B200ATTN_API void b200attn_debug_set_timeline_buffer(void* dev_ptr);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
