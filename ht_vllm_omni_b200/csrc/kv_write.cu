// kv_write.cu — Kernel 1: paged KV write with in-register FP8 (e4m3, satfinite) quantisation.
//
// Semantics follow the reference's reshape_and_cache path
//   ($VLLM/v1/attention/ops/triton_reshape_and_cache_flash.py:44-124 and the CUDA op
//    _C_cache_ops.reshape_and_cache_flash called at $VLLM/v1/attention/backends/flashinfer.py:1846-1855):
//   slot < 0 -> skip; block = slot / bs, off = slot % bs; value = x / scale (fp32, IEEE divide),
//   then round-to-nearest-even + saturate-to-finite to e4m3.  16-bit caches are a bit copy.
//
// Layout: one row = one (token, K|V, kv head) = D contiguous elements.  Each thread moves 16 elements:
// two 16-byte loads, and one 16-byte (fp8) or two 16-byte (16-bit) stores, so a row is written by
// D/16 consecutive lanes -> full 128-byte lines on the HND cache layout.
#include "common.cuh"

namespace b200attn {

struct KvWriteParams {
  const void* key;
  const void* value;
  int64_t key_stride_t, key_stride_h, value_stride_t, value_stride_h;
  void* k_cache;
  void* v_cache;
  int64_t stride_block, stride_slot, stride_head;
  const int64_t* slot_mapping;
  const float* k_scale;
  const float* v_scale;
  int scale_stride;
  int T, Hkv, D, block_size;
  FastDiv d_lanes, d_heads, d_bs;   // threads per row (D / 16), kv heads, block size
};

template <int IN_DT>
__device__ __forceinline__ void unpack8(const uint4& raw, float (&f)[8]) {
  using E = Elem<IN_DT>;
  const typename E::type* e = reinterpret_cast<const typename E::type*>(&raw);
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] = E::to_float(e[i]);
}

__device__ __forceinline__ uint32_t quant4(const float* f, float scale) {
  // x / scale with IEEE rounding (never the approximate divide), then cvt.rn.satfinite.e4m3x2.f32
  float2 lo = make_float2(__fdiv_rn(f[0], scale), __fdiv_rn(f[1], scale));
  float2 hi = make_float2(__fdiv_rn(f[2], scale), __fdiv_rn(f[3], scale));
  uint32_t a = __nv_cvt_float2_to_fp8x2(lo, __NV_SATFINITE, __NV_E4M3);
  uint32_t b = __nv_cvt_float2_to_fp8x2(hi, __NV_SATFINITE, __NV_E4M3);
  return a | (b << 16);
}

template <int IN_DT, bool FP8>
__global__ void __launch_bounds__(256) kv_write_kernel(const KvWriteParams p) {
  // 32-bit index math with multiply-shift divisions (the host rejects launches with >= 2^31 threads)
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t row, lane_in_row;
  p.d_lanes.divmod(gid, row, lane_in_row);
  const int col = (int)lane_in_row << 4;
  const uint32_t rows = (uint32_t)p.T * 2 * p.Hkv;
  // Dependent launch: let the decode kernel that follows become resident right away (it prefetches K/V tiles that
  // this kernel cannot touch and only then waits for us); everything above is index math, wait for the producer of
  // key/value here.
  pdl_launch_dependents();
  pdl_wait();
  if (row >= rows) return;
  uint32_t tkv, hh;
  p.d_heads.divmod(row, tkv, hh);
  const int h = (int)hh, kv = (int)(tkv & 1), t = (int)(tkv >> 1);
  // the slot, the two data vectors and the scale are independent loads: issue them together (one memory round trip on the
  // critical path between two decode kernels instead of two); a padding token's key/value row exists, reading it is harmless
  using E = Elem<IN_DT>;
  const typename E::type* src =
      kv ? reinterpret_cast<const typename E::type*>(p.value) + (int64_t)t * p.value_stride_t + (int64_t)h * p.value_stride_h
         : reinterpret_cast<const typename E::type*>(p.key) + (int64_t)t * p.key_stride_t + (int64_t)h * p.key_stride_h;
  const int64_t slot = __ldg(p.slot_mapping + t);
  const uint4 r0 = __ldg(reinterpret_cast<const uint4*>(src + col));
  const uint4 r1 = __ldg(reinterpret_cast<const uint4*>(src + col + 8));
  float scale = 1.0f;
  if constexpr (FP8) scale = __ldg((kv ? p.v_scale : p.k_scale) + (int64_t)h * p.scale_stride);
  if (slot < 0) return;  // padding token
  uint32_t blk32, off32;
  p.d_bs.divmod((uint32_t)slot, blk32, off32);   // slots are < num_blocks * block_size < 2^32
  const int64_t blk = blk32, off = off32;
  const int64_t dst_off = blk * p.stride_block + off * p.stride_slot + (int64_t)h * p.stride_head + col;

  if constexpr (FP8) {
    float f[16];
    unpack8<IN_DT>(r0, *reinterpret_cast<float(*)[8]>(&f[0]));
    unpack8<IN_DT>(r1, *reinterpret_cast<float(*)[8]>(&f[8]));
    uint4 o;
    o.x = quant4(f + 0, scale);
    o.y = quant4(f + 4, scale);
    o.z = quant4(f + 8, scale);
    o.w = quant4(f + 12, scale);
    uint8_t* dst = reinterpret_cast<uint8_t*>(kv ? p.v_cache : p.k_cache) + dst_off;
    *reinterpret_cast<uint4*>(dst) = o;
  } else {
    typename E::type* dst = reinterpret_cast<typename E::type*>(kv ? p.v_cache : p.k_cache) + dst_off;
    *reinterpret_cast<uint4*>(dst) = r0;
    *reinterpret_cast<uint4*>(dst + 8) = r1;
  }
}

int launch_with_pdl(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args);

}  // namespace b200attn

using namespace b200attn;

extern "C" int b200attn_kv_write(const void* key, const void* value, int64_t key_stride_t, int64_t key_stride_h,
                                 int64_t value_stride_t, int64_t value_stride_h, void* k_cache, void* v_cache,
                                 int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                                 const int64_t* slot_mapping, int32_t num_tokens, int32_t num_kv_heads,
                                 int32_t head_dim, int32_t block_size, int32_t in_dtype, int32_t kv_dtype,
                                 const float* k_scale, const float* v_scale, int32_t scale_stride, void* stream) {
  if (num_tokens == 0) return B200ATTN_OK;
  if (!key || !value || !k_cache || !v_cache || !slot_mapping) {
    set_error("kv_write: null pointer argument");
    return B200ATTN_EINVAL;
  }
  if (num_tokens < 0 || num_kv_heads <= 0 || block_size <= 0 || head_dim <= 0 || (head_dim % 16) != 0) {
    set_error("kv_write: bad shape T=%d Hkv=%d D=%d block_size=%d (D must be a multiple of 16)", num_tokens,
              num_kv_heads, head_dim, block_size);
    return B200ATTN_EINVAL;
  }
  if (in_dtype != B200ATTN_BF16 && in_dtype != B200ATTN_FP16) {
    set_error("kv_write: key/value dtype must be bf16 or fp16 (got %d)", in_dtype);
    return B200ATTN_EINVAL;
  }
  const bool fp8 = kv_dtype == B200ATTN_FP8_E4M3;
  if (!fp8 && kv_dtype != in_dtype) {
    set_error("kv_write: 16-bit cache dtype (%d) must equal key/value dtype (%d)", kv_dtype, in_dtype);
    return B200ATTN_EINVAL;
  }
  if (fp8 && (!k_scale || !v_scale)) {
    set_error("kv_write: fp8 cache needs k_scale / v_scale");
    return B200ATTN_EINVAL;
  }
  const int64_t es_out = fp8 ? 1 : 2;
  auto mis = [](const void* p, int64_t a) { return (reinterpret_cast<uintptr_t>(p) % a) != 0; };
  if (mis(key, 16) || mis(value, 16) || mis(k_cache, 16) || mis(v_cache, 16) || (key_stride_t % 8) || (key_stride_h % 8) ||
      (value_stride_t % 8) || (value_stride_h % 8) || ((stride_block * es_out) % 16) || ((stride_slot * es_out) % 16) ||
      ((stride_head * es_out) % 16)) {
    set_error("kv_write: pointers and strides must be 16-byte aligned");
    return B200ATTN_EINVAL;
  }
  KvWriteParams p;
  p.key = key; p.value = value;
  p.key_stride_t = key_stride_t; p.key_stride_h = key_stride_h;
  p.value_stride_t = value_stride_t; p.value_stride_h = value_stride_h;
  p.k_cache = k_cache; p.v_cache = v_cache;
  p.stride_block = stride_block; p.stride_slot = stride_slot; p.stride_head = stride_head;
  p.slot_mapping = slot_mapping; p.k_scale = k_scale; p.v_scale = v_scale; p.scale_stride = scale_stride;
  p.T = num_tokens; p.Hkv = num_kv_heads; p.D = head_dim; p.block_size = block_size;
  p.d_lanes.init((uint32_t)head_dim / 16); p.d_heads.init((uint32_t)num_kv_heads); p.d_bs.init((uint32_t)block_size);

  const int64_t threads = (int64_t)num_tokens * 2 * num_kv_heads * (head_dim / 16);
  if (threads >= (1ll << 31)) {
    set_error("kv_write: %lld work items exceed the 2^31 limit of one launch; split the token range", (long long)threads);
    return B200ATTN_EINVAL;
  }
  const int block = 256;
  const int64_t grid = (threads + block - 1) / block;
  const void* fn;
  if (fp8)
    fn = in_dtype == B200ATTN_BF16 ? (const void*)kv_write_kernel<B200ATTN_BF16, true> : (const void*)kv_write_kernel<B200ATTN_FP16, true>;
  else
    fn = in_dtype == B200ATTN_BF16 ? (const void*)kv_write_kernel<B200ATTN_BF16, false> : (const void*)kv_write_kernel<B200ATTN_FP16, false>;
  void* args[] = {&p};
  return launch_with_pdl(fn, dim3((unsigned)grid), dim3(block), 0, static_cast<cudaStream_t>(stream), args);
}

// ------------------------------------------------------------------------------------------------
// Per-(token, head) dynamic FP8 scales ("fp8_per_token_head").
// Reference: _reshape_cache_per_token_head ($VLLM/v1/attention/ops/triton_reshape_and_cache_flash.py:138-236), one program
// per (token, head):   scale = max(max|x| / 448, 1e-6);  byte = e4m3(clamp(x * (1 / scale), -448, 448));  scale stored in a
// float32 side cache [num_blocks][block_size][Hkv] (any strides).  The two divisions are float32 `/` in Triton, which the
// NVIDIA backend lowers to div.full.f32 (<= 2 ulp, not IEEE): B200ATTN_KV_DYN_IEEE_DIV selects div.rn instead (what numpy /
// the Triton interpreter compute, and what the oracle's goldens pin); the default reproduces the compiled reference.
// Mapping: a (token, K|V, head) row is handled by D/16 consecutive threads (16 elements each), the row's absmax is reduced
// with warp shuffles (D/16 divides 32, rows are aligned to it), one thread stores the scale.
// ------------------------------------------------------------------------------------------------
namespace b200attn {

struct KvWriteDynParams {
  KvWriteParams w;          // k_scale / v_scale / scale_stride unused
  float* k_scale_cache;
  float* v_scale_cache;
  int64_t sc_stride_block, sc_stride_slot, sc_stride_head;   // elements of float32
  int ieee_div;
};

__device__ __forceinline__ float div_full(float a, float b) {
  float r;
  asm("div.full.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

template <int IN_DT>
__global__ void __launch_bounds__(256) kv_write_dyn_kernel(const KvWriteDynParams q) {
  const KvWriteParams& p = q.w;
  const uint32_t lanes = (uint32_t)p.D >> 4;
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t row, lane_in_row;
  p.d_lanes.divmod(gid, row, lane_in_row);
  const int col = (int)lane_in_row << 4;
  const uint32_t rows = (uint32_t)p.T * 2 * p.Hkv;
  pdl_launch_dependents();
  pdl_wait();
  const bool live = row < rows;   // nobody leaves before the shuffles
  uint32_t tkv, hh;
  p.d_heads.divmod(live ? row : 0u, tkv, hh);
  const int h = (int)hh, kv = (int)(tkv & 1), t = (int)(tkv >> 1);
  using E = Elem<IN_DT>;
  const typename E::type* src =
      kv ? reinterpret_cast<const typename E::type*>(p.value) + (int64_t)t * p.value_stride_t + (int64_t)h * p.value_stride_h
         : reinterpret_cast<const typename E::type*>(p.key) + (int64_t)t * p.key_stride_t + (int64_t)h * p.key_stride_h;
  const int64_t slot = __ldg(p.slot_mapping + t);
  float f[16];
  if (live) {
    const uint4 r0 = __ldg(reinterpret_cast<const uint4*>(src + col));
    const uint4 r1 = __ldg(reinterpret_cast<const uint4*>(src + col + 8));
    unpack8<IN_DT>(r0, *reinterpret_cast<float(*)[8]>(&f[0]));
    unpack8<IN_DT>(r1, *reinterpret_cast<float(*)[8]>(&f[8]));
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) f[i] = 0.f;
  }
  float amax = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) amax = fmaxf(amax, fabsf(f[i]));
  for (uint32_t off = 1; off < lanes; off <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, off));
  if (!live || slot < 0) return;
  const float scale = fmaxf(q.ieee_div ? __fdiv_rn(amax, 448.0f) : div_full(amax, 448.0f), 1e-6f);
  const float inv = q.ieee_div ? __fdiv_rn(1.0f, scale) : div_full(1.0f, scale);
  uint32_t blk32, off32;
  p.d_bs.divmod((uint32_t)slot, blk32, off32);
  const int64_t blk = blk32, off = off32;
  if (col == 0) (kv ? q.v_scale_cache : q.k_scale_cache)[blk * q.sc_stride_block + off * q.sc_stride_slot + (int64_t)h * q.sc_stride_head] = scale;
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 lo = make_float2(fminf(fmaxf(f[4 * i + 0] * inv, -448.f), 448.f), fminf(fmaxf(f[4 * i + 1] * inv, -448.f), 448.f));
    float2 hi = make_float2(fminf(fmaxf(f[4 * i + 2] * inv, -448.f), 448.f), fminf(fmaxf(f[4 * i + 3] * inv, -448.f), 448.f));
    w[i] = (uint32_t)__nv_cvt_float2_to_fp8x2(lo, __NV_SATFINITE, __NV_E4M3) | ((uint32_t)__nv_cvt_float2_to_fp8x2(hi, __NV_SATFINITE, __NV_E4M3) << 16);
  }
  uint8_t* dst = reinterpret_cast<uint8_t*>(kv ? p.v_cache : p.k_cache) + blk * p.stride_block + off * p.stride_slot + (int64_t)h * p.stride_head + col;
  *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
}

}  // namespace b200attn

extern "C" int b200attn_kv_write_per_token_head(const void* key, const void* value, int64_t key_stride_t, int64_t key_stride_h,
                                                int64_t value_stride_t, int64_t value_stride_h, void* k_cache, void* v_cache,
                                                int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                                                float* k_scale_cache, float* v_scale_cache, int64_t scale_stride_block,
                                                int64_t scale_stride_slot, int64_t scale_stride_head, const int64_t* slot_mapping,
                                                int32_t num_tokens, int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                                                int32_t in_dtype, int32_t flags, void* stream) {
  if (num_tokens == 0) return B200ATTN_OK;
  if (!key || !value || !k_cache || !v_cache || !slot_mapping || !k_scale_cache || !v_scale_cache) {
    set_error("kv_write_per_token_head: null pointer argument");
    return B200ATTN_EINVAL;
  }
  if (num_tokens < 0 || num_kv_heads <= 0 || block_size <= 0 || head_dim < 16 || head_dim > 512 || (head_dim & (head_dim - 1)) != 0) {
    set_error("kv_write_per_token_head: bad shape T=%d Hkv=%d D=%d block_size=%d (D must be a power of two in 16..512)", num_tokens,
              num_kv_heads, head_dim, block_size);
    return B200ATTN_EINVAL;
  }
  if (in_dtype != B200ATTN_BF16 && in_dtype != B200ATTN_FP16) {
    set_error("kv_write_per_token_head: key/value dtype must be bf16 or fp16 (got %d)", in_dtype);
    return B200ATTN_EINVAL;
  }
  auto mis = [](const void* p, int64_t a) { return (reinterpret_cast<uintptr_t>(p) % a) != 0; };
  if (mis(key, 16) || mis(value, 16) || mis(k_cache, 16) || mis(v_cache, 16) || mis(k_scale_cache, 4) || mis(v_scale_cache, 4) ||
      (key_stride_t % 8) || (key_stride_h % 8) || (value_stride_t % 8) || (value_stride_h % 8) || (stride_block % 16) ||
      (stride_slot % 16) || (stride_head % 16)) {
    set_error("kv_write_per_token_head: pointers and strides must be 16-byte aligned");
    return B200ATTN_EINVAL;
  }
  const int64_t threads = (int64_t)num_tokens * 2 * num_kv_heads * (head_dim / 16);
  if (threads >= (1ll << 31)) {
    set_error("kv_write_per_token_head: %lld work items exceed the 2^31 limit of one launch", (long long)threads);
    return B200ATTN_EINVAL;
  }
  KvWriteDynParams q;
  KvWriteParams& p = q.w;
  p.key = key; p.value = value;
  p.key_stride_t = key_stride_t; p.key_stride_h = key_stride_h;
  p.value_stride_t = value_stride_t; p.value_stride_h = value_stride_h;
  p.k_cache = k_cache; p.v_cache = v_cache;
  p.stride_block = stride_block; p.stride_slot = stride_slot; p.stride_head = stride_head;
  p.slot_mapping = slot_mapping; p.k_scale = nullptr; p.v_scale = nullptr; p.scale_stride = 0;
  p.T = num_tokens; p.Hkv = num_kv_heads; p.D = head_dim; p.block_size = block_size;
  p.d_lanes.init((uint32_t)head_dim / 16); p.d_heads.init((uint32_t)num_kv_heads); p.d_bs.init((uint32_t)block_size);
  q.k_scale_cache = k_scale_cache; q.v_scale_cache = v_scale_cache;
  q.sc_stride_block = scale_stride_block; q.sc_stride_slot = scale_stride_slot; q.sc_stride_head = scale_stride_head;
  q.ieee_div = (flags & B200ATTN_KV_DYN_IEEE_DIV) ? 1 : 0;
  const int block = 256;
  const void* fn = in_dtype == B200ATTN_BF16 ? (const void*)kv_write_dyn_kernel<B200ATTN_BF16> : (const void*)kv_write_dyn_kernel<B200ATTN_FP16>;
  void* args[] = {&q};
  return launch_with_pdl(fn, dim3((unsigned)((threads + block - 1) / block)), dim3(block), 0, static_cast<cudaStream_t>(stream), args);
}
