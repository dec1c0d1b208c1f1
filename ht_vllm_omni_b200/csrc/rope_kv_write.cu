// rope_kv_write.cu — fused rotary embedding + paged KV write (the step before the path).
//
// Replaces, in ONE launch, the pair the reference issues per layer and step:
//   RotaryEmbedding.forward(positions, q, k)      ($VLLM/model_executor/layers/rotary_embedding/base.py:130-180,
//                                                   common.py:144-183; CUDA op csrc/pos_encoding_kernels.cu)
//   reshape_and_cache_flash(k, v, ...)            ($VLLM/v1/attention/backends/flashinfer.py:1846-1855)
// behind vLLM's hook AttentionImpl.do_rope_and_kv_cache_update ($VLLM/v1/attention/backend.py:815-840).  K is read once
// instead of read-rotated-written-read; the rotated q and k are also stored back in place, as the unfused op does.
//
// Arithmetic is the reference's:  o1 = rnd(x1 cos - x2 sin),  o2 = rnd(x2 cos + x1 sin)  in fp32 with one rounding to the
// activation dtype (the compiled CUDA op), or with every elementary op rounded (B200ATTN_ROPE_ROUND_EACH_OP, torch-native),
// with (x1, x2) = the two halves of the rotary span (neox) or even / odd elements (GPT-J).  The cache bytes are then
// those of kv_write.cu: fp32 IEEE divide by the head's scale, cvt.rn.satfinite.e4m3.
//
// Mapping: a (token, head) row of D elements is handled by D/16 consecutive threads, each owning two 8-element chunks:
//   neox rotary thread j:  [8j, 8j+8) and its partners [rot/2 + 8j, +8)          (cos/sin index 8j..8j+7)
//   GPT-J rotary thread j: [16j, 16j+16) = 8 adjacent pairs                      (cos/sin index 8j..8j+7)
//   pass-through thread (elements >= rotary_dim, and every V row): 16 contiguous elements, no rotation.
// All global accesses are 16-byte vectors (8-byte for an fp8 cache chunk).
#include "common.cuh"

namespace b200attn {

struct RopeKvParams {
  void* query;
  void* key;
  const void* value;
  int64_t q_stride_t, q_stride_h, key_stride_t, key_stride_h, value_stride_t, value_stride_h;
  const int64_t* positions;
  int64_t pos_axis_stride;   // 0: positions [T]; else positions [3][T] (M-RoPE: temporal / height / width), this many elements apart
  const void* cos_sin;   // [max_pos][rot]: cos (rot/2) then sin (rot/2), activation dtype
  const void* q_norm_w;  // nullable: RMSNorm weight [D] applied to every q head before the rotation (Qwen3 q_norm)
  const void* k_norm_w;  // nullable: likewise for k (k_norm)
  float rms_eps;
  void* k_cache;
  void* v_cache;
  int64_t stride_block, stride_slot, stride_head;
  const int64_t* slot_mapping;
  const float* k_scale;
  const float* v_scale;
  int scale_stride;
  int T, Hq, Hkv, D, rot, block_size, is_neox, each_op;
  FastDiv d_lanes, d_units, d_bs;         // threads per row (D / 16), heads per token (Hq + 2 Hkv), block size
  int sec_t, sec_h, sec_w, interleaved;   // M-RoPE sections (frequency counts per axis, sum = rot / 2); all 0 = plain RoPE
};

template <int DT>
__device__ __forceinline__ float rnd(float x) {
  using E = Elem<DT>;
  return E::to_float(E::from_float(x));
}

// The two roundings the reference has: its compiled CUDA op (csrc/pos_encoding_kernels.cu, what a GPU run executes)
// computes in fp32 and rounds once -- measured on the box, tests/tools/rope_probe.py: 0 mismatches in 524288 elements;
// products of two 16-bit floats are exact in fp32, so FMA contraction cannot change the result -- while forward_native
// (torch bf16 / fp16 tensor arithmetic) rounds every elementary operation.
template <int DT>
__device__ __forceinline__ float rot1(float x1, float x2, float c, float s, int each_op) {
  return each_op ? rnd<DT>(rnd<DT>(x1 * c) - rnd<DT>(x2 * s)) : rnd<DT>(x1 * c - x2 * s);
}
template <int DT>
__device__ __forceinline__ float rot2(float x1, float x2, float c, float s, int each_op) {
  return each_op ? rnd<DT>(rnd<DT>(x2 * c) + rnd<DT>(x1 * s)) : rnd<DT>(x2 * c + x1 * s);
}

template <int DT>
__device__ __forceinline__ void load8(const void* base, int64_t off, float (&f)[8]) {
  using E = Elem<DT>;
  const uint4 raw = *reinterpret_cast<const uint4*>(reinterpret_cast<const typename E::type*>(base) + off);
  const typename E::type* e = reinterpret_cast<const typename E::type*>(&raw);
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] = E::to_float(e[i]);
}

template <int DT>
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  using E = Elem<DT>;
  uint4 raw;
  typename E::type* e = reinterpret_cast<typename E::type*>(&raw);
#pragma unroll
  for (int i = 0; i < 8; ++i) e[i] = E::from_float(f[i]);
  return raw;
}

__device__ __forceinline__ uint2 quant8(const float (&f)[8], float scale) {
  uint32_t w[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const float2 lo = make_float2(__fdiv_rn(f[4 * i + 0], scale), __fdiv_rn(f[4 * i + 1], scale));
    const float2 hi = make_float2(__fdiv_rn(f[4 * i + 2], scale), __fdiv_rn(f[4 * i + 3], scale));
    w[i] = (uint32_t)__nv_cvt_float2_to_fp8x2(lo, __NV_SATFINITE, __NV_E4M3) |
           ((uint32_t)__nv_cvt_float2_to_fp8x2(hi, __NV_SATFINITE, __NV_E4M3) << 16);
  }
  return make_uint2(w[0], w[1]);
}

// M-RoPE: which position axis (0 temporal, 1 height, 2 width) drives rotary frequency `f`
// ($VLLM/model_executor/layers/rotary_embedding/mrope.py: forward_native :286-298 chunked [TTT..HHH..WWW],
// apply_interleaved_rope :190-198 [THWTHW...TT])
__device__ __forceinline__ int mrope_axis(const RopeKvParams& p, int f) {
  if (p.sec_t + p.sec_h + p.sec_w == 0) return 0;
  if (p.interleaved) {
    const int m = f % 3;
    if (m == 1 && f < 3 * p.sec_h) return 1;
    if (m == 2 && f < 3 * p.sec_w) return 2;
    return 0;
  }
  return f < p.sec_t ? 0 : (f < p.sec_t + p.sec_h ? 1 : 2);
}

// EXT = the launch uses the fused RMSNorm and / or M-RoPE stages; the plain instantiation carries none of their code (48
// registers and 5 resident CTAs per SM instead of 74 and 3, which costs 60 % of the throughput at prefill sizes).
template <int DT, bool FP8, bool EXT>
__global__ void __launch_bounds__(256, EXT ? 4 : 5) rope_kv_write_kernel(const RopeKvParams p) {
  using E = Elem<DT>;
  // 32-bit index math (the host rejects launches with >= 2^31 threads): 64-bit divisions would cost more than the rotation
  const uint32_t lanes = (uint32_t)p.D >> 4;
  const uint32_t units = (uint32_t)(p.Hq + 2 * p.Hkv);
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t row, ju;
  p.d_lanes.divmod(gid, row, ju);
  const int j = (int)ju;
  pdl_launch_dependents();
  pdl_wait();
  // With a fused RMSNorm the D/16 threads of a row reduce its sum of squares by warp shuffles (D/16 divides 32 and rows are
  // aligned to it), so no thread may leave before that point; without it the early exits below are taken as before.
  const bool any_norm = EXT && (p.q_norm_w != nullptr || p.k_norm_w != nullptr);
  bool live = row < (uint32_t)p.T * units;
  if (!any_norm && !live) return;
  uint32_t tu, uu;
  p.d_units.divmod(live ? row : 0u, tu, uu);
  const int t = (int)tu, u = (int)uu;
  const int kind = u < p.Hq ? 0 : (u < p.Hq + p.Hkv ? 1 : 2);   // q, k, v
  const int h = kind == 0 ? u : (kind == 1 ? u - p.Hq : u - p.Hq - p.Hkv);
  // slot, position and the data vectors are independent loads, issued together (see kv_write.cu)
  const int64_t slot = __ldg(p.slot_mapping + t);
  const int64_t pos = __ldg(p.positions + t);

  const int half = p.rot >> 1;
  const bool rotary = kind != 2 && j < (p.rot >> 4);
  const void* norm_w = !EXT ? nullptr : (kind == 0 ? p.q_norm_w : (kind == 1 ? p.k_norm_w : nullptr));
  int a0, a1;
  if (rotary && p.is_neox) { a0 = 8 * j; a1 = half + 8 * j; }
  else if (rotary) { a0 = 16 * j; a1 = a0 + 8; }
  else { a0 = (kind == 2 ? 0 : p.rot) + 16 * (j - (kind == 2 ? 0 : (p.rot >> 4))); a1 = a0 + 8; }
  if (kind == 0 && !rotary && norm_w == nullptr) live = false;   // q elements beyond the rotary span stay as they are
  if (!any_norm && !live) return;

  void* base = kind == 0 ? p.query : (kind == 1 ? p.key : const_cast<void*>(p.value));
  const int64_t roff = kind == 0 ? (int64_t)t * p.q_stride_t + (int64_t)h * p.q_stride_h
                     : kind == 1 ? (int64_t)t * p.key_stride_t + (int64_t)h * p.key_stride_h
                                 : (int64_t)t * p.value_stride_t + (int64_t)h * p.value_stride_h;
  float c0[8], c1[8];
  if (live) {
    load8<DT>(base, roff + a0, c0);
    load8<DT>(base, roff + a1, c1);
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) c0[i] = c1[i] = 0.f;
  }
  if (kind == 2 && slot < 0) live = false;   // padding token: nothing to write
  if (!any_norm && !live) return;

  if constexpr (EXT) if (any_norm) {
    // ---- RMSNorm over the head ($VLLM/model_executor/layers/layernorm.py forward_native / csrc/layernorm_kernels.cu):
    //   y = rnd(rnd(x * rsqrt(mean(x^2) + eps)) * w)   -- fp32 statistics, the normalised value rounded to the activation
    // dtype BEFORE the multiplication by the weight, which is itself a 16-bit multiply (one more rounding)
    float ss = 0.f;
    if (live && norm_w != nullptr) {
#pragma unroll
      for (int i = 0; i < 8; ++i) ss += c0[i] * c0[i] + c1[i] * c1[i];
    }
    for (uint32_t off = 1; off < lanes; off <<= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
    if (!live) return;
    if (norm_w != nullptr) {
      const float inv = rsqrtf(ss / (float)p.D + p.rms_eps);
      float w0[8], w1[8];
      load8<DT>(norm_w, a0, w0);
      load8<DT>(norm_w, a1, w1);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        c0[i] = rnd<DT>(rnd<DT>(c0[i] * inv) * w0[i]);
        c1[i] = rnd<DT>(rnd<DT>(c1[i] * inv) * w1[i]);
      }
    }
  }

  if (rotary) {
    float cs[8], sn[8];
    if (!EXT || p.pos_axis_stride == 0) {
      load8<DT>(p.cos_sin, pos * p.rot + 8 * j, cs);
      load8<DT>(p.cos_sin, pos * p.rot + half + 8 * j, sn);
    } else if constexpr (EXT) {
      // M-RoPE: frequency f takes cos / sin from the cache row of ITS axis' position; a chunk of 8 frequencies can touch all
      // three rows (interleaved layout), each fetched only when one of its frequencies needs it
      int ax[8];
      int need = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) { ax[i] = mrope_axis(p, 8 * j + i); need |= 1 << ax[i]; }
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        if (!(need & (1 << a))) continue;
        const int64_t pa = a == 0 ? pos : __ldg(p.positions + (int64_t)a * p.pos_axis_stride + t);
        float ca[8], sa[8];
        load8<DT>(p.cos_sin, pa * p.rot + 8 * j, ca);
        load8<DT>(p.cos_sin, pa * p.rot + half + 8 * j, sa);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (ax[i] == a) { cs[i] = ca[i]; sn[i] = sa[i]; }
      }
    }
    if (p.is_neox) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float x1 = c0[i], x2 = c1[i];
        c0[i] = rot1<DT>(x1, x2, cs[i], sn[i], p.each_op);
        c1[i] = rot2<DT>(x1, x2, cs[i], sn[i], p.each_op);
      }
    } else {
      float e[16];
#pragma unroll
      for (int i = 0; i < 8; ++i) { e[i] = c0[i]; e[8 + i] = c1[i]; }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float x1 = e[2 * i], x2 = e[2 * i + 1];
        e[2 * i] = rot1<DT>(x1, x2, cs[i], sn[i], p.each_op);
        e[2 * i + 1] = rot2<DT>(x1, x2, cs[i], sn[i], p.each_op);
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) { c0[i] = e[i]; c1[i] = e[8 + i]; }
    }
  }
  if (rotary || (kind != 2 && norm_w != nullptr)) {
    // rotated (and normalised) q / k go back in place, like torch.ops._C.rotary_embedding / fused_qk_norm_rope
    typename E::type* dst = reinterpret_cast<typename E::type*>(base) + roff;
    *reinterpret_cast<uint4*>(dst + a0) = pack8<DT>(c0);
    *reinterpret_cast<uint4*>(dst + a1) = pack8<DT>(c1);
  }
  if (kind == 0 || slot < 0) return;

  uint32_t blk32, off32;
  p.d_bs.divmod((uint32_t)slot, blk32, off32);   // slots are < num_blocks * block_size < 2^32
  const int64_t blk = blk32, off = off32;
  const int64_t dst_off = blk * p.stride_block + off * p.stride_slot + (int64_t)h * p.stride_head;
  if constexpr (FP8) {
    const float scale = __ldg((kind == 1 ? p.k_scale : p.v_scale) + (int64_t)h * p.scale_stride);
    uint8_t* dst = reinterpret_cast<uint8_t*>(kind == 1 ? p.k_cache : p.v_cache) + dst_off;
    *reinterpret_cast<uint2*>(dst + a0) = quant8(c0, scale);
    *reinterpret_cast<uint2*>(dst + a1) = quant8(c1, scale);
  } else {
    typename E::type* dst = reinterpret_cast<typename E::type*>(kind == 1 ? p.k_cache : p.v_cache) + dst_off;
    *reinterpret_cast<uint4*>(dst + a0) = pack8<DT>(c0);
    *reinterpret_cast<uint4*>(dst + a1) = pack8<DT>(c1);
  }
}

int launch_with_pdl(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args);

}  // namespace b200attn

using namespace b200attn;

static int rope_impl(void* query, int64_t q_stride_t, int64_t q_stride_h, int32_t num_q_heads, void* key,
                     const void* value, int64_t key_stride_t, int64_t key_stride_h, int64_t value_stride_t,
                     int64_t value_stride_h, const int64_t* positions, int64_t positions_axis_stride, int32_t sec_t, int32_t sec_h,
                     int32_t sec_w, const void* cos_sin_cache,
                     int32_t rotary_dim, int32_t rope_flags, const void* q_norm_weight, const void* k_norm_weight, float rms_eps,
                     void* k_cache, void* v_cache, int64_t stride_block,
                     int64_t stride_slot, int64_t stride_head, const int64_t* slot_mapping,
                     int32_t num_tokens, int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                     int32_t in_dtype, int32_t kv_dtype, const float* k_scale, const float* v_scale,
                     int32_t scale_stride, void* stream) {
  if (num_tokens == 0) return B200ATTN_OK;
  if (positions_axis_stride != 0) {
    if (sec_t < 0 || sec_h < 0 || sec_w < 0 || sec_t + sec_h + sec_w != rotary_dim / 2) {
      set_error("rope_kv_write: mrope_section (%d, %d, %d) must be non-negative and sum to rotary_dim / 2 = %d", sec_t, sec_h, sec_w,
                rotary_dim / 2);
      return B200ATTN_EINVAL;
    }
  } else {
    sec_t = sec_h = sec_w = 0;
  }
  if ((q_norm_weight || k_norm_weight) && (head_dim > 512 || (head_dim & (head_dim - 1)) != 0)) {
    set_error("rope_kv_write: the fused RMSNorm needs a power-of-two head_dim <= 512 (got %d)", head_dim);
    return B200ATTN_EINVAL;
  }
  if ((q_norm_weight && (reinterpret_cast<uintptr_t>(q_norm_weight) % 16)) || (k_norm_weight && (reinterpret_cast<uintptr_t>(k_norm_weight) % 16))) {
    set_error("rope_kv_write: norm weights must be 16-byte aligned");
    return B200ATTN_EINVAL;
  }
  if (!query || !key || !value || !positions || !cos_sin_cache || !k_cache || !v_cache || !slot_mapping) {
    set_error("rope_kv_write: null pointer argument");
    return B200ATTN_EINVAL;
  }
  if (num_tokens < 0 || num_q_heads <= 0 || num_kv_heads <= 0 || block_size <= 0 || head_dim <= 0 || (head_dim % 16) != 0) {
    set_error("rope_kv_write: bad shape T=%d Hq=%d Hkv=%d D=%d block_size=%d (D must be a multiple of 16)", num_tokens,
              num_q_heads, num_kv_heads, head_dim, block_size);
    return B200ATTN_EINVAL;
  }
  if (rotary_dim <= 0 || rotary_dim > head_dim || (rotary_dim % 16) != 0) {
    set_error("rope_kv_write: rotary_dim %d must be a positive multiple of 16 and <= head_dim %d", rotary_dim, head_dim);
    return B200ATTN_EINVAL;
  }
  if (in_dtype != B200ATTN_BF16 && in_dtype != B200ATTN_FP16) {
    set_error("rope_kv_write: q/k/v dtype must be bf16 or fp16 (got %d)", in_dtype);
    return B200ATTN_EINVAL;
  }
  const bool fp8 = kv_dtype == B200ATTN_FP8_E4M3;
  if (!fp8 && kv_dtype != in_dtype) {
    set_error("rope_kv_write: 16-bit cache dtype (%d) must equal q/k/v dtype (%d)", kv_dtype, in_dtype);
    return B200ATTN_EINVAL;
  }
  if (fp8 && (!k_scale || !v_scale)) {
    set_error("rope_kv_write: fp8 cache needs k_scale / v_scale");
    return B200ATTN_EINVAL;
  }
  const int64_t es_out = fp8 ? 1 : 2;
  auto mis = [](const void* p, int64_t a) { return (reinterpret_cast<uintptr_t>(p) % a) != 0; };
  if (mis(query, 16) || mis(key, 16) || mis(value, 16) || mis(cos_sin_cache, 16) || mis(k_cache, 16) || mis(v_cache, 16) ||
      (q_stride_t % 8) || (q_stride_h % 8) || (key_stride_t % 8) || (key_stride_h % 8) || (value_stride_t % 8) ||
      (value_stride_h % 8) || ((stride_block * es_out) % 16) || ((stride_slot * es_out) % 16) ||
      ((stride_head * es_out) % 16)) {
    set_error("rope_kv_write: pointers and strides must be 16-byte aligned");
    return B200ATTN_EINVAL;
  }
  RopeKvParams p;
  p.query = query; p.key = key; p.value = value;
  p.q_stride_t = q_stride_t; p.q_stride_h = q_stride_h;
  p.key_stride_t = key_stride_t; p.key_stride_h = key_stride_h;
  p.value_stride_t = value_stride_t; p.value_stride_h = value_stride_h;
  p.positions = positions; p.cos_sin = cos_sin_cache;
  p.pos_axis_stride = positions_axis_stride;
  p.sec_t = sec_t; p.sec_h = sec_h; p.sec_w = sec_w;
  p.interleaved = (rope_flags & B200ATTN_ROPE_MROPE_INTERLEAVED) ? 1 : 0;
  p.q_norm_w = q_norm_weight; p.k_norm_w = k_norm_weight; p.rms_eps = rms_eps;
  p.k_cache = k_cache; p.v_cache = v_cache;
  p.stride_block = stride_block; p.stride_slot = stride_slot; p.stride_head = stride_head;
  p.slot_mapping = slot_mapping; p.k_scale = k_scale; p.v_scale = v_scale; p.scale_stride = scale_stride;
  p.T = num_tokens; p.Hq = num_q_heads; p.Hkv = num_kv_heads; p.D = head_dim; p.rot = rotary_dim;
  p.block_size = block_size;
  p.d_lanes.init((uint32_t)head_dim / 16); p.d_units.init((uint32_t)(num_q_heads + 2 * num_kv_heads)); p.d_bs.init((uint32_t)block_size);
  p.is_neox = (rope_flags & B200ATTN_ROPE_NEOX) ? 1 : 0;
  p.each_op = (rope_flags & B200ATTN_ROPE_ROUND_EACH_OP) ? 1 : 0;

  const int64_t threads = (int64_t)num_tokens * (num_q_heads + 2 * num_kv_heads) * (head_dim / 16);
  const int block = 256;
  const int64_t grid = (threads + block - 1) / block;
  if (threads >= (1ll << 31)) {
    set_error("rope_kv_write: %lld work items exceed the 2^31 limit of one launch; split the token range", (long long)threads);
    return B200ATTN_EINVAL;
  }
  const void* fn;
  const bool ext = q_norm_weight != nullptr || k_norm_weight != nullptr || positions_axis_stride != 0;
#define B200_ROPE_FN(DT, F8) (ext ? (const void*)rope_kv_write_kernel<DT, F8, true> : (const void*)rope_kv_write_kernel<DT, F8, false>)
  if (fp8)
    fn = in_dtype == B200ATTN_BF16 ? B200_ROPE_FN(B200ATTN_BF16, true) : B200_ROPE_FN(B200ATTN_FP16, true);
  else
    fn = in_dtype == B200ATTN_BF16 ? B200_ROPE_FN(B200ATTN_BF16, false) : B200_ROPE_FN(B200ATTN_FP16, false);
#undef B200_ROPE_FN
  void* args[] = {&p};
  return launch_with_pdl(fn, dim3((unsigned)grid), dim3(block), 0, static_cast<cudaStream_t>(stream), args);
}

extern "C" int b200attn_rope_kv_write(void* query, int64_t q_stride_t, int64_t q_stride_h, int32_t num_q_heads, void* key,
                                      const void* value, int64_t key_stride_t, int64_t key_stride_h, int64_t value_stride_t,
                                      int64_t value_stride_h, const int64_t* positions, const void* cos_sin_cache,
                                      int32_t rotary_dim, int32_t rope_flags, void* k_cache, void* v_cache, int64_t stride_block,
                                      int64_t stride_slot, int64_t stride_head, const int64_t* slot_mapping,
                                      int32_t num_tokens, int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                                      int32_t in_dtype, int32_t kv_dtype, const float* k_scale, const float* v_scale,
                                      int32_t scale_stride, void* stream) {
  return rope_impl(query, q_stride_t, q_stride_h, num_q_heads, key, value, key_stride_t, key_stride_h, value_stride_t, value_stride_h,
                   positions, 0, 0, 0, 0, cos_sin_cache, rotary_dim, rope_flags, nullptr, nullptr, 0.f, k_cache, v_cache, stride_block,
                   stride_slot, stride_head, slot_mapping, num_tokens, num_kv_heads, head_dim, block_size, in_dtype, kv_dtype, k_scale,
                   v_scale, scale_stride, stream);
}

extern "C" int b200attn_norm_rope_kv_write(void* query, int64_t q_stride_t, int64_t q_stride_h, int32_t num_q_heads, void* key,
                                           const void* value, int64_t key_stride_t, int64_t key_stride_h, int64_t value_stride_t,
                                           int64_t value_stride_h, const int64_t* positions, int64_t positions_axis_stride,
                                           int32_t mrope_t, int32_t mrope_h, int32_t mrope_w, const void* cos_sin_cache,
                                           int32_t rotary_dim, int32_t rope_flags, const void* q_norm_weight,
                                           const void* k_norm_weight, float rms_eps, void* k_cache, void* v_cache,
                                           int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                                           const int64_t* slot_mapping, int32_t num_tokens, int32_t num_kv_heads, int32_t head_dim,
                                           int32_t block_size, int32_t in_dtype, int32_t kv_dtype, const float* k_scale,
                                           const float* v_scale, int32_t scale_stride, void* stream) {
  return rope_impl(query, q_stride_t, q_stride_h, num_q_heads, key, value, key_stride_t, key_stride_h, value_stride_t, value_stride_h,
                   positions, positions_axis_stride, mrope_t, mrope_h, mrope_w, cos_sin_cache, rotary_dim, rope_flags, q_norm_weight,
                   k_norm_weight, rms_eps, k_cache, v_cache, stride_block, stride_slot, stride_head, slot_mapping, num_tokens,
                   num_kv_heads, head_dim, block_size, in_dtype, kv_dtype, k_scale, v_scale, scale_stride, stream);
}
