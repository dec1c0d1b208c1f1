// decode_simt.cu — split-KV paged decode for head sizes the tensor-core kernels do not tile (head_dim 64 and 256; the
// TMA / tcgen05 kernels are built around 128-byte rows).  Same semantics as decode.cu (kernel_unified_attention,
// $VLLM/v1/attention/ops/triton_unified_attention.py:440-643; FlashInfer's scale convention, flashinfer.py:1396-1404), same C
// entry point (b200attn_decode routes here by head_dim), CUDA cores only.
//
// Why it exists: the Qwen3-TTS talker config class has no head_dim field
// (vllm_omni/model_executor/models/qwen3_tts/configuration_qwen3_tts.py:376-398), so a random-init talker from class defaults
// resolves to hidden / heads = 64 (BASELINE configs[0] "D 64 variant"); FlashInfer serves 64 / 128 / 256
// ($VLLM/v1/attention/backends/flashinfer.py:403-405).  These shapes are latency-bound (C1: 2 MiB of KV per layer), so the
// kernel is written for correctness and coalescing, not for the HBM roofline.
//
// Mapping: a token's K (or V) row of D elements is read by D/8 consecutive threads, 8 elements each (8-byte loads for fp8,
// 16-byte for 16-bit): whole 64- / 128- / 256- / 512-byte rows per request.  A thread keeps its 8 query elements and 8 output
// accumulators for up to 8 query heads of the kv head in registers; logits are reduced over the D/8 threads by shuffles; the
// 128 / (D/8) token groups of a CTA run independent online softmaxes that are merged through shared memory at the end.
// Grid = (sequence x kv head x head pass) x S splits of the context; partials go to the caller's workspace and
// simt_merge_kernel folds them (S = 1 writes the output directly).
#include "decode_common.cuh"

namespace b200attn {

constexpr int kSimtThreads = 128;
constexpr int kSimtHeads = 8;   // query heads per pass

struct SimtParams {
  const void* q;
  int64_t q_stride_t, q_stride_h;
  void* out;
  int64_t out_stride_t, out_stride_h;
  const uint8_t* k_cache;
  const uint8_t* v_cache;
  int64_t stride_block, stride_slot, stride_head;   // elements
  const int32_t* block_table;
  int64_t bt_stride;
  const int32_t* seq_lens;
  const int32_t* row_map;
  const float* k_scale;
  const float* v_scale;
  int scale_stride;
  float* lse_out;
  float* ws_o;     // [unit][S][8][D]
  float* ws_ml;    // [unit][S][8][2]
  float sm_scale_log2;
  int B, Hq, Hkv, G, passes, S, block_size;
};

template <bool FP8, int QDT>
__device__ __forceinline__ void load8(const uint8_t* base, int64_t elem_off, float (&f)[8]) {
  if constexpr (FP8) {
    const uint2 raw = *reinterpret_cast<const uint2*>(base + elem_off);
    const uint32_t w[2] = {raw.x, raw.y};
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const uint32_t lo = e4m3x2_to_f16x2_lo(w[i]), hi = e4m3x2_to_f16x2_hi(w[i]);
      const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&lo));
      const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&hi));
      f[4 * i + 0] = a.x; f[4 * i + 1] = a.y; f[4 * i + 2] = b.x; f[4 * i + 3] = b.y;
    }
  } else {
    const uint4 raw = *reinterpret_cast<const uint4*>(base + elem_off * 2);
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = q_to_float<QDT>(w[i] & 0xffff);
      f[2 * i + 1] = q_to_float<QDT>(w[i] >> 16);
    }
  }
}

template <int D, bool FP8, int QDT>
__global__ void __launch_bounds__(kSimtThreads) decode_simt_kernel(const SimtParams p) {
  constexpr int TPT = D / 8;                 // threads per token row
  constexpr int GROUPS = kSimtThreads / TPT; // token groups per CTA
  __shared__ float sm_ml[GROUPS][kSimtHeads][2];
  __shared__ float sm_o[GROUPS][kSimtHeads][D];
  pdl_launch_dependents();
  pdl_wait();
  const int unit = blockIdx.x;               // (sequence, kv head, pass)
  const int sp = blockIdx.y;
  const int pass = unit % p.passes;
  const int kvh = (unit / p.passes) % p.Hkv;
  const int s = unit / (p.passes * p.Hkv);
  const int L = max(__ldg(p.seq_lens + s), 0);
  const int h0 = pass * kSimtHeads;          // first head of the kv head's group handled here
  const int nh = min(kSimtHeads, p.G - h0);
  const int grp = threadIdx.x / TPT, sub = threadIdx.x % TPT;
  const float ninf = -INFINITY;
  using E = Elem<QDT>;
  if (L == 0) {
    if (p.lse_out != nullptr && sp == 0) {   // rows without tokens: (0, -inf) when the caller combines by LSE
      for (int i = threadIdx.x; i < nh * D; i += kSimtThreads) {
        const int hq = kvh * p.G + h0 + i / D;
        reinterpret_cast<typename E::type*>(p.out)[(int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h + i % D] = E::from_float(0.f);
      }
      for (int i = threadIdx.x; i < nh; i += kSimtThreads) p.lse_out[(int64_t)s * p.Hq + kvh * p.G + h0 + i] = ninf;
    }
    return;
  }
  // this split's token range: equal runs of 64-token tiles
  const int tiles = (L + 63) / 64;
  const int per = (tiles + p.S - 1) / p.S;
  const int tok_lo = min(L, sp * per * 64), tok_hi = min(L, (sp + 1) * per * 64);

  float qv[kSimtHeads][8], acc[kSimtHeads][8], m[kSimtHeads], l[kSimtHeads], cs[kSimtHeads];
  const float kscale = FP8 ? __ldg(p.k_scale + (int64_t)kvh * p.scale_stride) : 1.0f;
#pragma unroll
  for (int h = 0; h < kSimtHeads; ++h) {
    m[h] = ninf; l[h] = 0.f; cs[h] = p.sm_scale_log2 * kscale;
#pragma unroll
    for (int i = 0; i < 8; ++i) { acc[h][i] = 0.f; qv[h][i] = 0.f; }
    if (h < nh) {
      const uint16_t* src = reinterpret_cast<const uint16_t*>(p.q) + (int64_t)s * p.q_stride_t + (int64_t)(kvh * p.G + h0 + h) * p.q_stride_h + sub * 8;
      const uint4 raw = *reinterpret_cast<const uint4*>(src);
      const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) { qv[h][2 * i] = q_to_float<QDT>(w[i] & 0xffff); qv[h][2 * i + 1] = q_to_float<QDT>(w[i] >> 16); }
    }
  }
  const int row = p.row_map ? __ldg(p.row_map + s) : s;
  const int32_t* bt = p.block_table + (int64_t)row * p.bt_stride;
  for (int base = tok_lo; base < tok_hi; base += GROUPS) {   // uniform trip count per CTA: every lane takes part in the shuffles
    const int t = base + grp;
    const bool valid = t < tok_hi;
    float kf[8], vf[8];
    if (valid) {
      const int blk = __ldg(bt + t / p.block_size);
      const int64_t off = (int64_t)blk * p.stride_block + (int64_t)(t % p.block_size) * p.stride_slot + (int64_t)kvh * p.stride_head + sub * 8;
      load8<FP8, QDT>(p.k_cache, off, kf);
      load8<FP8, QDT>(p.v_cache, off, vf);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) kf[i] = vf[i] = 0.f;
    }
#pragma unroll
    for (int h = 0; h < kSimtHeads; ++h) {
      if (h < nh) {
        float sdot = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) sdot += qv[h][i] * kf[i];
#pragma unroll
        for (int o = 1; o < TPT; o <<= 1) sdot += __shfl_xor_sync(0xffffffffu, sdot, o);
        if (!valid) continue;
        const float x = sdot * cs[h];
        const float mn = fmaxf(m[h], x);
        const float corr = fast_exp2(m[h] - mn);     // (-inf) - finite -> 0
        const float pr = fast_exp2(x - mn);
        m[h] = mn;
        l[h] = l[h] * corr + pr;
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[h][i] = acc[h][i] * corr + pr * vf[i];
      }
    }
  }
  // ---- merge the CTA's token groups ----
#pragma unroll
  for (int h = 0; h < kSimtHeads; ++h) {
    if (sub == 0) { sm_ml[grp][h][0] = m[h]; sm_ml[grp][h][1] = l[h]; }
  }
  __syncthreads();
#pragma unroll
  for (int h = 0; h < kSimtHeads; ++h) {
    float mg = ninf;
    for (int g = 0; g < GROUPS; ++g) mg = fmaxf(mg, sm_ml[g][h][0]);
    const float sc = (m[h] == ninf) ? 0.f : fast_exp2(m[h] - mg);
#pragma unroll
    for (int i = 0; i < 8; ++i) sm_o[grp][h][sub * 8 + i] = acc[h][i] * sc;
  }
  __syncthreads();
  const float vscale = FP8 ? __ldg(p.v_scale + (int64_t)kvh * p.scale_stride) : 1.0f;
  for (int i = threadIdx.x; i < nh * D; i += kSimtThreads) {
    const int h = i / D, d = i % D;
    float mg = ninf;
    for (int g = 0; g < GROUPS; ++g) mg = fmaxf(mg, sm_ml[g][h][0]);
    float lg = 0.f, o = 0.f;
    for (int g = 0; g < GROUPS; ++g) {
      const float mw = sm_ml[g][h][0];
      lg += (mw == ninf) ? 0.f : sm_ml[g][h][1] * fast_exp2(mw - mg);
      o += sm_o[g][h][d];
    }
    const int hq = kvh * p.G + h0 + h;
    if (p.S == 1) {
      reinterpret_cast<typename E::type*>(p.out)[(int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h + d] = E::from_float(o * (vscale / lg));
      if (p.lse_out != nullptr && d == 0) p.lse_out[(int64_t)s * p.Hq + hq] = (mg + fast_log2(lg)) * 0.6931471805599453f;
    } else {
      const size_t slot = ((size_t)unit * p.S + sp) * kSimtHeads + h;
      p.ws_o[slot * D + d] = o;
      if (d == 0) { p.ws_ml[slot * 2] = mg; p.ws_ml[slot * 2 + 1] = lg; }
    }
  }
}

// fold the S partials of every (sequence, kv head, pass, head): one CTA of D threads per (unit, head)
template <int D, int QDT>
__global__ void __launch_bounds__(D) simt_merge_kernel(const SimtParams p) {
  pdl_launch_dependents();
  pdl_wait();
  const int unit = blockIdx.x / kSimtHeads, h = blockIdx.x % kSimtHeads;
  const int pass = unit % p.passes;
  const int kvh = (unit / p.passes) % p.Hkv;
  const int s = unit / (p.passes * p.Hkv);
  const int h0 = pass * kSimtHeads;
  if (h >= min(kSimtHeads, p.G - h0)) return;
  if (max(__ldg(p.seq_lens + s), 0) == 0) return;
  const int d = threadIdx.x;
  const float ninf = -INFINITY;
  float mg = ninf;
  for (int sp = 0; sp < p.S; ++sp) mg = fmaxf(mg, p.ws_ml[(((size_t)unit * p.S + sp) * kSimtHeads + h) * 2]);
  float lg = 0.f, o = 0.f;
  for (int sp = 0; sp < p.S; ++sp) {
    const size_t slot = ((size_t)unit * p.S + sp) * kSimtHeads + h;
    const float mw = p.ws_ml[slot * 2];
    const float w = (mw == ninf) ? 0.f : fast_exp2(mw - mg);
    lg += p.ws_ml[slot * 2 + 1] * w;
    o += p.ws_o[slot * D + d] * w;
  }
  using E = Elem<QDT>;
  const int hq = kvh * p.G + h0 + h;
  const float vscale = p.v_scale ? __ldg(p.v_scale + (int64_t)kvh * p.scale_stride) : 1.0f;
  reinterpret_cast<typename E::type*>(p.out)[(int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h + d] = E::from_float(o * (vscale / lg));
  if (p.lse_out != nullptr && d == 0) p.lse_out[(int64_t)s * p.Hq + hq] = (mg + fast_log2(lg)) * 0.6931471805599453f;
}

int launch_with_pdl(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args);

template <int D>
static int launch_simt_d(SimtParams& p, bool fp8, int q_dtype, cudaStream_t st) {
  const void* fn;
  const void* mg;
  if (q_dtype == B200ATTN_BF16) {
    fn = fp8 ? (const void*)decode_simt_kernel<D, true, B200ATTN_BF16> : (const void*)decode_simt_kernel<D, false, B200ATTN_BF16>;
    mg = (const void*)simt_merge_kernel<D, B200ATTN_BF16>;
  } else {
    fn = fp8 ? (const void*)decode_simt_kernel<D, true, B200ATTN_FP16> : (const void*)decode_simt_kernel<D, false, B200ATTN_FP16>;
    mg = (const void*)simt_merge_kernel<D, B200ATTN_FP16>;
  }
  const int units = p.B * p.Hkv * p.passes;
  void* args[] = {&p};
  int rc = launch_with_pdl(fn, dim3(units, p.S), dim3(kSimtThreads), 0, st, args);
  if (rc || p.S == 1) return rc;
  return launch_with_pdl(mg, dim3(units * kSimtHeads), dim3(D), 0, st, args);
}

// head_dim 64 / 256 behind b200attn_decode.  `workspace` is the caller's split-KV workspace (b200attn_workspace_bytes).
int launch_decode_simt(const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out, int64_t out_stride_t, int64_t out_stride_h,
                       const void* k_cache, const void* v_cache, int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                       const int32_t* block_table, int64_t bt_stride, const int32_t* seq_lens, int B, int Hq, int Hkv, int D,
                       int block_size, int q_dtype, int kv_dtype, float sm_scale, const float* k_scale, const float* v_scale,
                       int scale_stride, int num_ctas, void* workspace, size_t workspace_bytes, float* lse_out,
                       const int32_t* row_map, cudaStream_t st) {
  if (D != 64 && D != 256) {
    set_error("decode: head_dim %d unsupported (64, 128 or 256)", D);
    return B200ATTN_EINVAL;
  }
  const bool fp8 = kv_dtype == B200ATTN_FP8_E4M3;
  const int64_t es = fp8 ? 1 : 2;
  if ((reinterpret_cast<uintptr_t>(k_cache) % 16) || (reinterpret_cast<uintptr_t>(v_cache) % 16) || ((stride_block * es) % 16) ||
      ((stride_slot * es) % 16) || ((stride_head * es) % 16) || (reinterpret_cast<uintptr_t>(q) % 16) || (q_stride_t % 8) || (q_stride_h % 8)) {
    set_error("decode: cache / query pointers and strides must be 16-byte aligned");
    return B200ATTN_EINVAL;
  }
  SimtParams p;
  p.q = q; p.q_stride_t = q_stride_t; p.q_stride_h = q_stride_h;
  p.out = out; p.out_stride_t = out_stride_t; p.out_stride_h = out_stride_h;
  p.k_cache = reinterpret_cast<const uint8_t*>(k_cache); p.v_cache = reinterpret_cast<const uint8_t*>(v_cache);
  p.stride_block = stride_block; p.stride_slot = stride_slot; p.stride_head = stride_head;
  p.block_table = block_table; p.bt_stride = bt_stride; p.seq_lens = seq_lens; p.row_map = row_map;
  p.k_scale = fp8 ? k_scale : nullptr; p.v_scale = fp8 ? v_scale : nullptr; p.scale_stride = scale_stride;
  p.lse_out = lse_out;
  p.sm_scale_log2 = sm_scale * 1.4426950408889634f;
  p.B = B; p.Hq = Hq; p.Hkv = Hkv; p.G = Hq / Hkv; p.passes = (p.G + kSimtHeads - 1) / kSimtHeads; p.block_size = block_size;
  const int units = B * Hkv * p.passes;
  // splits: fill ~2 CTAs per SM slot of the caller's grid, bounded by the workspace (one slot = 8 heads x (D + 2) floats)
  const size_t slot_bytes = (size_t)kSimtHeads * (D + 2) * sizeof(float);
  int S = (2 * num_ctas + units - 1) / units;
  const size_t cap = workspace_bytes / slot_bytes / (size_t)units;
  if ((size_t)S > cap) S = (int)cap;
  if (S > 64) S = 64;
  if (S < 1) S = 1;
  p.S = S;
  p.ws_o = reinterpret_cast<float*>(workspace);
  p.ws_ml = p.ws_o + (size_t)units * S * kSimtHeads * D;
  return D == 64 ? launch_simt_d<64>(p, fp8, q_dtype, st) : launch_simt_d<256>(p, fp8, q_dtype, st);
}

}  // namespace b200attn
