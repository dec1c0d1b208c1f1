// decode_common.cuh — pieces shared by the two split-KV decode kernels (decode.cu: HMMA consumers; decode_tc.cu:
// tcgen05/TMEM): launch parameters, the debug timeline stamp and the in-kernel merge of split-KV partials.
#pragma once
#include "common.cuh"

namespace b200attn {

struct DecodeParams {
  const void* q;
  int64_t q_stride_t, q_stride_h;
  void* out;
  int64_t out_stride_t, out_stride_h;
  const int32_t* block_table;
  int64_t bt_stride;
  const int32_t* seq_lens;
  const int32_t* row_map;  // nullable: block-table row of decode row s (query-token rows of one request share a row)
  int32_t* plan;           // read-only except the arrival counters of the in-kernel merge
  float* ws_o;
  float* ws_ml;
  const float* k_scale;
  const float* v_scale;
  float* lse_out;
  float sm_scale_log2;  // sm_scale * log2(e)
  int scale_stride;
  int B, Hq, Hv;        // Hv = virtual kv heads = Hkv * gsplit
  int G;                // query heads per virtual kv head (<= 16)
  int gsplit;           // virtual heads per physical kv head
  int cpb;              // 16-token chunks per cache block (block_size / 16)
  int fresh;            // trailing positions of a row the preceding kernel may have written (< 0: unknown, no early prefetch)
  int fused_merge;      // 1: the last CTA to finish a split (sequence, kv head) pair merges its partials in-kernel
  unsigned long long* dbg;  // debug timeline buffer [grid][8] (globaltimer ns) or nullptr
  int debug;            // bit0: consumers skip the math (pure streaming ceiling); never set in production
};

constexpr int kD = 128;

__device__ __forceinline__ void dbg_stamp(const DecodeParams& p, int slot) {
  if (p.dbg != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.dbg[(size_t)blockIdx.x * 8 + slot] = t;
  }
}

template <int QDT>
__device__ __forceinline__ float q_to_float(uint16_t raw) {
  if constexpr (QDT == B200ATTN_BF16) return __uint_as_float((uint32_t)raw << 16);
  else return __half2float(__ushort_as_half(raw));
}

// In-kernel merge (reduce_segments semantics, $VLLM/v1/attention/ops/triton_unified_attention.py:645-733): called by the
// 128 epilogue threads of a CTA right after they stored a partial (o, m, l) of the pair (s, hv) to the workspace.  The
// last CTA to deliver a partial of the pair reduces all of them and writes `out`.  GH = heads held per thread (8 or 16),
// red_o / red_ml = smem scratch of >= GH*4 and >= 2*GH floats, hand = 3 ints of smem.
template <int QDT, int GH>
__device__ __forceinline__ void fused_merge_tail(const DecodeParams& p, int s, int hv, int L, int pfx, float vscale, int tid,
                                                 float* red_o, float* red_ml, volatile int* hand) {
  const float ninf = -INFINITY;
  {
      named_bar_sync(1, kConsumerWarps * 32);  // all partial stores of this CTA are issued ...
      if (tid == 0) {
        const int tpc = __ldg(p.plan + 1);
        const int nt = (L + kTileTokens - 1) / kTileTokens;
        const long long x0 = (long long)pfx * p.Hv + (long long)hv * nt;
        const int c_lo = (int)(x0 / tpc), c_hi = (int)((x0 + nt - 1) / tpc);
        const int slot_lo = (x0 == (long long)c_lo * tpc) ? 0 : 1;
        int* cnt = p.plan + kPlanCounterOffset + c_lo * 2 + slot_lo;
        const int np = c_hi - c_lo + 1;
        __threadfence();  // ... and (cumulatively, through the barrier) made visible device-wide before the arrival counts
        const int old = atomicAdd(cnt, 1);
        const bool last = (old == np - 1);
        if (last) {
          *cnt = 0;         // every partial of the pair has arrived: re-arm the counter for the next launch
          __threadfence();  // acquire side of the hand-shake
        }
        hand[0] = last ? np : 0;
        hand[1] = c_lo;
        hand[2] = slot_lo;
      }
      named_bar_sync(1, kConsumerWarps * 32);
      const int np = hand[0];
      if (np > 0) {
        const int c_lo = hand[1], slot_lo = hand[2];
        auto part = [&](int i, int head) { return ((size_t)(c_lo + i) * 2 + (i == 0 ? slot_lo : 0)) * kMaxGroup + head; };
        constexpr int CH = GH == 8 ? 4 : 2;   // partials merged per round with every load in flight at once
        const int d = tid;
        using E = Elem<QDT>;
        if (np <= CH) {
          // fast path (the common case: a pair cut by one to three CTA boundaries).  One round trip to L2: issue the
          // partial-output loads first, then the (m, l) loads, then combine.
          float ov[CH][GH];
#pragma unroll
          for (int i = 0; i < CH; ++i)
#pragma unroll
            for (int h = 0; h < GH; ++h) ov[i][h] = (i < np && h < p.G) ? ldcg_f32(p.ws_o + part(i, h) * kD + d) : 0.f;
          {
            const int head = tid >> 3, i = tid & 7;   // 8 lanes per head, lane i <-> partial i
            float2 ml = make_float2(ninf, 0.f);
            if (head < p.G && i < np) ml = ldcg_f32x2(p.ws_ml + part(i, head) * 2);
            float mg = ml.x;
#pragma unroll
            for (int off = 1; off < 8; off <<= 1) mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, off));
            const float wgt = (ml.x == ninf) ? 0.f : fast_exp2(ml.x - mg);
            float lg = ml.y * wgt;
#pragma unroll
            for (int off = 1; off < 8; off <<= 1) lg += __shfl_xor_sync(0xffffffffu, lg, off);
            if (head < p.G && i < CH) red_o[head * CH + i] = wgt * (vscale / lg);   // normalised weight of partial i
            if (head < p.G && i == 0) red_ml[head] = (mg + fast_log2(lg)) * 0.6931471805599453f;
          }
          named_bar_sync(1, kConsumerWarps * 32);
#pragma unroll
          for (int h = 0; h < GH; ++h) {
            if (h < p.G) {
              float acc = 0.f;
#pragma unroll
              for (int i = 0; i < CH; ++i) acc += red_o[h * CH + i] * ov[i][h];   // weights of absent partials are 0
              const int hq = hv * p.G + h;
              typename E::type* op = reinterpret_cast<typename E::type*>(p.out) + (int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h;
              op[d] = E::from_float(acc);
              if (p.lse_out != nullptr && d == 0) p.lse_out[(int64_t)s * p.Hq + hq] = red_ml[h];
            }
          }
        } else {
          // general path: any number of partials (small batches with very long contexts)
          {
            const int head = tid >> 3, sub = tid & 7;
            float mg = ninf, lg = 0.f;
            if (head < p.G)
              for (int i = sub; i < np; i += 8) mg = fmaxf(mg, ldcg_f32(p.ws_ml + part(i, head) * 2));
#pragma unroll
            for (int off = 1; off < 8; off <<= 1) mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, off));
            if (head < p.G)
              for (int i = sub; i < np; i += 8) {
                const float2 ml = ldcg_f32x2(p.ws_ml + part(i, head) * 2);
                lg += (ml.x == ninf) ? 0.f : ml.y * fast_exp2(ml.x - mg);
              }
#pragma unroll
            for (int off = 1; off < 8; off <<= 1) lg += __shfl_xor_sync(0xffffffffu, lg, off);
            if (sub == 0 && head < p.G) {
              red_ml[head * 2 + 0] = mg;
              red_ml[head * 2 + 1] = lg;
            }
          }
          named_bar_sync(1, kConsumerWarps * 32);
          float acc[GH];
#pragma unroll
          for (int h = 0; h < GH; ++h) acc[h] = 0.f;
          for (int i = 0; i < np; ++i) {
            float ov[GH], mv[GH];
#pragma unroll
            for (int h = 0; h < GH; ++h) {
              ov[h] = h < p.G ? ldcg_f32(p.ws_o + part(i, h) * kD + d) : 0.f;
              mv[h] = h < p.G ? ldcg_f32(p.ws_ml + part(i, h) * 2) : ninf;
            }
#pragma unroll
            for (int h = 0; h < GH; ++h) acc[h] += (mv[h] == ninf ? 0.f : fast_exp2(mv[h] - red_ml[(h < p.G ? h : 0) * 2])) * ov[h];
          }
#pragma unroll
          for (int h = 0; h < GH; ++h) {
            if (h < p.G) {
              const float mg = red_ml[h * 2 + 0], lg = red_ml[h * 2 + 1];
              const int hq = hv * p.G + h;
              typename E::type* op = reinterpret_cast<typename E::type*>(p.out) + (int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h;
              op[d] = E::from_float(acc[h] * (vscale / lg));
              if (p.lse_out != nullptr && d == 0) p.lse_out[(int64_t)s * p.Hq + hq] = (mg + fast_log2(lg)) * 0.6931471805599453f;
            }
          }
        }
      }
  }
}

}  // namespace b200attn
