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
  int box_chunks;       // 16-token chunks one TMA box covers (tcgen05 kernel: min(cpb, 4) when a power of two, else 1)
  FastDiv d_cpb, d_gsplit;   // multiply-shift forms of the two launch-constant divisors (the producer's first-TMA path ran three
                             // integer-divide sequences otherwise)
  int fresh;            // trailing positions of a row the preceding kernel may have written (< 0: unknown, no early prefetch)
  int fused_merge;      // 1: the last CTA to finish a split (sequence, kv head) pair merges its partials in-kernel
  int cluster;          // thread-block cluster size of this launch (1: none); partials of a pair whose CTAs share a cluster
                        // are folded through distributed shared memory instead of the L2 workspace
  // Fused KV write (pure-decode steps; b200attn_decode_kv_write): the CTA that streams the tile holding a row's LAST position
  // first quantises and stores that row's new key / value (token index == row index) to the cache, then reads the tile.
  // new_key == nullptr: the cache was written by an earlier kernel (b200attn_decode).
  const void* new_key;
  const void* new_value;
  int64_t nk_stride_t, nk_stride_h, nv_stride_t, nv_stride_h;
  const int64_t* slot_mapping;
  uint8_t* kc_base;
  uint8_t* vc_base;
  int64_t c_stride_block, c_stride_slot, c_stride_head;   // elements (= bytes for the fp8 cache)
  int block_size;
  // Per-(token, head) dynamic FP8 scales ("fp8_per_token_head"): float32 side caches [block][head][slot] (slot stride 1);
  // nullptr = static per-tensor / per-head scales (k_scale / v_scale above)
  const float* ks_cache;
  const float* vs_cache;
  int64_t sc_stride_block, sc_stride_head;   // float32 elements
  unsigned long long* dbg;  // debug timeline buffer [grid][16] (globaltimer ns) or nullptr
  int debug;            // bit0: consumers skip the math (pure streaming ceiling); never set in production
};

constexpr int kD = 128;
constexpr int kDbgSlots = 16;

__device__ __forceinline__ void dbg_stamp(const DecodeParams& p, int slot) {
  if (p.dbg != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.dbg[(size_t)blockIdx.x * kDbgSlots + slot] = t;
  }
}

template <int QDT>
__device__ __forceinline__ float q_to_float(uint16_t raw) {
  if constexpr (QDT == B200ATTN_BF16) return __uint_as_float((uint32_t)raw << 16);
  else return __half2float(__ushort_as_half(raw));
}

// In-kernel merge (reduce_segments semantics, $VLLM/v1/attention/ops/triton_unified_attention.py:645-733) of the split-KV
// partials of one (sequence, virtual kv head) pair.  Called by the 128 epilogue threads of a CTA at the end of a segment that
// does not cover its whole sequence; thread `tid` holds, for head dim d = tid and every head h < G, the CTA's unnormalised
// partial acc[h] = sum_j 2^(x_j - mg[h]) v_j[d] together with (mg[h], lg[h]).
//   1. one thread draws a ticket from the pair's arrival counter;
//   2. every CTA but the last stores its partial to the workspace and bumps the pair's "published" counter with a release;
//   3. the last CTA never stores its own partial: it waits until the others are published (normally long ago), loads their
//      partials (one L2 round trip, all loads in flight at once) and combines them with its registers.
// Both counters live in the plan buffer (zeroed by the plan kernel, re-armed here by the last CTA).  mrg = smem scratch of
// 64 floats, hand = 3 ints of smem.
// Where the pieces of a split (sequence, virtual kv head) pair live: CTAs c_lo..c_hi, the first one's workspace slot.
struct MergeLoc {
  int c_lo, c_hi, slot_lo, np;
};
__device__ __forceinline__ MergeLoc merge_locate(int Hv, int tpc, int hv, int L, int pfx) {
  MergeLoc m;
  const int nt = (L + kTileTokens - 1) / kTileTokens;
  const long long x0 = (long long)pfx * Hv + (long long)hv * nt;
  m.c_lo = (int)(x0 / tpc);
  m.c_hi = (int)((x0 + nt - 1) / tpc);
  m.slot_lo = (x0 == (long long)m.c_lo * tpc) ? 0 : 1;
  m.np = m.c_hi - m.c_lo + 1;
  return m;
}

// Ticket of the in-kernel merge (one thread): where the pair's counters are and in which order this CTA arrived.
struct MergeTicket {
  int* cnt;       // arrival counter; the "published" counter sits kPlanCounterWords / 2 words further
  int ticket, np, c_lo, slot_lo;
};
__device__ __forceinline__ MergeTicket merge_draw_ticket(const DecodeParams& p, int hv, int L, int pfx) {
  MergeTicket t;
  const int tpc = __ldg(p.plan + 1);
  const int nt = (L + kTileTokens - 1) / kTileTokens;
  const long long x0 = (long long)pfx * p.Hv + (long long)hv * nt;
  t.c_lo = (int)(x0 / tpc);
  const int c_hi = (int)((x0 + nt - 1) / tpc);
  t.slot_lo = (x0 == (long long)t.c_lo * tpc) ? 0 : 1;
  t.cnt = p.plan + kPlanCounterOffset + t.c_lo * 2 + t.slot_lo;
  t.np = c_hi - t.c_lo + 1;
  t.ticket = atomicAdd(t.cnt, 1);
  return t;
}
// release: every store this thread has observed (its own, and through barriers the other epilogue threads') is visible
// to whoever acquires the bumped counter
__device__ __forceinline__ void merge_publish(int* pub) {
  asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(pub) : "memory");
}

// Request ring between the epilogue threads and the publish-helper warp (smem): ring[0] = number of requests pushed so far
// (low 32 bits), ring[1 + (i & 7)] = request i (address of a "published" counter, or 0 = "no more requests").  A CTA cuts at
// most two (sequence, kv head) pairs, so at most three requests are ever pushed and the ring cannot overflow.
__device__ __forceinline__ void pub_ring_push(volatile unsigned long long* ring, unsigned long long req) {
  const uint32_t n = (uint32_t)ring[0];
  ring[1 + (n & 7)] = req;
  const uint32_t a = smem_u32(const_cast<unsigned long long*>(ring));
  asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" ::"r"(a), "r"(n + 1) : "memory");
}
__device__ __forceinline__ uint32_t pub_ring_count(volatile unsigned long long* ring) {
  uint32_t n;
  asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(n) : "r"(smem_u32(const_cast<unsigned long long*>(ring))) : "memory");
  return n;
}

// `tk` is thread 0's ticket (drawn any time before; other threads pass anything).  kAsyncPublish: instead of fencing inline
// (one memory round trip, several microseconds while the CTA's own KV stream saturates HBM), thread 0 leaves the address of
// the "published" counter in the request ring `pub_req`; a helper warp polls the ring and does the release + signal.
template <int QDT, int GH, bool kAsyncPublish>
__device__ __forceinline__ void fused_merge_tail(const DecodeParams& p, int s, int hv, float vscale, int tid, int pslot,
                                                 const MergeTicket& tk, const float (&acc)[GH], const float (&mg)[GH],
                                                 const float (&lg)[GH], float* mrg, volatile int* hand,
                                                 volatile unsigned long long* pub_req) {
  const float ninf = -INFINITY;
  if (tid == 0) {
    hand[0] = (tk.ticket == tk.np - 1) ? tk.np : 0;
    hand[1] = tk.c_lo;
    hand[2] = tk.slot_lo;
  }
  named_bar_sync(1, kConsumerWarps * 32);
  const int np = hand[0], c_lo = hand[1], slot_lo = hand[2];
  if (np == 0) {
    // ---- not the last: publish the partial ----
#pragma unroll
    for (int h = 0; h < GH; ++h) {
      if (h < p.G) {
        const size_t pidx = ((size_t)blockIdx.x * 2 + pslot) * kMaxGroup + h;
        __stcg(p.ws_o + pidx * kD + tid, acc[h]);
        if (tid == 0) __stcg(reinterpret_cast<float2*>(p.ws_ml + pidx * 2), make_float2(mg[h], lg[h]));
      }
    }
    named_bar_sync(1, kConsumerWarps * 32);   // every thread's stores are issued; the release is cumulative over them
    if (tid == 0) {
      if constexpr (kAsyncPublish) {
        // hand the release to the helper warp: request ring slot, then the sequence word it polls (CTA-scope release)
        pub_ring_push(pub_req, reinterpret_cast<unsigned long long>(tk.cnt + kPlanCounterWords / 2));
      } else {
        merge_publish(tk.cnt + kPlanCounterWords / 2);
      }
    }
    return;
  }
  // ---- the last CTA of the pair: wait for the others' partials, then reduce ----
  if (tid == 0) {
    int* cnt = tk.cnt;
    int* pub = cnt + kPlanCounterWords / 2;
    int seen;
    do {
      asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(pub) : "memory");
    } while (seen < np - 1);
    *cnt = 0;   // re-arm both counters for the next launch
    *pub = 0;
  }
  named_bar_sync(1, kConsumerWarps * 32);
  const int own = (int)blockIdx.x - c_lo;
  auto part = [&](int i, int h) { return ((size_t)(c_lo + i) * 2 + (i == 0 ? slot_lo : 0)) * kMaxGroup + h; };
  constexpr int CH = GH == 8 ? 4 : 2;   // partials per round: their loads are all in flight at once (one L2 round trip)
  float ms[GH], ls[GH], os[GH];
#pragma unroll
  for (int h = 0; h < GH; ++h) { ms[h] = ninf; ls[h] = 0.f; os[h] = 0.f; }   // partials are folded in index order (deterministic)
  float2* ml_s = reinterpret_cast<float2*>(mrg);   // [CH][GH] (m, l) of the round, loaded once and shared through smem
  for (int i0 = 0; i0 < np; i0 += CH) {
    float ov[CH][GH];
#pragma unroll
    for (int j = 0; j < CH; ++j)
#pragma unroll
      for (int h = 0; h < GH; ++h) {
        const int i = i0 + j;
        ov[j][h] = (i == own) ? acc[h] : ((i < np && h < p.G) ? ldcg_f32(p.ws_o + part(i, h) * kD + tid) : 0.f);
      }
    if (tid < CH * GH) {
      const int j = tid / GH, h = tid % GH, i = i0 + j;
      float2 v = make_float2(ninf, 0.f);
      if (i == own) {
#pragma unroll
        for (int hh = 0; hh < GH; ++hh)
          if (hh == h) v = make_float2(mg[hh], lg[hh]);
      } else if (i < np && h < p.G) {
        v = ldcg_f32x2(p.ws_ml + part(i, h) * 2);
      }
      ml_s[tid] = v;
    }
    named_bar_sync(1, kConsumerWarps * 32);
#pragma unroll
    for (int j = 0; j < CH; ++j)
#pragma unroll
      for (int h = 0; h < GH; ++h) {
        const float2 mv = ml_s[j * GH + h];
        const float mn = fmaxf(ms[h], mv.x);
        const float a = (ms[h] == ninf) ? 0.f : fast_exp2(ms[h] - mn);
        const float b = (mv.x == ninf) ? 0.f : fast_exp2(mv.x - mn);
        os[h] = os[h] * a + ov[j][h] * b;
        ls[h] = ls[h] * a + mv.y * b;
        ms[h] = mn;
      }
    if (i0 + CH < np) named_bar_sync(1, kConsumerWarps * 32);   // the next round overwrites ml_s
  }
  using E = Elem<QDT>;
#pragma unroll
  for (int h = 0; h < GH; ++h) {
    if (h < p.G) {
      const int hq = hv * p.G + h;
      typename E::type* op = reinterpret_cast<typename E::type*>(p.out) + (int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h;
      op[tid] = E::from_float(os[h] * (vscale / ls[h]));
      if (p.lse_out != nullptr && tid == 0) p.lse_out[(int64_t)s * p.Hq + hq] = (ms[h] + fast_log2(ls[h])) * 0.6931471805599453f;
    }
  }
}

}  // namespace b200attn
