// decode.cu — Kernel 2: persistent split-KV paged flash-decoding for sm_100a.
//
// Reference semantics: the decode branch of AttentionImpl.forward, i.e. kernel_unified_attention
// ($VLLM/v1/attention/ops/triton_unified_attention.py:440-643: paged gather by block_table, S = scale*Q.K^T,
// seq_len mask, online softmax, P.V, split-KV partial store) with the FlashInfer scale convention
// ($VLLM/v1/attention/backends/flashinfer.py:1396-1404: k_scale folded into the logit scale, v_scale applied
// to the output).
//
// Design:
//   * Work = flat list of (sequence, kv head, 64-token tile); the plan kernel cuts it into equal contiguous
//     ranges, one per persistent CTA, so ragged batches stay balanced.  A CTA walks its range as "segments"
//     (maximal runs inside one (sequence, kv head)).
//   * Warp 4 is the producer: it issues TMA tensor loads (UTMALDG, 128B swizzle, L2 evict-first) of 16-token x
//     128-byte boxes straight out of the paged cache into a ring of smem stages, and a bulk copy of the G query
//     rows of each segment.  mbarrier full/empty pairs, no __syncthreads in the steady state.
//   * Warps 0-3 are consumers; each owns one 16-token chunk of every stage:
//       S(16 heads x 16 tok) = Q.K^T on mma.sync m16n8k16 (heads are the M rows, padded to 16),
//       online softmax in registers (log2 domain), P reused as the B operand of O^T(128 x heads) += V^T.P^T.
//     FP8 K/V bytes are expanded e4m3 -> f16 in registers (exact, F2FP.UNPACK) right before the HMMA; the
//     byte gather that transposes V uses PRMT.  16-bit K/V feed the HMMA directly (V through LDSM.T).
//   * Segment end: the 4 warps combine their (m, l, O) through smem; a segment that covers the whole
//     sequence is normalised and stored to `out`, otherwise the partial goes to the workspace for merge.cu.
#include "decode_common.cuh"

namespace b200attn {

#ifndef B200_FP8_CTAS_NT1
#define B200_FP8_CTAS_NT1 2
#endif
#ifndef B200_KV16_CTAS_NT1
#define B200_KV16_CTAS_NT1 2
#endif
constexpr int kQSlotBytes = kMaxGroup * kD * 2;  // 4 KiB: 16 query rows
constexpr int kRedStride = kD + 4;               // floats; +4 keeps rows 16-byte aligned and staggers banks

template <int KVT>
struct KvCfg {
  static constexpr bool kFp8 = (KVT == B200ATTN_FP8_E4M3);
  static constexpr int kChunkBytes = kChunkTokens * kD * (kFp8 ? 1 : 2);  // 2 KiB / 4 KiB
  static constexpr int kStageBytes = kChunkBytes * kConsumerWarps * 2;    // K + V: 16 KiB / 32 KiB
};

// smem carve-up, shared by host (size) and device (offsets)
constexpr int kDynStageBytes = 2 * kTileTokens * 4;   // per stage: 64 K scales + 64 V scales (float32)
constexpr int kDynMaxStages = 8;

template <int KVT, int NT, bool DYN = false>
struct Smem {
  static constexpr int kStageBytes = KvCfg<KVT>::kStageBytes;
  static constexpr int kDynBytes = DYN ? kDynMaxStages * kDynStageBytes : 0;   // per-token scale ring (fp8_per_token_head)
  static constexpr int kRedOBytes = kConsumerWarps * NT * 8 * kRedStride * 4;
  static constexpr int kRedMlBytes = kConsumerWarps * kMaxGroup * 2 * 4;
  static constexpr int kDescBytes = 3 * 32;   // 2 segment descriptors + merge hand-off words
  static constexpr int kBarBytes = 256;
  static constexpr int kFixedBytes = 2 * kQSlotBytes + kRedOBytes + kRedMlBytes + kDescBytes + kBarBytes + kDynBytes;
  // fp8: 2 CTAs / SM (113 KiB, 6 x 16 KiB stages).  16-bit KV, G <= 8: also 2 CTAs / SM with 2 x 32 KiB stages each -- one CTA
  // per SM (6 stages) left the 4 consumer warps latency-bound at 75 % of HBM, two reach the copy peak (profiles/r01_sweep_*).
  // 16-bit KV with G > 8 (234 registers) stays at 1 CTA / SM.
  static constexpr int kCtas = KvCfg<KVT>::kFp8 ? (NT == 1 ? B200_FP8_CTAS_NT1 : 2) : (NT == 1 ? B200_KV16_CTAS_NT1 : 1);
  static constexpr int kBudget = kCtas == 3 ? 75 * 1024 : (kCtas == 2 ? 113 * 1024 : 227 * 1024);
  static constexpr int kStagesRaw = (kBudget - kFixedBytes - 1024) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kTotalBytes = kStages * kStageBytes + kFixedBytes + 1024;  // +1024: manual alignment
  static_assert(kStages >= 2, "pipeline too shallow");
  static_assert(kTotalBytes <= kBudget, "smem budget exceeded");
};

template <int KVT, int QDT, int NT, bool DYN = false>
__global__ void __launch_bounds__(kDecodeThreads, Smem<KVT, NT, DYN>::kCtas)
decode_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v, const DecodeParams p) {
  using Cfg = KvCfg<KVT>;
  using SM = Smem<KVT, NT, DYN>;
  static_assert(!DYN || Cfg::kFp8, "per-token scales belong to the fp8 cache");
  constexpr bool kFp8 = Cfg::kFp8;
  constexpr bool kMmaBf16 = !kFp8 && (KVT == B200ATTN_BF16);  // HMMA operand type (fp8 is expanded to f16)
  constexpr int kStages = SM::kStages;
  constexpr int kChunkBytes = Cfg::kChunkBytes;
  constexpr int kStageBytes = Cfg::kStageBytes;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t stages_base = smem_base;
  const uint32_t q_base = stages_base + kStages * kStageBytes;
  const uint32_t redo_base = q_base + 2 * kQSlotBytes;
  const uint32_t redml_base = redo_base + SM::kRedOBytes;
  const uint32_t desc_base = redml_base + SM::kRedMlBytes;
  const uint32_t bar_base = desc_base + SM::kDescBytes;
  const uint32_t dyn_base = bar_base + SM::kBarBytes;   // [stage][K | V][64 tokens] float32 (DYN only)
  // generic pointers for the small reduction buffers
  uint8_t* smem_al = smem_raw + (smem_base - smem_u32(smem_raw));
  float* red_o = reinterpret_cast<float*>(smem_al + (redo_base - smem_base));
  float* red_ml = reinterpret_cast<float*>(smem_al + (redml_base - smem_base));
  volatile int* desc = reinterpret_cast<volatile int*>(smem_al + (desc_base - smem_base));
  auto full_bar = [&](uint32_t s) { return bar_base + s * 8; };
  auto empty_bar = [&](uint32_t s) { return bar_base + (kStages + s) * 8; };
  auto qfull_bar = [&](uint32_t s) { return bar_base + (2 * kStages + s) * 8; };
  auto qempty_bar = [&](uint32_t s) { return bar_base + (2 * kStages + 2 + s) * 8; };

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) dbg_stamp(p, 0);
  // the CTA's plan entry: requested first, so its latency hides behind the barrier set-up
  const int32_t* ent = p.plan + kPlanHeaderWords + (size_t)blockIdx.x * kPlanEntryWords;
  const int4 cs = __ldg(reinterpret_cast<const int4*>(ent));
  const int4 cx = __ldg(reinterpret_cast<const int4*>(ent) + 1);   // seq_len, block-table row, tile prefix of the first segment; tiles per CTA
  const int4 chk = __ldg(reinterpret_cast<const int4*>(ent) + 2);
  const int first_ids = warp == kConsumerWarps ? __ldg(ent + kPlanEntryIds + lane) : 0;   // producer: block ids of the first 8 tiles

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), kConsumerWarps);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(qfull_bar(s), 1);
      mbar_init(qempty_bar(s), kConsumerWarps);
    }
    mbar_fence_init();
    prefetch_tensormap(&tm_k);
    prefetch_tensormap(&tm_v);
  }
  __syncthreads();

  // Programmatic dependent launch: let the kernel after us (the next layer's KV write) get resident early, then run the
  // part of our prologue that does not depend on the preceding kernel.  What may be touched BEFORE griddepcontrol.wait:
  // the plan, seq_lens and block_table (step-level metadata, older than the preceding kernel) and K/V tiles that lie
  // entirely below L - fresh (`fresh` = trailing positions the preceding kernel may have written; < 0 = unknown).
  // Q, the newest tokens, `out`, the workspace and the merge counters are only touched after the wait.
  pdl_launch_dependents();

  if (chk.x != (int)gridDim.x || chk.y != p.B) __trap();   // plan / launch mismatch
  // rows without any token: (o = 0, lse = -inf) when the caller wants the LSE (see decode_tc.cu); idle CTAs take part
  if (p.lse_out != nullptr && warp < kConsumerWarps) {
    using E = Elem<QDT>;
    pdl_wait();
    for (int s = blockIdx.x; s < p.B; s += gridDim.x) {
      if (__ldg(p.seq_lens + s) > 0) continue;
      for (int i = threadIdx.x; i < p.Hq * kD; i += kConsumerWarps * 32) {
        const int hq = i / kD, d = i % kD;
        reinterpret_cast<typename E::type*>(p.out)[(int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h + d] = E::from_float(0.f);
      }
      for (int hq = threadIdx.x; hq < p.Hq; hq += kConsumerWarps * 32) p.lse_out[(int64_t)s * p.Hq + hq] = -INFINITY;
    }
  }
  if (cs.x < 0 || cs.w <= 0) return;  // idle CTA (fewer tiles than CTAs)
  if (threadIdx.x == 0) dbg_stamp(p, 1);

  if (warp == kConsumerWarps) {
    // =========================================== producer ===========================================
    const uint64_t pol = policy_evict_first();
    int s = cs.x, hv = cs.y, tile = cs.z, remaining = cs.w;
    uint32_t stage = 0, phase = 0, sg = 0;
    const uint16_t* qptr = reinterpret_cast<const uint16_t*>(p.q);
    bool waited = false;
    int early_left = p.fresh >= 0 ? kStages : 0;  // tiles we may still issue before the wait (ring is empty at start)
    while (remaining > 0) {
      // the first segment's metadata comes with the plan entry (no dependent loads before the first TMA)
      const int L = sg == 0 ? cx.x : max(__ldg(p.seq_lens + s), 0);
      const int nt = (L + kTileTokens - 1) / kTileTokens;
      const int n = min(remaining, nt - tile);
      const int kvh = (int)p.d_gsplit.div((uint32_t)hv);
      const int row = sg == 0 ? cx.y : (p.row_map ? __ldg(p.row_map + s) : s);
      const int32_t* bt = p.block_table + (int64_t)row * p.bt_stride;
      const int pfx = sg == 0 ? cx.z : __ldg(p.plan + kPlanPrefixOffset + s);
      const uint32_t slot = sg & 1;
      bool q_issued = false;
      int blk = 0;
      for (int tt = 0; tt <= n; ++tt) {
        // ---- (once per segment, as soon as we are past the wait) query rows + segment descriptor ----
        const bool early_ok = !waited && early_left > 0 && tt < n && (tile + tt + 1) * kTileTokens <= L - p.fresh;
        if (!early_ok && !waited) {
          pdl_wait();
          waited = true;
        }
        if (waited && !q_issued) {
          mbar_wait(qempty_bar(slot), ((sg >> 1) & 1) ^ 1);
          if (lane == 0) {
            volatile int* d = desc + slot * 8;
            d[0] = s; d[1] = hv; d[2] = tile; d[3] = n; d[4] = L; d[5] = pfx;
            mbar_arrive_expect_tx(qfull_bar(slot), (uint32_t)p.G * kD * 2);
          }
          __syncwarp();
          if (lane < p.G) {
            const uint16_t* src = qptr + (int64_t)s * p.q_stride_t + (int64_t)(hv * p.G + lane) * p.q_stride_h;
            bulk_load_1d(q_base + slot * kQSlotBytes + lane * (kD * 2), src, kD * 2, qfull_bar(slot));
          }
          q_issued = true;
        }
        if (tt == n) break;
        if (early_ok) --early_left;
        // ---- K/V tile ----
        const int tl = tile + tt;
        if ((tt & 7) == 0) {  // refill 32 block ids (8 tiles) with one coalesced load
          const int ci = (tl << 2) + lane;
          blk = (sg == 0 && tt == 0) ? first_ids : ((ci * kChunkTokens < L) ? __ldg(bt + p.d_cpb.div((uint32_t)ci)) : 0);
        }
        mbar_wait(empty_bar(stage), phase ^ 1);
        const int tok0 = tl * kTileTokens;
        const int nvalid = min(kConsumerWarps, (L - tok0 + kChunkTokens - 1) / kChunkTokens);
        if (lane == 0) mbar_arrive_expect_tx(full_bar(stage), (uint32_t)nvalid * 2 * (kChunkBytes + (DYN ? kChunkTokens * 4 : 0)));
        __syncwarp();
        const int c = (lane >> 1) & 3;
        const int kv = lane & 1;
        const int b = __shfl_sync(0xffffffffu, blk, ((tt & 7) << 2) + c);
        if constexpr (DYN) {
          // the chunk's 16 K scales and 16 V scales: 64 contiguous bytes each in the [block][head][slot] side caches
          if (lane >= 8 && lane < 16 && c < nvalid) {
            const int ci = (tl << 2) + c;
            const int row0 = (int)((uint32_t)ci - p.d_cpb.div((uint32_t)ci) * (uint32_t)p.cpb) * kChunkTokens;
            const float* src = (kv ? p.vs_cache : p.ks_cache) + (int64_t)b * p.sc_stride_block + (int64_t)kvh * p.sc_stride_head + row0;
            bulk_load_1d(dyn_base + stage * kDynStageBytes + kv * (kTileTokens * 4) + c * (kChunkTokens * 4), src, kChunkTokens * 4, full_bar(stage));
          }
        }
        if (lane < 8 && c < nvalid) {
          const int ci = (tl << 2) + c;
          const int row0 = (int)((uint32_t)ci - p.d_cpb.div((uint32_t)ci) * (uint32_t)p.cpb) * kChunkTokens;
          const uint32_t dst = stages_base + stage * kStageBytes + (kv ? kConsumerWarps * kChunkBytes : 0) + c * kChunkBytes;
          const CUtensorMap* tm = kv ? &tm_v : &tm_k;
          if constexpr (kFp8) {
            tma_load_4d(dst, tm, full_bar(stage), 0, row0, kvh, b, pol);
          } else {
            tma_load_4d(dst, tm, full_bar(stage), 0, row0, kvh, b, pol);
            tma_load_4d(dst + 2048, tm, full_bar(stage), 64, row0, kvh, b, pol);
          }
        }
        if (sg == 0 && tt == 0 && lane == 0) dbg_stamp(p, 2);
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
      remaining -= n;
      tile += n;
      ++sg;
      if (tile == nt) {
        tile = 0;
        if (++hv == p.Hv) {
          hv = 0;
          do { ++s; } while (s < p.B && remaining > 0 && __ldg(p.seq_lens + s) <= 0);
          if (s >= p.B) break;
        }
      }
    }
    if (!waited) pdl_wait();
    return;
  }

  // Consumers never run ahead of the wait: they block on the Q barrier, which the producer arms after it.
  // ============================================= consumers =============================================
  // MMA roles (mma.sync m16n8k16, fp32 accumulate):
  //   S^T(16 tok x 8 heads) = K(16 tok x d) . Q^T(d x 8 heads)     A = K rows straight from the swizzled tile
  //   O^T(16 d x 8 heads)  += V^T(16 d x 16 tok) . P^T(16 tok x 8 heads)
  // so the GQA group sits on the N side (8 or 2x8 heads, no padding rows) and a thread owns the softmax statistics
  // of exactly the heads (2t, 2t+1) whose O^T columns it accumulates.  P^T goes from the S^T accumulator layout to the
  // B-operand layout with two movmatrix.trans per 8 heads.
  pdl_wait();
  const int w = warp;
  const int g = lane >> 2;
  const int t = lane & 3;
  const int tid = threadIdx.x;  // 0..127
  int remaining = cs.w;
  uint32_t stage = 0, phase = 0, sg = 0;
  const float ninf = -INFINITY;
  constexpr float kLazy = 8.0f;  // rescale O only when a row maximum grows by more than 2^8 (P <= 256 fits f16/bf16)

  while (remaining > 0) {
    const uint32_t slot = sg & 1;
    mbar_wait(qfull_bar(slot), (sg >> 1) & 1);
    const int s = desc[slot * 8 + 0];
    const int hv = desc[slot * 8 + 1];
    const int tile0 = desc[slot * 8 + 2];
    const int n = desc[slot * 8 + 3];
    const int L = desc[slot * 8 + 4];
    const int pfx = desc[slot * 8 + 5];
    const int kvh = (int)p.d_gsplit.div((uint32_t)hv);

    // ---- Q fragments (B operand: n = head g (+8), k = permuted head-dim) ----
    uint32_t qb[NT][8][2];
    float crow[NT][2];  // logit scale (log2 domain) of heads 8a+2t and 8a+2t+1
    {
      const float kscale = kFp8 ? __ldg(p.k_scale + (int64_t)kvh * p.scale_stride) : 1.0f;
      const uint32_t qs = q_base + slot * kQSlotBytes;
#pragma unroll
      for (int a = 0; a < NT; ++a) {
        const int head = a * 8 + g;
        const bool ok = head < p.G;
        uint2 r[8];
#pragma unroll
        for (int st = 0; st < 8; ++st) {
          const int dbase = kFp8 ? (32 * t + 4 * st) : (64 * (st >> 2) + 16 * t + 4 * (st & 3));
          r[st] = ok ? lds64(qs + head * (kD * 2) + dbase * 2) : make_uint2(0u, 0u);
        }
        float f = 1.0f;  // factor folded back into the logit scale
        if constexpr (kFp8 && QDT == B200ATTN_BF16) {
          // bf16 -> f16 with an exact per-head power-of-two pre-scale (keeps any finite bf16 row inside f16 range)
          float amax = 0.f;
#pragma unroll
          for (int st = 0; st < 8; ++st)
            amax = fmaxf(amax, fmaxf(fmaxf(fabsf(q_to_float<QDT>(r[st].x & 0xffff)), fabsf(q_to_float<QDT>(r[st].x >> 16))),
                                     fmaxf(fabsf(q_to_float<QDT>(r[st].y & 0xffff)), fabsf(q_to_float<QDT>(r[st].y >> 16)))));
          amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
          amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
          const uint32_t eb = (__float_as_uint(amax) >> 23) & 0xffu;
          // inv = 2^(8-e) puts the row maximum in [256, 512); f = 2^(e-8) undoes it on the logits.  Rows whose maximum
          // is below 2^-111 (biased exponent < 16) are left alone: their logits are ~0 either way.
          const float inv = eb >= 16 ? __uint_as_float((262u - eb) << 23) : 1.0f;
          f = eb >= 16 ? __uint_as_float((eb - 8u) << 23) : 1.0f;
#pragma unroll
          for (int st = 0; st < 8; ++st) {
            qb[a][st][0] = pack_f32x2<false>(q_to_float<QDT>(r[st].x & 0xffff) * inv, q_to_float<QDT>(r[st].x >> 16) * inv);
            qb[a][st][1] = pack_f32x2<false>(q_to_float<QDT>(r[st].y & 0xffff) * inv, q_to_float<QDT>(r[st].y >> 16) * inv);
          }
        } else {
#pragma unroll
          for (int st = 0; st < 8; ++st) { qb[a][st][0] = r[st].x; qb[a][st][1] = r[st].y; }
        }
        // the statistics / accumulators of heads (2t, 2t+1) live in this thread: fetch their factors from lanes g = head
        const float c = f * p.sm_scale_log2 * kscale;
        crow[a][0] = __shfl_sync(0xffffffffu, c, (2 * t) << 2);
        crow[a][1] = __shfl_sync(0xffffffffu, c, (2 * t + 1) << 2);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(qempty_bar(slot));

    float m_run[NT][2], l_run[NT][2];
    float o[NT][8][4];
#pragma unroll
    for (int a = 0; a < NT; ++a) {
      m_run[a][0] = m_run[a][1] = ninf;
      l_run[a][0] = l_run[a][1] = 0.f;
#pragma unroll
      for (int b = 0; b < 8; ++b)
#pragma unroll
        for (int c = 0; c < 4; ++c) o[a][b][c] = 0.f;
    }

    for (int tt = 0; tt < n; ++tt) {
      mbar_wait(full_bar(stage), phase);
      if (sg == 0 && tt == 0 && tid == 0) dbg_stamp(p, 3);
      const int tok0 = (tile0 + tt) * kTileTokens + w * kChunkTokens;
      if (tok0 < L && !(p.debug & 1)) {
        const uint32_t kbase = stages_base + stage * kStageBytes + w * kChunkBytes;
        const uint32_t vbase = kbase + kConsumerWarps * kChunkBytes;
        const bool partial = tok0 + kChunkTokens > L;
        // per-token scales of tokens g and g + 8 of this warp's chunk; V's carry 2^6 so that p * scale stays clear of f16
        // subnormals (undone by the constant v_scale = 2^-6 the host passes); positions past L must contribute exactly 0
        float ks_lo = 1.f, ks_hi = 1.f, vs_lo = 0.f, vs_hi = 0.f;
        if constexpr (DYN) {
          const uint32_t sb0 = dyn_base + stage * kDynStageBytes + (w * kChunkTokens + g) * 4;
          float a0, a1, b0, b1;
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(a0) : "r"(sb0));
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(a1) : "r"(sb0 + 32));
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(b0) : "r"(sb0 + kTileTokens * 4));
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(b1) : "r"(sb0 + kTileTokens * 4 + 32));
          ks_lo = a0; ks_hi = a1;
          vs_lo = (tok0 + g < L) ? b0 * 64.0f : 0.f;
          vs_hi = (tok0 + g + 8 < L) ? b1 * 64.0f : 0.f;
        }

        // ---------------- S^T = K.Q^T : two accumulator chains (even / odd k-steps) per 8 heads ----------------
        float sa[NT][4], sb[NT][4];
#pragma unroll
        for (int a = 0; a < NT; ++a)
#pragma unroll
          for (int c = 0; c < 4; ++c) sa[a][c] = sb[a][c] = 0.f;
        if constexpr (kFp8) {
          const uint32_t rA = kbase + g * 128, rB = kbase + (g + 8) * 128;
          const uint32_t c0 = (uint32_t)((2 * t) ^ g) << 4, c1 = (uint32_t)((2 * t + 1) ^ g) << 4;
          const uint4 ka0 = lds128(rA + c0), ka1 = lds128(rA + c1);
          const uint4 kb0 = lds128(rB + c0), kb1 = lds128(rB + c1);
          const uint32_t kaw[8] = {ka0.x, ka0.y, ka0.z, ka0.w, ka1.x, ka1.y, ka1.z, ka1.w};
          const uint32_t kbw[8] = {kb0.x, kb0.y, kb0.z, kb0.w, kb1.x, kb1.y, kb1.z, kb1.w};
#pragma unroll
          for (int st = 0; st < 8; ++st) {
            const uint32_t a0 = e4m3x2_to_f16x2_lo(kaw[st]), a2 = e4m3x2_to_f16x2_hi(kaw[st]);
            const uint32_t a1 = e4m3x2_to_f16x2_lo(kbw[st]), a3 = e4m3x2_to_f16x2_hi(kbw[st]);
#pragma unroll
            for (int a = 0; a < NT; ++a) {
              if (st & 1) mma_16816<false>(sb[a], a0, a1, a2, a3, qb[a][st][0], qb[a][st][1]);
              else mma_16816<false>(sa[a], a0, a1, a2, a3, qb[a][st][0], qb[a][st][1]);
            }
          }
        } else {
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              const uint32_t co = (uint32_t)((2 * t + i) ^ g) << 4;
              const uint4 ka = lds128(kbase + hf * 2048 + g * 128 + co);
              const uint4 kb = lds128(kbase + hf * 2048 + (g + 8) * 128 + co);
              const int st = hf * 4 + i * 2;
#pragma unroll
              for (int a = 0; a < NT; ++a) {
                mma_16816<kMmaBf16>(sa[a], ka.x, kb.x, ka.y, kb.y, qb[a][st][0], qb[a][st][1]);
                mma_16816<kMmaBf16>(sb[a], ka.z, kb.z, ka.w, kb.w, qb[a][st + 1][0], qb[a][st + 1][1]);
              }
            }
          }
        }

        // ---------------- online softmax (log2 domain, lazy rescale) ----------------
        // thread (g,t): x[a][0] = S[head 8a+2t][tok g], [1] = S[head 8a+2t+1][tok g], [2]/[3] = same heads, tok g+8
        uint32_t pb[NT][2];
#pragma unroll
        for (int a = 0; a < NT; ++a) {
          float x0 = (sa[a][0] + sb[a][0]) * crow[a][0], x1 = (sa[a][1] + sb[a][1]) * crow[a][1];
          float x2 = (sa[a][2] + sb[a][2]) * crow[a][0], x3 = (sa[a][3] + sb[a][3]) * crow[a][1];
          if constexpr (DYN) { x0 *= ks_lo; x1 *= ks_lo; x2 *= ks_hi; x3 *= ks_hi; }   // K's per-token scale (triton_unified_attention.py:506-510)
          if (partial) {
            if (tok0 + g >= L) { x0 = ninf; x1 = ninf; }
            if (tok0 + g + 8 >= L) { x2 = ninf; x3 = ninf; }
          }
          float mx0 = fmaxf(x0, x2), mx1 = fmaxf(x1, x3);
#pragma unroll
          for (int off = 4; off < 32; off <<= 1) {
            mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, off));
            mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, off));
          }
          const bool need = (mx0 > m_run[a][0] + kLazy) || (mx1 > m_run[a][1] + kLazy);  // -inf + 8 = -inf: first chunk
          if (__any_sync(0xffffffffu, need)) {
            const float n0 = fmaxf(m_run[a][0], mx0), n1 = fmaxf(m_run[a][1], mx1);
            const float al0 = fast_exp2(m_run[a][0] - n0), al1 = fast_exp2(m_run[a][1] - n1);  // (-inf) - finite -> 0
            m_run[a][0] = n0; m_run[a][1] = n1;
            l_run[a][0] *= al0; l_run[a][1] *= al1;
#pragma unroll
            for (int mt = 0; mt < 8; ++mt) {
              o[a][mt][0] *= al0; o[a][mt][1] *= al1; o[a][mt][2] *= al0; o[a][mt][3] *= al1;
            }
          }
          const float p0 = fast_exp2(x0 - m_run[a][0]), p1 = fast_exp2(x1 - m_run[a][1]);
          const float p2 = fast_exp2(x2 - m_run[a][0]), p3 = fast_exp2(x3 - m_run[a][1]);
          l_run[a][0] += p0 + p2;
          l_run[a][1] += p1 + p3;
          // [tok g][heads 2t,2t+1] --movmatrix.trans--> [head g][toks 2t,2t+1] = B fragment of P^T
          if constexpr (DYN) {   // V's per-token scale rides on P (triton_unified_attention.py:540-543); l_run keeps the unscaled p
            pb[a][0] = movmatrix_trans(pack_f32x2<kMmaBf16>(p0 * vs_lo, p1 * vs_lo));
            pb[a][1] = movmatrix_trans(pack_f32x2<kMmaBf16>(p2 * vs_hi, p3 * vs_hi));
          } else {
            pb[a][0] = movmatrix_trans(pack_f32x2<kMmaBf16>(p0, p1));
            pb[a][1] = movmatrix_trans(pack_f32x2<kMmaBf16>(p2, p3));
          }
        }

        // ---------------- O^T(d x heads) += V^T(d x tok) . P^T(tok x heads) ----------------
        if constexpr (kFp8) {
          // thread owns the 16 head-dims [16g, 16g+16) of tokens {2t, 2t+1, 2t+8, 2t+9}
          const uint32_t cg = (uint32_t)g;
          uint4 v0 = lds128(vbase + (2 * t) * 128 + ((cg ^ (uint32_t)(2 * t)) << 4));
          uint4 v1 = lds128(vbase + (2 * t + 1) * 128 + ((cg ^ (uint32_t)(2 * t + 1)) << 4));
          uint4 v2 = lds128(vbase + (2 * t + 8) * 128 + ((cg ^ (uint32_t)(2 * t)) << 4));
          uint4 v3 = lds128(vbase + (2 * t + 9) * 128 + ((cg ^ (uint32_t)(2 * t + 1)) << 4));
          if (partial) {  // never let stale cache bytes (possibly NaN patterns) meet p = 0
            const int j0 = tok0 + 2 * t;
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);
            if (j0 >= L) v0 = z;
            if (j0 + 1 >= L) v1 = z;
            if (j0 + 8 >= L) v2 = z;
            if (j0 + 9 >= L) v3 = z;
          }
          const uint32_t w0[4] = {v0.x, v0.y, v0.z, v0.w}, w1[4] = {v1.x, v1.y, v1.z, v1.w};
          const uint32_t w2[4] = {v2.x, v2.y, v2.z, v2.w}, w3[4] = {v3.x, v3.y, v3.z, v3.w};
#pragma unroll
          for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int hb = 0; hb < 2; ++hb) {
              const uint32_t sel = hb ? 0x7362u : 0x5140u;
              const uint32_t x01 = __byte_perm(w0[r], w1[r], sel);  // [tok0.d0, tok1.d0, tok0.d1, tok1.d1]
              const uint32_t x23 = __byte_perm(w2[r], w3[r], sel);
              const uint32_t a0 = e4m3x2_to_f16x2_lo(x01), a1 = e4m3x2_to_f16x2_hi(x01);
              const uint32_t a2 = e4m3x2_to_f16x2_lo(x23), a3 = e4m3x2_to_f16x2_hi(x23);
#pragma unroll
              for (int a = 0; a < NT; ++a) mma_16816<false>(o[a][2 * r + hb], a0, a1, a2, a3, pb[a][0], pb[a][1]);
            }
          }
        } else {
          const int mi = lane >> 3, rr = lane & 7;
          const uint32_t tokrow = (uint32_t)((mi >> 1) * 8 + rr);
          uint32_t keep_lo = 0xffffffffu, keep_hi = 0xffffffffu;  // a0/a1 (tok 2t,2t+1), a2/a3 (tok 2t+8,2t+9)
          if (partial) {
            const int j0 = tok0 + 2 * t;
            keep_lo = (j0 < L ? 0x0000ffffu : 0u) | (j0 + 1 < L ? 0xffff0000u : 0u);
            keep_hi = (j0 + 8 < L ? 0x0000ffffu : 0u) | (j0 + 9 < L ? 0xffff0000u : 0u);
          }
#pragma unroll
          for (int mt = 0; mt < 8; ++mt) {
            const uint32_t chunk = (uint32_t)(2 * (mt & 3) + (mi & 1));
            const uint32_t addr = vbase + (mt >> 2) * 2048 + tokrow * 128 + ((chunk ^ (uint32_t)rr) << 4);
            uint32_t a0, a1, a2, a3;
            ldmatrix_x4_trans(addr, a0, a1, a2, a3);
            a0 &= keep_lo; a1 &= keep_lo; a2 &= keep_hi; a3 &= keep_hi;
#pragma unroll
            for (int a = 0; a < NT; ++a) mma_16816<kMmaBf16>(o[a][mt], a0, a1, a2, a3, pb[a][0], pb[a][1]);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_bar(stage));
      if (++stage == kStages) { stage = 0; phase ^= 1; }
    }

    if (tid == 0 && remaining == n) dbg_stamp(p, 4);
    // ---------------- segment end: combine the 4 warps' (m, l, O) ----------------
#pragma unroll
    for (int a = 0; a < NT; ++a) {
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        float l = l_run[a][hh];
#pragma unroll
        for (int off = 4; off < 32; off <<= 1) l += __shfl_xor_sync(0xffffffffu, l, off);
        if (g == 0) {
          red_ml[(w * kMaxGroup + a * 8 + 2 * t + hh) * 2 + 0] = m_run[a][hh];
          red_ml[(w * kMaxGroup + a * 8 + 2 * t + hh) * 2 + 1] = l;
        }
      }
    }
    named_bar_sync(1, kConsumerWarps * 32);
#pragma unroll
    for (int a = 0; a < NT; ++a) {
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int head = a * 8 + 2 * t + hh;
        float mg = ninf;
#pragma unroll
        for (int ww = 0; ww < kConsumerWarps; ++ww) mg = fmaxf(mg, red_ml[(ww * kMaxGroup + head) * 2]);
        const float mine = red_ml[(w * kMaxGroup + head) * 2];
        const float sc = (mine == ninf) ? 0.f : fast_exp2(mine - mg);
        float* dst = red_o + ((w * NT * 8) + head) * kRedStride;
#pragma unroll
        for (int mt = 0; mt < 8; ++mt) {
          const int dlo = kFp8 ? (16 * g + 2 * mt) : (16 * mt + g);
          const int dhi = kFp8 ? (dlo + 1) : (dlo + 8);
          dst[dlo] = o[a][mt][hh] * sc;
          dst[dhi] = o[a][mt][2 + hh] * sc;
        }
      }
    }
    named_bar_sync(1, kConsumerWarps * 32);
    const bool whole = (tile0 == 0) && ((tile0 + n) * kTileTokens >= L);
    const float vscale = kFp8 ? __ldg(p.v_scale + (int64_t)kvh * p.scale_stride) : 1.0f;
    {
      const int pslot = sg == 0 ? 0 : 1;
      const int d = tid;  // 128 consumer threads <-> 128 head dims
      constexpr int GH = NT * 8;
      float accv[GH], mgv[GH], lgv[GH];
#pragma unroll
      for (int head = 0; head < GH; ++head) {
        float mg = ninf;
#pragma unroll
        for (int ww = 0; ww < kConsumerWarps; ++ww) mg = fmaxf(mg, red_ml[(ww * kMaxGroup + head) * 2]);
        float lg = 0.f, acc = 0.f;
#pragma unroll
        for (int ww = 0; ww < kConsumerWarps; ++ww) {
          const float mw = red_ml[(ww * kMaxGroup + head) * 2];
          lg += (mw == ninf) ? 0.f : red_ml[(ww * kMaxGroup + head) * 2 + 1] * fast_exp2(mw - mg);
          acc += red_o[((ww * NT * 8) + head) * kRedStride + d];
        }
        accv[head] = acc; mgv[head] = mg; lgv[head] = lg;
        if (head < p.G) {
          if (whole) {
            const int hq = hv * p.G + head;
            using E = Elem<QDT>;
            typename E::type* op = reinterpret_cast<typename E::type*>(p.out) + (int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h;
            op[d] = E::from_float(acc * (vscale / lg));
            if (p.lse_out != nullptr && d == 0)
              p.lse_out[(int64_t)s * p.Hq + hq] = (mg + fast_log2(lg)) * 0.6931471805599453f;
          } else if (!p.fused_merge) {
            const size_t pidx = ((size_t)blockIdx.x * 2 + pslot) * kMaxGroup + head;
            __stcg(p.ws_o + pidx * kD + d, acc);
            if (d == 0) __stcg(reinterpret_cast<float2*>(p.ws_ml + pidx * 2), make_float2(mg, lg));
          }
        }
      }
      // in-kernel merge: the last CTA to finish a piece of this (sequence, kv head) pair reduces all of them
      if (!whole && p.fused_merge) {
        MergeTicket tk = {};
        if (tid == 0) tk = merge_draw_ticket(p, hv, L, pfx);
        fused_merge_tail<QDT, GH, false>(p, s, hv, vscale, tid, pslot, tk, accv, mgv, lgv, red_o, desc + 16, nullptr);
      }
    }
    named_bar_sync(1, kConsumerWarps * 32);
    remaining -= n;
    ++sg;
  }
  if (tid == 0) dbg_stamp(p, 5);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
int encode_kv_tensor_map(CUtensorMap* tm, const void* base, int kv_dtype, int64_t stride_block, int64_t stride_slot,
                         int64_t stride_head, int64_t num_blocks, int block_size, int Hkv, int D, int box_tokens = kChunkTokens);
int launch_with_pdl(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args);
int option_force_simt();
int option_debug();
int option_tc();
int option_fused_write();
int option_big_box();
int launch_decode_tc(const CUtensorMap& tk, const CUtensorMap& tv, DecodeParams& p, int grid, int q_dtype, cudaStream_t stream);
int launch_decode_simt(const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out, int64_t out_stride_t, int64_t out_stride_h,
                       const void* k_cache, const void* v_cache, int64_t stride_block, int64_t stride_slot, int64_t stride_head,
                       const int32_t* block_table, int64_t bt_stride, const int32_t* seq_lens, int B, int Hq, int Hkv, int D,
                       int block_size, int q_dtype, int kv_dtype, float sm_scale, const float* k_scale, const float* v_scale,
                       int scale_stride, int num_ctas, void* workspace, size_t workspace_bytes, float* lse_out,
                       const int32_t* row_map, cudaStream_t st);
unsigned long long* option_debug_buffer();

// constant scales of the per-token-head mode: K's logit factor is 1, V's output factor undoes the 2^6 carried on P
__device__ float g_dyn_static_scales[2] = {1.0f, 1.0f / 64.0f};

template <int KVT, int QDT, int NT, bool DYN = false>
static int launch_decode(const CUtensorMap& tk, const CUtensorMap& tv, const DecodeParams& p, int grid, cudaStream_t stream) {
  using SM = Smem<KVT, NT, DYN>;
  auto fn = decode_kernel<KVT, QDT, NT, DYN>;
  static bool configured[64] = {};  // per instantiation and device (the attribute belongs to the context)
  int dev = 0;
  if (check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return B200ATTN_ECUDA;
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    if (check_cuda(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::kTotalBytes),
                   "cudaFuncSetAttribute(decode smem)"))
      return B200ATTN_ECUDA;
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  void* args[] = {(void*)&tk, (void*)&tv, (void*)&p};
  return launch_with_pdl((const void*)fn, dim3(grid), dim3(kDecodeThreads), SM::kTotalBytes, stream, args);
}

int decode_smem_bytes(int kv_dtype, int nt) {
  if (kv_dtype == B200ATTN_FP8_E4M3) return nt == 2 ? Smem<B200ATTN_FP8_E4M3, 2>::kTotalBytes : Smem<B200ATTN_FP8_E4M3, 1>::kTotalBytes;
  return nt == 2 ? Smem<B200ATTN_BF16, 2>::kTotalBytes : Smem<B200ATTN_BF16, 1>::kTotalBytes;
}

}  // namespace b200attn

using namespace b200attn;

// per-(token, head) scale side caches (b200attn_decode_per_token_head)
struct DynScales {
  const float* ks_cache;
  const float* vs_cache;
  int64_t stride_block, stride_slot, stride_head;
};

// the new token of every row, for the decode launch that also writes the cache (b200attn_decode_kv_write)
struct NewTokens {
  const void* key;
  const void* value;
  int64_t key_stride_t, key_stride_h, value_stride_t, value_stride_h;
  const int64_t* slot_mapping;
};

static int decode_impl(const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out, int64_t out_stride_t,
                       int64_t out_stride_h, const void* k_cache, const void* v_cache, int64_t stride_block,
                       int64_t stride_slot, int64_t stride_head, int64_t num_blocks, const int32_t* block_table,
                       int64_t block_table_stride, const int32_t* seq_lens, int32_t num_seqs, int32_t num_q_heads,
                       int32_t num_kv_heads, int32_t head_dim, int32_t block_size, int32_t q_dtype, int32_t kv_dtype,
                       float sm_scale, const float* k_scale, const float* v_scale, int32_t scale_stride,
                       const void* plan, int32_t num_ctas, void* workspace, size_t workspace_bytes,
                       float* lse_out, const int32_t* row_map, int32_t fresh_tokens, int32_t flags, void* stream,
                       const NewTokens* nw, const DynScales* dyn = nullptr) {
  if (num_seqs == 0) return B200ATTN_OK;
  if (!q || !out || !k_cache || !v_cache || !block_table || !seq_lens || !plan || !workspace) {
    set_error("decode: null pointer argument");
    return B200ATTN_EINVAL;
  }
  if (block_size <= 0 || block_size % kChunkTokens != 0 || block_size > 256) {
    set_error("decode: block_size %d must be a multiple of 16 and <= 256", block_size);
    return B200ATTN_EINVAL;
  }
  if (num_kv_heads <= 0 || num_q_heads <= 0 || num_q_heads % num_kv_heads != 0) {
    set_error("decode: Hq=%d must be a positive multiple of Hkv=%d", num_q_heads, num_kv_heads);
    return B200ATTN_EINVAL;
  }
  if (q_dtype != B200ATTN_BF16 && q_dtype != B200ATTN_FP16) {
    set_error("decode: q dtype must be bf16 or fp16");
    return B200ATTN_EINVAL;
  }
  const bool fp8 = kv_dtype == B200ATTN_FP8_E4M3;
  if (!fp8 && kv_dtype != q_dtype) {
    set_error("decode: 16-bit kv dtype (%d) must equal q dtype (%d)", kv_dtype, q_dtype);
    return B200ATTN_EINVAL;
  }
  if (fp8 && dyn == nullptr && (!k_scale || !v_scale)) {
    set_error("decode: fp8 cache needs k_scale / v_scale");
    return B200ATTN_EINVAL;
  }
  const int Gfull = num_q_heads / num_kv_heads;
  int gsplit = 1;
  while (Gfull / gsplit > kMaxGroup || (Gfull % gsplit) != 0) {
    ++gsplit;
    if (gsplit > Gfull) break;
  }
  if (gsplit > Gfull || Gfull / gsplit > kMaxGroup) {
    set_error("decode: cannot tile group size %d into passes of <= %d heads", Gfull, kMaxGroup);
    return B200ATTN_EINVAL;
  }
  const int G = Gfull / gsplit;
  if (head_dim != kD) {
    // head_dim 64 / 256: the CUDA-core kernel (decode_simt.cu); the tensor-core kernels tile 128-byte rows only
    if (dyn != nullptr) {
      set_error("decode_per_token_head: head_dim %d unsupported (only 128)", head_dim);
      return B200ATTN_EINVAL;
    }
    if (fp8 && (!k_scale || !v_scale)) {
      set_error("decode: fp8 cache needs k_scale / v_scale");
      return B200ATTN_EINVAL;
    }
    if (Gfull > 64) {
      set_error("decode: %d query heads per kv head unsupported", Gfull);
      return B200ATTN_EINVAL;
    }
    if (nw != nullptr) {
      int rcw = b200attn_kv_write(nw->key, nw->value, nw->key_stride_t, nw->key_stride_h, nw->value_stride_t, nw->value_stride_h,
                                  const_cast<void*>(k_cache), const_cast<void*>(v_cache), stride_block, stride_slot, stride_head,
                                  nw->slot_mapping, num_seqs, num_kv_heads, head_dim, block_size, q_dtype, kv_dtype, k_scale, v_scale,
                                  scale_stride, stream);
      if (rcw) return rcw;
    }
    return launch_decode_simt(q, q_stride_t, q_stride_h, out, out_stride_t, out_stride_h, k_cache, v_cache, stride_block, stride_slot,
                              stride_head, block_table, block_table_stride, seq_lens, num_seqs, num_q_heads, num_kv_heads, head_dim,
                              block_size, q_dtype, kv_dtype, sm_scale, k_scale, v_scale, scale_stride, num_ctas > 0 ? num_ctas : 1,
                              workspace, workspace_bytes, lse_out, row_map, static_cast<cudaStream_t>(stream));
  }
  if (num_ctas <= 0 || num_ctas > kMaxGrid) {
    set_error("decode: num_ctas %d out of range (1..%d)", num_ctas, kMaxGrid);
    return B200ATTN_EINVAL;
  }
  const size_t need = (ws_o_floats(num_ctas, kD) + ws_ml_floats(num_ctas)) * sizeof(float);
  if (workspace_bytes < need) {
    set_error("decode: workspace too small (%zu < %zu bytes)", workspace_bytes, need);
    return B200ATTN_ENOSPACE;
  }
  auto mis = [](const void* ptr, int64_t a) { return (reinterpret_cast<uintptr_t>(ptr) % a) != 0; };
  if (mis(q, 16) || (q_stride_t % 8) || (q_stride_h % 8) || mis(out, 2)) {
    set_error("decode: q must be 16-byte aligned with strides that are multiples of 8 elements");
    return B200ATTN_EINVAL;
  }

  // The tcgen05 kernel takes one TMA box per min(block_size, 64) tokens (block_size 32 / >= 64: half / a quarter of the
  // UTMALDG instructions per byte; the swizzled smem image is the same as that of consecutive 16-token boxes).  64 is the
  // limit because a piece of a row may start on any 64-token plan tile, so a 128-token box could straddle two cache blocks.
  const bool use_tc = dyn == nullptr && fp8 && G <= 8 && option_tc();
  const int cpb = block_size / kChunkTokens;
  const int box_chunks = (use_tc && option_big_box() && (cpb & (cpb - 1)) == 0) ? (cpb < 4 ? cpb : 4) : 1;
  CUtensorMap tk, tv;
  int rc = encode_kv_tensor_map(&tk, k_cache, kv_dtype, stride_block, stride_slot, stride_head, num_blocks, block_size, num_kv_heads, kD,
                                box_chunks * kChunkTokens);
  if (rc) return rc;
  rc = encode_kv_tensor_map(&tv, v_cache, kv_dtype, stride_block, stride_slot, stride_head, num_blocks, block_size, num_kv_heads, kD,
                            box_chunks * kChunkTokens);
  if (rc) return rc;

  DecodeParams p;
  p.q = q; p.q_stride_t = q_stride_t; p.q_stride_h = q_stride_h;
  p.out = out; p.out_stride_t = out_stride_t; p.out_stride_h = out_stride_h;
  p.block_table = block_table; p.bt_stride = block_table_stride;
  p.seq_lens = seq_lens;
  p.plan = reinterpret_cast<int32_t*>(const_cast<void*>(plan));
  p.row_map = row_map;
  p.fresh = fresh_tokens;
  p.fused_merge = (flags & B200ATTN_DECODE_FUSED_MERGE) ? 1 : 0;
  p.ws_o = reinterpret_cast<float*>(workspace);
  p.ws_ml = p.ws_o + ws_o_floats(num_ctas, kD);
  p.k_scale = k_scale; p.v_scale = v_scale; p.scale_stride = scale_stride;
  p.lse_out = lse_out;
  p.sm_scale_log2 = sm_scale * 1.4426950408889634f;
  p.B = num_seqs; p.Hq = num_q_heads; p.Hv = num_kv_heads * gsplit; p.G = G; p.gsplit = gsplit;
  p.cpb = cpb; p.box_chunks = box_chunks;
  p.d_cpb.init((uint32_t)p.cpb); p.d_gsplit.init((uint32_t)gsplit);
  p.cluster = 1;
  p.debug = option_debug();
  p.dbg = option_debug_buffer();
  p.new_key = nullptr; p.new_value = nullptr; p.slot_mapping = nullptr;
  p.nk_stride_t = p.nk_stride_h = p.nv_stride_t = p.nv_stride_h = 0;
  p.kc_base = nullptr; p.vc_base = nullptr;
  p.c_stride_block = stride_block; p.c_stride_slot = stride_slot; p.c_stride_head = stride_head;
  p.block_size = block_size;
  p.ks_cache = nullptr; p.vs_cache = nullptr; p.sc_stride_block = 0; p.sc_stride_head = 0;

  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dyn != nullptr) {
    // fp8_per_token_head: the HMMA consumers apply K's scale to the logits and V's to P (decode.cu, DYN instantiation)
    if (!fp8) {
      set_error("decode_per_token_head: needs an fp8 cache");
      return B200ATTN_EINVAL;
    }
    if (!dyn->ks_cache || !dyn->vs_cache || dyn->stride_slot != 1 || (dyn->stride_block % 4) || (dyn->stride_head % 4) ||
        (reinterpret_cast<uintptr_t>(dyn->ks_cache) % 16) || (reinterpret_cast<uintptr_t>(dyn->vs_cache) % 16)) {
      set_error("decode_per_token_head: scale caches must be 16-byte aligned float32 [block][head][slot] views (slot stride 1, "
                "block / head strides multiples of 4)");
      return B200ATTN_EINVAL;
    }
    float* cst = nullptr;
    if (check_cuda(cudaGetSymbolAddress(reinterpret_cast<void**>(&cst), g_dyn_static_scales), "cudaGetSymbolAddress")) return B200ATTN_ECUDA;
    p.k_scale = cst; p.v_scale = cst + 1; p.scale_stride = 0;
    p.ks_cache = dyn->ks_cache; p.vs_cache = dyn->vs_cache;
    p.sc_stride_block = dyn->stride_block; p.sc_stride_head = dyn->stride_head;
    if (G > 8) return q_dtype == B200ATTN_BF16 ? launch_decode<B200ATTN_FP8_E4M3, B200ATTN_BF16, 2, true>(tk, tv, p, num_ctas, st)
                                               : launch_decode<B200ATTN_FP8_E4M3, B200ATTN_FP16, 2, true>(tk, tv, p, num_ctas, st);
    return q_dtype == B200ATTN_BF16 ? launch_decode<B200ATTN_FP8_E4M3, B200ATTN_BF16, 1, true>(tk, tv, p, num_ctas, st)
                                    : launch_decode<B200ATTN_FP8_E4M3, B200ATTN_FP16, 1, true>(tk, tv, p, num_ctas, st);
  }
  if (nw != nullptr) {
    // In-kernel write: the tcgen05 kernel, rows == tokens (no row_map: with several query tokens per request a later row
    // reads what an earlier row's CTA writes).  Every other shape runs the KV-write kernel first -- same result, two launches.
    if (use_tc && row_map == nullptr && option_fused_write()) {
      p.new_key = nw->key; p.new_value = nw->value; p.slot_mapping = nw->slot_mapping;
      p.nk_stride_t = nw->key_stride_t; p.nk_stride_h = nw->key_stride_h;
      p.nv_stride_t = nw->value_stride_t; p.nv_stride_h = nw->value_stride_h;
      p.kc_base = reinterpret_cast<uint8_t*>(const_cast<void*>(k_cache));
      p.vc_base = reinterpret_cast<uint8_t*>(const_cast<void*>(v_cache));
      if (p.fresh == 0) p.fresh = 1;   // the row's last position is written by this launch: never read it before the wait
    } else {
      rc = b200attn_kv_write(nw->key, nw->value, nw->key_stride_t, nw->key_stride_h, nw->value_stride_t, nw->value_stride_h,
                             const_cast<void*>(k_cache), const_cast<void*>(v_cache), stride_block, stride_slot, stride_head,
                             nw->slot_mapping, num_seqs, num_kv_heads, head_dim, block_size, q_dtype, kv_dtype, k_scale, v_scale,
                             scale_stride, stream);
      if (rc) return rc;
      if (p.fresh == 0) p.fresh = 1;
    }
  }
  // fp8 KV with at most 8 query heads per kv head: both GEMMs on tcgen05 (decode_tc.cu); everything else: HMMA consumers
  if (use_tc) return launch_decode_tc(tk, tv, p, num_ctas, q_dtype, st);
  const int nt = G > 8 ? 2 : 1;
#define B200_LAUNCH(KVT, QDT)                                                       \
  (nt == 2 ? launch_decode<KVT, QDT, 2>(tk, tv, p, num_ctas, st) : launch_decode<KVT, QDT, 1>(tk, tv, p, num_ctas, st))
  if (fp8) return q_dtype == B200ATTN_BF16 ? B200_LAUNCH(B200ATTN_FP8_E4M3, B200ATTN_BF16) : B200_LAUNCH(B200ATTN_FP8_E4M3, B200ATTN_FP16);
  if (kv_dtype == B200ATTN_BF16) return B200_LAUNCH(B200ATTN_BF16, B200ATTN_BF16);
  return B200_LAUNCH(B200ATTN_FP16, B200ATTN_FP16);
#undef B200_LAUNCH
}

extern "C" int b200attn_decode(const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out, int64_t out_stride_t,
                               int64_t out_stride_h, const void* k_cache, const void* v_cache, int64_t stride_block,
                               int64_t stride_slot, int64_t stride_head, int64_t num_blocks, const int32_t* block_table,
                               int64_t block_table_stride, const int32_t* seq_lens, int32_t num_seqs, int32_t num_q_heads,
                               int32_t num_kv_heads, int32_t head_dim, int32_t block_size, int32_t q_dtype, int32_t kv_dtype,
                               float sm_scale, const float* k_scale, const float* v_scale, int32_t scale_stride,
                               const void* plan, int32_t num_ctas, void* workspace, size_t workspace_bytes,
                               float* lse_out, const int32_t* row_map, int32_t fresh_tokens, int32_t flags, void* stream) {
  return decode_impl(q, q_stride_t, q_stride_h, out, out_stride_t, out_stride_h, k_cache, v_cache, stride_block, stride_slot,
                     stride_head, num_blocks, block_table, block_table_stride, seq_lens, num_seqs, num_q_heads, num_kv_heads,
                     head_dim, block_size, q_dtype, kv_dtype, sm_scale, k_scale, v_scale, scale_stride, plan, num_ctas, workspace,
                     workspace_bytes, lse_out, row_map, fresh_tokens, flags, stream, nullptr);
}

extern "C" int b200attn_decode_kv_write(const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out, int64_t out_stride_t,
                                        int64_t out_stride_h, const void* key, const void* value, int64_t key_stride_t,
                                        int64_t key_stride_h, int64_t value_stride_t, int64_t value_stride_h,
                                        const int64_t* slot_mapping, void* k_cache, void* v_cache, int64_t stride_block,
                                        int64_t stride_slot, int64_t stride_head, int64_t num_blocks, const int32_t* block_table,
                                        int64_t block_table_stride, const int32_t* seq_lens, int32_t num_seqs,
                                        int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                                        int32_t q_dtype, int32_t kv_dtype, float sm_scale, const float* k_scale,
                                        const float* v_scale, int32_t scale_stride, const void* plan, int32_t num_ctas,
                                        void* workspace, size_t workspace_bytes, float* lse_out, const int32_t* row_map,
                                        int32_t fresh_tokens, int32_t flags, void* stream) {
  if (num_seqs == 0) return B200ATTN_OK;
  if (!key || !value || !slot_mapping) {
    set_error("decode_kv_write: null key / value / slot_mapping");
    return B200ATTN_EINVAL;
  }
  auto mis = [](const void* ptr, int64_t a) { return (reinterpret_cast<uintptr_t>(ptr) % a) != 0; };
  if (mis(key, 16) || mis(value, 16) || (key_stride_t % 8) || (key_stride_h % 8) || (value_stride_t % 8) || (value_stride_h % 8)) {
    set_error("decode_kv_write: key / value must be 16-byte aligned with strides that are multiples of 8 elements");
    return B200ATTN_EINVAL;
  }
  NewTokens nw = {key, value, key_stride_t, key_stride_h, value_stride_t, value_stride_h, slot_mapping};
  return decode_impl(q, q_stride_t, q_stride_h, out, out_stride_t, out_stride_h, k_cache, v_cache, stride_block, stride_slot,
                     stride_head, num_blocks, block_table, block_table_stride, seq_lens, num_seqs, num_q_heads, num_kv_heads,
                     head_dim, block_size, q_dtype, kv_dtype, sm_scale, k_scale, v_scale, scale_stride, plan, num_ctas, workspace,
                     workspace_bytes, lse_out, row_map, fresh_tokens, flags, stream, &nw);
}

extern "C" int b200attn_decode_per_token_head(const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out, int64_t out_stride_t,
                                              int64_t out_stride_h, const void* k_cache, const void* v_cache, int64_t stride_block,
                                              int64_t stride_slot, int64_t stride_head, int64_t num_blocks,
                                              const float* k_scale_cache, const float* v_scale_cache, int64_t scale_stride_block,
                                              int64_t scale_stride_slot, int64_t scale_stride_head, const int32_t* block_table,
                                              int64_t block_table_stride, const int32_t* seq_lens, int32_t num_seqs,
                                              int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim, int32_t block_size,
                                              int32_t q_dtype, float sm_scale, const void* plan, int32_t num_ctas, void* workspace,
                                              size_t workspace_bytes, float* lse_out, const int32_t* row_map, int32_t fresh_tokens,
                                              int32_t flags, void* stream) {
  DynScales dyn = {k_scale_cache, v_scale_cache, scale_stride_block, scale_stride_slot, scale_stride_head};
  return decode_impl(q, q_stride_t, q_stride_h, out, out_stride_t, out_stride_h, k_cache, v_cache, stride_block, stride_slot,
                     stride_head, num_blocks, block_table, block_table_stride, seq_lens, num_seqs, num_q_heads, num_kv_heads,
                     head_dim, block_size, q_dtype, B200ATTN_FP8_E4M3, sm_scale, nullptr, nullptr, 0, plan, num_ctas, workspace,
                     workspace_bytes, lse_out, row_map, fresh_tokens, flags, stream, nullptr, &dyn);
}
