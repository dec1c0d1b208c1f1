// prefill_tc.cu — varlen causal flash-attention forward over the paged FP8 (e4m3) KV cache for rows with q_len > 1
// (chunked prefill, mixed batches, spec-decode verification): SURVEY.md §8(f) rank 2.
//
// Reference semantics: the prefill branch of AttentionImpl.forward, i.e. kernel_unified_attention with q_len > 1
// ($VLLM/v1/attention/ops/triton_unified_attention.py:286-303 (causal length of a query row = context_len + q_pos + 1),
// 440-545 (paged gather, S = scale*Q.K^T, mask, online softmax, P.V)) and FlashInfer's
// trtllm_batch_context_with_kv_cache ($VLLM/v1/attention/backends/flashinfer.py:1500-1689); scale convention as in
// decode (k_scale folded into the logit scale, v_scale applied to the output, flashinfer.py:1396-1404).
//
// One CTA = one (request, kv head, 128-row query tile); a query tile packs TPT = 128 / G tokens x G query heads of the kv
// head (GQA rows share the K/V stream).  Both GEMMs run on the 5th-generation tensor cores, accumulators in TMEM:
//     S[128 rows x 128 tok]  = (Qhi + Qlo)[128 x 128 d] . K_tile^T       8 x tcgen05.mma kind::f8f6f4 M128 N128 K32
//     O[128 rows x 128 d]   += (Phi + Plo)[128 x 128 tok] . V_tile        8 x tcgen05.mma M128 N128 K32 (B = V, MN-major)
// The 16-bit query rows are split into two e4m3 terms after an exact per-row power-of-two pre-scale that puts the row
// maximum in [128, 256): q = hi + lo is exact for every element >= 2^-9 of the row maximum (below that the absolute error is
// <= 2^-10, i.e. 2^-17 of the maximum); the probabilities p' = 2^(x - m + 4) in (0, 256] are split the same way.  K and V
// tiles are consumed exactly as TMA wrote them (128-byte swizzle), as in decode_tc.cu.
//   warps 0-7  softmax: two threads per query row (TMEM lane), one per 64-column half of the S tile: one TMEM read of the
//              half row into registers, row maximum exchanged through smem, online softmax with a lazily raised row maximum,
//              exp2, hi / lo split (hi by integer rounding of the fp32 bits, lo by one cvt per pair), P written to smem as
//              the A operand of the second GEMM.  Two warps per SM sub-partition keep the issue slots and the MUFU busy.
//   warp 8     producer: TMA tensor loads of 16-token x 128-byte boxes of the paged cache into a K ring and a V ring.
//   warp 9     MMA issuer (one lane), TMEM allocation.
#include <cstdio>
#include <cstdlib>

#include "decode_common.cuh"

namespace b200attn {

#ifndef B200_PF_PTMEM
#define B200_PF_PTMEM 1   // P (hi, lo) as the TMEM-resident A operand of the second GEMM (0: staged in shared memory)
#endif
#ifndef B200_PF_HALF2
#define B200_PF_HALF2 1   // probabilities in f16x2 from the exponent on (0: the fp32 / integer-rounding path of the first version;
                          // 2: exponent in fp32, hi from the fp32 pair -- 5 % slower than 1)
#endif
#ifndef B200_PF_SPLIT
#define B200_PF_SPLIT 2   // softmax threads per query row (2 or 4; 4 = 16 softmax warps measured 8 % slower: 707 vs 656 us on the 8k prefill)
#endif
constexpr int kPfSplit = B200_PF_SPLIT;
constexpr int kPfCols = 128 / kPfSplit;       // S columns / head dims per softmax thread
constexpr int kPfSoftmaxThreads = 128 * kPfSplit;
constexpr int kPfSoftmaxWarps = kPfSoftmaxThreads / 32;
constexpr int kPfThreads = kPfSoftmaxThreads + 64;
constexpr int kPfProducerWarp = kPfSoftmaxWarps, kPfMmaWarp = kPfSoftmaxWarps + 1;
static_assert(kPfSplit == 2 || kPfSplit == 4, "two or four threads per query row");
constexpr int kPfTile = 128;                  // kv tokens per tile = query rows per tile
#ifndef B200_PF_CTAS2
#define B200_PF_CTAS2 0   // 1: two CTAs per SM, S single-buffered with P aliased onto it (256 TMEM columns), 2-stage K / V rings --
                          // measured 46 % SLOWER (8k prefill 667 -> 974 us): kept as an experiment switch, see DESIGN.md 3.5
#endif
#if B200_PF_CTAS2 && !B200_PF_PTMEM
#error "B200_PF_CTAS2 needs the TMEM-resident P operand"
#endif
constexpr int kPfCtasPerSm = B200_PF_CTAS2 ? 2 : 1;
constexpr int kPfStages = B200_PF_CTAS2 ? 2 : 3;
constexpr int kPfSBufs = B200_PF_CTAS2 ? 1 : 2;           // S (and P) buffers in TMEM
constexpr int kPfTileBytes = kPfTile * kD;    // 16 KiB (fp8)
constexpr int kPfBoxBytes = kChunkTokens * kD;
// TMEM columns.  One CTA per SM: S0 [0,128) S1 [128,256) O [256,384) P0 P1 [384,512).  Two CTAs per SM: S [0,128), P (hi 32 +
// lo 32 columns) written over S once every softmax thread has its logits in registers, O [128,256): the chain
// QK(j) -> softmax(j) -> PV(j) -> QK(j+1) of one CTA is serial and the SM's other CTA fills its gaps (and hides the ~12k
// cycles of per-CTA prologue + epilogue: query load and split, pipeline fill, output store).
constexpr int kPfTmemCols = B200_PF_CTAS2 ? 256 : 512;
constexpr int kPfColS = 0, kPfColO = B200_PF_CTAS2 ? 128 : 256, kPfColP = B200_PF_CTAS2 ? 0 : 384;
constexpr float kPfMargin = 4.0f;             // the row maximum is raised when a logit exceeds it by 2^4
constexpr float kPfShift = 4.0f;              // p' = p * 2^4 <= 2^8 < 448

struct PfSmem {
  static constexpr int kQ = 0;                                   // Qhi, Qlo
  static constexpr int kK = kQ + 2 * kPfTileBytes;
  static constexpr int kV = kK + kPfStages * kPfTileBytes;
  static constexpr int kP = kV + kPfStages * kPfTileBytes;       // [2 buffers][hi, lo]
  static constexpr int kXch = kP + (B200_PF_PTMEM ? 0 : 4 * kPfTileBytes);             // float [2 slots][kPfSplit parts][128 rows]: row max / sum exchange
  static constexpr int kBars = kXch + 2 * kPfSplit * 128 * 4;
  static constexpr int kNumBars = 4 * kPfStages + 10;
  static constexpr int kTmemPtr = kBars + kNumBars * 8;
  static constexpr int kTotal = kTmemPtr + 16;
  static_assert(kTotal <= 227 * 1024 / kPfCtasPerSm - 1024, "smem budget exceeded");
};

struct PrefillParams {
  const void* q;
  int64_t q_stride_t, q_stride_h;
  void* out;
  int64_t out_stride_t, out_stride_h;
  const int32_t* block_table;      // rows of the prefill requests (row r = request req0 + r is the caller's business: pointer is pre-offset)
  int64_t bt_stride;
  const int32_t* seq_lens;         // [num_reqs] total length (context + this step's tokens), pre-offset
  const int32_t* query_start_loc;  // [num_reqs + 1] token offsets, pre-offset; tokens are indexed relative to q / out as given
  const int32_t* tile_prefix;      // [num_reqs + 1] exclusive scan of ceil(q_len / TPT)
  const float* k_scale;
  const float* v_scale;
  float sm_scale_log2;
  int scale_stride;
  int num_reqs, Hq, Hv, G, gsplit, cpb, tpt;
  int tok0;                        // query_start_loc[0] of the slice (token index subtracted from every offset)
  unsigned long long* dbg;         // per-CTA wait-cycle counters [grid][16] (debug builds of the launch only)
  int variant;                     // debug instantiation only, bit flags: 1 = softmax warps skip their arithmetic, 2 = no PV MMAs, 4 = hi terms only, 8 = no QK MMAs
};

__device__ __forceinline__ void pf_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void pf_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void pf_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void pf_mma_f8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// A operand read from TENSOR MEMORY (row = TMEM lane, 4 e4m3 elements per 32-bit column, K ascending with the column)
__device__ __forceinline__ void pf_mma_f8_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void pf_st16_nowait(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void pf_st_nowait(uint32_t taddr, const uint32_t (&r)[16]) { pf_st16_nowait(taddr, r); }
__device__ __forceinline__ void pf_st_nowait(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void pf_st_nowait(uint32_t taddr, const uint32_t (&r)[4]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1,%2,%3,%4};" ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]) : "memory");
}
__device__ __forceinline__ void pf_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void pf_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,"
      "%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void pf_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,"
      "%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]),
        "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]),
        "r"(r[30]), "r"(r[31])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
// shared-memory matrix descriptor (same encoding as decode_tc.cu)
__device__ __forceinline__ uint64_t pf_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46) | (2ull << 61);   // 128-byte swizzle
}
// kind::f8f6f4, e4m3 x e4m3 -> f32, M = 128, N = 128
constexpr uint32_t kPfIdesc = (1u << 4) | ((128u >> 3) << 17) | ((128u >> 4) << 24);
constexpr uint32_t kPfIdescQK = kPfIdesc;                 // A (Q operand) K-major, B (K tile) K-major
constexpr uint32_t kPfIdescPV = kPfIdesc | (1u << 16);    // A (P operand) K-major, B (V tile) MN-major

__device__ __forceinline__ uint32_t pf_f32x2_to_e4m3x2(float lo, float hi) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float2 pf_e4m3x2_to_f32x2(uint32_t two_bytes) {
  uint32_t h2;
  asm("{\n\t.reg .b16 t;\n\tcvt.u16.u32 t, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, t;\n\t}" : "=r"(h2) : "r"(two_bytes));
  return __half22float2(*reinterpret_cast<__half2*>(&h2));
}
// x = hi + lo with hi, lo e4m3: the two bytes of each term for a pair of values
__device__ __forceinline__ void pf_split_pair(float x0, float x1, uint32_t& hi2, uint32_t& lo2) {
  hi2 = pf_f32x2_to_e4m3x2(x0, x1);
  const float2 h = pf_e4m3x2_to_f32x2(hi2);
  lo2 = pf_f32x2_to_e4m3x2(x0 - h.x, x1 - h.y);
}
// Four non-negative values p (each <= 256): hi = p rounded (half up) to 4 significant bits by integer arithmetic on the
// fp32 bits -- an e4m3 value whose byte is the rebased exponent | 3 mantissa bits -- or 0 below the e4m3 normal range
// (2^-6); lo = e4m3(p - hi) by cvt.  Returns the four hi bytes and the four lo bytes, packed.
__device__ __forceinline__ void pf_split4_pos(const float (&pv)[4], uint32_t& hi4, uint32_t& lo4) {
  uint32_t t[4];
  float r[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint32_t hb = (__float_as_uint(pv[i]) + 0x00080000u) & 0xfff00000u;
    hb = hb >= 0x3c800000u ? hb : 0u;
    r[i] = pv[i] - __uint_as_float(hb);
    t[i] = hb ? hb - 0x3c000000u : 0u;      // e4m3 byte at bits [20, 28)
  }
  hi4 = (t[0] >> 20) | (t[1] >> 12) | (t[2] >> 4) | (t[3] << 4);
  lo4 = pf_f32x2_to_e4m3x2(r[0], r[1]) | (pf_f32x2_to_e4m3x2(r[2], r[3]) << 16);
}
// byte offset of (row r, 16-byte chunk c) inside a 128-row x 128-byte operand tile laid out like a TMA box with 128-byte swizzle
__device__ __forceinline__ uint32_t pf_swz(int r, int c) { return (uint32_t)r * 128u + (uint32_t)((c ^ (r & 7)) << 4); }

template <int QDT, bool kDbg>
__global__ void __launch_bounds__(kPfThreads, kPfCtasPerSm)
prefill_tc_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v, const PrefillParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = smem_u32(smem_raw);
  if (base & 1023u) __trap();
  const uint32_t q_base = base + PfSmem::kQ, k_base = base + PfSmem::kK, v_base = base + PfSmem::kV, p_base = base + PfSmem::kP;
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(smem_raw + PfSmem::kTmemPtr);
  const uint32_t bar0 = base + PfSmem::kBars;
  auto kfull = [&](uint32_t i) { return bar0 + i * 8; };
  auto kempty = [&](uint32_t i) { return bar0 + (kPfStages + i) * 8; };
  auto vfull = [&](uint32_t i) { return bar0 + (2 * kPfStages + i) * 8; };
  auto vempty = [&](uint32_t i) { return bar0 + (3 * kPfStages + i) * 8; };
  const uint32_t bar1 = bar0 + 4 * kPfStages * 8;
  const uint32_t qfull = bar1;
  auto sfull = [&](uint32_t i) { return bar1 + (1 + i) * 8; };
  auto sempty = [&](uint32_t i) { return bar1 + (3 + i) * 8; };
  auto pfull = [&](uint32_t i) { return bar1 + (5 + i) * 8; };
  auto pempty = [&](uint32_t i) { return bar1 + (7 + i) * 8; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // debug instantiation: cycles spent in each wait, per role (slot meanings in tests/tools/prefill_profile.py)
  long long acc[4] = {0, 0, 0, 0};
  const long long t_begin = kDbg ? clock64() : 0;
  auto wait_t = [&](uint32_t bar, uint32_t par, int slot) {
    if (kDbg) {
      const long long t0 = clock64();
      mbar_wait(bar, par);
      acc[slot] += clock64() - t0;
    } else {
      mbar_wait(bar, par);
    }
  };

  // ---- which (request, query tile, kv head) is this CTA?  heaviest units (late tiles) first ----
  const int total_tiles = __ldg(p.tile_prefix + p.num_reqs);
  const int unit = (int)blockIdx.x;
  const bool idle = unit >= total_tiles * p.Hv;
  int r = 0, qt = 0, hv = 0;
  if (!idle) {
    const int tg = total_tiles - 1 - unit / p.Hv;
    hv = unit % p.Hv;
    int lo = 0, hi = p.num_reqs;   // largest r with tile_prefix[r] <= tg
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (__ldg(p.tile_prefix + mid) <= tg) lo = mid; else hi = mid;
    }
    r = lo;
    qt = tg - __ldg(p.tile_prefix + r);
  }
  if (threadIdx.x == 0) {
    for (int i = 0; i < kPfStages; ++i) {
      mbar_init(kfull(i), 1); mbar_init(kempty(i), 1);
      mbar_init(vfull(i), 1); mbar_init(vempty(i), 1);
    }
    mbar_init(qfull, kPfSoftmaxWarps);
    for (int i = 0; i < 2; ++i) {
      mbar_init(sfull(i), 1); mbar_init(sempty(i), kPfSoftmaxWarps);
      mbar_init(pfull(i), kPfSoftmaxWarps); mbar_init(pempty(i), 1);
    }
    mbar_fence_init();
    prefetch_tensormap(&tm_k);
    prefetch_tensormap(&tm_v);
  }
  if (warp == kPfMmaWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(base + PfSmem::kTmemPtr), "r"(kPfTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  pf_fence_before();
  __syncthreads();
  pf_fence_after();
  const uint32_t tmem = *tmem_ptr_smem;
  pdl_launch_dependents();
  pdl_wait();   // the KV write of this step (and whoever produced q) precedes us

  if (!idle) {
    const int q0 = __ldg(p.query_start_loc + r) - p.tok0, q1 = __ldg(p.query_start_loc + r + 1) - p.tok0;
    const int q_len = q1 - q0;
    const int L = max(__ldg(p.seq_lens + r), 0);
    const int ctx = L - q_len;                                  // tokens already in the cache before this step's chunk
    const int t_lo = qt * p.tpt, t_hi = min(q_len, t_lo + p.tpt);   // query tokens of this tile
    const int rows = (t_hi - t_lo) * p.G;
    const int kv_end = max(0, min(L, ctx + t_hi));              // the last row of the tile sees [0, ctx + t_hi)
    const int n_kv = (kv_end + kPfTile - 1) / kPfTile;
    const int kvh = hv / p.gsplit;
    if (kDbg && p.dbg != nullptr && threadIdx.x == 0) p.dbg[(size_t)blockIdx.x * 16 + 14] = (unsigned long long)n_kv;

    if (warp == kPfProducerWarp) {
      // =========================================== producer ===========================================
      const uint64_t pol = policy_evict_normal();   // K/V tiles are re-read by the other query tiles of the request
      const int32_t* bt = p.block_table + (int64_t)r * p.bt_stride;
      for (int j = 0; j < n_kv; ++j) {
        const int tok0 = j * kPfTile;
        const int nbox = (min(kPfTile, kv_end - tok0) + kChunkTokens - 1) / kChunkTokens;
        const int c = lane & 7;
        const int ci = (tok0 >> 4) + c;
        const int b = (c < nbox) ? __ldg(bt + ci / p.cpb) : 0;
        const int row0 = (ci % p.cpb) * kChunkTokens;
        const uint32_t st = j % kPfStages, par = ((j / kPfStages) & 1) ^ 1;
        wait_t(kempty(st), par, 0);
        if (lane == 0) mbar_arrive_expect_tx(kfull(st), (uint32_t)nbox * kPfBoxBytes);
        __syncwarp();
        if (lane < 8 && c < nbox) tma_load_4d(k_base + st * kPfTileBytes + c * kPfBoxBytes, &tm_k, kfull(st), 0, row0, kvh, b, pol);
        wait_t(vempty(st), par, 1);
        if (lane == 0) mbar_arrive_expect_tx(vfull(st), (uint32_t)nbox * kPfBoxBytes);
        __syncwarp();
        if (lane >= 8 && lane < 16 && c < nbox) tma_load_4d(v_base + st * kPfTileBytes + c * kPfBoxBytes, &tm_v, vfull(st), 0, row0, kvh, b, pol);
      }
    } else if (warp == kPfMmaWarp) {
      // =========================================== MMA issuer ===========================================
      auto issue_qk = [&](int j) {
        const uint32_t b = j % kPfSBufs, st = j % kPfStages;
        wait_t(kfull(st), (j / kPfStages) & 1, 0);
        wait_t(sempty(b), ((j / kPfSBufs) & 1) ^ 1, 1);
        pf_fence_after();
        if (lane == 0) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (kDbg && (p.variant & 8)) break;
            const uint64_t db = pf_desc(k_base + st * kPfTileBytes + k * 32, 16, 1024);
            pf_mma_f8(tmem + kPfColS + b * 128, pf_desc(q_base + k * 32, 16, 1024), db, kPfIdescQK, k > 0);
            if (!kDbg || !(p.variant & 4)) pf_mma_f8(tmem + kPfColS + b * 128, pf_desc(q_base + kPfTileBytes + k * 32, 16, 1024), db, kPfIdescQK, 1u);
          }
          pf_commit(kempty(st));
          pf_commit(sfull(b));
        }
        __syncwarp();
      };
      auto issue_pv = [&](int j) {
        const uint32_t b = j % kPfSBufs, st = j % kPfStages;
        wait_t(vfull(st), (j / kPfStages) & 1, 2);
        wait_t(pfull(b), (j / kPfSBufs) & 1, 3);
        pf_fence_after();
        if (lane == 0) {
          const int nk = (min(kPfTile, kv_end - j * kPfTile) + 31) >> 5;
          for (int k = 0; k < nk; ++k) {
            if (kDbg && (p.variant & 2)) break;
            const uint64_t db = pf_desc(v_base + st * kPfTileBytes + k * 4096, 0, 1024);
#if B200_PF_PTMEM
            // P lives in TMEM: buffer b = 64 columns (32 hi + 32 lo) from kPfColP on, a 32-token K step = 8 columns
            pf_mma_f8_ts(tmem + kPfColO, tmem + kPfColP + b * 64 + k * 8, db, kPfIdescPV, (j > 0 || k > 0) ? 1u : 0u);
            if (!kDbg || !(p.variant & 4)) pf_mma_f8_ts(tmem + kPfColO, tmem + kPfColP + b * 64 + 32 + k * 8, db, kPfIdescPV, 1u);
#else
            pf_mma_f8(tmem + kPfColO, pf_desc(p_base + (2 * b) * kPfTileBytes + k * 32, 16, 1024), db, kPfIdescPV, (j > 0 || k > 0) ? 1u : 0u);
            if (!kDbg || !(p.variant & 4)) pf_mma_f8(tmem + kPfColO, pf_desc(p_base + (2 * b + 1) * kPfTileBytes + k * 32, 16, 1024), db, kPfIdescPV, 1u);
#endif
          }
          pf_commit(vempty(st));
          pf_commit(pempty(b));
        }
        __syncwarp();
      };
      if (n_kv > 0) {
        mbar_wait(qfull, 0);
        issue_qk(0);
        for (int j = 0; j < n_kv; ++j) {
#if B200_PF_CTAS2
          issue_pv(j);                         // P(j) sits where S(j) was: the next QK may only follow this PV (the tensor
          if (j + 1 < n_kv) issue_qk(j + 1);   // pipe executes in issue order)
#else
          if (j + 1 < n_kv) issue_qk(j + 1);   // QK of the next tile runs while the softmax warps work on this one
          issue_pv(j);
#endif
        }
      }
    } else {
      // =========================================== softmax warps ===========================================
      const int tid = threadIdx.x;                       // 0..kPfSoftmaxThreads-1
      const int row = tid & 127, part = tid >> 7;        // query row of the tile (= TMEM lane) and which kPfCols columns of it
      const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
      const float ninf = -INFINITY;
      const bool row_ok = row < rows;
      const int t = t_lo + row / p.G, g = row % p.G;     // query token (within the request's chunk) and head of this row
      const int limit = row_ok ? min(L, ctx + t + 1) : 0;   // this row attends to cache positions [0, limit)
      using E = Elem<QDT>;
      float* xch = reinterpret_cast<float*>(smem_raw + PfSmem::kXch);
      int xs = 0;
      // the row's other threads' values combined (maximum or sum): write, barrier over the softmax threads, read
      // (alternating slots)
      auto exchange = [&](float v, bool sum) {
        xch[(xs * kPfSplit + part) * 128 + row] = v;
        const long long t0 = kDbg ? clock64() : 0;
        named_bar_sync(1, kPfSoftmaxThreads);
        if (kDbg) acc[2] += clock64() - t0;
        float o = sum ? 0.f : -INFINITY;
#pragma unroll
        for (int i = 1; i < kPfSplit; ++i) {
          const float x = xch[(xs * kPfSplit + ((part + i) & (kPfSplit - 1))) * 128 + row];
          o = sum ? o + x : fmaxf(o, x);
        }
        xs ^= 1;
        return o;
      };
      // ---- the row's query (this thread: 64 of the 128 head dims): exact power-of-two pre-scale, hi / lo split ----
      float crow;
      {
        const uint16_t* src = reinterpret_cast<const uint16_t*>(p.q) + (int64_t)(q0 + t) * p.q_stride_t +
                              (int64_t)(hv * p.G + g) * p.q_stride_h + part * kPfCols;
        uint4 raw[kPfCols / 8];
#pragma unroll
        for (int c = 0; c < kPfCols / 8; ++c) raw[c] = row_ok ? __ldcg(reinterpret_cast<const uint4*>(src) + c) : make_uint4(0u, 0u, 0u, 0u);
        float amax = 0.f;
#pragma unroll
        for (int c = 0; c < kPfCols / 8; ++c) {
          const uint32_t w[4] = {raw[c].x, raw[c].y, raw[c].z, raw[c].w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
            amax = fmaxf(amax, fmaxf(fabsf(q_to_float<QDT>(w[i] & 0xffff)), fabsf(q_to_float<QDT>(w[i] >> 16))));
        }
        amax = fmaxf(amax, exchange(amax, false));
        const uint32_t eb = (__float_as_uint(amax) >> 23) & 0xffu;
        const bool norm = eb >= 16 && eb < 255;
        const float inv = norm ? __uint_as_float((261u - eb) << 23) : 1.0f;   // row maximum -> [128, 256)
        const float fct = norm ? __uint_as_float((eb - 7u) << 23) : 1.0f;
        const float kscale = __ldg(p.k_scale + (int64_t)kvh * p.scale_stride);
        crow = fct * p.sm_scale_log2 * kscale;
#pragma unroll
        for (int c = 0; c < kPfCols / 16; ++c) {   // 16 head dims -> one 16-byte chunk of each operand row
          uint32_t hi[8], lo[8];
#pragma unroll
          for (int hlf = 0; hlf < 2; ++hlf) {
            const uint4 v = raw[2 * c + hlf];
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
              pf_split_pair(q_to_float<QDT>(w[i] & 0xffff) * inv, q_to_float<QDT>(w[i] >> 16) * inv, hi[hlf * 4 + i], lo[hlf * 4 + i]);
          }
          const uint32_t a = q_base + pf_swz(row, part * (kPfCols / 16) + c);
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(hi[0] | (hi[1] << 16)), "r"(hi[2] | (hi[3] << 16)),
                       "r"(hi[4] | (hi[5] << 16)), "r"(hi[6] | (hi[7] << 16)) : "memory");
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a + kPfTileBytes), "r"(lo[0] | (lo[1] << 16)), "r"(lo[2] | (lo[3] << 16)),
                       "r"(lo[4] | (lo[5] << 16)), "r"(lo[6] | (lo[7] << 16)) : "memory");
        }
        pf_fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(qfull);
      }

      float m_run = ninf, l_run = 0.f;
      for (int j = 0; j < n_kv; ++j) {
        const uint32_t b = j % kPfSBufs, bpar = (j / kPfSBufs) & 1;
        const int col0 = j * kPfTile + part * kPfCols;       // cache position of this thread's first column
        const bool need_mask = col0 + kPfCols > limit;
        wait_t(sfull(b), bpar, 0);
        pf_fence_after();
        if (kDbg && (p.variant & 1)) {
          pf_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(sempty(b));
          wait_t(pempty(b), bpar ^ 1, 1);
          pf_fence_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(pfull(b));
          continue;
        }
        // ---- the half row of logits: one TMEM read, kept in registers ----
        uint32_t sr[kPfCols];
#pragma unroll
        for (int cc = 0; cc < kPfCols / 32; ++cc)
          pf_ld32(tmem + lane_base + kPfColS + b * 128 + part * kPfCols + cc * 32, *reinterpret_cast<uint32_t(*)[32]>(&sr[cc * 32]));
        pf_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(sempty(b));               // S is in registers: the QK of tile j + 2 may overwrite it
        float mx = ninf;
        if (need_mask) {
#pragma unroll
          for (int i = 0; i < kPfCols; ++i) {
            if (col0 + i >= limit) sr[i] = 0xff800000u;      // -inf (crow > 0); also replaces whatever stale K rows produced
            mx = fmaxf(mx, __uint_as_float(sr[i]));
          }
        } else {
#pragma unroll
          for (int i = 0; i < kPfCols; ++i) mx = fmaxf(mx, __uint_as_float(sr[i]));
        }
        mx *= crow;
        mx = fmaxf(mx, exchange(mx, false));
        // ---- raise the running maximum lazily (always on the first tile); O is rescaled in TMEM ----
        const bool raise = mx > m_run + kPfMargin;
        if (__any_sync(0xffffffffu, raise)) {
          const float m_new = raise ? mx : m_run;
          const float fac = (m_run == ninf) ? 0.f : fast_exp2(m_run - m_new);
          l_run *= fac;
          m_run = m_new;
          if (j > 0) {
            mbar_wait(pempty((j - 1) % kPfSBufs), ((j - 1) / kPfSBufs) & 1);   // every PV issued so far has completed
            pf_fence_after();
#pragma unroll 1
            for (int cc = 0; cc < kPfCols / 32; ++cc) {
              uint32_t o[32];
              pf_ld32(tmem + lane_base + kPfColO + part * kPfCols + cc * 32, o);
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * fac);
              pf_st32(tmem + lane_base + kPfColO + part * kPfCols + cc * 32, o);
            }
            pf_fence_before();
          }
        }
        // ---- p' = 2^(x - m + 4), row sum, hi / lo split, A-operand rows of the second GEMM ----
        wait_t(pempty(b), bpar ^ 1, 1);   // the PV that last read this P buffer has completed
        const float off = kPfShift - m_run;
#if B200_PF_PTMEM && !B200_PF_CTAS2
        uint32_t hw[kPfCols / 4], lw[kPfCols / 4];  // this thread's kPfCols tokens of the row: hi bytes, lo bytes
#endif
#pragma unroll
        for (int c = 0; c < kPfCols / 16; ++c) {    // 16 tokens -> one 16-byte chunk of each operand row
          uint32_t h4[4], l4[4];
#if B200_PF_HALF2
          // two elements per instruction from the exponent on: t = x crow + off in fp32, packed to f16x2 (|t| < 16 wherever p'
          // matters: spacing <= 2^-7, i.e. p' good to 0.3 %, the precision the 8-bit hi + lo operand keeps anyway), ex2.f16x2,
          // hi = e4m3(p'), lo = e4m3(p' - hi), and the row sum carried as f16x2 partial sums of 16 values (exact to 2^-11 each)
          uint32_t sacc = 0u;
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            uint32_t hb[2], lb[2], pp[2];
#pragma unroll
            for (int q2 = 0; q2 < 2; ++q2) {
              const float t0 = fmaf(__uint_as_float(sr[c * 16 + w * 4 + 2 * q2]), crow, off);
              const float t1 = fmaf(__uint_as_float(sr[c * 16 + w * 4 + 2 * q2 + 1]), crow, off);
              uint32_t p2, back, d2;
              uint16_t hi2, lo2;
#if B200_PF_HALF2 == 2
              // exponent in fp32 (ex2.approx.f16x2 is two scalar MUFUs plus byte permutes anyway), hi straight from the fp32
              // pair, p' packed to f16x2 only for the lo term and the row sum
              const float p0 = fast_exp2(t0), p1 = fast_exp2(t1);
              asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(hi2) : "f"(p1), "f"(p0));
              asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(p2) : "f"(p1), "f"(p0));
#else
              uint32_t t2;
              asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(t2) : "f"(t1), "f"(t0));
              asm("ex2.approx.f16x2 %0, %1;" : "=r"(p2) : "r"(t2));
              asm("cvt.rn.satfinite.e4m3x2.f16x2 %0, %1;" : "=h"(hi2) : "r"(p2));
#endif
              asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(back) : "h"(hi2));
              asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(d2) : "r"(p2), "r"(back));
              asm("cvt.rn.satfinite.e4m3x2.f16x2 %0, %1;" : "=h"(lo2) : "r"(d2));
              hb[q2] = hi2; lb[q2] = lo2; pp[q2] = p2;
            }
            h4[w] = hb[0] | (hb[1] << 16);
            l4[w] = lb[0] | (lb[1] << 16);
            uint32_t s2;
            asm("add.rn.f16x2 %0, %1, %2;" : "=r"(s2) : "r"(pp[0]), "r"(pp[1]));
            asm("add.rn.f16x2 %0, %1, %2;" : "=r"(sacc) : "r"(sacc), "r"(s2));
          }
          {
            const float2 sf = __half22float2(*reinterpret_cast<const __half2*>(&sacc));
            l_run += sf.x + sf.y;
          }
#else
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            float pv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              pv[i] = fast_exp2(fmaf(__uint_as_float(sr[c * 16 + w * 4 + i]), crow, off));
              l_run += pv[i];
            }
            pf_split4_pos(pv, h4[w], l4[w]);
          }
#endif
#if B200_PF_PTMEM && B200_PF_CTAS2
          // straight to tensor memory, 16 tokens at a time: no hi / lo staging registers (the two-CTA build has 96 per thread)
          pf_st_nowait(tmem + lane_base + kPfColP + b * 64 + part * (kPfCols / 4) + c * 4, h4);
          pf_st_nowait(tmem + lane_base + kPfColP + b * 64 + 32 + part * (kPfCols / 4) + c * 4, l4);
#elif B200_PF_PTMEM
#pragma unroll
          for (int w = 0; w < 4; ++w) { hw[c * 4 + w] = h4[w]; lw[c * 4 + w] = l4[w]; }
#else
          const uint32_t a = p_base + (2 * b) * kPfTileBytes + pf_swz(row, part * (kPfCols / 16) + c);
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(h4[0]), "r"(h4[1]), "r"(h4[2]), "r"(h4[3]) : "memory");
          asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a + kPfTileBytes), "r"(l4[0]), "r"(l4[1]), "r"(l4[2]), "r"(l4[3]) : "memory");
#endif
        }
#if B200_PF_PTMEM
        // the A operand of the second GEMM goes to tensor memory: no shared-memory write here and no shared-memory read by the MMA
#if !B200_PF_CTAS2
        pf_st_nowait(tmem + lane_base + kPfColP + b * 64 + part * (kPfCols / 4), hw);
        pf_st_nowait(tmem + lane_base + kPfColP + b * 64 + 32 + part * (kPfCols / 4), lw);
#endif
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        pf_fence_before();
#endif
        {
          // rows of the V tile past the valid count that the last 32-token MMA step still reads: stale smem or cache slots
          // beyond seq_len (possibly NaN patterns) must not meet p = 0
          const int v = min(kPfTile, kv_end - j * kPfTile);
          if (v & 31) {
            const uint32_t st = j % kPfStages;
            mbar_wait(vfull(st), (j / kPfStages) & 1);
            const int nchunk = (32 - (v & 31)) * 8;
            for (int idx = tid; idx < nchunk; idx += kPfSoftmaxThreads)
              asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(v_base + st * kPfTileBytes + (v + (idx >> 3)) * 128 + (idx & 7) * 16), "r"(0u) : "memory");
          }
        }
        pf_fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(pfull(b));
      }
      // ---- epilogue: O row / l -> out ----
      l_run += exchange(l_run, true);
      if (n_kv > 0) {
        mbar_wait(pempty((n_kv - 1) % kPfSBufs), ((n_kv - 1) / kPfSBufs) & 1);
        pf_fence_after();
      }
      const float vscale = __ldg(p.v_scale + (int64_t)kvh * p.scale_stride);
      const float onorm = (n_kv > 0 && l_run > 0.f) ? vscale / l_run : 0.f;
      typename E::type* op = reinterpret_cast<typename E::type*>(p.out) + (int64_t)(q0 + t) * p.out_stride_t +
                             (int64_t)(hv * p.G + g) * p.out_stride_h + part * kPfCols;
#pragma unroll 1
      for (int cc = 0; cc < kPfCols / 32; ++cc) {
        uint32_t o[32];
        if (n_kv > 0) {
          pf_ld32(tmem + lane_base + kPfColO + part * kPfCols + cc * 32, o);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = 0u;
        }
        if (row_ok) {
#pragma unroll
          for (int v4 = 0; v4 < 4; ++v4) {
            uint4 pk;
            uint32_t* w = reinterpret_cast<uint32_t*>(&pk);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const typename E::type e0 = E::from_float(__uint_as_float(o[v4 * 8 + 2 * i]) * onorm);
              const typename E::type e1 = E::from_float(__uint_as_float(o[v4 * 8 + 2 * i + 1]) * onorm);
              w[i] = (uint32_t)(*reinterpret_cast<const uint16_t*>(&e0)) | ((uint32_t)(*reinterpret_cast<const uint16_t*>(&e1)) << 16);
            }
            *reinterpret_cast<uint4*>(op + cc * 32 + v4 * 8) = pk;
          }
        }
      }
      pf_fence_before();
    }
  }

  if (kDbg && p.dbg != nullptr && lane == 0 && (warp == 0 || warp == kPfProducerWarp || warp == kPfMmaWarp)) {
    unsigned long long* d = p.dbg + (size_t)blockIdx.x * 16;
    const int o = warp == 0 ? 0 : (warp == kPfProducerWarp ? 4 : 8);
    for (int i = 0; i < 4; ++i) d[o + i] = (unsigned long long)acc[i];
    if (warp == 0) { d[12] = (unsigned long long)(clock64() - t_begin); d[13] = idle ? 0ull : 1ull; }
  }
  pf_fence_before();
  __syncthreads();
  if (warp == kPfMmaWarp) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kPfTmemCols) : "memory");
  }
}

// tile_prefix[r] = sum_{r' < r} ceil(q_len_r' / tpt) over requests with q_len >= min_q_len; one warp-sized CTA is plenty (num_reqs <= a few thousand)
__global__ void __launch_bounds__(256) prefill_plan_kernel(const int32_t* __restrict__ query_start_loc, int num_reqs, int tpt,
                                                           int min_q_len, int32_t* __restrict__ tile_prefix) {
  __shared__ int warp_sums[8];
  __shared__ int carry_s;
  pdl_wait();
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < num_reqs; base += 256) {
    const int i = base + tid;
    int nt = 0;
    if (i < num_reqs) {
      const int ql = query_start_loc[i + 1] - query_start_loc[i];
      nt = (ql > 0 && ql >= min_q_len) ? (ql + tpt - 1) / tpt : 0;   // shorter requests belong to the decode kernel
    }
    int v = nt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, v, off);
      if (lane >= off) v += u;
    }
    if (lane == 31) warp_sums[wid] = v;
    __syncthreads();
    int woff = 0;
    for (int k = 0; k < wid; ++k) woff += warp_sums[k];
    const int carry = carry_s;
    if (i < num_reqs) tile_prefix[i] = carry + woff + v - nt;
    __syncthreads();
    if (tid == 255) carry_s = carry + woff + v;
    __syncthreads();
  }
  if (tid == 0) tile_prefix[num_reqs] = carry_s;
}

int encode_kv_tensor_map(CUtensorMap* tm, const void* base, int kv_dtype, int64_t stride_block, int64_t stride_slot,
                         int64_t stride_head, int64_t num_blocks, int block_size, int Hkv, int D, int box_tokens = kChunkTokens);
int launch_with_pdl(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args);
unsigned long long* option_debug_buffer();
int option_debug();

}  // namespace b200attn

using namespace b200attn;

// query heads per virtual kv head (<= kMaxGroup) and the number of virtual heads a kv head is cut into
static bool prefill_group(int Hq, int Hkv, int* G, int* gsplit) {
  if (Hkv <= 0 || Hq <= 0 || Hq % Hkv) return false;
  const int Gfull = Hq / Hkv;
  int gs = 1;
  while (Gfull / gs > kMaxGroup || (Gfull % gs) != 0) {
    if (++gs > Gfull) return false;
  }
  *G = Gfull / gs;
  *gsplit = gs;
  return true;
}

extern "C" int b200attn_prefill_tile_tokens(int32_t num_q_heads, int32_t num_kv_heads) {
  int G, gs;
  if (!prefill_group(num_q_heads, num_kv_heads, &G, &gs)) return 0;
  return kPfTile / G;
}

extern "C" int b200attn_prefill_plan(const int32_t* query_start_loc, int32_t num_reqs, int32_t num_q_heads, int32_t num_kv_heads,
                                     int32_t min_query_len, int32_t* tile_prefix, void* stream) {
  if (num_reqs == 0) return B200ATTN_OK;
  if (!query_start_loc || !tile_prefix || num_reqs < 0) {
    set_error("prefill_plan: null pointer argument");
    return B200ATTN_EINVAL;
  }
  int tpt = b200attn_prefill_tile_tokens(num_q_heads, num_kv_heads);
  if (tpt <= 0) {
    set_error("prefill_plan: bad head counts Hq=%d Hkv=%d", num_q_heads, num_kv_heads);
    return B200ATTN_EINVAL;
  }
  void* args[] = {(void*)&query_start_loc, &num_reqs, &tpt, &min_query_len, (void*)&tile_prefix};
  return launch_with_pdl((const void*)prefill_plan_kernel, dim3(1), dim3(256), 0, static_cast<cudaStream_t>(stream), args);
}

extern "C" int b200attn_prefill(const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out, int64_t out_stride_t,
                                int64_t out_stride_h, const void* k_cache, const void* v_cache, int64_t stride_block,
                                int64_t stride_slot, int64_t stride_head, int64_t num_blocks, const int32_t* block_table,
                                int64_t block_table_stride, const int32_t* seq_lens, const int32_t* query_start_loc,
                                int32_t num_reqs, int32_t num_tokens, int32_t first_token, int32_t num_q_heads, int32_t num_kv_heads,
                                int32_t head_dim, int32_t block_size, int32_t q_dtype, int32_t kv_dtype, float sm_scale,
                                const float* k_scale, const float* v_scale, int32_t scale_stride, const int32_t* tile_prefix,
                                void* stream) {
  if (num_reqs == 0 || num_tokens == 0) return B200ATTN_OK;
  if (!q || !out || !k_cache || !v_cache || !block_table || !seq_lens || !query_start_loc || !tile_prefix) {
    set_error("prefill: null pointer argument");
    return B200ATTN_EINVAL;
  }
  if (head_dim != kD) {
    set_error("prefill: head_dim %d unsupported (only 128)", head_dim);
    return B200ATTN_EINVAL;
  }
  if (kv_dtype != B200ATTN_FP8_E4M3) {
    set_error("prefill: only the fp8_e4m3 KV cache has a tensor-core prefill kernel (16-bit caches go through decode rows)");
    return B200ATTN_EINVAL;
  }
  if (!k_scale || !v_scale) {
    set_error("prefill: fp8 cache needs k_scale / v_scale");
    return B200ATTN_EINVAL;
  }
  if (q_dtype != B200ATTN_BF16 && q_dtype != B200ATTN_FP16) {
    set_error("prefill: q dtype must be bf16 or fp16");
    return B200ATTN_EINVAL;
  }
  if (block_size <= 0 || block_size % kChunkTokens != 0 || block_size > 256) {
    set_error("prefill: block_size %d must be a multiple of 16 and <= 256", block_size);
    return B200ATTN_EINVAL;
  }
  const int tpt = b200attn_prefill_tile_tokens(num_q_heads, num_kv_heads);
  if (tpt <= 0) {
    set_error("prefill: bad head counts Hq=%d Hkv=%d", num_q_heads, num_kv_heads);
    return B200ATTN_EINVAL;
  }
  auto mis = [](const void* ptr, int64_t a) { return (reinterpret_cast<uintptr_t>(ptr) % a) != 0; };
  if (mis(q, 16) || (q_stride_t % 8) || (q_stride_h % 8) || mis(out, 16) || (out_stride_t % 8) || (out_stride_h % 8)) {
    set_error("prefill: q / out must be 16-byte aligned with strides that are multiples of 8 elements");
    return B200ATTN_EINVAL;
  }
  int G = 1, gsplit = 1;
  prefill_group(num_q_heads, num_kv_heads, &G, &gsplit);
  CUtensorMap tk, tv;
  int rc = encode_kv_tensor_map(&tk, k_cache, kv_dtype, stride_block, stride_slot, stride_head, num_blocks, block_size, num_kv_heads, kD);
  if (rc) return rc;
  rc = encode_kv_tensor_map(&tv, v_cache, kv_dtype, stride_block, stride_slot, stride_head, num_blocks, block_size, num_kv_heads, kD);
  if (rc) return rc;
  PrefillParams p;
  p.q = q; p.q_stride_t = q_stride_t; p.q_stride_h = q_stride_h;
  p.out = out; p.out_stride_t = out_stride_t; p.out_stride_h = out_stride_h;
  p.block_table = block_table; p.bt_stride = block_table_stride;
  p.seq_lens = seq_lens; p.query_start_loc = query_start_loc; p.tile_prefix = tile_prefix;
  p.k_scale = k_scale; p.v_scale = v_scale; p.scale_stride = scale_stride;
  p.sm_scale_log2 = sm_scale * 1.4426950408889634f;
  p.num_reqs = num_reqs; p.Hq = num_q_heads; p.Hv = num_kv_heads * gsplit; p.G = G; p.gsplit = gsplit;
  p.cpb = block_size / kChunkTokens; p.tpt = tpt; p.tok0 = first_token;
  // upper bound on the number of query tiles that needs no device read: every request wastes less than one tile
  const long long max_tiles = (long long)(num_tokens + tpt - 1) / tpt + num_reqs;
  const long long grid = max_tiles * p.Hv;
  if (grid > 2147483647LL) {
    set_error("prefill: grid too large");
    return B200ATTN_EINVAL;
  }
  p.dbg = option_debug_buffer();
  p.variant = option_debug();
  const bool dbg = p.dbg != nullptr;
  const void* fn = q_dtype == B200ATTN_BF16
                       ? (dbg ? (const void*)prefill_tc_kernel<B200ATTN_BF16, true> : (const void*)prefill_tc_kernel<B200ATTN_BF16, false>)
                       : (dbg ? (const void*)prefill_tc_kernel<B200ATTN_FP16, true> : (const void*)prefill_tc_kernel<B200ATTN_FP16, false>);
  static bool configured[64][4] = {};
  int dev = 0;
  if (check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return B200ATTN_ECUDA;
  const int qi = (q_dtype == B200ATTN_BF16 ? 0 : 1) + (dbg ? 2 : 0);
  if (dev < 0 || dev >= 64 || !configured[dev][qi]) {
    if (check_cuda(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, PfSmem::kTotal), "cudaFuncSetAttribute(prefill smem)"))
      return B200ATTN_ECUDA;
    if (kPfCtasPerSm > 1 &&
        check_cuda(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared),
                   "cudaFuncSetAttribute(prefill carveout)"))
      return B200ATTN_ECUDA;
    if (dev >= 0 && dev < 64) configured[dev][qi] = true;
  }
  void* args[] = {(void*)&tk, (void*)&tv, (void*)&p};
  return launch_with_pdl(fn, dim3((unsigned)grid), dim3(kPfThreads), PfSmem::kTotal, static_cast<cudaStream_t>(stream), args);
}
