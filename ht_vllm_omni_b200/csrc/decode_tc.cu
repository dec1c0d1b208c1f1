// decode_tc.cu — Kernel 2, tcgen05 flavour: persistent split-KV paged flash-decoding over an FP8 (e4m3) KV cache with
// both GEMMs on the 5th-generation tensor cores (tcgen05.mma kind::f8f6f4, accumulators in TMEM) — no FP8->F16
// conversion of K/V on the CUDA cores at all.
//
// Reference semantics are those of decode.cu (kernel_unified_attention, $VLLM/v1/attention/ops/triton_unified_attention.py:
// 440-643, FlashInfer scale convention $VLLM/v1/attention/backends/flashinfer.py:1396-1404); the work plan, the partial
// (o, m, l) workspace format and the in-kernel merge are shared with decode.cu, so both kernels are interchangeable
// behind b200attn_decode.
//
// Per CTA (192 threads, 2 CTAs / SM):
//   warp 4   producer: TMA tensor loads (UTMALDG, 128-B swizzle) of 16-token x 128-byte boxes out of the paged cache
//            into two rings of 128-token tiles (K ring, V ring; 16 KiB per tile), bulk copy of the query rows.
//   warp 5   MMA issuer (one lane): per 128-token tile
//              S^T[128 tok x 16]  = K_tile[128 x 128 d] . Qop^T      4 x tcgen05.mma M128 N16 K32  (A, B K-major)
//              O^T[128 d  x 16]  += V_tile^T[128 d x 128 tok] . P^T  <=4 x tcgen05.mma M128 N16 K32 (A, B MN-major)
//            The K/V tiles are consumed exactly as TMA wrote them (the V tile is simply described MN-major).
//   warps 0-3 softmax: thread = one token row of S^T (tcgen05.ld 32x32b), online softmax in the log2 domain against a
//            CTA-wide lazily updated reference maximum, P written back to smem as e4m3 for the second GEMM.
// FP8 operands without losing the 16-bit activations: the bf16 query is split EXACTLY into two e4m3 terms
// q*2^k = hi + lo (8 significant bits = 4 + 4) that occupy N columns h and 8+h of the same MMA, and the probabilities are
// split the same way (p*16 = hi + lo/16), so the column pairs are recombined in fp32 after the MMA.
#include <cstdio>
#include <cstdlib>

#include "decode_common.cuh"

namespace b200attn {

constexpr int kTcThreads = 224;               // 4 softmax warps + producer + MMA issuer + publish helper
constexpr int kTcTile = 128;                 // tokens per MMA tile (= 2 plan tiles of 64)
#ifndef B200_TC_STAGES
#define B200_TC_STAGES 3
#endif
constexpr int kTcStages = B200_TC_STAGES;    // depth of the K ring and of the V ring (2: three CTAs fit per SM, see below)
constexpr int kTcTileBytes = kTcTile * kD;   // 16 KiB (fp8)
constexpr int kTcBoxBytes = kChunkTokens * kD;
constexpr int kTcQRawBytes = 8 * kD * 2;     // 8 query rows of 16-bit elements
constexpr int kTcQopBytes = 16 * kD;         // 16 rows (8 hi + 8 lo) of e4m3
constexpr int kTcPBytes = kTcTile * 16;      // [token][8 hi + 8 lo] e4m3
constexpr int kTcScratchFloats = 176;         // 156..173: publish request ring (9 x 8 bytes)
constexpr int kTcRecvBytes = 8 * kD * 4 + 8 * 2 * 4;   // one (o[8][128], (m, l)[8]) partial pushed by the next CTA of the cluster
constexpr int kTcTmemCols = 64;              // S0 [0,16) S1 [16,32) O [32,48)
constexpr float kTcMargin = 4.0f;            // the reference maximum is raised when a logit exceeds it by 2^4
constexpr float kTcPShift = 4.0f;            // p' = p * 2^4, so p' <= 2^8 < 448 (e4m3 max)

struct TcSmem {
  static constexpr int kK = 0;
  static constexpr int kV = kK + kTcStages * kTcTileBytes;
  static constexpr int kQRaw = kV + kTcStages * kTcTileBytes;
  static constexpr int kQop = kQRaw + kTcQRawBytes;
  static constexpr int kP = kQop + 2 * kTcQopBytes;
  static constexpr int kScratch = kP + 2 * kTcPBytes;
  static constexpr int kPlan = kScratch + kTcScratchFloats * 4;     // the CTA's plan entry (first 32 words)
  static constexpr int kRecv = kPlan + 128;                        // partial received over distributed shared memory
  static constexpr int kBars = kRecv + kTcRecvBytes;
  static constexpr int kNumBars = 4 * kTcStages + 17;             // + 1: "partial received" (armed by the cluster peer)
  static constexpr int kTmemPtr = kBars + kNumBars * 8;
  static constexpr int kTotal = kTmemPtr + 16;
  // The grid is 2 CTAs per SM, but a CTA is sized so that THREE fit (75 KiB, <= 112 registers): with programmatic
  // dependent launch the first CTA of the next kernel on this stream becomes resident beside the two running ones, runs
  // its prologue and prefetches its first K/V tiles, so launch latency, metadata fetch and the merge tail of one kernel
  // overlap the streaming of its neighbours.
  static constexpr int kCtasPerSm = kTcStages <= 2 ? 3 : 2;
  static_assert(kTotal <= (kCtasPerSm == 3 ? 75 : 113) * 1024, "smem budget exceeded");
};

// ---------------------------------------------------------------------------------------------
// tcgen05 wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void tc_mma_f8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrives (count 1) on an mbarrier of this CTA once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// Shared-memory matrix descriptor (SM100 "version 1"): start address, leading / stride byte offsets (all >> 4), layout.
constexpr uint64_t kLayoutNone = 0, kLayoutSw128 = 2;
__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint64_t layout) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46) | (layout << 61);
}
// Instruction descriptor, kind::f8f6f4, e4m3 x e4m3 -> f32, M = 128, N = 16.
constexpr uint32_t kIdescBase = (1u << 4) | ((16u >> 3) << 17) | ((128u >> 4) << 24);
constexpr uint32_t kIdescQK = kIdescBase;                              // A (K tile) K-major, B (Qop) K-major
constexpr uint32_t kIdescPV = kIdescBase | (1u << 15) | (1u << 16);   // A (V tile) MN-major, B (P^T) MN-major

__device__ __forceinline__ uint32_t f32x2_to_e4m3x2(float lo, float hi) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float2 e4m3x2_to_f32x2(uint32_t two_bytes) {
  uint32_t h2;
  asm("{\n\t.reg .b16 t;\n\tcvt.u16.u32 t, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, t;\n\t}" : "=r"(h2) : "r"(two_bytes));
  return __half22float2(*reinterpret_cast<__half2*>(&h2));
}
// x = hi + lo / lo_scale with hi, lo e4m3: returns the two bytes for a pair of values
__device__ __forceinline__ void split_pair(float x0, float x1, float lo_scale, uint32_t& hi2, uint32_t& lo2) {
  hi2 = f32x2_to_e4m3x2(x0, x1);
  const float2 h = e4m3x2_to_f32x2(hi2);
  lo2 = f32x2_to_e4m3x2((x0 - h.x) * lo_scale, (x1 - h.y) * lo_scale);
}

// ---------------------------------------------------------------------------------------------
// The work of a CTA: its contiguous range of 64-token plan tiles, walked segment by segment ((sequence, kv head) runs)
// and, inside a segment, MMA tile by MMA tile.  Every warp role walks the same sequence on its own.
// ---------------------------------------------------------------------------------------------
struct TcCursor {
  int s, hv, tile, remaining;       // walk state (plan units)
  int seg_s, seg_hv, seg_tile0, seg_n, seg_L, seg_end;
  int u, m;                         // MMA tile index inside the segment, MMA tiles in the segment
  int sg;                           // segment ordinal in this CTA
  int L0;                           // seq_len of the first segment's row (from the plan entry: no dependent load)
  __device__ __forceinline__ void init(const int4& cs, int L_first) { s = cs.x; hv = cs.y; tile = cs.z; remaining = cs.w; u = 0; m = 0; sg = -1; L0 = L_first; }
  __device__ __forceinline__ bool next_segment(const DecodeParams& p) {
    if (remaining <= 0 || s >= p.B) return false;
    const int L = sg < 0 ? L0 : max(__ldg(p.seq_lens + s), 0);
    const int nt = (L + kTileTokens - 1) / kTileTokens;
    const int n = min(remaining, nt - tile);
    seg_s = s; seg_hv = hv; seg_tile0 = tile; seg_n = n; seg_L = L;
    seg_end = min((tile + n) * kTileTokens, L);
    m = (n + 1) >> 1; u = -1; ++sg;
    remaining -= n;
    tile += n;
    if (tile == nt) {
      tile = 0;
      if (++hv == p.Hv) {
        hv = 0;
        do { ++s; } while (s < p.B && remaining > 0 && __ldg(p.seq_lens + s) <= 0);
      }
    }
    return true;
  }
  // advance to the next MMA tile; false when the CTA's range is exhausted
  __device__ __forceinline__ bool next_tile(const DecodeParams& p) {
    if (u + 1 < m) { ++u; return true; }
    if (!next_segment(p)) return false;
    u = 0;
    return true;
  }
  __device__ __forceinline__ int tok0() const { return seg_tile0 * kTileTokens + u * kTcTile; }
  __device__ __forceinline__ int valid() const { return min(kTcTile, seg_end - tok0()); }
  __device__ __forceinline__ bool first() const { return u == 0; }
  __device__ __forceinline__ bool last() const { return u == m - 1; }
};

// Two CTAs per SM: 128 registers per thread.  (144 would remove the last 24-byte spill, but a CTA's 7 warps land 2-2-2-1 on
// the four scheduler partitions and 4 warps x 144 x 32 registers exceed one partition's 16 K registers: measured, the grid
// then runs one CTA per SM, 30.5 -> 51.8 us on c2.)
template <int QDT, bool kDbg>
__global__ void __launch_bounds__(kTcThreads, TcSmem::kCtasPerSm)
decode_tc_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v, const DecodeParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];   // 128-byte-swizzled tiles need 1024-byte alignment
  const uint32_t base = smem_u32(smem_raw);
  if (base & 1023u) __trap();
  uint8_t* gen = smem_raw;
  const uint32_t k_base = base + TcSmem::kK, v_base = base + TcSmem::kV;
  const uint32_t qraw_base = base + TcSmem::kQRaw, qop_base = base + TcSmem::kQop, p_base = base + TcSmem::kP;
  float* scratch = reinterpret_cast<float*>(gen + TcSmem::kScratch);
  volatile uint32_t* tmem_ptr_smem = reinterpret_cast<volatile uint32_t*>(gen + TcSmem::kTmemPtr);
  const uint32_t bar0 = base + TcSmem::kBars;
  auto kfull = [&](uint32_t i) { return bar0 + i * 8; };
  auto kempty = [&](uint32_t i) { return bar0 + (kTcStages + i) * 8; };
  auto vfull = [&](uint32_t i) { return bar0 + (2 * kTcStages + i) * 8; };
  auto vempty = [&](uint32_t i) { return bar0 + (3 * kTcStages + i) * 8; };
  const uint32_t bar1 = bar0 + 4 * kTcStages * 8;
  auto qopfull = [&](uint32_t i) { return bar1 + (4 + i) * 8; };
  auto qopempty = [&](uint32_t i) { return bar1 + (6 + i) * 8; };
  auto sfull = [&](uint32_t i) { return bar1 + (8 + i) * 8; };
  auto sempty = [&](uint32_t i) { return bar1 + (10 + i) * 8; };
  auto pfull = [&](uint32_t i) { return bar1 + (12 + i) * 8; };
  auto pempty = [&](uint32_t i) { return bar1 + (14 + i) * 8; };
  const uint32_t recv_bar = bar1 + 16 * 8;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (kDbg && threadIdx.x == 0) {
    dbg_stamp(p, 0);
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if (p.dbg != nullptr && p.cluster > 1) p.dbg[(size_t)blockIdx.x * kDbgSlots + 6] = smid;   // (slot 6 = publish time without clusters)
  }

  // ---------------------------------------------------------------------------------------------------------------------
  // Prologue.  Every role requests the CTA's plan entry itself, first thing (one 16-byte-aligned broadcast load per warp,
  // all in flight together), so nobody waits for anybody's load.  The producer warp is on the critical path (entry -> first
  // TMA): it initialises only the ring barriers it arms itself and ARRIVES on the set-up barrier without waiting; the other
  // six warps (barrier init, TMEM allocation) sync on it.
  // ---------------------------------------------------------------------------------------------------------------------
  const int32_t* ent = p.plan + kPlanHeaderWords + (size_t)blockIdx.x * kPlanEntryWords;
  int e1 = 0;   // producer: block id `lane` of the CTA's first 32 chunks
  if (warp == 4) e1 = __ldg(ent + kPlanEntryIds + lane);
  const int4 cs = __ldg(reinterpret_cast<const int4*>(ent));        // seq, kv head, first tile, tile count
  const int4 cx = __ldg(reinterpret_cast<const int4*>(ent) + 1);    // seq_len, block-table row, tile prefix, tiles per CTA
  const int4 chk = __ldg(reinterpret_cast<const int4*>(ent) + 2);   // grid, batch the plan was cut for
  constexpr int kSetupBar = 2, kSetupThreads = kTcThreads;
  if (warp == 4) {
    if (lane == 0) {
      for (int i = 0; i < kTcStages; ++i) {
        mbar_init(kfull(i), 1); mbar_init(kempty(i), 1);
        mbar_init(vfull(i), 1); mbar_init(vempty(i), 1);
      }
      mbar_fence_init();
      prefetch_tensormap(&tm_k);
      prefetch_tensormap(&tm_v);
    }
    __syncwarp();
    named_bar_arrive(kSetupBar, kSetupThreads);   // "the ring barriers exist": does not wait for the other warps
  } else {
    if (threadIdx.x == 0) {
      for (int i = 0; i < 2; ++i) {
        mbar_init(qopfull(i), 4); mbar_init(qopempty(i), 1);
        mbar_init(sfull(i), 1); mbar_init(sempty(i), 4);
        mbar_init(pfull(i), 4); mbar_init(pempty(i), 1);
      }
      mbar_init(recv_bar, 1);
      *reinterpret_cast<volatile unsigned long long*>(scratch + 156) = 0ull;
      mbar_fence_init();
    }
    if (warp == 5) {  // TMEM: one warp allocates (and frees at the end)
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(base + TcSmem::kTmemPtr), "r"(kTcTmemCols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    named_bar_sync(kSetupBar, kSetupThreads);
    tc_fence_after();
  }
  // "my barriers are initialised"; waited for only by CTAs that push to a peer.  The producer warp arrives AFTER it has issued
  // its first tile's TMA loads: a cluster-scope release is a fence, and nothing the peer waits for is produced by this warp.
  if (p.cluster > 1 && warp != 4) cluster_arrive_release();
  pdl_launch_dependents();
  if (chk.x != (int)gridDim.x || chk.y != p.B) __trap();   // plan / launch mismatch
  const int L_first = cx.x;
  const bool idle = cs.x < 0 || cs.w <= 0;   // fewer tiles than CTAs

  if (warp == 4) {
    // =========================================== producer ===========================================
    bool cl_arrived = false;
    if (!idle) {
    const uint64_t pol = policy_evict_first();
    TcCursor cur; cur.init(cs, L_first);
    const int row_first = cx.y;
    bool waited = false;
    int early_left = p.fresh >= 0 ? kTcStages : 0;
    uint32_t it = 0;       // MMA tile ordinal in this CTA
    int blk = 0, blk_next = 0;
    const int32_t* bt = nullptr;
    int kvh = 0;
    bool nt_loaded = false, nt_stored = false;   // fused KV write: the row's new token is in registers / in the cache
    int64_t nt_slot = -1;
    uint4 nt_raw = make_uint4(0u, 0u, 0u, 0u);
    float nt_scale = 1.f;
    if (kDbg && lane == 0) dbg_stamp(p, 8);
    while (cur.next_tile(p)) {
      const int tok0 = cur.tok0();
      if (cur.first()) {
        const int row = cur.sg == 0 ? row_first : (p.row_map ? __ldg(p.row_map + cur.seg_s) : cur.seg_s);
        bt = p.block_table + (int64_t)row * p.bt_stride;
        kvh = (int)p.d_gsplit.div((uint32_t)cur.seg_hv);
        if (cur.sg == 0) {
          blk = e1;   // the plan entry carries the first 32 block ids: no dependent load before the first TMA
        } else {
          const int ci = (tok0 >> 4) + lane;
          blk = (ci * kChunkTokens < cur.seg_L) ? __ldg(bt + p.d_cpb.div((uint32_t)ci)) : 0;
        }
      } else if ((cur.u & 3) == 0) {
        blk = blk_next;
      }
      if ((cur.u & 3) == 1 && (cur.u | 3) + 1 < cur.m) {  // ids of MMA tiles [4k+4, 4k+8): requested three tiles ahead of their use
        const int ci = (tok0 >> 4) + 24 + lane;
        blk_next = (ci * kChunkTokens < cur.seg_L) ? __ldg(bt + p.d_cpb.div((uint32_t)ci)) : 0;
      }
      // K/V tiles that lie entirely below L - fresh cannot have been touched by the kernel before us: they are requested
      // BEFORE griddepcontrol.wait (programmatic dependent launch), at most one ring fill
      const bool early_ok = !waited && early_left > 0 && tok0 + kTcTile <= cur.seg_L - p.fresh;
      if (!early_ok && !waited) { pdl_wait(); waited = true; }
      if (early_ok) --early_left;
      // ---- fused KV write: the piece that ends its row owns the row's new token.  Lanes 0-15 carry 8 key elements each,
      // lanes 16-31 the value's.  The loads are requested at the first tile after the wait above (key / value come from the
      // kernel before us), the bytes go to the cache one tile later -- mid-stream, where the ring hides the producer's stall --
      // or at the latest right before the TMA of the tile that holds position L - 1 reads them back
      // (generic-proxy store -> fence.proxy.async -> async-proxy load, same warp).
      if (p.new_key != nullptr && cur.seg_end == cur.seg_L) {
        if (cur.first()) { nt_loaded = false; nt_stored = false; }
        bool just_loaded = false;
        if (!nt_loaded && waited) {
          const int hl = lane & 15;
          const uint16_t* src = lane < 16
              ? reinterpret_cast<const uint16_t*>(p.new_key) + (int64_t)cur.seg_s * p.nk_stride_t + (int64_t)kvh * p.nk_stride_h + hl * 8
              : reinterpret_cast<const uint16_t*>(p.new_value) + (int64_t)cur.seg_s * p.nv_stride_t + (int64_t)kvh * p.nv_stride_h + hl * 8;
          nt_slot = __ldg(p.slot_mapping + cur.seg_s);
          nt_raw = ldg_nc_v4(src);
          nt_scale = __ldg((lane < 16 ? p.k_scale : p.v_scale) + (int64_t)kvh * p.scale_stride);
          nt_loaded = true;
          just_loaded = true;
        }
        if (nt_loaded && !nt_stored && (!just_loaded || cur.last())) {
          if (nt_slot >= 0) {
            float f[8];
            f[0] = q_to_float<QDT>(nt_raw.x & 0xffff); f[1] = q_to_float<QDT>(nt_raw.x >> 16);
            f[2] = q_to_float<QDT>(nt_raw.y & 0xffff); f[3] = q_to_float<QDT>(nt_raw.y >> 16);
            f[4] = q_to_float<QDT>(nt_raw.z & 0xffff); f[5] = q_to_float<QDT>(nt_raw.z >> 16);
            f[6] = q_to_float<QDT>(nt_raw.w & 0xffff); f[7] = q_to_float<QDT>(nt_raw.w >> 16);
            // x / scale in fp32 with IEEE rounding, then round-to-nearest-even + saturate-to-finite: kv_write.cu's arithmetic
            const uint32_t w0 = f32x2_to_e4m3x2(__fdiv_rn(f[0], nt_scale), __fdiv_rn(f[1], nt_scale)) |
                                (f32x2_to_e4m3x2(__fdiv_rn(f[2], nt_scale), __fdiv_rn(f[3], nt_scale)) << 16);
            const uint32_t w1 = f32x2_to_e4m3x2(__fdiv_rn(f[4], nt_scale), __fdiv_rn(f[5], nt_scale)) |
                                (f32x2_to_e4m3x2(__fdiv_rn(f[6], nt_scale), __fdiv_rn(f[7], nt_scale)) << 16);
            const int64_t nblk = nt_slot / p.block_size, noff = nt_slot - nblk * p.block_size;
            uint8_t* dst = (lane < 16 ? p.kc_base : p.vc_base) + nblk * p.c_stride_block + noff * p.c_stride_slot +
                           (int64_t)kvh * p.c_stride_head + (lane & 15) * 8;
            asm volatile("st.global.v2.u32 [%0], {%1, %2};" ::"l"(dst), "r"(w0), "r"(w1) : "memory");
          }
          asm volatile("fence.proxy.async.global;" ::: "memory");
          __syncwarp();
          asm volatile("fence.proxy.async.global;" ::: "memory");
          nt_stored = true;
        }
      }
      const int per = p.box_chunks;   // 16-token chunks per TMA box: whole boxes are loaded (the cache block is allocated whole)
      const int nbox = ((cur.valid() + kChunkTokens - 1) / kChunkTokens + per - 1) & ~(per - 1);
      const uint32_t st = it % kTcStages, par = ((it / kTcStages) & 1) ^ 1;
      const int c = lane & 7;
      const bool issue = c < nbox && (c & (per - 1)) == 0;
      const int b = __shfl_sync(0xffffffffu, blk, ((cur.u & 3) << 3) + c);
      const int ci = (tok0 >> 4) + c;
      const int row0 = (int)((uint32_t)ci - p.d_cpb.div((uint32_t)ci) * (uint32_t)p.cpb) * kChunkTokens;
      mbar_wait(kempty(st), par);
      if (kDbg && it == 0 && lane == 0) dbg_stamp(p, 14);
      if (lane == 0) mbar_arrive_expect_tx(kfull(st), (uint32_t)nbox * kTcBoxBytes);
      __syncwarp();
      if (lane < 8 && issue) tma_load_4d(k_base + st * kTcTileBytes + c * kTcBoxBytes, &tm_k, kfull(st), 0, row0, kvh, b, pol);
      if (kDbg && it == 0 && lane == 0) dbg_stamp(p, 15);
      mbar_wait(vempty(st), par);
      if (lane == 0) mbar_arrive_expect_tx(vfull(st), (uint32_t)nbox * kTcBoxBytes);
      __syncwarp();
      if (lane >= 8 && lane < 16 && issue) tma_load_4d(v_base + st * kTcTileBytes + c * kTcBoxBytes, &tm_v, vfull(st), 0, row0, kvh, b, pol);
      if (kDbg && it == 0 && lane == 0) dbg_stamp(p, 2);
      if (it == 0 && p.cluster > 1) { cluster_arrive_release(); cl_arrived = true; }
      ++it;
    }
    if (!waited) pdl_wait();
    }
    if (p.cluster > 1 && !cl_arrived) cluster_arrive_release();   // idle CTA / empty range: still owed to the cluster barrier
  } else {
    const uint32_t tmem = *tmem_ptr_smem;
    if (kDbg && threadIdx.x == 0) dbg_stamp(p, 1);
    // rows without any token (padding of a CUDA-graph batch, or a sequence-split rank that holds no position of a short
    // context): the partial is (o = 0, lse = -inf), written here so a consumer that combines partials by LSE never meets a
    // stale buffer.  Only when the caller asked for the LSE; idle CTAs take part.
    if (p.lse_out != nullptr && warp < 4) {
      using E = Elem<QDT>;
      pdl_wait();
      for (int s = blockIdx.x; s < p.B; s += gridDim.x) {
        if (__ldg(p.seq_lens + s) > 0) continue;
        for (int i = threadIdx.x; i < p.Hq * kD; i += 128) {
          const int hq = i / kD, d = i % kD;
          reinterpret_cast<typename E::type*>(p.out)[(int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h + d] = E::from_float(0.f);
        }
        for (int hq = threadIdx.x; hq < p.Hq; hq += 128) p.lse_out[(int64_t)s * p.Hq + hq] = -INFINITY;
      }
    }
    if (idle) {
      // nothing to do: fall through to the teardown (TMEM is given back there)
    } else if (warp == 5) {
      // =========================================== MMA issuer ===========================================
      pdl_wait();
      TcCursor cq, cp;
      cq.init(cs, L_first); cp.init(cs, L_first);
      uint32_t iq = 0, ip = 0;  // tile ordinals of the QK and PV streams
      auto issue_qk = [&]() {
        const uint32_t slot = cq.sg & 1, b = iq & 1, st = iq % kTcStages;
        if (cq.first()) mbar_wait(qopfull(slot), (cq.sg >> 1) & 1);
        mbar_wait(kfull(st), (iq / kTcStages) & 1);
        mbar_wait(sempty(b), ((iq >> 1) & 1) ^ 1);
        tc_fence_after();
        if (lane == 0) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t da = tc_desc(k_base + st * kTcTileBytes + k * 32, 16, 1024, kLayoutSw128);
            const uint64_t db = tc_desc(qop_base + slot * kTcQopBytes + k * 256, 128, 1024, kLayoutNone);
            tc_mma_f8(tmem + b * 16, da, db, kIdescQK, k > 0);
          }
          tc_commit(kempty(st));
          tc_commit(sfull(b));
          if (cq.last()) tc_commit(qopempty(slot));
          if (kDbg && iq == 0) dbg_stamp(p, 10);
        }
        __syncwarp();
        ++iq;
      };
      auto issue_pv = [&]() {
        const uint32_t b = ip & 1, st = ip % kTcStages;
        mbar_wait(vfull(st), (ip / kTcStages) & 1);
        mbar_wait(pfull(b), (ip >> 1) & 1);
        tc_fence_after();
        if (lane == 0) {
          const int nk = (cp.valid() + 31) >> 5;
          for (int k = 0; k < nk; ++k) {
            const uint64_t da = tc_desc(v_base + st * kTcTileBytes + k * 4096, 0, 1024, kLayoutSw128);
            const uint64_t db = tc_desc(p_base + b * kTcPBytes + k * 512, 128, 0, kLayoutNone);
            tc_mma_f8(tmem + 32, da, db, kIdescPV, (k > 0 || !cp.first()) ? 1u : 0u);
          }
          tc_commit(vempty(st));
          tc_commit(pempty(b));
        }
        __syncwarp();
        ++ip;
      };
      if (cq.next_tile(p)) issue_qk();
      while (cp.next_tile(p)) {
        // QK of the next tile runs while the softmax warps work on this one -- inside a segment.  Across a segment boundary
        // the order flips: the next segment's Qop is only written after the softmax warps have seen this segment's last PV.
        const bool have = cq.next_tile(p);
        if (have && !cq.first()) issue_qk();
        issue_pv();
        if (have && cq.first()) issue_qk();
      }
    } else if (warp == 6) {
      // =========================================== publish helper ===========================================
      // Takes the release fence + counter bump of the in-kernel merge off the softmax warps: a memory round trip costs
      // microseconds while the CTA's own KV stream saturates HBM, and the softmax warps would stall the whole pipeline on it.
      volatile unsigned long long* ring = reinterpret_cast<volatile unsigned long long*>(scratch + 156);
      for (uint32_t k = 0;; ++k) {
        while (pub_ring_count(ring) <= k) __nanosleep(64);
        const unsigned long long req = ring[1 + (k & 7)];
        if (req == 0) break;   // the softmax warps are done
        if (lane == 0) merge_publish(reinterpret_cast<int*>(req));
        __syncwarp();
      }
    } else {
      // =========================================== softmax warps ===========================================
      pdl_wait();
      const int w = warp;
      const int tid = threadIdx.x;  // 0..127 = row of the 128-row MMA tiles (token for S^T, head dim for O^T)
      const uint32_t lane_base = (uint32_t)(w * 32) << 16;   // TMEM lanes this warp may touch
      float* cs_sm = scratch;                                  // [2][8]  logit scale per head (by Q slot)
      float* wmax = scratch + 16;                              // [4][8]  per-warp tile maxima
      volatile int* flags = reinterpret_cast<volatile int*>(scratch + 48);  // [2][4] "a logit left the window" votes
      float* red_l = scratch + 56;                             // [4][8]  per-warp row sums
      float* mrg = scratch + 88;                               // merge scratch (64 floats)
      volatile int* hand = reinterpret_cast<volatile int*>(scratch + 152);
      volatile unsigned long long* pub_req = reinterpret_cast<volatile unsigned long long*>(scratch + 156);   // request ring [1 + 8]
      const float2* recv_ml = reinterpret_cast<const float2*>(gen + TcSmem::kRecv + 8 * kD * 4);
      const int tpc = cx.w;
      MergeTicket tk = {};
      const float ninf = -INFINITY;
      TcCursor cur; cur.init(cs, L_first);
      uint32_t it = 0;
      float m_ref[8], l_part[8], crow[8];
      bool stamped = false, cluster_ready = false, seg_recv = false, seg_push = false;
      int seg_mode = 0;
      float seg_vscale = 1.f;

      // ---------------- per segment: split the 16-bit query rows into e4m3 hi / lo operand rows (Qop) ----------------
      // Runs one segment AHEAD of the tile loop (cursor `qc`): the next segment's operand is published before this segment's
      // epilogue, so the MMA warp can issue the next QK right behind the last PV.
      TcCursor qc; qc.init(cs, L_first);
      const int qh = tid >> 4, qpart = tid & 15;   // 16 threads per head, 8 consecutive head dims each
      // the query rows of segment qc: every thread copies ITS 16 bytes global -> smem asynchronously (no registers held, no
      // other thread involved), requested one tile ahead of the split below so the latency hides behind the tile's softmax
      auto request_q = [&]() {
        if (qh < p.G) {
          const uint16_t* src = reinterpret_cast<const uint16_t*>(p.q) + (int64_t)qc.seg_s * p.q_stride_t +
                                (int64_t)(qc.seg_hv * p.G + qh) * p.q_stride_h + qpart * 8;
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(qraw_base + tid * 16), "l"(src) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
      };
      auto prep_q = [&]() {
        const uint32_t slot = qc.sg & 1;
        const int kvh = (int)p.d_gsplit.div((uint32_t)qc.seg_hv);
        const int h = qh, part = qpart;
        float kscale = 0.f;
        if (part == 0) kscale = __ldg(p.k_scale + (int64_t)kvh * p.scale_stride);   // in flight while the query arrives
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        uint4 raw = make_uint4(0u, 0u, 0u, 0u);
        if (h < p.G) raw = lds128(qraw_base + tid * 16);   // the 16 bytes this thread copied itself
        float f[8];
        f[0] = q_to_float<QDT>(raw.x & 0xffff); f[1] = q_to_float<QDT>(raw.x >> 16);
        f[2] = q_to_float<QDT>(raw.y & 0xffff); f[3] = q_to_float<QDT>(raw.y >> 16);
        f[4] = q_to_float<QDT>(raw.z & 0xffff); f[5] = q_to_float<QDT>(raw.z >> 16);
        f[6] = q_to_float<QDT>(raw.w & 0xffff); f[7] = q_to_float<QDT>(raw.w >> 16);
        float amax = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(f[i]));
#pragma unroll
        for (int off = 1; off < 16; off <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, off));
        // exact power-of-two pre-scale that puts the row maximum in [128, 256); undone on the logits
        const uint32_t eb = (__float_as_uint(amax) >> 23) & 0xffu;
        const bool norm = eb >= 16 && eb < 255;
        const float inv = norm ? __uint_as_float((261u - eb) << 23) : 1.0f;
        const float fct = norm ? __uint_as_float((eb - 7u) << 23) : 1.0f;
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split_pair(f[2 * i] * inv, f[2 * i + 1] * inv, 16.0f, hi[i], lo[i]);
        mbar_wait(qopempty(slot), ((qc.sg >> 1) & 1) ^ 1);   // the MMAs of segment sg-2 are done with this slot
        // K-major, no swizzle: 8x16-byte core matrices; byte (row r, d) at (r/8)*1024 + (d/16)*128 + (r%8)*16 + d%16
        const uint32_t qa = qop_base + slot * kTcQopBytes + (part >> 1) * 128 + h * 16 + (part & 1) * 8;
        asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(qa), "r"(hi[0] | (hi[1] << 16)), "r"(hi[2] | (hi[3] << 16)) : "memory");
        asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(qa + 1024), "r"(lo[0] | (lo[1] << 16)), "r"(lo[2] | (lo[3] << 16)) : "memory");
        if (part == 0) cs_sm[slot * 8 + h] = (h < p.G) ? fct * p.sm_scale_log2 * kscale : 0.f;
        fence_proxy_async_smem();
        named_bar_sync(1, 128);
        if (lane == 0) mbar_arrive(qopfull(slot));
      };
      bool q_pending = false;   // a request for segment qc is in flight
      if (qc.next_segment(p)) { request_q(); prep_q(); }
      if (kDbg && tid == 0) dbg_stamp(p, 9);

      while (cur.next_tile(p)) {
        if (cur.first()) {
          const uint32_t slot = cur.sg & 1;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            crow[i] = cs_sm[slot * 8 + i];
            m_ref[i] = ninf;
            l_part[i] = 0.f;
          }
          // how this segment ends is known now: decide it here, off the tail's critical path
          //   0 whole sequence -> final output     1 partial to the workspace (merge kernel follows)
          //   2 in-kernel merge through L2 (ticket / publish)     3 fold along the cluster chain (distributed shared memory)
          seg_vscale = __ldg(p.v_scale + (int64_t)((int)p.d_gsplit.div((uint32_t)cur.seg_hv)) * p.scale_stride);
          const bool whole_seg = (cur.seg_tile0 == 0) && ((cur.seg_tile0 + cur.seg_n) * kTileTokens >= cur.seg_L);
          seg_mode = 0;
          if (!whole_seg) {
            seg_mode = 1;
            if (p.fused_merge) {
              const int pfx = cur.sg == 0 ? cx.z : __ldg(p.plan + kPlanPrefixOffset + cur.seg_s);
              const MergeLoc ml = merge_locate(p.Hv, tpc, cur.seg_hv, cur.seg_L, pfx);
              const bool in_cluster = p.cluster > 1 && (ml.c_lo / p.cluster == ml.c_hi / p.cluster);
              seg_mode = in_cluster ? 3 : 2;
              seg_recv = in_cluster && (int)blockIdx.x < ml.c_hi;   // a piece behind us pushes its fold to us
              seg_push = in_cluster && (int)blockIdx.x > ml.c_lo;   // we push ours to the piece before us
              if (seg_recv && tid == 0) mbar_arrive_expect_tx(recv_bar, kTcRecvBytes);   // the push is kTcRecvBytes of st.async
            }
          }
        }
        // The merge ticket of a cut segment is drawn a few tiles BEFORE its end, so the round trip of the atomic hides behind
        // the remaining tiles (any order of arrival is valid: the last to draw merges, and it waits for the others' partials).
        // Pairs whose pieces all live in this CTA's cluster need no ticket: they are folded through distributed shared memory.
        if (tid == 0 && seg_mode == 2 && cur.u == max(cur.m - 4, 0)) {
          const int pfx = cur.sg == 0 ? cx.z : __ldg(p.plan + kPlanPrefixOffset + cur.seg_s);
          tk = merge_draw_ticket(p, cur.seg_hv, cur.seg_L, pfx);
        }
        // the peer's "partial received" barrier must exist before we arrive on it: its set-up signal (cluster barrier) has
        // normally completed microseconds ago; taking the wait here keeps even that out of the tail
        if (seg_mode == 3 && seg_push && !cluster_ready && (cur.u == 1 || cur.last())) { cluster_wait_acquire(); cluster_ready = true; }

        if (cur.last() && qc.next_segment(p)) { request_q(); q_pending = true; }   // next segment's query rows: one tile ahead

        // ---------------- one 128-token tile: S^T row of this thread's token ----------------
        const uint32_t b = it & 1;
        mbar_wait(sfull(b), (it >> 1) & 1);
        tc_fence_after();
        if (kDbg && !stamped && tid == 0) { dbg_stamp(p, 3); stamped = true; }
        uint32_t r[16];
        tc_ld16(tmem + lane_base + b * 16, r);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(sempty(b));
        const bool ok = cur.tok0() + tid < cur.seg_end;
        float x[8];
        bool exceed = false;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          x[i] = ok ? (__uint_as_float(r[i]) + __uint_as_float(r[8 + i]) * 0.0625f) * crow[i] : ninf;
          exceed |= x[i] > m_ref[i] + kTcMargin;
        }
        const int any = __any_sync(0xffffffffu, exceed);
        if (lane == 0) flags[b * 4 + w] = any;
        named_bar_sync(1, 128);
        if (flags[b * 4 + 0] | flags[b * 4 + 1] | flags[b * 4 + 2] | flags[b * 4 + 3]) {
          // ---------------- raise the CTA-wide reference maximum (always on a segment's first tile, rare later) ----------------
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float mx = x[i];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
            if (lane == 0) wmax[w * 8 + i] = mx;
          }
          named_bar_sync(1, 128);
          float fac[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float t = fmaxf(fmaxf(wmax[i], wmax[8 + i]), fmaxf(wmax[16 + i], wmax[24 + i]));
            const float mn = fmaxf(m_ref[i], t);
            fac[i] = (m_ref[i] == ninf) ? 0.f : fast_exp2(m_ref[i] - mn);
            m_ref[i] = mn;
            l_part[i] *= fac[i];
          }
          if (!cur.first()) {
            // O^T lives in TMEM: wait for every PV issued so far, then scale this thread's row (head dim) of it
            mbar_wait(pempty((it - 1) & 1), ((it - 1) >> 1) & 1);
            tc_fence_after();
            uint32_t o[16];
            tc_ld16(tmem + lane_base + 32, o);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              o[i] = __float_as_uint(__uint_as_float(o[i]) * fac[i]);
              o[8 + i] = __float_as_uint(__uint_as_float(o[8 + i]) * fac[i]);
            }
            tc_st16(tmem + lane_base + 32, o);
            tc_fence_before();
          }
        }
        // p' = 2^(x - m_ref + 4) in (0, 256], split into e4m3 hi + lo/16
        float pr[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          pr[i] = fast_exp2(x[i] - m_ref[i] + kTcPShift);
          l_part[i] += pr[i];
        }
        uint32_t ph[4], pl[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) split_pair(pr[2 * i], pr[2 * i + 1], 16.0f, ph[i], pl[i]);
        mbar_wait(pempty(b), ((it >> 1) & 1) ^ 1);   // PV of tile it-2 no longer reads this P buffer
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(p_base + b * kTcPBytes + tid * 16), "r"(ph[0] | (ph[1] << 16)),
                     "r"(ph[2] | (ph[3] << 16)), "r"(pl[0] | (pl[1] << 16)), "r"(pl[2] | (pl[3] << 16))
                     : "memory");
        {
          // rows of the V tile past the valid count that the last 32-token MMA step still reads: stale smem or cache slots
          // beyond seq_len (possibly NaN patterns) must not meet p = 0
          const int v = cur.valid();
          if (v & 31) {
            const uint32_t st = it % kTcStages;
            mbar_wait(vfull(st), (it / kTcStages) & 1);
            const int nchunk = (32 - (v & 31)) * 8;
            for (int idx = tid; idx < nchunk; idx += 128)
              asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(v_base + st * kTcTileBytes + (v + (idx >> 3)) * 128 + (idx & 7) * 16), "r"(0u) : "memory");
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(pfull(b));
        ++it;

        if (cur.last() && q_pending) { prep_q(); q_pending = false; }   // publish the next segment's operand before this epilogue
        if (cur.last()) {
          // ---------------- segment end: O^T row of this thread's head dim, row sums, output or partial ----------------
          if (kDbg && tid == 0 && cur.remaining <= 0) dbg_stamp(p, 4);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float v = l_part[i];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
            if (lane == 0) red_l[w * 8 + i] = v;
          }
          mbar_wait(pempty((it - 1) & 1), ((it - 1) >> 1) & 1);   // the segment's last PV has completed
          tc_fence_after();
          uint32_t o[16];
          tc_ld16(tmem + lane_base + 32, o);
          tc_fence_before();
          named_bar_sync(1, 128);
          if (kDbg && tid == 0 && cur.remaining <= 0) dbg_stamp(p, 11);
          const int s = cur.seg_s, hv = cur.seg_hv, L = cur.seg_L;
          const int kvh = (int)p.d_gsplit.div((uint32_t)hv);
          const float vscale = seg_vscale;
          const int pslot = cur.sg == 0 ? 0 : 1;
          const int d = tid;
          using E = Elem<QDT>;
          float accv[8], mgv[8], lgv[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            lgv[i] = (red_l[i] + red_l[8 + i]) + (red_l[16 + i] + red_l[24 + i]);
            accv[i] = __uint_as_float(o[i]) + __uint_as_float(o[8 + i]) * 0.0625f;
            mgv[i] = m_ref[i] - kTcPShift;   // p' = 2^(x - mg)
          }
          auto store_final = [&]() {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (i < p.G) {
                const int hq = hv * p.G + i;
                typename E::type* op = reinterpret_cast<typename E::type*>(p.out) + (int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h;
                op[d] = E::from_float(accv[i] * (vscale / lgv[i]));
                if (p.lse_out != nullptr && d == 0) p.lse_out[(int64_t)s * p.Hq + hq] = (mgv[i] + fast_log2(lgv[i])) * 0.6931471805599453f;
              }
            }
          };
          if (seg_mode == 0) {
            store_final();
          } else if (seg_mode == 1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (i < p.G) {
                const size_t pidx = ((size_t)blockIdx.x * 2 + pslot) * kMaxGroup + i;
                __stcg(p.ws_o + pidx * kD + d, accv[i]);
                if (d == 0) __stcg(reinterpret_cast<float2*>(p.ws_ml + pidx * 2), make_float2(mgv[i], lgv[i]));
              }
            }
          } else {
            if (seg_mode == 3) {
              // ---- the pair's pieces all sit in this cluster: fold them along the chain c_hi -> ... -> c_lo through
              // distributed shared memory (one push + one mbarrier arrive per hop; no L2 round trips, no atomics) ----
              const int me = (int)blockIdx.x;
              if (seg_recv) {   // receive the fold of the pieces behind us (its bytes were announced at the segment's start)
                while (!mbar_try_wait_acq_cluster(recv_bar, 0)) {
                }
                if (kDbg && tid == 0) dbg_stamp(p, 12);
                float rov[8];
                {
                  const uint4 r0 = lds128(base + TcSmem::kRecv + d * 32), r1 = lds128(base + TcSmem::kRecv + d * 32 + 16);
                  rov[0] = __uint_as_float(r0.x); rov[1] = __uint_as_float(r0.y); rov[2] = __uint_as_float(r0.z); rov[3] = __uint_as_float(r0.w);
                  rov[4] = __uint_as_float(r1.x); rov[5] = __uint_as_float(r1.y); rov[6] = __uint_as_float(r1.z); rov[7] = __uint_as_float(r1.w);
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                  if (i < p.G) {
                    const float2 rml = recv_ml[i];
                    const float ro = rov[i];
                    const float mn = fmaxf(mgv[i], rml.x);
                    const float a = (mgv[i] == ninf) ? 0.f : fast_exp2(mgv[i] - mn);
                    const float bsc = (rml.x == ninf) ? 0.f : fast_exp2(rml.x - mn);
                    accv[i] = accv[i] * a + ro * bsc;
                    lgv[i] = lgv[i] * a + rml.y * bsc;
                    mgv[i] = mn;
                  }
                }
              }
              if (seg_push) {   // push to the piece before us
                // st.async: every store carries its own completion to the peer's "partial received" barrier (which expects
                // kTcRecvBytes), so nobody here fences, syncs or waits for the stores to be performed
                const uint32_t peer = cluster_ctarank() - 1;
                const uint32_t ro = mapa_shared(base + TcSmem::kRecv, peer);
                const uint32_t rb = mapa_shared(recv_bar, peer);
                st_async_f32x4(ro + (uint32_t)d * 32, accv[0], accv[1], accv[2], accv[3], rb);        // layout [head dim][8 heads]
                st_async_f32x4(ro + (uint32_t)d * 32 + 16, accv[4], accv[5], accv[6], accv[7], rb);
#pragma unroll
                for (int i = 0; i < 8; ++i)
                  if (d == i) st_async_f32x2(ro + 8 * kD * 4 + (uint32_t)i * 8, mgv[i], lgv[i], rb);
                if (kDbg && tid == 0) dbg_stamp(p, 13);
              } else {
                store_final();
              }
            } else {
              unsigned long long tm0 = 0, tm1 = 0;
              if (kDbg && p.dbg != nullptr && tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tm0));
              fused_merge_tail<QDT, 8, true>(p, s, hv, vscale, tid, pslot, tk, accv, mgv, lgv, mrg, hand, pub_req);
              if (kDbg && p.dbg != nullptr && tid == 0) {
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tm1));
                p.dbg[(size_t)blockIdx.x * kDbgSlots + (hand[0] > 0 ? 7 : 6)] += tm1 - tm0;   // 6: publish calls, 7: merge calls
              }
            }
          }
          named_bar_sync(1, 128);
        }
      }
      if (kDbg && tid == 0) dbg_stamp(p, 5);
      if (tid == 0) pub_ring_push(pub_req, 0ull);   // tell the publish helper to leave
    }
  }

  // teardown: every role is done with TMEM (the softmax warps waited for the last PV)
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    const uint32_t tmem = *tmem_ptr_smem;
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTcTmemCols) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
int launch_with_pdl_cluster(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args, int cluster);
int option_cluster();

template <int QDT, bool kDbg>
static int launch_tc(const CUtensorMap& tk, const CUtensorMap& tv, DecodeParams& p, int grid, cudaStream_t stream) {
  auto fn = decode_tc_kernel<QDT, kDbg>;
  static bool configured[64] = {};   // per device: the attribute belongs to the context
  int dev = 0;
  if (check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return B200ATTN_ECUDA;
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    if (check_cuda(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, TcSmem::kTotal), "cudaFuncSetAttribute(decode_tc smem)"))
      return B200ATTN_ECUDA;
    (void)cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (getenv("B200ATTN_VERBOSE")) {
      int nb = 0;
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fn, kTcThreads, TcSmem::kTotal);
      fprintf(stderr, "[b200attn] decode_tc: %d B smem / CTA, %d stages, occupancy %d CTAs / SM\n", TcSmem::kTotal, kTcStages, nb);
    }
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  // clusters of 2 CTAs when the grid allows: the two halves of a (sequence, kv head) pair exchange their partials through
  // distributed shared memory
  const int cl = option_cluster();
  p.cluster = (cl > 1 && grid % cl == 0) ? cl : 1;
  void* args[] = {(void*)&tk, (void*)&tv, (void*)&p};
  return launch_with_pdl_cluster((const void*)fn, dim3(grid), dim3(kTcThreads), TcSmem::kTotal, stream, args, p.cluster);
}

// fp8 KV, group size <= 8 per (virtual) kv head
int launch_decode_tc(const CUtensorMap& tk, const CUtensorMap& tv, DecodeParams& p, int grid, int q_dtype, cudaStream_t stream) {
  if (p.dbg != nullptr)
    return q_dtype == B200ATTN_BF16 ? launch_tc<B200ATTN_BF16, true>(tk, tv, p, grid, stream) : launch_tc<B200ATTN_FP16, true>(tk, tv, p, grid, stream);
  return q_dtype == B200ATTN_BF16 ? launch_tc<B200ATTN_BF16, false>(tk, tv, p, grid, stream) : launch_tc<B200ATTN_FP16, false>(tk, tv, p, grid, stream);
}

}  // namespace b200attn
