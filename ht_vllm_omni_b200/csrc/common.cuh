// common.cuh — shared device helpers for libb200attn (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200attn.h"

namespace b200attn {

// ---------------------------------------------------------------------------------------------
// Geometry of the split-KV decode pipeline
// ---------------------------------------------------------------------------------------------
constexpr int kChunkTokens = 16;                              // one TMA box / one consumer warp
constexpr int kConsumerWarps = 4;                             // chunks per tile
constexpr int kTileTokens = kChunkTokens * kConsumerWarps;    // 64 = B200ATTN_TILE_TOKENS
constexpr int kDecodeThreads = (kConsumerWarps + 1) * 32;     // + 1 producer warp
constexpr int kCtasPerSm = 2;
constexpr int kMaxGrid = 1024;                                // plan capacity (CTAs)
constexpr int kMaxGroup = B200ATTN_MAX_GROUP;                 // q heads per kv head per CTA pass
static_assert(kTileTokens == B200ATTN_TILE_TOKENS, "header / kernel tile mismatch");

// Plan buffer layout (int32 words), ABI version 3:
//   [0] total_tiles  [1] tiles_per_cta  [2] grid  [3] num_seqs  [4] num_kv_heads (virtual)  [5] chunks per cache block  [6..7] 0
//   [8 .. 8+kMaxGrid*64)        one 256-byte ENTRY per persistent CTA, everything its first TMA needs in one coalesced load:
//        [0] seq (-1: idle) [1] kv head (virtual) [2] first 64-token tile [3] tile count
//        [4] seq_len of that row [5] block-table row (row_map[seq] or seq) [6] tile prefix of seq [7] tiles_per_cta
//        [8] grid [9] num_seqs (both checked against the launch: a plan / launch mismatch traps instead of reading garbage)
//        [32..64) block ids of the first 32 16-token chunks of the CTA's first segment (chunk = tile * 4 + j)
//   [.. +kMaxGrid*4)            counters of the in-kernel merge, one (ticket, published) pair per (first CTA, slot) of a
//                               split (sequence, kv head) pair; zeroed by the plan kernel, self-resetting afterwards
//   [.. +B+1)                   tile_prefix[num_seqs+1] (exclusive scan of ceil(L/64))
constexpr int kPlanHeaderWords = 8;
constexpr int kPlanEntryWords = 64;
constexpr int kPlanEntryIds = 32;
constexpr int kPlanCtaWords = kMaxGrid * kPlanEntryWords;
constexpr int kPlanCounterOffset = kPlanHeaderWords + kPlanCtaWords;
constexpr int kPlanCounterWords = kMaxGrid * 4;   // [0, 2*kMaxGrid): arrival tickets, [2*kMaxGrid, 4*kMaxGrid): published partials
constexpr int kPlanPrefixOffset = kPlanCounterOffset + kPlanCounterWords;

// Workspace layout: ws_o [grid][2 slots][16 heads][D] fp32, then ws_ml [grid][2][16][2] fp32.
__host__ __device__ inline size_t ws_o_floats(int grid, int D) { return (size_t)grid * 2 * kMaxGroup * D; }
__host__ __device__ inline size_t ws_ml_floats(int grid) { return (size_t)grid * 2 * kMaxGroup * 2; }

// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// 4-D tiled TMA load, global -> shared, completion on an mbarrier, with an L2 eviction policy.
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0,
                                            int c1, int c2, int c3, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(dst),
      "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(policy)
      : "memory");
}
// 1-D bulk copy, global -> shared (16-byte aligned, size multiple of 16).
__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint64_t policy_evict_normal() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void prefetch_tensormap(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
}

// Thread-block cluster / distributed shared memory
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t local_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void st_cluster_f32x2(uint32_t addr, float a, float b) {
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(a), "f"(b) : "memory");
}
// asynchronous stores into another CTA's shared memory that carry their own completion: each one adds its byte count to the
// transaction count of an mbarrier in the DESTINATION CTA, so the receiver's wait on that barrier is the only ordering
// needed -- no fence or release by the sender (its threads never wait for the stores to be performed)
__device__ __forceinline__ void st_async_f32x4(uint32_t addr, float a, float b, float c, float d, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(addr), "f"(a),
               "f"(b), "f"(c), "f"(d), "r"(remote_bar)
               : "memory");
}
__device__ __forceinline__ void st_async_f32x2(uint32_t addr, float a, float b, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];" ::"r"(addr), "f"(a), "f"(b),
               "r"(remote_bar)
               : "memory");
}
// arrive (count 1) on an mbarrier in another CTA of the cluster; release at cluster scope orders this thread's earlier
// (and, through CTA barriers, its CTA's earlier) distributed-shared-memory stores before it
__device__ __forceinline__ void mbar_arrive_remote_release(uint32_t remote_bar) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote_bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_acq_cluster(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void cluster_arrive_release() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait_acquire() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

// Programmatic dependent launch (no-ops when the kernel was launched without the attribute).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
// non-blocking arrival on a named barrier other warps bar.sync on (`threads` = total participants, arrivers included)
__device__ __forceinline__ void named_bar_arrive(int id, int threads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
// L2-coherent loads that the compiler must not reorder or merge (the in-kernel merge issues them in a chosen order)
__device__ __forceinline__ float ldcg_f32(const float* p) {
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float2 ldcg_f32x2(const float* p) {
  float2 v;
  asm volatile("ld.global.cg.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

// transpose an 8x8 matrix of b16 held one row-pair per thread: in: thread (g,t) = M[g][2t,2t+1]; out: M[2t,2t+1][g]
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}

// two e4m3 bytes (low / high half of a 32-bit word) -> packed f16x2 (exact)
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2_lo(uint32_t w) {
  uint32_t r;
  asm("{\n\t.reg .b16 lo, hi;\n\tmov.b32 {lo, hi}, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, lo;\n\t}" : "=r"(r) : "r"(w));
  return r;
}
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2_hi(uint32_t w) {
  uint32_t r;
  asm("{\n\t.reg .b16 lo, hi;\n\tmov.b32 {lo, hi}, %1;\n\tcvt.rn.f16x2.e4m3x2 %0, hi;\n\t}" : "=r"(r) : "r"(w));
  return r;
}

// D(16x8, f32) += A(16x16) * B(16x8); kBf16 selects bf16 vs f16 operands.
template <bool kBf16>
__device__ __forceinline__ void mma_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                          uint32_t b0, uint32_t b1) {
  if constexpr (kBf16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
}

template <bool kBf16>
__device__ __forceinline__ uint32_t pack_f32x2(float lo, float hi) {
  uint32_t r;
  if constexpr (kBf16) {
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  } else {
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  }
  return r;
}

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_log2(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 16-bit element <-> float for the two activation dtypes
template <int DT>
struct Elem;
template <>
struct Elem<B200ATTN_BF16> {
  using type = __nv_bfloat16;
  static __device__ __forceinline__ float to_float(type v) { return __bfloat162float(v); }
  static __device__ __forceinline__ type from_float(float v) { return __float2bfloat16_rn(v); }
};
template <>
struct Elem<B200ATTN_FP16> {
  using type = __half;
  static __device__ __forceinline__ float to_float(type v) { return __half2float(v); }
  static __device__ __forceinline__ type from_float(float v) { return __float2half_rn(v); }
};

// Division of a 32-bit index by a launch-constant divisor without the integer-divide sequence (which goes through the XU
// pipe: ~20 instructions per quotient; the element-wise kernels did three per 16 bytes, profiles/r01_ncu_rope_kv_write_v8.txt
// 33 % pipe_xu).  Granlund-Montgomery round-up method, exact for every 32-bit n: for d = 2^k a shift, otherwise
//   L = ceil(log2 d), m = floor(2^32 (2^L - d) / d) + 1,  q = (t + ((n - t) >> 1)) >> (L - 1)  with  t = umulhi(m, n).
struct FastDiv {
  uint32_t d, m, s;   // s = L - 1 (or log2 d for powers of two, with m = 0)
  __host__ void init(uint32_t div) {
    d = div ? div : 1;
    uint32_t L = 0;
    while ((1ull << L) < d) ++L;
    if ((1ull << L) == d) { m = 0; s = L; return; }
    m = (uint32_t)(((1ull << 32) * ((1ull << L) - d)) / d + 1);
    s = L - 1;
  }
  __device__ __forceinline__ uint32_t div(uint32_t n) const {
    if (m == 0) return n >> s;
    const uint32_t t = __umulhi(m, n);
    return (t + ((n - t) >> 1)) >> s;
  }
  __device__ __forceinline__ void divmod(uint32_t n, uint32_t& q, uint32_t& r) const { q = div(n); r = n - q * d; }
};

// host-side error plumbing (abi.cu)
void set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);

}  // namespace b200attn
