// peer_exchange.cu — the two exchange steps of a sequence-split replica group (tp > Hkv, SURVEY.md 8e remedy 1) done with
// one-way PUSHES over NVLink peer memory, no collective, no barrier and no memory fence:
//
//   peer_gather_heads_kernel   before the decode: every rank pushes its own query heads to every peer and unpacks what the
//                              peers pushed into its local gathered-query buffer.  Replaces the NCCL all-gather of the
//                              group's query heads ($VLLM/v1/attention/backends/flashinfer.py:1714-1738,
//                              get_dcp_group().all_gather(query)).
//   peer_combine_kernel        after the decode: every rank pushes, for each peer, the slice of its partial (o, lse) that
//                              covers THAT peer's query heads, and merges its own slice with the slices it receives by
//                              log-sum-exp.  Replaces cp_lse_ag_out_rs (all-gather of lse + reduce-scatter of o) /
//                              merge_attn_states ($VLLM/v1/attention/ops/merge_attn_states.py:7-101).
//
// Transport: the "LL" protocol -- every 8-byte store carries 4 bytes of payload and the 4-byte launch count (epoch) of the
// exchange, and 8-byte stores are atomic over NVLink, so the receiver simply polls each word of its LOCAL receive buffer until
// the epoch matches: the data IS the flag.  No __threadfence_system / release / acquire (measured: a fence-and-flag version
// of these kernels cost 11-15 us per launch on B200 NVLink, almost all of it in MEMBAR.SYS), no second round trip; a step
// costs one NVLink crossing.  Payload doubles on the wire (32 KiB -> 64 KiB per peer and step at thinker shapes), which is
// irrelevant at 900 GB/s.  Epochs are launch counts kept in device memory (read after griddepcontrol.wait, bumped by the last
// CTA), so nothing is ever reset and CUDA-graph replays work; receive buffers alternate by epoch parity, so a rank that runs
// one launch ahead never overwrites words a peer is still reading.  The buffers are symmetric allocations; their base
// pointers arrive as DEVICE arrays of per-rank pointers (torch symmetric memory `buffer_ptrs_dev`).
#include "common.cuh"

namespace b200attn {

constexpr int kPeerCtas = B200ATTN_PEER_CTAS;
constexpr int kPeerMaxRanks = 8;

__device__ __forceinline__ void st_ll(uint2* p, uint32_t data, uint32_t flag) {
  asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(data), "r"(flag) : "memory");
}
// spin until the word carries this launch's epoch; a peer that never arrives must not hang the GPU: trap after ~4 s
__device__ __forceinline__ uint32_t ld_ll(const uint2* p, uint32_t flag) {
  uint32_t d, f;
  long long t0 = 0;
  for (int spin = 0;; ++spin) {
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(d), "=r"(f) : "l"(p) : "memory");
    if (f == flag) return d;
    if (spin == 64) t0 = clock64();
    if (spin > 64 && (spin & 1023) == 0 && clock64() - t0 > 8000000000ll) __trap();
  }
}

// two words polled together (their loads overlap)
__device__ __forceinline__ void ld_ll2(const uint2* p0, const uint2* p1, uint32_t flag, uint32_t& d0, uint32_t& d1) {
  uint32_t f0, f1;
  long long t0 = 0;
  for (int spin = 0;; ++spin) {
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(d0), "=r"(f0) : "l"(p0) : "memory");
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(d1), "=r"(f1) : "l"(p1) : "memory");
    if (f0 == flag && f1 == flag) return;
    if (spin == 64) t0 = clock64();
    if (spin > 64 && (spin & 1023) == 0 && clock64() - t0 > 8000000000ll) __trap();
  }
}

// epoch of this launch (same on every rank: every rank issues the same sequence of calls); bumped by the last CTA to finish
__device__ __forceinline__ uint32_t epoch_begin(const int* epoch) {
  return (uint32_t)(*reinterpret_cast<const volatile int*>(epoch)) + 1u;
}
__device__ __forceinline__ void epoch_end(int* epoch, uint32_t e) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(epoch + 1, 1) == (int)gridDim.x - 1) {
      epoch[1] = 0;
      __threadfence();
      *reinterpret_cast<volatile int*>(epoch) = (int)e;
    }
  }
}

struct GatherParams {
  const void* local;           // [n][Hl][D] of 16-bit elements, any token / head stride
  int64_t stride_t, stride_h;
  void* gathered;              // LOCAL [n][R * Hl][D]
  uint2* const* ll_ptrs;       // per rank: [2 parities][R senders][cap_words]
  int64_t cap_words;
  int* epoch;                  // local: [0] completed launches, [1] CTAs done in the running launch
  int n, Hl, D, rank, R;
};

__global__ void __launch_bounds__(256) peer_gather_heads_kernel(const GatherParams p) {
  pdl_launch_dependents();
  pdl_wait();   // the local query heads come from the kernel before us; the epoch word from our own previous launch
  const uint32_t e = epoch_begin(p.epoch);
  const int cta = blockIdx.x;
  const int per = (p.n + gridDim.x - 1) / gridDim.x;
  const int t0 = cta * per, t1 = min(p.n, t0 + per);
  const int wpr = p.D >> 1;                       // 32-bit words per head row
  const int items = max(0, t1 - t0) * p.Hl * wpr;
  const int Hg = p.Hl * p.R;
  const int64_t slot = (int64_t)(e & 1) * p.R * p.cap_words;
  // ---- push my heads to every peer (and straight into my own gathered buffer) ----
  for (int i = threadIdx.x; i < items; i += blockDim.x) {
    const int w = i % wpr, h = (i / wpr) % p.Hl, t = t0 + i / (wpr * p.Hl);
    const uint32_t x = *reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint16_t*>(p.local) + (int64_t)t * p.stride_t + (int64_t)h * p.stride_h + w * 2);
    const int64_t widx = ((int64_t)t * p.Hl + h) * wpr + w;
    for (int j = 0; j < p.R; ++j)
      if (j != p.rank) st_ll(p.ll_ptrs[j] + slot + (int64_t)p.rank * p.cap_words + widx, x, e);
    reinterpret_cast<uint32_t*>(p.gathered)[((int64_t)t * Hg + p.rank * p.Hl + h) * wpr + w] = x;
  }
  // ---- unpack what the peers pushed to me ----
  const uint2* mine = p.ll_ptrs[p.rank] + slot;
  for (int j = 0; j < p.R; ++j) {
    if (j == p.rank) continue;
    for (int i = threadIdx.x; i < items; i += blockDim.x) {
      const int w = i % wpr, h = (i / wpr) % p.Hl, t = t0 + i / (wpr * p.Hl);
      const uint32_t x = ld_ll(mine + (int64_t)j * p.cap_words + ((int64_t)t * p.Hl + h) * wpr + w, e);
      reinterpret_cast<uint32_t*>(p.gathered)[((int64_t)t * Hg + j * p.Hl + h) * wpr + w] = x;
    }
  }
  epoch_end(p.epoch, e);
}

struct CombineParams {
  const void* o_part;          // [n][R * Hl][D] local partial (this rank's slice of the context, all heads of the group)
  const float* lse_part;       // [n][R * Hl] natural-log LSE, -inf = no token on this rank
  uint2* const* ll_ptrs;       // per rank: [2 parities][R senders][cap_words]; a sender's region = o words [cap][Hl][D/2], then lse [cap][Hl]
  int64_t cap_words;
  int* epoch;
  void* out;                   // [n][Hl][D]
  int64_t out_stride_t, out_stride_h;
  int n, cap, Hl, D, rank, R;
};

template <int DT>
__global__ void __launch_bounds__(256) peer_combine_kernel(const CombineParams p) {
  using E = Elem<DT>;
  pdl_launch_dependents();
  pdl_wait();   // the local partial comes from the decode kernel before us
  const uint32_t e = epoch_begin(p.epoch);
  const int cta = blockIdx.x;
  const int per = (p.n + gridDim.x - 1) / gridDim.x;
  const int t0 = cta * per, t1 = min(p.n, t0 + per);
  const int nt = max(0, t1 - t0);
  const int wpr = p.D >> 1;
  const int Hg = p.Hl * p.R;
  const int64_t slot = (int64_t)(e & 1) * p.R * p.cap_words;
  const int64_t lse_base = (int64_t)p.cap * p.Hl * wpr;   // word offset of the lse area inside a sender's region
  // ---- push: for every peer j, my partial of ITS heads ----
  for (int j = 0; j < p.R; ++j) {
    if (j == p.rank) continue;
    uint2* dst = p.ll_ptrs[j] + slot + (int64_t)p.rank * p.cap_words;
    for (int i = threadIdx.x; i < nt * p.Hl * wpr; i += blockDim.x) {
      const int w = i % wpr, h = (i / wpr) % p.Hl, t = t0 + i / (wpr * p.Hl);
      const uint32_t x = *reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint16_t*>(p.o_part) + ((int64_t)t * Hg + j * p.Hl + h) * p.D + w * 2);
      st_ll(dst + ((int64_t)t * p.Hl + h) * wpr + w, x, e);
    }
    for (int i = threadIdx.x; i < nt * p.Hl; i += blockDim.x) {
      const int h = i % p.Hl, t = t0 + i / p.Hl;
      st_ll(dst + lse_base + (int64_t)t * p.Hl + h, __float_as_uint(p.lse_part[(int64_t)t * Hg + j * p.Hl + h]), e);
    }
  }
  // ---- merge my heads: own slice (local) + what the peers pushed (polled word by word) ----
  const uint2* mine = p.ll_ptrs[p.rank] + slot;
  for (int i = threadIdx.x; i < nt * p.Hl * wpr; i += blockDim.x) {
    const int w = i % wpr, h = (i / wpr) % p.Hl, t = t0 + i / (wpr * p.Hl);
    float lse[kPeerMaxRanks];
    uint32_t ov[kPeerMaxRanks];
#pragma unroll
    for (int j = 0; j < kPeerMaxRanks; ++j) {
      if (j < p.R) {
        if (j == p.rank) {
          lse[j] = p.lse_part[(int64_t)t * Hg + j * p.Hl + h];
          ov[j] = *reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint16_t*>(p.o_part) + ((int64_t)t * Hg + j * p.Hl + h) * p.D + w * 2);
        } else {
          const uint2* src = mine + (int64_t)j * p.cap_words;
          uint32_t a, b;
          ld_ll2(src + lse_base + (int64_t)t * p.Hl + h, src + ((int64_t)t * p.Hl + h) * wpr + w, e, a, b);   // both in flight together
          lse[j] = __uint_as_float(a);
          ov[j] = b;
        }
      }
    }
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < kPeerMaxRanks; ++j)
      if (j < p.R) mx = fmaxf(mx, lse[j]);
    float den = 0.f, a0 = 0.f, a1 = 0.f;
#pragma unroll
    for (int j = 0; j < kPeerMaxRanks; ++j) {
      if (j < p.R) {
        const float wgt = (lse[j] == -INFINITY) ? 0.f : __expf(lse[j] - mx);
        den += wgt;
        const typename E::type* ee = reinterpret_cast<const typename E::type*>(&ov[j]);
        a0 += wgt * E::to_float(ee[0]);
        a1 += wgt * E::to_float(ee[1]);
      }
    }
    uint32_t res;
    typename E::type* r = reinterpret_cast<typename E::type*>(&res);
    r[0] = E::from_float(den > 0.f ? a0 / den : 0.f);
    r[1] = E::from_float(den > 0.f ? a1 / den : 0.f);
    *reinterpret_cast<uint32_t*>(reinterpret_cast<typename E::type*>(p.out) + (int64_t)t * p.out_stride_t + (int64_t)h * p.out_stride_h + w * 2) = res;
  }
  epoch_end(p.epoch, e);
}

int launch_with_pdl(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args);

static int peer_grid(int n) { return n < kPeerCtas ? (n > 0 ? n : 1) : kPeerCtas; }

}  // namespace b200attn

using namespace b200attn;

static int peer_args_ok(const char* what, int n, int Hl, int D, int dtype, int rank, int R) {
  if (n < 0 || Hl <= 0 || D <= 0 || (D % 8) != 0) {
    set_error("%s: bad shape n=%d heads_local=%d head_dim=%d (head_dim must be a multiple of 8)", what, n, Hl, D);
    return B200ATTN_EINVAL;
  }
  if (dtype != B200ATTN_BF16 && dtype != B200ATTN_FP16) {
    set_error("%s: dtype must be bf16 or fp16", what);
    return B200ATTN_EINVAL;
  }
  if (R < 1 || R > kPeerMaxRanks || rank < 0 || rank >= R) {
    set_error("%s: rank %d / %d ranks out of range (1..%d ranks)", what, rank, R, kPeerMaxRanks);
    return B200ATTN_EINVAL;
  }
  return 0;
}

extern "C" size_t b200attn_peer_ll_bytes(int32_t rows_capacity, int32_t heads_local, int32_t head_dim, int32_t num_ranks, int32_t with_lse) {
  const size_t words = (size_t)rows_capacity * heads_local * (head_dim / 2) + (with_lse ? (size_t)rows_capacity * heads_local : 0);
  return 2 * (size_t)num_ranks * words * 8;   // [2 parities][senders][words] of 8 bytes
}

extern "C" int b200attn_peer_gather_heads(const void* local, int64_t stride_t, int64_t stride_h, int32_t n, int32_t rows_capacity,
                                          int32_t heads_local, int32_t head_dim, int32_t dtype, void* gathered,
                                          void* const* ll_ptrs, int32_t* epoch, int32_t rank, int32_t num_ranks, void* stream) {
  if (n == 0) return B200ATTN_OK;
  if (!local || !gathered || !ll_ptrs || !epoch) {
    set_error("peer_gather_heads: null pointer argument");
    return B200ATTN_EINVAL;
  }
  int rc = peer_args_ok("peer_gather_heads", n, heads_local, head_dim, dtype, rank, num_ranks);
  if (rc) return rc;
  if (n > rows_capacity) {
    set_error("peer_gather_heads: %d rows exceed the buffers' capacity %d", n, rows_capacity);
    return B200ATTN_ENOSPACE;
  }
  if ((reinterpret_cast<uintptr_t>(local) % 4) || (reinterpret_cast<uintptr_t>(gathered) % 4) || (stride_t % 2) || (stride_h % 2)) {
    set_error("peer_gather_heads: pointers and strides must be 4-byte aligned");
    return B200ATTN_EINVAL;
  }
  GatherParams p;
  p.local = local; p.stride_t = stride_t; p.stride_h = stride_h;
  p.gathered = gathered;
  p.ll_ptrs = reinterpret_cast<uint2* const*>(ll_ptrs);
  p.cap_words = (int64_t)rows_capacity * heads_local * (head_dim / 2);
  p.epoch = epoch;
  p.n = n; p.Hl = heads_local; p.D = head_dim; p.rank = rank; p.R = num_ranks;
  void* args[] = {&p};
  return launch_with_pdl((const void*)peer_gather_heads_kernel, dim3(peer_grid(n)), dim3(256), 0, static_cast<cudaStream_t>(stream), args);
}

extern "C" int b200attn_peer_combine(const void* o_part, const float* lse_part, int32_t n, int32_t rows_capacity,
                                     int32_t heads_local, int32_t head_dim, int32_t dtype, void* const* ll_ptrs, int32_t* epoch,
                                     int32_t rank, int32_t num_ranks, void* out, int64_t out_stride_t, int64_t out_stride_h,
                                     void* stream) {
  if (n == 0) return B200ATTN_OK;
  if (!o_part || !lse_part || !ll_ptrs || !epoch || !out) {
    set_error("peer_combine: null pointer argument");
    return B200ATTN_EINVAL;
  }
  int rc = peer_args_ok("peer_combine", n, heads_local, head_dim, dtype, rank, num_ranks);
  if (rc) return rc;
  if (n > rows_capacity) {
    set_error("peer_combine: %d rows exceed the receive buffers' capacity %d", n, rows_capacity);
    return B200ATTN_ENOSPACE;
  }
  if ((reinterpret_cast<uintptr_t>(o_part) % 4) || (reinterpret_cast<uintptr_t>(out) % 4) || (out_stride_t % 2) || (out_stride_h % 2)) {
    set_error("peer_combine: pointers and strides must be 4-byte aligned");
    return B200ATTN_EINVAL;
  }
  CombineParams p;
  p.o_part = o_part; p.lse_part = lse_part;
  p.ll_ptrs = reinterpret_cast<uint2* const*>(ll_ptrs);
  p.cap_words = (int64_t)rows_capacity * heads_local * (head_dim / 2) + (int64_t)rows_capacity * heads_local;
  p.epoch = epoch;
  p.out = out; p.out_stride_t = out_stride_t; p.out_stride_h = out_stride_h;
  p.n = n; p.cap = rows_capacity; p.Hl = heads_local; p.D = head_dim; p.rank = rank; p.R = num_ranks;
  void* args[] = {&p};
  const void* fn = dtype == B200ATTN_BF16 ? (const void*)peer_combine_kernel<B200ATTN_BF16> : (const void*)peer_combine_kernel<B200ATTN_FP16>;
  return launch_with_pdl(fn, dim3(peer_grid(n)), dim3(256), 0, static_cast<cudaStream_t>(stream), args);
}
