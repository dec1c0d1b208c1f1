// peer_exchange.cu — the two exchange steps of a sequence-split replica group (tp > Hkv, SURVEY.md 8e remedy 1) done with
// PUSHES over NVLink peer memory and per-CTA flags instead of collectives:
//
//   peer_gather_heads_kernel   before the decode: every rank stores its own query heads into the gathered-query buffer of
//                              EVERY rank of the group (its own included), then raises a flag at each peer and waits for the
//                              peers' flags.  Replaces the NCCL all-gather of the group's query heads
//                              ($VLLM/v1/attention/backends/flashinfer.py:1714-1738, get_dcp_group().all_gather(query)).
//   peer_combine_kernel        after the decode: every rank stores, for each peer, the slice of its partial (o, lse) that
//                              covers THAT peer's query heads into the peer's receive buffer, raises the flag, waits for the
//                              peers' flags, and merges its own slice with the received ones by log-sum-exp.  Replaces
//                              cp_lse_ag_out_rs (all-gather of lse + reduce-scatter of o) / merge_attn_states
//                              ($VLLM/v1/attention/ops/merge_attn_states.py:7-101).
//
// A push is one-way: stores travel to the peer without anyone waiting for a reply, the flag follows them (system-scope
// fence + release store), and the receiver only ever reads LOCAL memory after an acquire on its own flag word -- so each step
// costs one NVLink crossing instead of the barrier + two load round trips of the pull variant (b200attn_merge_lse_peers).
// Each CTA owns a contiguous chunk of tokens and its own flag words and epoch counter, so CTAs never synchronise with each
// other; flag values are launch counts kept in device memory (monotonic), so nothing is reset and CUDA-graph replays work.
// The buffers are symmetric allocations (one per rank, same layout); their base pointers arrive as DEVICE arrays of
// per-rank pointers (torch symmetric memory `buffer_ptrs_dev`).
#include "common.cuh"

namespace b200attn {

constexpr int kPeerCtas = B200ATTN_PEER_CTAS;   // flag words / epoch counters per (op, sender)
constexpr int kPeerMaxRanks = 8;

__device__ __forceinline__ void st_release_sys(int* p, int v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
  int v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_cv_v4(const void* p) {   // bypass L1: the line was written by a peer over NVLink
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_cv_f32(const float* p) {
  float v;
  asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

// Raise this CTA's flag at every peer, then wait until every peer's flag for this CTA has reached `epoch`.
// A peer that never arrives (a crashed rank) must not hang the GPU: after ~4 s of polling the kernel traps.
__device__ __forceinline__ void signal_and_wait(int* const* flag_ptrs, int64_t flag_off, int rank, int R, int cta, int epoch) {
  __threadfence_system();   // this thread's stores to the peers are performed before the flag below can be seen
  __syncthreads();
  const int j = threadIdx.x;
  if (j < R && j != rank) {
    st_release_sys(flag_ptrs[j] + flag_off + rank * kPeerCtas + cta, epoch);
    const int* mine = flag_ptrs[rank] + flag_off + j * kPeerCtas + cta;
    const long long t0 = clock64();
    while (ld_acquire_sys(mine) - epoch < 0) {
      __nanosleep(64);
      if (clock64() - t0 > 8000000000ll) __trap();
    }
  }
  __syncthreads();
}

struct GatherParams {
  const void* local;           // [n][Hl][D] of 16-bit elements, any token / head stride
  int64_t stride_t, stride_h;
  void* const* gather_ptrs;    // per rank: [n][R * Hl][D]
  int64_t gather_off;          // elements
  int* const* flag_ptrs;
  int64_t flag_off;
  int* epoch;                  // local, [kPeerCtas]
  int n, Hl, D, rank, R;
};

__global__ void __launch_bounds__(256) peer_gather_heads_kernel(const GatherParams p) {
  __shared__ int s_epoch;
  pdl_launch_dependents();
  pdl_wait();   // the local query heads come from the kernel before us
  const int cta = blockIdx.x;
  if (threadIdx.x == 0) s_epoch = ++p.epoch[cta];
  __syncthreads();
  const int epoch = s_epoch;
  const int per = (p.n + gridDim.x - 1) / gridDim.x;
  const int t0 = cta * per, t1 = min(p.n, t0 + per);
  const int vec = p.D >> 3;                     // 16-byte vectors per head row
  const int items = max(0, t1 - t0) * p.Hl * vec;
  const int Hg = p.Hl * p.R;
  for (int i = threadIdx.x; i < items; i += blockDim.x) {
    const int v = i % vec, h = (i / vec) % p.Hl, t = t0 + i / (vec * p.Hl);
    const uint4 x = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.local) + (int64_t)t * p.stride_t + (int64_t)h * p.stride_h + v * 8);
    const int64_t dst = p.gather_off + ((int64_t)t * Hg + p.rank * p.Hl + h) * p.D + v * 8;
#pragma unroll 1
    for (int j = 0; j < p.R; ++j) *reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.gather_ptrs[j]) + dst) = x;
  }
  signal_and_wait(p.flag_ptrs, p.flag_off, p.rank, p.R, cta, epoch);
}

struct CombineParams {
  const void* o_part;          // [n][R * Hl][D] local partial (this rank's slice of the context, all heads of the group)
  const float* lse_part;       // [n][R * Hl] natural-log LSE, -inf = no token on this rank
  void* const* recv_o_ptrs;    // per rank: [R senders][cap][Hl][D]
  float* const* recv_lse_ptrs; // per rank: [R senders][cap][Hl]
  int64_t recv_o_slot, recv_lse_slot; // elements per slot: the kernel alternates between two slots by launch parity, so a
                                      // rank that runs one launch ahead never writes the slot a peer is still merging
  int* const* flag_ptrs;
  int64_t flag_off;
  int* epoch;
  void* out;                   // [n][Hl][D]
  int64_t out_stride_t, out_stride_h;
  int n, cap, Hl, D, rank, R;
};

template <int DT>
__global__ void __launch_bounds__(256) peer_combine_kernel(const CombineParams p) {
  using E = Elem<DT>;
  __shared__ int s_epoch;
  pdl_launch_dependents();
  pdl_wait();   // the local partial comes from the decode kernel before us
  const int cta = blockIdx.x;
  if (threadIdx.x == 0) s_epoch = ++p.epoch[cta];
  __syncthreads();
  const int epoch = s_epoch;
  const int64_t recv_o_off = (int64_t)(epoch & 1) * p.recv_o_slot, recv_lse_off = (int64_t)(epoch & 1) * p.recv_lse_slot;
  const int per = (p.n + gridDim.x - 1) / gridDim.x;
  const int t0 = cta * per, t1 = min(p.n, t0 + per);
  const int nt = max(0, t1 - t0);
  const int vec = p.D >> 3;
  const int Hg = p.Hl * p.R;
  // ---- push: for every peer j, my partial of ITS heads ----
  for (int j = 0; j < p.R; ++j) {
    if (j == p.rank) continue;
    uint16_t* ro = reinterpret_cast<uint16_t*>(p.recv_o_ptrs[j]) + recv_o_off + (int64_t)p.rank * p.cap * p.Hl * p.D;
    float* rl = p.recv_lse_ptrs[j] + recv_lse_off + (int64_t)p.rank * p.cap * p.Hl;
    for (int i = threadIdx.x; i < nt * p.Hl * vec; i += blockDim.x) {
      const int v = i % vec, h = (i / vec) % p.Hl, t = t0 + i / (vec * p.Hl);
      const uint4 x = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.o_part) + ((int64_t)t * Hg + j * p.Hl + h) * p.D + v * 8);
      *reinterpret_cast<uint4*>(ro + ((int64_t)t * p.Hl + h) * p.D + v * 8) = x;
    }
    for (int i = threadIdx.x; i < nt * p.Hl; i += blockDim.x) {
      const int h = i % p.Hl, t = t0 + i / p.Hl;
      rl[(int64_t)t * p.Hl + h] = p.lse_part[(int64_t)t * Hg + j * p.Hl + h];
    }
  }
  signal_and_wait(p.flag_ptrs, p.flag_off, p.rank, p.R, cta, epoch);
  // ---- merge my heads: own slice (local) + what the peers pushed (local receive buffer) ----
  const uint16_t* my_o = reinterpret_cast<const uint16_t*>(p.recv_o_ptrs[p.rank]) + recv_o_off;
  const float* my_l = p.recv_lse_ptrs[p.rank] + recv_lse_off;
  for (int i = threadIdx.x; i < nt * p.Hl * vec; i += blockDim.x) {
    const int v = i % vec, h = (i / vec) % p.Hl, t = t0 + i / (vec * p.Hl);
    float lse[kPeerMaxRanks];
    uint4 ov[kPeerMaxRanks];
#pragma unroll
    for (int j = 0; j < kPeerMaxRanks; ++j) {
      if (j < p.R) {
        if (j == p.rank) {
          lse[j] = p.lse_part[(int64_t)t * Hg + j * p.Hl + h];
          ov[j] = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint16_t*>(p.o_part) + ((int64_t)t * Hg + j * p.Hl + h) * p.D + v * 8);
        } else {
          lse[j] = ld_cv_f32(my_l + ((int64_t)j * p.cap + t) * p.Hl + h);
          ov[j] = ld_cv_v4(my_o + (((int64_t)j * p.cap + t) * p.Hl + h) * p.D + v * 8);
        }
      }
    }
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < kPeerMaxRanks; ++j)
      if (j < p.R) mx = fmaxf(mx, lse[j]);
    float den = 0.f, acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
#pragma unroll
    for (int j = 0; j < kPeerMaxRanks; ++j) {
      if (j < p.R) {
        const float wgt = (lse[j] == -INFINITY) ? 0.f : __expf(lse[j] - mx);
        den += wgt;
        const typename E::type* e = reinterpret_cast<const typename E::type*>(&ov[j]);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += wgt * E::to_float(e[k]);
      }
    }
    uint4 res;
    typename E::type* r = reinterpret_cast<typename E::type*>(&res);
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = E::from_float(den > 0.f ? acc[k] / den : 0.f);
    *reinterpret_cast<uint4*>(reinterpret_cast<typename E::type*>(p.out) + (int64_t)t * p.out_stride_t + (int64_t)h * p.out_stride_h + v * 8) = res;
  }
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

extern "C" int b200attn_peer_gather_heads(const void* local, int64_t stride_t, int64_t stride_h, int32_t n, int32_t heads_local,
                                          int32_t head_dim, int32_t dtype, void* const* gather_ptrs, int64_t gather_offset,
                                          int32_t* const* flag_ptrs, int64_t flag_offset, int32_t* epoch, int32_t rank,
                                          int32_t num_ranks, void* stream) {
  if (n == 0) return B200ATTN_OK;
  if (!local || !gather_ptrs || !flag_ptrs || !epoch) {
    set_error("peer_gather_heads: null pointer argument");
    return B200ATTN_EINVAL;
  }
  int rc = peer_args_ok("peer_gather_heads", n, heads_local, head_dim, dtype, rank, num_ranks);
  if (rc) return rc;
  if ((reinterpret_cast<uintptr_t>(local) % 16) || (stride_t % 8) || (stride_h % 8) || (gather_offset % 8)) {
    set_error("peer_gather_heads: pointers, strides and offsets must be 16-byte aligned");
    return B200ATTN_EINVAL;
  }
  GatherParams p;
  p.local = local; p.stride_t = stride_t; p.stride_h = stride_h;
  p.gather_ptrs = gather_ptrs; p.gather_off = gather_offset;
  p.flag_ptrs = flag_ptrs; p.flag_off = flag_offset; p.epoch = epoch;
  p.n = n; p.Hl = heads_local; p.D = head_dim; p.rank = rank; p.R = num_ranks;
  void* args[] = {&p};
  return launch_with_pdl((const void*)peer_gather_heads_kernel, dim3(peer_grid(n)), dim3(256), 0, static_cast<cudaStream_t>(stream), args);
}

extern "C" int b200attn_peer_combine(const void* o_part, const float* lse_part, int32_t n, int32_t rows_capacity,
                                     int32_t heads_local, int32_t head_dim, int32_t dtype, void* const* recv_o_ptrs,
                                     int64_t recv_o_slot_elems, float* const* recv_lse_ptrs, int64_t recv_lse_slot_elems,
                                     int32_t* const* flag_ptrs, int64_t flag_offset, int32_t* epoch, int32_t rank,
                                     int32_t num_ranks, void* out, int64_t out_stride_t, int64_t out_stride_h, void* stream) {
  if (n == 0) return B200ATTN_OK;
  if (!o_part || !lse_part || !recv_o_ptrs || !recv_lse_ptrs || !flag_ptrs || !epoch || !out) {
    set_error("peer_combine: null pointer argument");
    return B200ATTN_EINVAL;
  }
  int rc = peer_args_ok("peer_combine", n, heads_local, head_dim, dtype, rank, num_ranks);
  if (rc) return rc;
  if (n > rows_capacity) {
    set_error("peer_combine: %d rows exceed the receive buffers' capacity %d", n, rows_capacity);
    return B200ATTN_ENOSPACE;
  }
  if ((reinterpret_cast<uintptr_t>(o_part) % 16) || (reinterpret_cast<uintptr_t>(out) % 16) || (out_stride_t % 8) || (out_stride_h % 8) ||
      (recv_o_slot_elems % 8)) {
    set_error("peer_combine: pointers, strides and offsets must be 16-byte aligned");
    return B200ATTN_EINVAL;
  }
  CombineParams p;
  p.o_part = o_part; p.lse_part = lse_part;
  p.recv_o_ptrs = recv_o_ptrs; p.recv_lse_ptrs = recv_lse_ptrs; p.recv_o_slot = recv_o_slot_elems; p.recv_lse_slot = recv_lse_slot_elems;
  p.flag_ptrs = flag_ptrs; p.flag_off = flag_offset; p.epoch = epoch;
  p.out = out; p.out_stride_t = out_stride_t; p.out_stride_h = out_stride_h;
  p.n = n; p.cap = rows_capacity; p.Hl = heads_local; p.D = head_dim; p.rank = rank; p.R = num_ranks;
  void* args[] = {&p};
  const void* fn = dtype == B200ATTN_BF16 ? (const void*)peer_combine_kernel<B200ATTN_BF16> : (const void*)peer_combine_kernel<B200ATTN_FP16>;
  return launch_with_pdl(fn, dim3(peer_grid(n)), dim3(256), 0, static_cast<cudaStream_t>(stream), args);
}
