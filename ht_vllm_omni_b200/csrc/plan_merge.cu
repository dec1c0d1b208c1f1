// plan_merge.cu — the split-KV work plan and Kernel 3 (merge of partial online-softmax results).
//
// Reference semantics of the merge: reduce_segments ($VLLM/v1/attention/ops/triton_unified_attention.py:645-733):
//   m = max_i m_i;  l = sum_i l_i * e^(m_i - m);  o = sum_i o_i * e^(m_i - m) / l
// and merge_attn_states ($VLLM/v1/attention/ops/merge_attn_states.py:7-101) for the LSE form.
// Ours works in the log2 domain (m_i are log2-scaled logits) and handles a variable number of partials per
// (sequence, kv head), located through the plan instead of a fixed segment count.
#include "common.cuh"

namespace b200attn {

// ------------------------------------------------------------------------------------------------
// plan: one CTA.  tile_prefix = exclusive scan of ceil(L_s / 64); flat work index x in
// [0, Hv * tile_prefix[B]) enumerates (s, hv, tile) with tile fastest, then hv, then s.
// CTA c owns x in [c * tpc, min((c + 1) * tpc, total)).
// ------------------------------------------------------------------------------------------------
constexpr int kPlanSmemSeqs = 4096;   // tile prefixes of up to this many rows are searched in shared memory

__global__ void __launch_bounds__(256) plan_kernel(const int32_t* __restrict__ seq_lens, int B, int Hv, int grid,
                                                   const int32_t* __restrict__ block_table, int64_t bt_stride,
                                                   const int32_t* __restrict__ row_map, int cpb, int32_t* __restrict__ plan) {
  __shared__ int warp_sums[8];
  __shared__ int carry_s;
  __shared__ int prefix_s[kPlanSmemSeqs + 1];
  __shared__ int4 ent_s[kMaxGrid];          // (first chunk, seq_len, block-table row, live) of every CTA, for the id gather
  // No early griddepcontrol.launch_dependents here: kernels that follow may read the plan BEFORE their own
  // griddepcontrol.wait (decode prefetches K/V tiles early), so they must not start until this grid has completed.
  pdl_wait();
  int32_t* prefix = plan + kPlanPrefixOffset;
  const bool in_smem = B <= kPlanSmemSeqs;   // the binary searches below then never touch global memory
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (int i = tid; i < kPlanCounterWords / 4; i += 256) reinterpret_cast<int4*>(plan + kPlanCounterOffset)[i] = make_int4(0, 0, 0, 0);
  if (tid == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < B; base += 256) {
    const int i = base + tid;
    int nt = 0;
    if (i < B) {
      const int L = seq_lens[i];
      nt = L > 0 ? (L + kTileTokens - 1) / kTileTokens : 0;
    }
    int v = nt;  // inclusive warp scan
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
    if (i < B) {
      prefix[i] = carry + woff + v - nt;  // exclusive
      if (in_smem) prefix_s[i] = carry + woff + v - nt;
    }
    __syncthreads();
    if (tid == 255) carry_s = carry + woff + v;
    __syncthreads();
  }
  const int total_seq_tiles = carry_s;
  if (tid == 0) {
    prefix[B] = total_seq_tiles;
    if (in_smem) prefix_s[B] = total_seq_tiles;
  }
  const long long total = (long long)total_seq_tiles * Hv;
  const int tpc = total > 0 ? (int)((total + grid - 1) / grid) : 1;
  if (tid == 0) {
    plan[0] = (int)total;
    plan[1] = tpc;
    plan[2] = grid;
    plan[3] = B;
    plan[4] = Hv;
    plan[5] = cpb;
    plan[6] = plan[7] = 0;
  }
  __syncthreads();  // prefix[] (global / shared) written by this CTA is visible to it after the barrier
  const int* pf = in_smem ? prefix_s : prefix;
  int32_t* ent = plan + kPlanHeaderWords;
  for (int c = tid; c < grid; c += 256) {
    const long long x = (long long)c * tpc;
    int4 e = make_int4(-1, 0, 0, 0);
    int L = 0, row = 0, pfx = 0;
    if (x < total) {
      // largest s with Hv * prefix[s] <= x  (prefix is non-decreasing; empty sequences repeat a value,
      // taking the largest index lands on the sequence that owns x)
      int lo = 0, hi = B;  // invariant: Hv*prefix[lo] <= x < Hv*prefix[hi]
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if ((long long)pf[mid] * Hv <= x) lo = mid; else hi = mid;
      }
      const int s = lo;
      pfx = pf[s];
      const int nt = pf[s + 1] - pfx;
      const long long r = x - (long long)pfx * Hv;
      e.x = s;
      e.y = (int)(r / nt);
      e.z = (int)(r % nt);
      const long long left = total - x;
      e.w = (int)(left < tpc ? left : tpc);
      L = max(seq_lens[s], 0);
      row = row_map ? row_map[s] : s;
    }
    int32_t* d = ent + (size_t)c * kPlanEntryWords;
    *reinterpret_cast<int4*>(d) = e;
    *reinterpret_cast<int4*>(d + 4) = make_int4(L, row, pfx, tpc);
    *reinterpret_cast<int4*>(d + 8) = make_int4(grid, B, 0, 0);
    ent_s[c] = make_int4(e.z * (kTileTokens / kChunkTokens), L, row, e.x >= 0 ? 1 : 0);
  }
  __syncthreads();
  // block ids of the first 32 16-token chunks of every CTA's first segment: the decode producer issues its first TMA
  // loads after ONE dependent load (this entry) instead of plan -> seq_lens -> row_map -> block_table.  One warp per CTA
  // entry and pass: 32 independent block-table loads, one coalesced 128-byte store.
  for (int c = wid; c < grid; c += 8) {
    const int4 e = ent_s[c];
    int id = 0;
    if (e.w) {
      const int ci = e.x + lane;
      if (ci * kChunkTokens < e.y) id = block_table[(int64_t)e.z * bt_stride + ci / cpb];
    }
    ent[(size_t)c * kPlanEntryWords + kPlanEntryIds + lane] = id;
  }
}

// ------------------------------------------------------------------------------------------------
// merge: one CTA per (sequence, virtual kv head) pair; G x D outputs.
// ------------------------------------------------------------------------------------------------
struct MergeParams {
  const int32_t* plan;
  const float* ws_o;
  const float* ws_ml;
  void* out;
  int64_t out_stride_t, out_stride_h;
  const float* v_scale;
  float* lse_out;
  int scale_stride;
  int Hq, G, gsplit, D;
};

// One CTA per (pair, head): the partials of a pair can be many (a long sequence in a small batch is cut across the whole
// grid), so pass 1 spreads them over all 128 threads and pass 2 over the 4 warps, each lane owning 4 head dims (float4,
// 8 independent loads in flight per thread).
template <int ODT>
__global__ void __launch_bounds__(128) merge_kernel(const MergeParams p) {
  extern __shared__ float wts[];  // [n_parts] normalised weights of this head
  __shared__ float red[4];
  __shared__ float4 osum[4][32];
  pdl_wait();
  pdl_launch_dependents();
  const int32_t* plan = p.plan;
  const int total = plan[0], tpc = plan[1], Hv = plan[4];
  if (total <= 0) return;
  const int pair = blockIdx.x, head = blockIdx.y;
  const int s = pair / Hv, hv = pair - s * Hv;
  const int32_t* prefix = plan + kPlanPrefixOffset;
  const int nt = prefix[s + 1] - prefix[s];
  if (nt == 0) return;
  const long long x0 = (long long)prefix[s] * Hv + (long long)hv * nt;
  const long long x1 = x0 + nt;
  const int c_lo = (int)(x0 / tpc), c_hi = (int)((x1 - 1) / tpc);
  if (c_lo == c_hi) return;  // whole pair lived in one CTA: decode already wrote the final output
  const int np = c_hi - c_lo + 1;
  const int slot_lo = (x0 == (long long)c_lo * tpc) ? 0 : 1;  // later CTAs start on this pair: slot 0
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  auto pidx_of = [&](int i) { return ((size_t)(c_lo + i) * 2 + (i == 0 ? slot_lo : 0)) * kMaxGroup + head; };

  // pass 1: global max / sum over the partials (block reduction), normalised weights to smem
  float mg = -INFINITY;
  for (int i = tid; i < np; i += 128) mg = fmaxf(mg, p.ws_ml[pidx_of(i) * 2]);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, off));
  if (lane == 0) red[wid] = mg;
  __syncthreads();
  mg = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  __syncthreads();
  float lg = 0.f;
  for (int i = tid; i < np; i += 128) {
    const float2 ml = *reinterpret_cast<const float2*>(p.ws_ml + pidx_of(i) * 2);
    const float wgt = fast_exp2(ml.x - mg);
    wts[i] = wgt;
    lg += ml.y * wgt;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) lg += __shfl_xor_sync(0xffffffffu, lg, off);
  if (lane == 0) red[wid] = lg;
  __syncthreads();
  lg = (red[0] + red[1]) + (red[2] + red[3]);
  const int kvh = hv / p.gsplit;
  const float vscale = p.v_scale ? p.v_scale[(int64_t)kvh * p.scale_stride] : 1.0f;
  const float onorm = vscale / lg;

  // pass 2: weighted sum; warp w folds partials w, w+4, ... in index order, lane = 4 consecutive head dims
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
  for (int i = wid; i < np; i += 4) {
    const float w = wts[i];
    const float4 v = __ldcg(reinterpret_cast<const float4*>(p.ws_o + pidx_of(i) * p.D) + lane);
    acc.x += w * v.x; acc.y += w * v.y; acc.z += w * v.z; acc.w += w * v.w;
  }
  osum[wid][lane] = acc;
  __syncthreads();
  const float* os = reinterpret_cast<const float*>(&osum[0][0]);
  const int d = tid;
  const float o = (os[d] + os[128 + d]) + (os[256 + d] + os[384 + d]);
  using E = Elem<ODT>;
  const int hq = hv * p.G + head;
  typename E::type* op = reinterpret_cast<typename E::type*>(p.out) + (int64_t)s * p.out_stride_t + (int64_t)hq * p.out_stride_h;
  op[d] = E::from_float(o * onorm);
  if (d == 0 && p.lse_out) p.lse_out[(int64_t)s * p.Hq + hq] = (mg + fast_log2(lg)) * 0.6931471805599453f;
}

// ------------------------------------------------------------------------------------------------
// LSE merge of P full attention results (sequence-split exchange across shards / ranks).
// One warp per (row, head); lanes stride the head dim.
// ------------------------------------------------------------------------------------------------
template <int DT>
__global__ void __launch_bounds__(128) merge_lse_kernel(const void* parts_o, const float* parts_lse, int P,
                                                        int64_t stride_o, int64_t stride_lse, void* out, float* lse_out,
                                                        int64_t rows /* n * Hq */, int D) {
  pdl_wait();
  pdl_launch_dependents();
  using E = Elem<DT>;
  const int64_t row = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float mx = -INFINITY;
  for (int i = 0; i < P; ++i) mx = fmaxf(mx, parts_lse[i * stride_lse + row]);
  float den = 0.f;
  for (int i = 0; i < P; ++i) {
    const float l = parts_lse[i * stride_lse + row];
    den += (l == -INFINITY) ? 0.f : __expf(l - mx);
  }
  const typename E::type* po = reinterpret_cast<const typename E::type*>(parts_o);
  typename E::type* oo = reinterpret_cast<typename E::type*>(out);
  for (int d = lane; d < D; d += 32) {
    float acc = 0.f;
    for (int i = 0; i < P; ++i) {
      const float l = parts_lse[i * stride_lse + row];
      const float wgt = (l == -INFINITY) ? 0.f : __expf(l - mx);
      acc += wgt * E::to_float(po[i * stride_o + row * D + d]);
    }
    oo[row * D + d] = E::from_float(den > 0.f ? acc / den : 0.f);
  }
  if (lane == 0 && lse_out) lse_out[row] = den > 0.f ? mx + __logf(den) : -INFINITY;
}

// Same combine with the P parts living in P different buffers -- the ranks' symmetric-memory buffers of a sequence-split
// group, read straight over NVLink (peer loads bypass the caches: the buffers are rewritten every other layer).  One warp
// per (row, head); lanes stride the head dim.
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

// Half a warp per (row, head): 16 lanes x 8 head dims (one 16-byte peer load per part and lane); all of a row's loads are
// issued before anything is consumed, so the merge costs two NVLink round trips (lse, then o) whatever P is.  D must be 128.
constexpr int kMaxPeerParts = 8;
template <int DT>
__global__ void __launch_bounds__(128) merge_lse_peers_kernel(const void* const* o_ptrs, const float* const* lse_ptrs, int P,
                                                              int64_t o_off, int64_t lse_off, void* out, float* lse_out,
                                                              int64_t rows, int D) {
  pdl_wait();
  pdl_launch_dependents();
  using E = Elem<DT>;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 4);
  if (row >= rows) return;
  const int l16 = threadIdx.x & 15;
  float lse[kMaxPeerParts];
  uint4 ov[kMaxPeerParts];
#pragma unroll
  for (int i = 0; i < kMaxPeerParts; ++i) {
    if (i < P) {
      lse[i] = __ldcv(lse_ptrs[i] + lse_off + row);
      ov[i] = ld_peer_v4(reinterpret_cast<const typename E::type*>(o_ptrs[i]) + o_off + row * D + l16 * 8);
    }
  }
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < kMaxPeerParts; ++i)
    if (i < P) mx = fmaxf(mx, lse[i]);
  float den = 0.f, acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxPeerParts; ++i) {
    if (i < P) {
      const float wgt = (lse[i] == -INFINITY) ? 0.f : __expf(lse[i] - mx);
      den += wgt;
      const typename E::type* e = reinterpret_cast<const typename E::type*>(&ov[i]);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += wgt * E::to_float(e[k]);
    }
  }
  uint4 res;
  typename E::type* r = reinterpret_cast<typename E::type*>(&res);
#pragma unroll
  for (int k = 0; k < 8; ++k) r[k] = E::from_float(den > 0.f ? acc[k] / den : 0.f);
  *reinterpret_cast<uint4*>(reinterpret_cast<typename E::type*>(out) + row * D + l16 * 8) = res;
  if (l16 == 0 && lse_out) lse_out[row] = den > 0.f ? mx + __logf(den) : -INFINITY;
}

int launch_with_pdl(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args);

static int split_group(int Hq, int Hkv, int* G, int* gsplit) {
  if (Hkv <= 0 || Hq <= 0 || Hq % Hkv) return -1;
  const int Gfull = Hq / Hkv;
  int gs = 1;
  while (gs <= Gfull && (Gfull / gs > kMaxGroup || (Gfull % gs) != 0)) ++gs;
  if (gs > Gfull) return -1;
  *G = Gfull / gs;
  *gsplit = gs;
  return 0;
}

}  // namespace b200attn

using namespace b200attn;

extern "C" size_t b200attn_plan_bytes(int32_t max_num_seqs) {
  if (max_num_seqs < 0) max_num_seqs = 0;
  return ((size_t)kPlanPrefixOffset + (size_t)max_num_seqs + 1 + 3) / 4 * 4 * sizeof(int32_t);
}

extern "C" size_t b200attn_workspace_bytes(int32_t head_dim, int32_t num_ctas) {
  if (num_ctas <= 0 || num_ctas > kMaxGrid) num_ctas = kMaxGrid;
  return (ws_o_floats(num_ctas, head_dim) + ws_ml_floats(num_ctas)) * sizeof(float);
}

extern "C" int b200attn_decode_plan(const int32_t* seq_lens, int32_t num_seqs, int32_t num_q_heads, int32_t num_kv_heads,
                                    int32_t num_ctas, const int32_t* block_table, int64_t block_table_stride,
                                    const int32_t* row_map, int32_t block_size, void* plan, size_t plan_bytes, void* stream) {
  if (!seq_lens || !plan || !block_table) {
    set_error("decode_plan: null pointer argument");
    return B200ATTN_EINVAL;
  }
  int G, gsplit;
  if (num_seqs < 0 || split_group(num_q_heads, num_kv_heads, &G, &gsplit)) {
    set_error("decode_plan: bad shape B=%d Hq=%d Hkv=%d", num_seqs, num_q_heads, num_kv_heads);
    return B200ATTN_EINVAL;
  }
  if (num_ctas <= 0 || num_ctas > kMaxGrid) {
    set_error("decode_plan: num_ctas %d out of range (1..%d)", num_ctas, kMaxGrid);
    return B200ATTN_EINVAL;
  }
  if (block_size <= 0 || block_size % kChunkTokens != 0) {
    set_error("decode_plan: block_size %d must be a positive multiple of %d", block_size, kChunkTokens);
    return B200ATTN_EINVAL;
  }
  if (plan_bytes < b200attn_plan_bytes(num_seqs)) {
    set_error("decode_plan: plan buffer too small (%zu < %zu)", plan_bytes, b200attn_plan_bytes(num_seqs));
    return B200ATTN_ENOSPACE;
  }
  int B = num_seqs, Hv = num_kv_heads * gsplit, grid = num_ctas, cpb = block_size / kChunkTokens;
  int32_t* pl = reinterpret_cast<int32_t*>(plan);
  void* args[] = {(void*)&seq_lens, &B, &Hv, &grid, (void*)&block_table, &block_table_stride, (void*)&row_map, &cpb, &pl};
  return launch_with_pdl((const void*)plan_kernel, dim3(1), dim3(256), 0, static_cast<cudaStream_t>(stream), args);
}

extern "C" int b200attn_merge(const void* plan, int32_t num_ctas, const void* workspace, void* out, int64_t out_stride_t,
                              int64_t out_stride_h, int32_t num_seqs, int32_t num_q_heads, int32_t num_kv_heads,
                              int32_t head_dim, int32_t out_dtype, const float* v_scale, int32_t scale_stride,
                              float* lse_out, void* stream) {
  if (num_seqs == 0) return B200ATTN_OK;
  if (!plan || !workspace || !out) {
    set_error("merge: null pointer argument");
    return B200ATTN_EINVAL;
  }
  int G, gsplit;
  if (split_group(num_q_heads, num_kv_heads, &G, &gsplit)) {
    set_error("merge: bad head counts Hq=%d Hkv=%d", num_q_heads, num_kv_heads);
    return B200ATTN_EINVAL;
  }
  if (num_ctas <= 0 || num_ctas > kMaxGrid) {
    set_error("merge: num_ctas %d out of range", num_ctas);
    return B200ATTN_EINVAL;
  }
  if (out_dtype != B200ATTN_BF16 && out_dtype != B200ATTN_FP16) {
    set_error("merge: out dtype must be bf16 or fp16");
    return B200ATTN_EINVAL;
  }
  if (head_dim != 128) {
    set_error("merge: head_dim %d unsupported (only 128)", head_dim);
    return B200ATTN_EINVAL;
  }
  MergeParams p;
  p.plan = reinterpret_cast<const int32_t*>(plan);
  p.ws_o = reinterpret_cast<const float*>(workspace);
  p.ws_ml = p.ws_o + ws_o_floats(num_ctas, head_dim);
  p.out = out; p.out_stride_t = out_stride_t; p.out_stride_h = out_stride_h;
  p.v_scale = v_scale; p.scale_stride = scale_stride; p.lse_out = lse_out;
  p.Hq = num_q_heads; p.G = G; p.gsplit = gsplit; p.D = head_dim;
  const int Hv = num_kv_heads * gsplit;
  const size_t smem = (size_t)num_ctas * sizeof(float);  // a pair spans at most num_ctas partials
  const void* fn = out_dtype == B200ATTN_BF16 ? (const void*)merge_kernel<B200ATTN_BF16> : (const void*)merge_kernel<B200ATTN_FP16>;
  if (smem > 48 * 1024) {
    if (check_cuda(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "cudaFuncSetAttribute(merge smem)"))
      return B200ATTN_ECUDA;
  }
  void* args[] = {&p};
  return launch_with_pdl(fn, dim3((unsigned)(num_seqs * Hv), (unsigned)G), dim3(128), smem, static_cast<cudaStream_t>(stream), args);
}

extern "C" int b200attn_merge_lse(const void* parts_o, const float* parts_lse, int32_t num_parts, int64_t part_stride_o,
                                  int64_t part_stride_lse, void* out, float* lse_out, int32_t n, int32_t num_q_heads,
                                  int32_t head_dim, int32_t dtype, void* stream) {
  if (n == 0) return B200ATTN_OK;
  if (!parts_o || !parts_lse || !out || num_parts <= 0) {
    set_error("merge_lse: null pointer or no parts");
    return B200ATTN_EINVAL;
  }
  if (dtype != B200ATTN_BF16 && dtype != B200ATTN_FP16) {
    set_error("merge_lse: dtype must be bf16 or fp16");
    return B200ATTN_EINVAL;
  }
  int64_t rows = (int64_t)n * num_q_heads;
  int P = num_parts, D = head_dim;
  const void* fn = dtype == B200ATTN_BF16 ? (const void*)merge_lse_kernel<B200ATTN_BF16> : (const void*)merge_lse_kernel<B200ATTN_FP16>;
  void* args[] = {(void*)&parts_o, (void*)&parts_lse, &P, &part_stride_o, &part_stride_lse, &out, &lse_out, &rows, &D};
  return launch_with_pdl(fn, dim3((unsigned)((rows + 3) / 4)), dim3(128), 0, static_cast<cudaStream_t>(stream), args);
}

extern "C" int b200attn_merge_lse_peers(const void* const* parts_o_ptrs, const float* const* parts_lse_ptrs, int32_t num_parts,
                                        int64_t o_offset, int64_t lse_offset, void* out, float* lse_out, int32_t n,
                                        int32_t num_q_heads, int32_t head_dim, int32_t dtype, void* stream) {
  if (n == 0) return B200ATTN_OK;
  if (!parts_o_ptrs || !parts_lse_ptrs || !out || num_parts <= 0) {
    set_error("merge_lse_peers: null pointer or no parts");
    return B200ATTN_EINVAL;
  }
  if (dtype != B200ATTN_BF16 && dtype != B200ATTN_FP16) {
    set_error("merge_lse_peers: dtype must be bf16 or fp16");
    return B200ATTN_EINVAL;
  }
  if (head_dim != 128 || num_parts > kMaxPeerParts) {
    set_error("merge_lse_peers: head_dim must be 128 and num_parts <= %d", kMaxPeerParts);
    return B200ATTN_EINVAL;
  }
  int64_t rows = (int64_t)n * num_q_heads;
  int P = num_parts, D = head_dim;
  const void* fn = dtype == B200ATTN_BF16 ? (const void*)merge_lse_peers_kernel<B200ATTN_BF16> : (const void*)merge_lse_peers_kernel<B200ATTN_FP16>;
  void* args[] = {(void*)&parts_o_ptrs, (void*)&parts_lse_ptrs, &P, &o_offset, &lse_offset, &out, &lse_out, &rows, &D};
  return launch_with_pdl(fn, dim3((unsigned)((rows + 7) / 8)), dim3(128), 0, static_cast<cudaStream_t>(stream), args);
}
