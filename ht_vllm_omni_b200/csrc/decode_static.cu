// decode_static.cu — b200attn_decode_static: plan + split-KV paged decode for callers that may not allocate device memory
// (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The plan buffer and the split-KV workspace that b200attn_decode_plan / b200attn_decode
// take from the caller live in ONE file-scope device array of fixed size, held by every GPU that loads the library until
// the process exits.  The kernels are the ones behind b200attn_decode_plan and b200attn_decode (fused in-kernel merge);
// only where their scratch comes from differs.
//
// The array's name is a host-side stand-in in host code, not a device address: its device address is looked up ONCE PER
// DEVICE with that device current (cudaGetSymbolAddress) and that pointer is what the kernels receive.
// All calls share the array.  Nothing in it has to survive between calls: the plan kernel rewrites the plan entries and
// zeroes the merge counters on every call, and the workspace partials are written before they are read.  Two calls that
// overlap on the GPU (different streams, or graph branches without an edge between them) RACE on it.
#include "decode_common.cuh"

namespace b200attn {

constexpr size_t kStaticPlanBytes =
    (((size_t)kPlanPrefixOffset + (size_t)B200ATTN_STATIC_MAX_SEQS + 1 + 3) / 4 * 4 * sizeof(int32_t) + 255) / 256 * 256;
// ws_o_floats(ctas, kD) + ws_ml_floats(ctas) of common.cuh, spelled out as a constant expression (checked at run time below)
constexpr size_t kStaticWsBytes =
    ((size_t)B200ATTN_STATIC_MAX_CTAS * 2 * kMaxGroup * kD + (size_t)B200ATTN_STATIC_MAX_CTAS * 2 * kMaxGroup * 2) * sizeof(float);
constexpr size_t kStaticBytes = kStaticPlanBytes + kStaticWsBytes;
static_assert(kStaticBytes <= (size_t)6 << 20, "static scratch is documented as at most 6 MiB");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

// device address of g_static_scratch on the CURRENT device (looked up once per device)
static int static_scratch_ptr(unsigned char** out) {
  static unsigned char* addr[64] = {};
  int dev = 0;
  if (check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return B200ATTN_ECUDA;
  if (dev >= 0 && dev < 64 && addr[dev] != nullptr) {
    *out = addr[dev];
    return 0;
  }
  unsigned char* ptr = nullptr;
  if (check_cuda(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_static_scratch), "cudaGetSymbolAddress(static scratch)"))
    return B200ATTN_ECUDA;
  if (ptr == nullptr) {
    set_error("decode_static: cudaGetSymbolAddress returned a null address");
    return B200ATTN_ECUDA;
  }
  if (dev >= 0 && dev < 64) addr[dev] = ptr;
  *out = ptr;
  return 0;
}

}  // namespace b200attn

using namespace b200attn;

extern "C" int b200attn_decode_static(const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out, int64_t out_stride_t,
                                      int64_t out_stride_h, const void* k_cache, const void* v_cache, int64_t stride_block,
                                      int64_t stride_slot, int64_t stride_head, int64_t num_blocks, const int32_t* block_table,
                                      int64_t block_table_stride, const int32_t* seq_lens, int32_t num_seqs, int32_t num_q_heads,
                                      int32_t num_kv_heads, int32_t head_dim, int32_t block_size, int32_t q_dtype, int32_t kv_dtype,
                                      float sm_scale, const float* k_scale, const float* v_scale, int32_t scale_stride,
                                      float* lse_out, int32_t fresh_tokens, void* stream) {
  if (head_dim != kD) {
    set_error("decode_static: head_dim %d not supported (the static scratch is laid out for head_dim %d)", head_dim, kD);
    return B200ATTN_EINVAL;
  }
  if (num_seqs < 0 || num_seqs > B200ATTN_STATIC_MAX_SEQS) {
    set_error("decode_static: num_seqs %d exceeds the static scratch's maximum of %d rows (use b200attn_decode with a caller-owned plan "
              "and workspace for larger batches)", num_seqs, B200ATTN_STATIC_MAX_SEQS);
    return B200ATTN_ENOSPACE;
  }
  if (num_seqs == 0) return 0;
  int ctas = b200attn_default_ctas(kv_dtype);
  if (ctas <= 0) {
    set_error("decode_static: no CUDA device");
    return B200ATTN_ECUDA;
  }
  if (ctas > B200ATTN_STATIC_MAX_CTAS) ctas = B200ATTN_STATIC_MAX_CTAS;   // a smaller grid is still a correct one
  if (b200attn_workspace_bytes(kD, ctas) > kStaticWsBytes || b200attn_plan_bytes(num_seqs) > kStaticPlanBytes) {
    set_error("decode_static: internal size mismatch (workspace %zu > %zu or plan %zu > %zu)", b200attn_workspace_bytes(kD, ctas),
              kStaticWsBytes, b200attn_plan_bytes(num_seqs), kStaticPlanBytes);
    return B200ATTN_ENOSPACE;
  }
  unsigned char* scratch = nullptr;
  int rc = static_scratch_ptr(&scratch);
  if (rc) return rc;
  void* plan = scratch;
  void* workspace = scratch + kStaticPlanBytes;
  rc = b200attn_decode_plan(seq_lens, num_seqs, num_q_heads, num_kv_heads, ctas, block_table, block_table_stride, nullptr, block_size,
                            plan, kStaticPlanBytes, stream);
  if (rc) return rc;
  return b200attn_decode(q, q_stride_t, q_stride_h, out, out_stride_t, out_stride_h, k_cache, v_cache, stride_block, stride_slot,
                         stride_head, num_blocks, block_table, block_table_stride, seq_lens, num_seqs, num_q_heads, num_kv_heads,
                         head_dim, block_size, q_dtype, kv_dtype, sm_scale, k_scale, v_scale, scale_stride, plan, ctas, workspace,
                         kStaticWsBytes, lse_out, nullptr, fresh_tokens, B200ATTN_DECODE_FUSED_MERGE, stream);
}
