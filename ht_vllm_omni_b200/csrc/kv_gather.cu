// kv_gather.cu — KV hand-off gather (the step after the path): one request's blocks out of EVERY
// layer's paged cache into one contiguous [layer][K|V][token][head][dim] buffer, in one launch.
//
// Replaces the per-layer Python loop of OmniKVTransferManager._extract_kv_cache
//   ($REF/vllm_omni/distributed/omni_connectors/kv_transfer_manager.py:224-301):
//     selected = blocks[valid_ids]; flat = selected.flatten(0, 1)[:seq_len]; flat.cpu().contiguous()
// i.e. token i of the result is slot i % block_size of block_ids[i / block_size], for i < min(seq_len, n_ids * block_size).
// Bytes are moved raw (an fp8 cache stays e4m3 bytes; scales travel in the metadata, as in the reference).  The host side
// filters out-of-range block ids exactly as the reference does (:269-270); the kernel still never dereferences an id outside
// [0, num_blocks) -- such a token is zero-filled.
//
// Pure copy, HBM-bound: one thread moves 16 bytes; a (token, head) row of D elements is read as one contiguous run from the
// cache (any layout: strides are parameters) and written contiguously.
#include "common.cuh"

namespace b200attn {

struct KvGatherParams {
  const void* const* k_planes;   // device array [num_layers] of K-plane base pointers
  const void* const* v_planes;
  int64_t stride_block, stride_slot, stride_head;   // BYTES
  int64_t num_blocks;
  const int32_t* block_ids;
  uint8_t* out;                  // [num_layers][2][n_tokens][Hkv][row_bytes]
  int num_layers, n_tokens, Hkv, block_size, row_bytes;
  FastDiv d_vec, d_heads, d_bs;   // vectors per row, kv heads, block size
};

__global__ void __launch_bounds__(256) kv_gather_kernel(const KvGatherParams p) {
  pdl_wait();
  pdl_launch_dependents();
  // blockIdx.y = layer * 2 + plane; blockIdx.x covers the plane's n_tokens * Hkv * (row_bytes / 16) vectors in 32-bit math
  // (the first version did three 64-bit divisions per 16 bytes and was issue-bound at 75 %: profiles/r01_ncu_kv_gather_v8.txt)
  const uint32_t vec_per_row = (uint32_t)p.row_bytes >> 4;
  const uint32_t per_plane = (uint32_t)p.n_tokens * p.Hkv * vec_per_row;
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= per_plane) return;
  uint32_t rowi, c, tok, h, bi, slot;
  p.d_vec.divmod(idx, rowi, c);
  p.d_heads.divmod(rowi, tok, h);
  const uint32_t kv = blockIdx.y & 1, layer = blockIdx.y >> 1;
  p.d_bs.divmod(tok, bi, slot);
  const int32_t blk = __ldg(p.block_ids + bi);
  uint4 v = make_uint4(0u, 0u, 0u, 0u);
  if (blk >= 0 && blk < p.num_blocks) {
    const uint8_t* base = reinterpret_cast<const uint8_t*>(kv ? p.v_planes[layer] : p.k_planes[layer]);
    v = __ldg(reinterpret_cast<const uint4*>(base + blk * p.stride_block + slot * p.stride_slot + h * p.stride_head) + c);
  }
  const uint64_t gid = (uint64_t)blockIdx.y * per_plane + idx;
  reinterpret_cast<uint4*>(p.out)[gid] = v;
}

int launch_with_pdl(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args);

}  // namespace b200attn

using namespace b200attn;

extern "C" int b200attn_kv_gather(const void* const* k_planes, const void* const* v_planes, int32_t num_layers,
                                  int64_t stride_block_bytes, int64_t stride_slot_bytes, int64_t stride_head_bytes,
                                  int64_t num_blocks, const int32_t* block_ids, int32_t n_tokens, int32_t num_kv_heads,
                                  int32_t block_size, int32_t row_bytes, void* out, void* stream) {
  if (n_tokens == 0 || num_layers == 0) return B200ATTN_OK;
  if (!k_planes || !v_planes || !block_ids || !out) {
    set_error("kv_gather: null pointer argument");
    return B200ATTN_EINVAL;
  }
  if (n_tokens < 0 || num_layers < 0 || num_kv_heads <= 0 || block_size <= 0 || row_bytes <= 0 || (row_bytes % 16) != 0) {
    set_error("kv_gather: bad shape layers=%d tokens=%d Hkv=%d block_size=%d row_bytes=%d (row must be a multiple of 16 bytes)",
              num_layers, n_tokens, num_kv_heads, block_size, row_bytes);
    return B200ATTN_EINVAL;
  }
  if ((stride_block_bytes % 16) || (stride_slot_bytes % 16) || (stride_head_bytes % 16) || (reinterpret_cast<uintptr_t>(out) % 16)) {
    set_error("kv_gather: strides and the output must be 16-byte aligned");
    return B200ATTN_EINVAL;
  }
  KvGatherParams p;
  p.k_planes = k_planes; p.v_planes = v_planes;
  p.stride_block = stride_block_bytes; p.stride_slot = stride_slot_bytes; p.stride_head = stride_head_bytes;
  p.num_blocks = num_blocks; p.block_ids = block_ids; p.out = reinterpret_cast<uint8_t*>(out);
  p.num_layers = num_layers; p.n_tokens = n_tokens; p.Hkv = num_kv_heads; p.block_size = block_size; p.row_bytes = row_bytes;
  p.d_vec.init((uint32_t)row_bytes / 16); p.d_heads.init((uint32_t)num_kv_heads); p.d_bs.init((uint32_t)block_size);
  const int64_t per_plane = (int64_t)n_tokens * num_kv_heads * (row_bytes / 16);
  const int64_t gx = (per_plane + 255) / 256;
  if (per_plane >= (1ll << 32) || gx >= (1ll << 31) || 2ll * num_layers > 65535) {
    set_error("kv_gather: %lld vectors per plane / %d layers exceed one launch; split the request", (long long)per_plane, num_layers);
    return B200ATTN_EINVAL;
  }
  void* args[] = {&p};
  return launch_with_pdl((const void*)kv_gather_kernel, dim3((unsigned)gx, (unsigned)(2 * num_layers)), dim3(256), 0, static_cast<cudaStream_t>(stream), args);
}
