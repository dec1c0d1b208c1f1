// abi.cu — host-side plumbing of libb200attn: error strings, options, launch helper (programmatic dependent
// launch), TMA tensor-map encoding through the driver entry point (no link-time libcuda dependency, so the
// library loads on a machine without a driver and fails loudly only when a kernel is requested).
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "common.cuh"

namespace b200attn {

static thread_local char g_err[512] = "";
static int g_opt_pdl = 1;
static int g_opt_force_simt = 0;
static int g_opt_debug = 0;
static int g_opt_tc = 1;
static int g_opt_big_box = 1;       // tcgen05 decode: one TMA box per min(block_size, 64) tokens instead of one per 16
static int g_opt_fused_write = 1;   // b200attn_decode_kv_write writes the new token inside the decode kernel when it can
static int g_opt_cluster = 2;   // thread-block cluster size of the tcgen05 decode kernel (1 = no clusters)
static unsigned long long* g_dbg_buf = nullptr;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  (void)cudaGetLastError();  // clear the sticky-less error so later calls report their own
  return B200ATTN_ECUDA;
}

int option_force_simt() { return g_opt_force_simt; }
int option_debug() { return g_opt_debug; }
int option_tc() { return g_opt_tc; }
int option_cluster() { return g_opt_cluster; }
int option_fused_write() { return g_opt_fused_write; }
int option_big_box() { return g_opt_big_box; }
unsigned long long* option_debug_buffer() { return g_dbg_buf; }

int launch_with_pdl_cluster(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args, int cluster) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (g_opt_pdl) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = (unsigned)cluster;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return check_cuda(cudaLaunchKernelExC(&cfg, func, args), "cudaLaunchKernelExC");
}

int launch_with_pdl(const void* func, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, void** args) {
  return launch_with_pdl_cluster(func, grid, block, smem, stream, args, 1);
}

// ---- cuTensorMapEncodeTiled via the runtime's driver entry point lookup ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;
static cudaError_t g_encode_err = cudaSuccess;

static void resolve_encode() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  g_encode_err = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (g_encode_err == cudaSuccess && qres == cudaDriverEntryPointSuccess) g_encode = reinterpret_cast<EncodeTiledFn>(fn);
}

// K or V plane of the paged cache as a 4-D tensor (D, block_size, Hkv, num_blocks), box = 16 tokens x 128 bytes,
// 128-byte swizzle so that both the row-wise K reads and the column-gathering V reads are bank-conflict free.
int encode_kv_tensor_map(CUtensorMap* tm, const void* base, int kv_dtype, int64_t stride_block, int64_t stride_slot,
                         int64_t stride_head, int64_t num_blocks, int block_size, int Hkv, int D, int box_tokens) {
  std::call_once(g_encode_once, resolve_encode);
  if (!g_encode) {
    if (g_encode_err != cudaSuccess) return check_cuda(g_encode_err, "cudaGetDriverEntryPoint(cuTensorMapEncodeTiled)");
    set_error("cuTensorMapEncodeTiled not available from this driver");
    return B200ATTN_ECUDA;
  }
  const bool fp8 = kv_dtype == B200ATTN_FP8_E4M3;
  const int64_t es = fp8 ? 1 : 2;
  if (num_blocks <= 0) {
    set_error("decode: num_blocks must be positive");
    return B200ATTN_EINVAL;
  }
  if ((reinterpret_cast<uintptr_t>(base) % 16) || ((stride_block * es) % 16) || ((stride_slot * es) % 16) || ((stride_head * es) % 16)) {
    set_error("decode: cache base and strides must be 16-byte aligned for TMA");
    return B200ATTN_EINVAL;
  }
  cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)block_size, (cuuint64_t)Hkv, (cuuint64_t)num_blocks};
  cuuint64_t strides[3] = {(cuuint64_t)(stride_slot * es), (cuuint64_t)(stride_head * es), (cuuint64_t)(stride_block * es)};
  cuuint32_t box[4] = {(cuuint32_t)(fp8 ? 128 : 64), (cuuint32_t)box_tokens, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUtensorMapDataType dt = fp8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_UINT16;
  CUresult r = g_encode(tm, dt, 4, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (dims %d x %d x %d x %lld)", (int)r, D, block_size, Hkv,
              (long long)num_blocks);
    return B200ATTN_ECUDA;
  }
  return 0;
}

}  // namespace b200attn

using namespace b200attn;

extern "C" int b200attn_abi_version(void) { return B200ATTN_ABI_VERSION; }

extern "C" const char* b200attn_last_error(void) { return g_err; }

extern "C" int b200attn_sm_count(void) {
  static int cached = -1;
  if (cached >= 0) return cached;
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) {
    (void)cudaGetLastError();
    return 0;
  }
  cached = n;
  return n;
}

extern "C" int b200attn_default_ctas(int32_t kv_dtype) {
  const int sms = b200attn_sm_count();
  if (sms <= 0) return 0;
  (void)kv_dtype;
  int n = sms * 2;   // every variant is sized for two resident CTAs per SM (decode.cu Smem<>)
  return n > kMaxGrid ? kMaxGrid : n;
}

extern "C" int b200attn_set_option(const char* name, int32_t value) {
  if (!name) return B200ATTN_EINVAL;
  int* slot = nullptr;
  if (!strcmp(name, "pdl")) slot = &g_opt_pdl;
  else if (!strcmp(name, "force_simt")) slot = &g_opt_force_simt;
  else if (!strcmp(name, "debug")) slot = &g_opt_debug;
  else if (!strcmp(name, "tc")) slot = &g_opt_tc;
  else if (!strcmp(name, "cluster")) slot = &g_opt_cluster;
  else if (!strcmp(name, "fused_write")) slot = &g_opt_fused_write;
  else if (!strcmp(name, "big_box")) slot = &g_opt_big_box;
  if (!slot) {
    set_error("set_option: unknown option '%s'", name);
    return B200ATTN_EINVAL;
  }
  const int prev = *slot;
  *slot = value;
  return prev;
}

// debug only: device buffer of [num_ctas][8] uint64 that the decode kernel fills with %globaltimer stamps
extern "C" void b200attn_debug_set_timeline_buffer(void* dev_ptr) {
  g_dbg_buf = reinterpret_cast<unsigned long long*>(dev_ptr);
}
