// Raw tcgen05.mma kind::f8f6f4 issue rate on one SM per shape / operand source (microbenchmark for DESIGN.md 3.5):
// one CTA per SM, one thread issues `iters` MMAs back to back into the same accumulator, clock64 around issue + completion.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build_variants/mma_rate tests/tools/cu/mma_rate.cu && build_variants/mma_rate
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo >> 4) & 0x3fffu) << 16) | ((uint64_t)((sbo >> 4) & 0x3fffu) << 32) |
         (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], [%1], %2, %3, p;\n\t}"
               ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

// mode bit 0: A from TMEM; bit 1: B MN-major; n = 128 or 256; ndst = number of distinct accumulators cycled through
__global__ void __launch_bounds__(128, 1) rate_kernel(int mode, int n, int iters, int ndst, int ctas_active, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint32_t tmem_ptr;
  __shared__ __align__(8) uint64_t bar;
  const uint32_t base = smem_u32(smem);
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x38383838u;   // e4m3 1.0
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_ptr)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_ptr;
  if ((int)blockIdx.x < ctas_active && threadIdx.x == 0) {
    const uint32_t idesc = (1u << 4) | (((uint32_t)n >> 3) << 17) | ((128u >> 4) << 24) | ((mode & 2) ? (1u << 16) : 0u);
    // descriptors precomputed, four MMAs per trip: the single issuing thread must not be the limit
    uint64_t da[4], db[4];
    uint32_t ta[4];
    for (int k = 0; k < 4; ++k) {
      da[k] = desc_sw128(base + k * 32, 16, 1024);
      db[k] = (mode & 2) ? desc_sw128(base + 32768 + k * 4096, 0, 1024) : desc_sw128(base + 32768 + k * 32, 16, 1024);
      ta[k] = tmem + 448 + k * 8;
    }
    const long long t0 = clock64();
    for (int i = 0; i < iters; i += 4) {
      const uint32_t d = tmem + (uint32_t)((i >> 3) & (ndst - 1)) * (uint32_t)n;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (mode & 1) mma_ts(d, ta[k], db[k], idesc, 1u);
        else mma_ss(d, da[k], db[k], idesc, 1u);
      }
    }
    const long long t1 = clock64();
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    uint32_t done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    const long long t2 = clock64();
    out[blockIdx.x * 2] = t1 - t0;
    out[blockIdx.x * 2 + 1] = t2 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

int main() {
  long long* out;
  cudaMalloc(&out, 148 * 2 * sizeof(long long));
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
  const int iters = 4096;
  for (int ctas : {1, 148})
    for (int n : {16, 32, 64, 128, 256})
      for (int mode = 0; mode < 4; ++mode)
        for (int ndst : {1, 2}) {
          if (n == 256 && ndst == 2 && (mode & 1)) continue;   // A operand columns would overlap the accumulators
          cudaMemset(out, 0, 148 * 2 * sizeof(long long));
          rate_kernel<<<148, 128, 96 * 1024>>>(mode, n, iters, ndst, ctas, out);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
          long long h[2];
          cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
          printf("active CTAs %3d  M128 N%d K32  A %s  B %s  accumulators %d: issue %.1f clk/MMA, complete %.1f clk/MMA (nominal fp8 %.0f)\n",
                 ctas, n, (mode & 1) ? "tmem" : "smem", (mode & 2) ? "MN-major" : "K-major ", ndst, (double)h[0] / iters, (double)h[1] / iters,
                 128.0 * n * 32 * 2 / 15472.0);
        }
  return 0;
}
