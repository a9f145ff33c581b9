// Micro-benchmark: issue rate / execution time of tcgen05.mma (kind::f16, M=128) from one thread
// as a function of N, shared-memory layout and operand reuse.  Build: nvcc -gencode arch=compute_100a,code=sm_100a
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ uint64_t desc_sw128(uint32_t addr, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ uint64_t desc_none(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// mode bits: 0 = vary A k-slice, 1 = vary B k-slice, 2 = vary A row offset (tap), 3 = no-swizzle layout
__global__ void bench(int N, int mode, int iters, long long* out, uint32_t tcols, int issuer) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "r"(tcols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = tslot;
  if (threadIdx.x == issuer) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
    const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;
    const bool nosw = mode & 8;
    const uint64_t da0 = nosw ? desc_none(a0, 4096, 128) : desc_sw128(a0, 1280);
    const uint64_t db0 = nosw ? desc_none(b0, N * 16, 128) : desc_sw128(b0, 1024);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const uint32_t ks = i & 3, tap = (i >> 2) % 9;
      uint64_t da = da0, db = db0;
      if (mode & 1) da += nosw ? ks * 512 : 2 * ks;
      if (mode & 2) db += nosw ? ks * (N * 2) : 2 * ks;
      if (mode & 4) da += ((tap / 3) * 10 + tap % 3) * 8;
      mma(tm, da, db, idesc, i > 0);
    }
    long long t1 = clock64();
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    uint32_t done = 0;
    while (!done) {
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                   : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    }
    long long t2 = clock64();
    out[0] = t1 - t0;
    out[1] = t2 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(tcols) : "memory");
}
int main(int argc, char** argv) {
  const int threads = argc > 1 ? atoi(argv[1]) : 128;
  const uint32_t tcols = argc > 2 ? atoi(argv[2]) : 256;
  const int issuer = argc > 3 ? atoi(argv[3]) : 0;
  const int smem_kb = argc > 4 ? atoi(argv[4]) : 200;
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_kb * 1024);
  const int iters = 720;
  const int Ns[] = {16, 64, 128, 256};
  for (int nosw = 0; nosw < 1; ++nosw)
    for (int N : Ns)
      for (int mode = 3; mode < 4; ++mode) {
        bench<<<1, threads, smem_kb * 1024>>>(N, mode | (nosw << 3), iters, d, tcols, issuer);
        long long h[2];
        cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        printf("%s N=%3d varyA_k=%d varyB_k=%d varyA_tap=%d : issue %6.1f clk/mma, complete %6.1f clk/mma\n",
               nosw ? "noswz" : "sw128", N, mode & 1, (mode >> 1) & 1, (mode >> 2) & 1, (double)h[0] / iters, (double)h[1] / iters);
      }
  return 0;
}
