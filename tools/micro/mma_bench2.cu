// Micro-benchmark 2: the conv_tc K loop in isolation (5x1 taps, SPLIT=3: Ahi*[Bhi|Blo] N=2BN + Alo*Bhi N=BN per
// 16-channel k-step, SW128 operands, aligned taps) with optional background activity of the other warp roles:
//   bit 0: 8 warps spin on an mbarrier (epilogue warps waiting for the accumulator)
//   bit 1: 8 warps stream st.shared.v2 into another A stage (producers)
//   bit 2: one thread streams 16 KB bulk copies global -> shared (weight loader)
//   bit 3: SPLIT=1 (one MMA per k-step)
//   bit 4: A SBO = 1280 instead of 1024;  bit 5: no tap variation;  bit 6: no B stage alternation;  bit 7: launch only 320 threads' worth of work (others exit)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -o mma_bench2 mma_bench2.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ uint64_t desc_sw128(uint32_t addr, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void mma_lh(uint32_t tmem_d, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "mov.b64 da, {%1, %2};\n\t"
      "mov.b64 db, {%3, %4};\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
      ::"r"(tmem_d), "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ bool elect() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ bool try_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  return done != 0;
}
__global__ void __launch_bounds__(576, 1) bench(int BN, int mode, int ksteps, const uint8_t* gsrc, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bars[6];
  __shared__ uint32_t tslot;
  __shared__ volatile int stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_done = smem_u32(&bars[0]), bar_spin = smem_u32(&bars[1]), bar_bulk = smem_u32(&bars[2]), bar_tap = smem_u32(&bars[3]), bar_rdy = smem_u32(&bars[4]);
  if (threadIdx.x == 0) {
    stop = 0;
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_done));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_spin));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_bulk));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_tap));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_rdy));
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_rdy) : "memory");   // phase 0 complete
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tslot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tm = tslot;
  // layout: A stage 0 [hi 20 KB | lo 20 KB], A stage 1 (40 KB), B ring 2 x (2*BN*128)
  const uint32_t a0 = smem_u32(smem), a_part = 20 * 1024, a1 = a0 + 2 * a_part, b0 = a1 + 2 * a_part;
  const uint32_t b_stage = 2u * BN * 128u;
  if (warp < 8) {
    if (mode & 1) { while (!try_wait(bar_spin, 0)) {} }
  } else if (warp == 8) {
    // clean issue path (same as conv_tc): whole warp runs the loops, one elected lane issues, descriptors as 32-bit words
    const bool split1 = mode & 8;
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | (8u << 24);
    const uint32_t idesc2 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 2) << 17) | (8u << 24);
    const uint64_t da0 = desc_sw128(0, (mode & 16) ? 1280 : 1024), db0 = desc_sw128(0, 1024);
    const uint32_t dhi_a = (uint32_t)(da0 >> 32), dhi_b = (uint32_t)(db0 >> 32);
    const uint32_t alo0 = (uint32_t)da0 + (a0 >> 4), blo0 = (uint32_t)db0 + (b0 >> 4);
    const uint32_t apart16 = a_part >> 4, bst16 = b_stage >> 4;
    const int groups = ksteps / 20;
    long long t0 = clock64();
    for (int g = 0; g < groups; ++g) {
      if (elect()) {
        uint32_t acc = g > 0;
        for (uint32_t tap = 0; tap < 5; ++tap) {
          const uint32_t alo = alo0 + ((mode & 32) ? 0 : tap * 64), blo = blo0 + ((mode & 64) ? 0 : (tap & 1) * bst16);
#pragma unroll
          for (uint32_t ks = 0; ks < 4; ++ks) {
            if (split1) {
              mma_lh(tm, alo + 2 * ks, dhi_a, blo + 2 * ks, dhi_b, idesc, ks == 0 ? acc : 1u);
            } else {
              mma_lh(tm, alo + 2 * ks, dhi_a, blo + 2 * ks, dhi_b, idesc2, ks == 0 ? acc : 1u);
              mma_lh(tm, alo + 2 * ks + apart16, dhi_a, blo + 2 * ks, dhi_b, idesc, 1);
            }
          }
          acc = 1;
          if (mode & 128) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_tap) : "memory");
          if (mode & 256) { while (!try_wait(bar_rdy, 0)) {} asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
        }
      }
      __syncwarp();
    }
    long long t1 = clock64();
    if (lane == 0) {
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_done) : "memory");
      while (!try_wait(bar_done, 0)) {}
      long long t2 = clock64();
      out[0] = t1 - t0;
      out[1] = t2 - t0;
      stop = 1;
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_spin) : "memory");
    }
  } else if (warp == 9) {
    if ((mode & 4) && lane == 0) {
      uint32_t ph = 0, it = 0;
      while (!stop) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_bulk), "r"(16384u) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(b0 + 2 * b_stage + (it & 1) * 16384u), "l"(gsrc + (size_t)(it & 63) * 16384), "r"(16384u), "r"(bar_bulk) : "memory");
        while (!try_wait(bar_bulk, ph)) {}
        ph ^= 1; ++it;
      }
    }
  } else {
    if (mode & 2) {
      const int tid = threadIdx.x - 320;
      uint8_t* dst = smem + 2 * a_part;
      uint32_t x = tid;
      while (!stop) {
#pragma unroll
        for (int u = 0; u < 12; ++u) {
          const uint32_t p = (tid >> 4) + 16 * u, q = tid & 15;
          const uint32_t off = p * 128u + (((q >> 1) ^ (p & 7)) << 4) + (q & 1) * 8u;
          if (p < 160) {
            *reinterpret_cast<uint2*>(dst + off) = make_uint2(x, x + u);
            *reinterpret_cast<uint2*>(dst + a_part + off) = make_uint2(x + 1, x + u);
          }
        }
        x += 7;
        __nanosleep(200);
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 8) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(512u) : "memory");
}
int main() {
  long long* d; cudaMalloc(&d, 16);
  uint8_t* g; cudaMalloc(&g, 64 * 16384); cudaMemset(g, 0, 64 * 16384);
  const int smem = 220 * 1024;
  cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int ksteps = 800;
  for (int BN : {64})
    for (int mode : {0, 128, 256, 384}) {
      if (2 * 40 * 1024 + 2 * 2 * BN * 128 + 32768 > smem) continue;
      bench<<<1, 576, smem>>>(BN, mode, ksteps, g, d);
      long long h[2];
      cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      printf("commit/tap=%d wait/tap=%d ", (mode >> 7) & 1, (mode >> 8) & 1);
      printf("BN=%3d split=%d spin=%d sts=%d bulk=%d : issue %6.1f clk/kstep, complete %6.1f clk/kstep\n", BN, (mode & 8) ? 1 : 3,
             mode & 1, (mode >> 1) & 1, (mode >> 2) & 1, (double)h[0] / ksteps, (double)h[1] / ksteps);
    }
  return 0;
}
