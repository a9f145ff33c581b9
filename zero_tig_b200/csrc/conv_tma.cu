// Full-resolution convolutions of the three Retinex networks (model/model.py:15-81) as a TMA-fed
// tcgen05 implicit GEMM over bf16 NHWC activations.
//
// Why a second tensor-core kernel: conv_tc.cu keeps fp32 activations and converts them in producer
// warps; at 1080p that kernel is bound by CUDA-core instruction issue and by tap-shifted operand views
// that straddle two swizzle atoms (ncu: 10 % tensor pipe, 22 % DRAM, profiles/r01_ncu_conv_tc_*.txt).
// Here the activations between layers are bf16 (the MMA consumed bf16 anyway), so
//   * one elected thread moves the input with TMA (cp.async.bulk.tensor.3d, SWIZZLE_128B, zero fill
//     outside the image = the convolution's zero padding): no producer warps, no conversions;
//   * the 16x8 pixel tile is loaded as kw copies shifted by one pixel, each a [rows][8 px][64 ch] box
//     whose 8-pixel rows are exactly one 1024 B swizzle atom: every filter tap (ky,kx) is the view
//     "copy kx, starting ky atoms in" - always atom aligned (58 clk per M128 K16 MMA instead of 114);
//   * weights (<= 72 KB) stay resident in shared memory, accumulators are double buffered in TMEM,
//     8 epilogue warps apply bias / activation / residual and write bf16 (or the small fp32 heads).
// GEMM view: M = 128 pixels (16 rows x 8), N = cout (<= 64), K = taps x 64 channels (cin <= 64).
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdlib.h>
#include <algorithm>
#include <map>
#include <mutex>
#include <tuple>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace zt {

constexpr int TM_TW = 8, TM_TH = 16;
constexpr int TM_THREADS = 320;           // warps 0-7 epilogue, 8 MMA issuer, 9 TMA / weight loader
constexpr int TM_W_MMA = 8, TM_W_LOAD = 9;
constexpr int TM_BAR_BYTES = 1024;
constexpr int TM_STG_BYTES = 8 * 2048;   // epilogue staging: 32 pixels x 64 B per epilogue warp

struct TmGeom {
  int thin;                               // single-copy mode with 32-byte pixel rows (cin <= 16): SWIZZLE_32B box of 16 channels,
                                          // so neither TMA nor the tensor core touches the 96 zero bytes of a 128-byte row
  int single, PW;                         // single = 1: ONE copy of the tile with its halo (PW = 8 + kw - 1 pixels wide), taps are
                                          // unaligned shifted views of it (as conv_tc.cu); 0: kw pixel-shifted, atom-aligned copies
  int BN, kh, kw, nks, rows;              // rows = TM_TH + kh - 1
  int tiles_x, S, NA, tmem_cols;
  uint32_t copy_bytes, a_stage, w_tap, w_bytes;
  int coal;                               // bf16 outputs: transpose the tile through shared memory so global stores / residual
                                          // loads touch 8 lines per instruction instead of 32 (TM_STG_BYTES after the weights)
};
struct TmArgs {
  const uint8_t* weight;                  // [tap][BN][128 B] bf16, swizzled
  const float* bias;
  void* out;
  int out_cs;
  const __nv_bfloat16* res;
  int res_cs;
  int H, W, cout, ph, pw;
  const float* head_w;                    // OUT = 3: fp32 [8][64] weights of a fused 1x1 head (zero padded), bias [8]
  const float* head_b;
  int head_n;                             // real head count (<= 8): the epilogue multiplies 4, 6 or 8 columns of head_w
};

__device__ __forceinline__ float tm_act(float v, int act) {
  if (act == ZT_ACT_RELU) return fmaxf(v, 0.f);
  if (act == ZT_ACT_LRELU02) return v > 0.f ? v : 0.2f * v;
  if (act == ZT_ACT_SIGMOID) return 1.f / (1.f + expf(-v));
  return v;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}

// OUT: 0 = bf16 out, 1 = bf16 out with residual (out = res + act(v)), 2 = fp32 out (cout <= 8),
//      3 = fp32 out of a fused 1x1 head: out[h] = head_b[h] + sum_c head_w[h][c] * act(v[c] + bias[c]), h < out_cs
//          (the 48 -> 3 / 48 -> 6 heads of the denoisers, model.py:27,43: the 48-channel activation never goes to HBM)
// two TMEM loads in flight, one wait (each tcgen05.ld + wait round trip is ~100 clk of epilogue latency)
__device__ __forceinline__ void tm_ld16_nowait(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
#define ZT_TM_R16(a) "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]), \
                     "+r"(a[8]), "+r"(a[9]), "+r"(a[10]), "+r"(a[11]), "+r"(a[12]), "+r"(a[13]), "+r"(a[14]), "+r"(a[15])
__device__ __forceinline__ void tm_ld_wait2(uint32_t (&a)[16], uint32_t (&b)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;" : ZT_TM_R16(a), ZT_TM_R16(b) :: "memory");
}
// (p0, p1) += x * (w0, w1) as one packed FFMA2 (two IEEE fp32 FMAs: the same bits as two FFMAs)
__device__ __forceinline__ void tm_fma2(float& p0, float& p1, float x, float w0, float w1) {
  asm("{\n\t.reg .b64 xx, ww, pp;\n\tmov.b64 xx, {%2, %2};\n\tmov.b64 ww, {%3, %4};\n\tmov.b64 pp, {%0, %1};\n\t"
      "fma.rn.f32x2 pp, xx, ww, pp;\n\tmov.b64 {%0, %1}, pp;\n\t}" : "+f"(p0), "+f"(p1) : "f"(x), "f"(w0), "f"(w1));
}

// The MMAs of one pixel tile of a 3x3 convolution, fully unrolled: 9 taps x NKS 16-channel k-steps.  The generic loop
// below (runtime kh / kw / nks read from the parameter struct) costs ~25 dependent scalar instructions per MMA in the
// single issuing thread - 225 clk per MMA on the 16-channel layers, 100 clk on the 64-channel ones, more than the
// MMA itself (ncu source page, profiles/r02_ncu_conv_tma.txt); here an MMA is two adds away from the previous one.
template <int NKS>
__device__ __forceinline__ void tm_issue_3x3(uint32_t tmem_d, uint32_t abase, uint32_t bbase, uint32_t tap_x16,
                                             uint32_t tap_y16, uint32_t wtap16, uint32_t dhiA, uint32_t dhi,
                                             uint32_t idesc) {
#pragma unroll
  for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const uint32_t alo = abase + (uint32_t)kx * tap_x16 + (uint32_t)ky * tap_y16;
      const uint32_t blo = bbase + (uint32_t)(ky * 3 + kx) * wtap16;
#pragma unroll
      for (int ks = 0; ks < NKS; ++ks)
        tc_mma_bf16_lh(tmem_d, alo + 2u * ks, dhiA, blo + 2u * ks, dhi, idesc, (ky | kx | ks) ? 1u : 0u);
    }
  }
}

template <int OUT, int ACT>
__global__ void __launch_bounds__(TM_THREADS, 1) conv_tma_kernel(const __grid_constant__ CUtensorMap tmap,
                                                                 const TmArgs a, const TmGeom g) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  // programmatic dependent launch: barrier init, TMEM allocation and the (constant) weight fetch run while the
  // preceding kernel drains; activations are touched only after griddepcontrol.wait
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + TM_BAR_BYTES - 16);
  uint8_t* smA = smem + TM_BAR_BYTES;
  uint8_t* smW = smA + (size_t)g.NA * g.a_stage;
  const uint32_t bar0 = smem_u32(smem);
  const int NA = g.NA;
  auto fullA = [&](uint32_t s) { return bar0 + 8u * s; };
  auto emptyA = [&](uint32_t s) { return bar0 + 8u * (NA + s); };
  const uint32_t fullW = bar0 + 8u * (2 * NA);
  auto accFull = [&](uint32_t s) { return bar0 + 8u * (2 * NA + 1 + s); };
  auto accEmpty = [&](uint32_t s) { return bar0 + 8u * (2 * NA + 3 + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = g.S;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NA; ++s) { mbar_init(fullA(s), 1); mbar_init(emptyA(s), 1); }
    mbar_init(fullW, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(accFull(s), 1); mbar_init(accEmpty(s), 256); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __shared__ __align__(16) float hw_s[OUT == 3 ? 8 * 64 : 4];
  __shared__ __align__(16) float part_s[OUT == 3 ? 2 * 4 * 32 * 8 : 4];      // double buffered by tile parity
  if (OUT == 3) {
    // constants, fetched before the dependency wait; stored channel-major so one channel's 8 weights are two LDS.128
    for (int i = threadIdx.x; i < 8 * 64; i += TM_THREADS) hw_s[(i & 63) * 8 + (i >> 6)] = __ldg(a.head_w + i);
  }
  if (warp == TM_W_MMA) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(tmem_slot)), "r"((uint32_t)g.tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 8) {
    // ===================== epilogue =====================
    // thread = one pixel (TMEM lane) x 32 consecutive output channels of its half
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int quad = warp & 3, half = warp >> 2;
    const int c0 = half * 32;
    const bool has_cols = c0 < g.BN;
    const bool second16 = c0 + 16 < g.BN;
    float bias[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) bias[i] = (a.bias && c0 + i < a.cout) ? __ldg(a.bias + c0 + i) : 0.f;
    const bool coal = OUT <= 1 && g.coal != 0;
    uint4* stg = reinterpret_cast<uint4*>(smW + g.w_bytes + (size_t)warp * 2048);
    uint32_t lt = 0;
    // tile coordinates advance incrementally (an integer division per tile sat on the epilogue's critical path)
    const int step_y = (int)gridDim.x / g.tiles_x, step_x = (int)gridDim.x - step_y * g.tiles_x;
    int tyi = (int)blockIdx.x / g.tiles_x, txi = (int)blockIdx.x - tyi * g.tiles_x;
    for (int t = blockIdx.x; t < S; t += gridDim.x, ++lt, tyi += step_y, txi += step_x) {
      if (txi >= g.tiles_x) { txi -= g.tiles_x; ++tyi; }
      const int oy = tyi * TM_TH + quad * 4 + (lane >> 3), ox = txi * TM_TW + (lane & 7);
      const bool ok = has_cols && oy < a.H && ox < a.W;
      const int64_t pix = (int64_t)oy * a.W + ox;
      const uint32_t buf = lt & 1;
      uint4 rv[4];
      // coalesced mapping: instruction i, lane l <-> pixel (tile row quad*4 + i, x = l >> 2), 16-byte chunk l & 3 of this half:
      // four lanes cover the 64 contiguous bytes of a pixel, a warp instruction touches 8 lines (thread = pixel: 32)
      const int cx = txi * TM_TW + (lane >> 2), cch = c0 + 8 * (lane & 3);
      const bool cok = coal && cx < a.W && cch < a.cout;
      if (OUT == 1 && coal) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int cy = tyi * TM_TH + quad * 4 + i;
          rv[i] = make_uint4(0u, 0u, 0u, 0u);
          if (cok && cy < a.H) rv[i] = __ldg(reinterpret_cast<const uint4*>(a.res + ((int64_t)cy * a.W + cx) * a.res_cs + cch));
        }
      } else if (OUT == 1) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          rv[j] = make_uint4(0u, 0u, 0u, 0u);
          if (ok && c0 + 8 * j < a.cout) rv[j] = __ldg(reinterpret_cast<const uint4*>(a.res + pix * a.res_cs + c0 + 8 * j));
        }
      }
      mbar_wait(accFull(buf), (lt >> 1) & 1);
      tc_fence_after();
      uint32_t v[32];
      if (has_cols) {
        const uint32_t taddr = tmem_base + buf * (uint32_t)g.BN + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0;
        uint32_t v0[16], v1[16];
        if (second16) {                                               // warp-uniform
          tm_ld16_nowait(taddr, v0);
          tm_ld16_nowait(taddr + 16, v1);
          tm_ld_wait2(v0, v1);
#pragma unroll
          for (int i = 0; i < 16; ++i) { v[i] = v0[i]; v[16 + i] = v1[i]; }
        } else {
          tmem_ld16(taddr, v0);
#pragma unroll
          for (int i = 0; i < 16; ++i) { v[i] = v0[i]; v[16 + i] = 0u; }
        }
      }
      // the accumulator is in registers: hand the TMEM buffer back before the stores
      tc_fence_before();
      mbar_arrive(accEmpty(buf));
      if (OUT == 3) {
        // fused 1x1 head: partial dot products over this thread's <= 32 channels, the two channel halves of a pixel
        // (warps q and q + 4, same lane) meet in shared memory under a 64-thread named barrier
        float part[8];
#pragma unroll
        for (int h = 0; h < 8; ++h) part[h] = 0.f;
        if (has_cols) {
          if (a.head_n > 6) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float x = tm_act(__uint_as_float(v[i]) + bias[i], ACT);
              const float4 w0 = *reinterpret_cast<const float4*>(hw_s + (c0 + i) * 8);
              const float4 w1 = *reinterpret_cast<const float4*>(hw_s + (c0 + i) * 8 + 4);
              tm_fma2(part[0], part[1], x, w0.x, w0.y);
              tm_fma2(part[2], part[3], x, w0.z, w0.w);
              tm_fma2(part[4], part[5], x, w1.x, w1.y);
              tm_fma2(part[6], part[7], x, w1.z, w1.w);
            }
          } else if (a.head_n > 4) {                                  // 5..6 heads (denoise_2: 6)
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float x = tm_act(__uint_as_float(v[i]) + bias[i], ACT);
              const float4 w0 = *reinterpret_cast<const float4*>(hw_s + (c0 + i) * 8);
              const float2 w1 = *reinterpret_cast<const float2*>(hw_s + (c0 + i) * 8 + 4);
              tm_fma2(part[0], part[1], x, w0.x, w0.y);
              tm_fma2(part[2], part[3], x, w0.z, w0.w);
              tm_fma2(part[4], part[5], x, w1.x, w1.y);
            }
          } else {                                                    // <= 4 heads (denoise_1: 3): half the weights
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              const float x = tm_act(__uint_as_float(v[i]) + bias[i], ACT);
              const float4 w0 = *reinterpret_cast<const float4*>(hw_s + (c0 + i) * 8);
              tm_fma2(part[0], part[1], x, w0.x, w0.y);
              tm_fma2(part[2], part[3], x, w0.z, w0.w);
            }
          }
        }
        float* ps = part_s + ((buf * 4 + quad) * 32 + lane) * 8;
        if (half == 1) {
          *reinterpret_cast<float4*>(ps) = make_float4(part[0], part[1], part[2], part[3]);
          *reinterpret_cast<float4*>(ps + 4) = make_float4(part[4], part[5], part[6], part[7]);
        }
        asm volatile("bar.sync %0, 64;" ::"r"(1 + quad) : "memory");
        if (half == 0 && oy < a.H && ox < a.W) {
          const float4 p0 = *reinterpret_cast<const float4*>(ps), p1 = *reinterpret_cast<const float4*>(ps + 4);
          float* o = reinterpret_cast<float*>(a.out) + pix * a.out_cs;
          *reinterpret_cast<float4*>(o) = make_float4(part[0] + p0.x + __ldg(a.head_b + 0), part[1] + p0.y + __ldg(a.head_b + 1),
                                                      part[2] + p0.z + __ldg(a.head_b + 2), part[3] + p0.w + __ldg(a.head_b + 3));
          if (a.out_cs > 4)
            *reinterpret_cast<float4*>(o + 4) = make_float4(part[4] + p1.x + __ldg(a.head_b + 4), part[5] + p1.y + __ldg(a.head_b + 5),
                                                            part[6] + p1.z + __ldg(a.head_b + 6), part[7] + p1.w + __ldg(a.head_b + 7));
        }
        // no second barrier: the next tile uses the other half of part_s, and the tile after it is written only after
        // that tile's barrier, which the reading warp reaches after these loads
        continue;
      }
      if (OUT <= 1 && coal) {
        if (!has_cols) continue;                                     // warp-uniform
        // staging rows of 64 B, 16-byte chunk position XOR ((row >> 1) & 3): conflict-free for both access patterns
        uint4* own = stg + lane * 4;
        const uint32_t swz = (lane >> 1) & 3;
        if (OUT == 1) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int pr = 8 * i + (lane >> 2);
            stg[pr * 4 + ((lane & 3) ^ ((pr >> 1) & 3))] = rv[i];
          }
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j) rv[j] = own[j ^ swz];
          __syncwarp();
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float r[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) r[e] = tm_act(__uint_as_float(v[8 * j + e]) + bias[8 * j + e], ACT);
          if (OUT == 1) {
            const uint32_t rw[4] = {rv[j].x, rv[j].y, rv[j].z, rv[j].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&rw[e]);
              r[2 * e] += __low2float(h);
              r[2 * e + 1] += __high2float(h);
            }
          }
          own[j ^ swz] = make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]), pack_bf16x2(r[4], r[5]),
                                    pack_bf16x2(r[6], r[7]));
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int pr = 8 * i + (lane >> 2), cy = tyi * TM_TH + quad * 4 + i;
          const uint4 q = stg[pr * 4 + ((lane & 3) ^ ((pr >> 1) & 3))];
          if (cok && cy < a.H)
            *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(a.out) + ((int64_t)cy * a.W + cx) * a.out_cs + cch) = q;
        }
        __syncwarp();                                                // the next tile overwrites the staging rows
        continue;
      }
      if (!ok) continue;
      if (OUT == 2) {
        float* o = reinterpret_cast<float*>(a.out) + pix * a.out_cs;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          if (4 * j < a.out_cs) {
            float r[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int c = 4 * j + e;
              r[e] = c < a.cout ? tm_act(__uint_as_float(v[c]) + bias[c], ACT) : 0.f;
            }
            *reinterpret_cast<float4*>(o + 4 * j) = make_float4(r[0], r[1], r[2], r[3]);
          }
        }
      } else {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(a.out) + pix * a.out_cs + c0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (c0 + 8 * j < a.cout) {
            float r[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) r[e] = tm_act(__uint_as_float(v[8 * j + e]) + bias[8 * j + e], ACT);
            if (OUT == 1) {
              const uint32_t rw[4] = {rv[j].x, rv[j].y, rv[j].z, rv[j].w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&rw[e]);
                r[2 * e] += __low2float(h);
                r[2 * e + 1] += __high2float(h);
              }
            }
            *reinterpret_cast<uint4*>(o + 8 * j) = make_uint4(pack_bf16x2(r[0], r[1]), pack_bf16x2(r[2], r[3]),
                                                              pack_bf16x2(r[4], r[5]), pack_bf16x2(r[6], r[7]));
          }
        }
      }
    }
  } else if (warp == TM_W_MMA) {
    // ===================== MMA issuer =====================
    const uint32_t BN = g.BN;
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((BN >> 3) << 17) | (8u << 24);
    const uint64_t desc0 = make_desc(0, 1024u);
    const uint32_t dhi = (uint32_t)(desc0 >> 32), dlo = (uint32_t)desc0;
    // single-copy mode: 8-pixel row groups of the patch are PW * 128 B apart, a tap is a row offset of (ky * PW + kx)
    // thin mode: the same views over 32-byte pixel rows (descriptor layout type SWIZZLE_32B, 8-pixel groups PW * 32 B apart)
    const uint64_t desc_thin = ((make_desc(0, (uint32_t)g.PW * 32u) & ~(7ull << 61)) | (6ull << 61));
    const uint32_t dhiA = g.thin ? (uint32_t)(desc_thin >> 32)
                                 : g.single ? (uint32_t)(make_desc(0, (uint32_t)g.PW * 128u) >> 32) : dhi;
    const uint32_t tap_y16 = g.thin ? (uint32_t)g.PW * 2u : g.single ? (uint32_t)g.PW * 8u : 64u;
    const uint32_t tap_x16 = g.thin ? 2u : g.single ? 8u : (g.copy_bytes >> 4);
    const uint32_t smA16 = smem_u32(smA) >> 4, smW16 = smem_u32(smW) >> 4;
    const uint32_t a_stage16 = g.a_stage >> 4, wtap16 = g.w_tap >> 4;
    const uint32_t kh = g.kh, kw = g.kw, nks = g.nks;
    const uint32_t issue3x3 = (kh == 3 && kw == 3 && (nks == 1 || nks == 3 || nks == 4)) ? nks : 0u;
    mbar_wait(fullW, 0);
    tc_fence_after();
    uint32_t lt = 0, sa = 0, pa = 0;
    for (int t = blockIdx.x; t < S; t += gridDim.x, ++lt) {
      const uint32_t buf = lt & 1;
      mbar_wait(accEmpty(buf), ((lt >> 1) & 1) ^ 1);
      mbar_wait(fullA(sa), pa);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t tmem_d = tmem_base + buf * BN;
        uint32_t acc = 0, blo = dlo + smW16;
        const uint32_t abase = dlo + smA16 + sa * a_stage16;
        if (issue3x3 == 1) tm_issue_3x3<1>(tmem_d, abase, blo, tap_x16, tap_y16, wtap16, dhiA, dhi, idesc);
        else if (issue3x3 == 3) tm_issue_3x3<3>(tmem_d, abase, blo, tap_x16, tap_y16, wtap16, dhiA, dhi, idesc);
        else if (issue3x3 == 4) tm_issue_3x3<4>(tmem_d, abase, blo, tap_x16, tap_y16, wtap16, dhiA, dhi, idesc);
        else
        for (uint32_t ky = 0; ky < kh; ++ky) {
          for (uint32_t kx = 0; kx < kw; ++kx, blo += wtap16) {
            const uint32_t alo = abase + kx * tap_x16 + ky * tap_y16;      // aligned copies: ky atoms of 1024 B
            for (uint32_t ks = 0; ks < nks; ++ks) {
              tc_mma_bf16_lh(tmem_d, alo + 2 * ks, dhiA, blo + 2 * ks, dhi, idesc, acc);
              acc = 1;
            }
          }
        }
        tc_commit(emptyA(sa));
        tc_commit(accFull(buf));
      }
      __syncwarp();
      if (++sa == (uint32_t)NA) { sa = 0; pa ^= 1; }
    }
  } else {
    // ===================== loader: weights once, then one TMA box per kx copy per tile =====================
    if (lane == 0) {
      tma_prefetch_desc(&tmap);
      if ((int)blockIdx.x < S) {
        mbar_arrive_expect_tx(fullW, g.w_bytes);
        const uint32_t ntaps = g.kh * g.kw;
        for (uint32_t it = 0; it < ntaps; ++it)
          bulk_g2s(smem_u32(smW + it * g.w_tap), a.weight + (size_t)it * g.w_tap, g.w_tap, fullW);
      }
      asm volatile("griddepcontrol.wait;" ::: "memory");
      uint32_t sa = 0, pa = 1;
      for (int t = blockIdx.x; t < S; t += gridDim.x) {
        const int tyi = t / g.tiles_x, txi = t - tyi * g.tiles_x;
        mbar_wait(emptyA(sa), pa);
        if (g.single) {
          mbar_arrive_expect_tx(fullA(sa), g.copy_bytes);
          tma_load_3d(smem_u32(smA + sa * g.a_stage), &tmap, 0, txi * TM_TW - a.pw, tyi * TM_TH - a.ph, fullA(sa));
        } else {
          mbar_arrive_expect_tx(fullA(sa), g.a_stage);
          for (int kx = 0; kx < g.kw; ++kx)
            tma_load_3d(smem_u32(smA + sa * g.a_stage + kx * g.copy_bytes), &tmap, 0, txi * TM_TW + kx - a.pw,
                        tyi * TM_TH - a.ph, fullA(sa));
        }
        if (++sa == (uint32_t)NA) { sa = 0; pa ^= 1; }
      }
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == TM_W_MMA) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)g.tmem_cols)
                 : "memory");
  }
}

// NCHW fp32 planes (+add) -> bf16 NHWC [HW][cs] (cs = 8 or 16), zero padded: the first-layer inputs
// (virtual torch.cat of model.py:313,326,334)
template <int CS>
__global__ void __launch_bounds__(256) pack_nhwc_bf16_kernel(const float* __restrict__ s0, const float* __restrict__ s1,
                                                             const float* __restrict__ s2, const float* __restrict__ s3,
                                                             int c_each, int nsrc, float add,
                                                             __nv_bfloat16* __restrict__ out, int64_t HW) {
  const float* src[4] = {s0, s1, s2, s3};
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += (int64_t)gridDim.x * blockDim.x) {
    float v[CS];
#pragma unroll
    for (int c = 0; c < CS; ++c) v[c] = 0.f;
    int ch = 0;
    for (int s = 0; s < nsrc; ++s)
      for (int c = 0; c < c_each; ++c, ++ch) {
        const float x = __ldg(src[s] + c * HW + p) + add;
#pragma unroll
        for (int k = 0; k < CS; ++k)
          if (k == ch) v[k] = x;
      }
    uint4* o = reinterpret_cast<uint4*>(out + p * CS);
#pragma unroll
    for (int j = 0; j < CS / 8; ++j)
      o[j] = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                        pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
  }
}

// fast path: three channels per source, four consecutive pixels per thread (float4 plane loads, 4*CS*2 contiguous
// bytes written per thread), channel placement resolved at compile time
template <int CS, int NSRC>
__global__ void __launch_bounds__(256, 3) pack_nhwc_bf16_x4_kernel(const float* __restrict__ s0, const float* __restrict__ s1,
                                                                const float* __restrict__ s2, const float* __restrict__ s3,
                                                                float add, __nv_bfloat16* __restrict__ out, int64_t HW) {
  __shared__ uint4 stage[8][32 * (4 * CS / 8)];
  pdl_enter();
  const float* src[4] = {s0, s1, s2, s3};
  const int64_t n4 = HW >> 2;
  // n4 is a multiple of 32 (checked by the host): whole warps iterate together (shared-memory transpose below)
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 v[NSRC * 3];
#pragma unroll
    for (int s = 0; s < NSRC; ++s)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        v[s * 3 + c] = __ldg(reinterpret_cast<const float4*>(src[s] + c * HW) + i);
    // Each thread owns 4*CS*2 contiguous output bytes; stored directly a warp store would touch 32 lines with 16 B
    // each.  The warp's 32 * U uint4 go through shared memory (XOR-swizzled against bank conflicts) so that every
    // store instruction writes 512 contiguous bytes.
    constexpr int U = 4 * CS / 8;                       // uint4 per thread (4 or 8)
    uint4* tile = stage[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int px = 0; px < 4; ++px) {
      float r[CS];
#pragma unroll
      for (int c = 0; c < CS; ++c) {
        if (c < NSRC * 3) {
          const float4 t = v[c];
          r[c] = (px == 0 ? t.x : px == 1 ? t.y : px == 2 ? t.z : t.w) + add;
        } else {
          r[c] = 0.f;
        }
      }
#pragma unroll
      for (int j = 0; j < CS / 8; ++j) {
        const int u = px * (CS / 8) + j;
        tile[lane * U + (u ^ (lane & (U - 1)))] =
            make_uint4(pack_bf16x2(r[8 * j], r[8 * j + 1]), pack_bf16x2(r[8 * j + 2], r[8 * j + 3]),
                       pack_bf16x2(r[8 * j + 4], r[8 * j + 5]), pack_bf16x2(r[8 * j + 6], r[8 * j + 7]));
      }
    }
    __syncwarp();
    uint4* o = reinterpret_cast<uint4*>(out) + (i - lane) * U;   // the warp's 32 * U contiguous uint4
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const int e = k * 32 + lane;                      // element e belongs to lane e / U of this iteration
      const int t = e / U, u = e % U;
      o[e] = tile[t * U + (u ^ (t & (U - 1)))];
    }
    __syncwarp();
  }
}

// ---- host: tensor-map cache ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
    (void)cudaGetLastError();
  }
  return fn;
}
static std::mutex g_tm_mu;
static std::map<std::tuple<const void*, int, int, int, int, int>, CUtensorMap> g_tm_cache;

static int get_tensor_map(const void* base, int C, int W, int H, int rows, int box_w, int thin, CUtensorMap* out) {
  std::lock_guard<std::mutex> lk(g_tm_mu);
  auto key = std::make_tuple(base, C, W, H, rows, box_w + (thin ? 1000 : 0));
  auto it = g_tm_cache.find(key);
  if (it != g_tm_cache.end()) { *out = it->second; return ZT_OK; }
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("conv_bf16_tma: cuTensorMapEncodeTiled is unavailable (no CUDA driver)"); return ZT_ERR_CUDA; }
  const cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H};
  const cuuint64_t strides[2] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2};
  const cuuint32_t box[3] = {thin ? 16u : 64u, (cuuint32_t)box_w, (cuuint32_t)rows};
  const cuuint32_t estr[3] = {1u, 1u, 1u};
  CUtensorMap m;
  const CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, thin ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("conv_bf16_tma: cuTensorMapEncodeTiled failed (%d)", (int)r); return ZT_ERR_CUDA; }
  if (g_tm_cache.size() > 256) g_tm_cache.clear();
  g_tm_cache[key] = m;
  *out = m;
  return ZT_OK;
}

}  // namespace zt

extern "C" int zt_pack_nhwc_bf16(const float* s0, const float* s1, const float* s2, const float* s3, int c_each,
                                 int nsrc, float add, void* out, int cs, int64_t HW, void* stream) {
  using namespace zt;
  ZT_REQUIRE(s0 && out && nsrc >= 1 && nsrc <= 4 && c_each >= 1 && c_each * nsrc <= cs && HW > 0, "pack_nhwc_bf16: bad args");
  ZT_REQUIRE((nsrc < 2 || s1) && (nsrc < 3 || s2) && (nsrc < 4 || s3), "pack_nhwc_bf16: missing source");
  ZT_REQUIRE(cs == 8 || cs == 16, "pack_nhwc_bf16: cs must be 8 or 16 (got %d)", cs);
  const bool al16 = (((uintptr_t)s0 | (uintptr_t)s1 | (uintptr_t)s2 | (uintptr_t)s3) & 15) == 0;
  if (c_each == 3 && (HW & 127) == 0 && al16 && nsrc * 3 <= cs) {
    const int b4 = (int)std::min<int64_t>(ceil_div64(HW >> 2, 256), 64 * kNumSMs);   // one item per thread up to 4K frames
    cudaStream_t st = (cudaStream_t)stream;
    __nv_bfloat16* o = (__nv_bfloat16*)out;
    cudaError_t e = cudaErrorInvalidValue;
    if (cs == 8 && nsrc == 1) e = launch_pdl(pack_nhwc_bf16_x4_kernel<8, 1>, dim3(b4), dim3(256), 0, st, s0, s1, s2, s3, add, o, HW);
    else if (cs == 8 && nsrc == 2) e = launch_pdl(pack_nhwc_bf16_x4_kernel<8, 2>, dim3(b4), dim3(256), 0, st, s0, s1, s2, s3, add, o, HW);
    else if (cs == 16 && nsrc == 3) e = launch_pdl(pack_nhwc_bf16_x4_kernel<16, 3>, dim3(b4), dim3(256), 0, st, s0, s1, s2, s3, add, o, HW);
    else if (cs == 16 && nsrc == 4) e = launch_pdl(pack_nhwc_bf16_x4_kernel<16, 4>, dim3(b4), dim3(256), 0, st, s0, s1, s2, s3, add, o, HW);
    if (e == cudaSuccess) return check_launch("pack_nhwc_bf16_x4_kernel");
    (void)cudaGetLastError();      // combination without a fast path: generic kernel below
  }
  const int blocks = (int)std::min<int64_t>(ceil_div64(HW, 256), 16 * kNumSMs);
  if (cs == 8)
    pack_nhwc_bf16_kernel<8><<<blocks, 256, 0, (cudaStream_t)stream>>>(s0, s1, s2, s3, c_each, nsrc, add,
                                                                       (__nv_bfloat16*)out, HW);
  else
    pack_nhwc_bf16_kernel<16><<<blocks, 256, 0, (cudaStream_t)stream>>>(s0, s1, s2, s3, c_each, nsrc, add,
                                                                        (__nv_bfloat16*)out, HW);
  return check_launch("pack_nhwc_bf16_kernel");
}

extern "C" int zt_conv2d_bf16_tma(const ZtConvBf16* dp, void* stream) {
  using namespace zt;
  ZT_REQUIRE(dp != nullptr, "conv_bf16_tma: null descriptor");
  const ZtConvBf16& d = *dp;
  ZT_REQUIRE(d.in && d.weight && d.out, "conv_bf16_tma: null pointer");
  ZT_REQUIRE((d.kh == 1 || d.kh == 3) && d.kw == d.kh, "conv_bf16_tma: 1x1 or 3x3 only (got %dx%d)", d.kh, d.kw);
  ZT_REQUIRE(d.cin >= 8 && d.cin <= 64 && (d.cin & 7) == 0, "conv_bf16_tma: cin=%d must be a multiple of 8 in [8,64]", d.cin);
  ZT_REQUIRE(d.cout >= 1 && d.cout <= 64, "conv_bf16_tma: cout=%d must be <= 64", d.cout);
  ZT_REQUIRE(d.H > 0 && d.W > 0 && (int64_t)d.H * d.W < (1ll << 31), "conv_bf16_tma: bad dims");
  ZT_REQUIRE(((uintptr_t)d.in & 15) == 0 && ((uintptr_t)d.weight & 15) == 0 && ((uintptr_t)d.out & 15) == 0,
             "conv_bf16_tma: pointers must be 16-byte aligned");
  if (d.head_n) {
    ZT_REQUIRE(d.head_n >= 1 && d.head_n <= 8 && d.head_w && d.head_b && d.out_f32 && !d.res &&
                   (d.out_cs == 4 || d.out_cs == 8) && d.out_cs >= d.head_n,
               "conv_bf16_tma: fused head needs head_n <= 8, fp32 output with out_cs 4 or 8, no residual");
  } else if (d.out_f32) {
    ZT_REQUIRE(d.cout <= 8 && (d.out_cs == 4 || d.out_cs == 8) && d.out_cs >= d.cout && !d.res,
               "conv_bf16_tma: fp32 output needs cout <= 8, out_cs 4 or 8, no residual");
  } else {
    ZT_REQUIRE((d.cout & 7) == 0 && (d.out_cs & 7) == 0 && d.out_cs >= d.cout, "conv_bf16_tma: bf16 output needs cout, out_cs %% 8 == 0");
    ZT_REQUIRE(!d.res || ((d.res_cs & 7) == 0 && ((uintptr_t)d.res & 15) == 0), "conv_bf16_tma: bad residual");
  }
  TmGeom g;
  g.BN = ceil_div(d.cout, 16) * 16;
  g.kh = d.kh; g.kw = d.kw;
  g.nks = ceil_div(d.cin, 16);
  g.rows = TM_TH + d.kh - 1;
  g.tiles_x = ceil_div(d.W, TM_TW);
  g.S = g.tiles_x * ceil_div(d.H, TM_TH);
  // One copy with halo + unaligned tap views costs more per MMA (operand rows straddle swizzle atoms) but a third of
  // the TMA bytes: the thin layers (K = 16 per tap, or few output channels) are bound by the copies, the 64 -> 64
  // blocks by the MMAs.  ZT_TMA_SINGLE: -1 = never, else "use it when cin * BN <= value" (default 48 * 48).
  static const int single_thr = getenv("ZT_TMA_SINGLE") ? atoi(getenv("ZT_TMA_SINGLE")) : 48 * 48;
  g.single = (d.kw == 3 && single_thr >= 0 && d.cin * g.BN <= single_thr) ? 1 : 0;
  g.PW = TM_TW + d.kw - 1;
  // ZT_TMA_SW32=1: 32-byte pixel rows (SWIZZLE_32B box of 16 channels) for the 8 / 16-channel first layers.  Correct
  // (tests pass) but measured SLOWER than the zero-padded 128-byte rows: 83 vs 73 us for 3 -> 48 at 1080p - the
  // unaligned tap views cost more in the 32-byte layout and the layers are within 1.3-2x of the HBM floor anyway.
  static const int sw32 = getenv("ZT_TMA_SW32") ? atoi(getenv("ZT_TMA_SW32")) : 0;
  g.thin = (g.single && sw32 && d.cin <= 16) ? 1 : 0;
  if (g.single) {
    g.copy_bytes = (uint32_t)(g.rows * g.PW) * (g.thin ? 32u : 128u);     // bytes the TMA box writes
    g.a_stage = (g.copy_bytes + 1023u) / 1024u * 1024u;
  } else {
    g.copy_bytes = (uint32_t)g.rows * 1024u;
    g.a_stage = g.copy_bytes * d.kw;
  }
  g.w_tap = (uint32_t)g.BN * 128u;
  g.w_bytes = g.w_tap * d.kh * d.kw;
  g.tmem_cols = 32;
  while (g.tmem_cols < 2 * g.BN) g.tmem_cols <<= 1;
  // ZT_TMA_COALESCE=0 keeps the thread-per-pixel stores (every lane of a store instruction in a different 128-byte line)
  static const int coalesce = getenv("ZT_TMA_COALESCE") ? atoi(getenv("ZT_TMA_COALESCE")) : 1;
  g.coal = (coalesce && !d.out_f32 && !d.head_n) ? 1 : 0;
  const size_t stg_bytes = g.coal ? TM_STG_BYTES : 0;
  const size_t budget = 226 * 1024 - TM_BAR_BYTES - g.w_bytes - stg_bytes - (d.head_n ? 10 * 1024 : 0);  // static smem of OUT = 3
  g.NA = (int)std::min<size_t>(g.single ? 6 : 4, budget / g.a_stage);
  ZT_REQUIRE(g.NA >= 2, "conv_bf16_tma: tile does not fit shared memory");
  const size_t smem = TM_BAR_BYTES + (size_t)g.NA * g.a_stage + g.w_bytes + stg_bytes;
  CUtensorMap tmap;
  if (int e = get_tensor_map(d.in, d.cin, d.W, d.H, g.rows, g.single ? g.PW : TM_TW, g.thin, &tmap)) return e;
  TmArgs a;
  a.weight = (const uint8_t*)d.weight; a.bias = d.bias; a.out = d.out; a.out_cs = d.out_cs;
  a.res = (const __nv_bfloat16*)d.res; a.res_cs = d.res_cs;
  a.H = d.H; a.W = d.W; a.cout = d.cout; a.ph = d.kh / 2; a.pw = d.kw / 2;
  a.head_w = d.head_w; a.head_b = d.head_b; a.head_n = d.head_n;
  const int gx = std::min(kNumSMs, g.S);
  cudaStream_t st = (cudaStream_t)stream;
#define ZT_TM_LAUNCH(OUT, ACT)                                                                              \
  do {                                                                                                      \
    static bool attr_done = false;                                                                          \
    if (!attr_done) {                                                                                       \
      ZT_CUDA(cudaFuncSetAttribute(conv_tma_kernel<OUT, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                                   ((OUT) == 3 ? 216 : 227) * 1024));    /* OUT = 3 has 10 KB static */       \
      attr_done = true;                                                                                     \
    }                                                                                                       \
    ZT_CUDA(launch_pdl(conv_tma_kernel<OUT, ACT>, dim3(gx), dim3(TM_THREADS), smem, st, tmap, a, g));       \
    return check_launch("conv_tma_kernel");                                                                 \
  } while (0)
  if (d.head_n) {
    if (d.act == ZT_ACT_LRELU02) ZT_TM_LAUNCH(3, ZT_ACT_LRELU02);
    if (d.act == ZT_ACT_RELU) ZT_TM_LAUNCH(3, ZT_ACT_RELU);
  } else if (d.out_f32) {
    if (d.act == ZT_ACT_NONE) ZT_TM_LAUNCH(2, ZT_ACT_NONE);
    if (d.act == ZT_ACT_SIGMOID) ZT_TM_LAUNCH(2, ZT_ACT_SIGMOID);
  } else if (d.res) {
    if (d.act == ZT_ACT_RELU) ZT_TM_LAUNCH(1, ZT_ACT_RELU);
  } else {
    if (d.act == ZT_ACT_RELU) ZT_TM_LAUNCH(0, ZT_ACT_RELU);
    if (d.act == ZT_ACT_LRELU02) ZT_TM_LAUNCH(0, ZT_ACT_LRELU02);
  }
  set_error("conv_bf16_tma: unsupported output/activation combination (out_f32=%d res=%d act=%d)", d.out_f32,
            d.res != nullptr, d.act);
  return ZT_ERR_ARG;
}
