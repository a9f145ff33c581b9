// Convolution weight gradient on the tensor cores (train.py:129 loss.backward(): dW of every nn.Conv2d of the three
// Retinex networks, model/model.py:15-81):   dW[tap][ci][co] = sum over pixels p of X[p + tap][ci] * G[p][co].
//
// GEMM view: M = (tap, ci), N = co, K = pixels.  Both operands are NHWC (channels contiguous), i.e. "MN-major" for
// this GEMM: a 128-byte shared-memory row is ONE pixel (K index) with its 64 channels (M or N index), eight pixels of
// an image row are one 1024-byte swizzle atom - exactly what a TMA box (64 ch, 8 + kw - 1 px, 16 + kh - 1 rows,
// SWIZZLE_128B) writes.  tcgen05 consumes MN-major operands directly (instruction-descriptor bits 15 / 16), so
//   * the X patch (16 x 8 pixel tile + halo, zero filled outside the image = the convolution's padding) is loaded
//     once per tile and every filter tap is a shifted view of it (start address + (ky * PW + kx) * 128 B);
//   * two taps are stacked into one M = 128 MMA through the descriptor's leading-dimension offset (the second 64
//     rows start one tap further), five MMA groups cover the nine taps of a 3x3 filter (the tenth half is ignored);
//   * a K step is 16 pixels = two image rows of the tile (stride-dimension offset = patch row pitch);
//   * products are the three-term bf16 hi/lo split (Xhi*Ghi + Xlo*Ghi + Xhi*Glo, fp32-grade), accumulated in TMEM
//     over ALL tiles of a persistent CTA; one atomic flush of the 5 x 128 x 64 accumulator block per CTA at the end.
// The operands arrive as bf16 hi/lo pairs (zt_split_bf16_hilo); round 1's kernel did this contraction in fp32 FFMA
// at ~19 TFLOP/s and was half of the training step.
#include <cuda.h>
#include <cuda_bf16.h>
#include <string.h>
#include <algorithm>
#include "common.cuh"
#include "tc_ptx.cuh"
#include "tma_host.cuh"

namespace zt {

constexpr int WG_THREADS = 192;            // warps 0-3 final flush (TMEM lane quadrants), 4 MMA issuer, 5 TMA loader
constexpr int WG_W_MMA = 4, WG_W_LOAD = 5;
constexpr int WG_BAR_BYTES = 1024;
constexpr int WG_STAGES = 2;
constexpr uint32_t WG_G_PART = 16384;      // 128 pixels x 128 B
constexpr uint32_t WG_TMEM_COLS = 512;

struct WgArgs {
  float* dW;                 // [kh][kw][cin_p][w_cs] fp32, accumulated with atomics
  int H, W, kh, kw, ph, pw, cin, cout, cin_p, w_cs;
  int PW, tiles_x, tiles;
  uint32_t x_part, x_box, stage;
  int npairs, N;             // tap pairs (M = 128 each), N = ceil16(cout)
};

// MN-major SWIZZLE_128B shared-memory descriptor: LBO = byte offset between 64-element blocks along M/N,
// SBO = byte offset between 8-row groups along K (cute/atom/mma_traits_sm100.hpp, make_umma_desc<Major::MN>)
__device__ __forceinline__ uint64_t wg_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
         (1ull << 46) | (2ull << 61);
}

__global__ void __launch_bounds__(WG_THREADS, 1) wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmXh,
                                                                 const __grid_constant__ CUtensorMap tmXl,
                                                                 const __grid_constant__ CUtensorMap tmGh,
                                                                 const __grid_constant__ CUtensorMap tmGl, const WgArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const uint32_t bar0 = smem_u32(smem);
  auto full = [&](uint32_t s) { return bar0 + 8u * s; };
  auto empty = [&](uint32_t s) { return bar0 + 8u * (WG_STAGES + s); };
  const uint32_t accFull = bar0 + 8u * (2 * WG_STAGES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 512);
  uint8_t* stage0 = smem + WG_BAR_BYTES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < WG_STAGES; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
    mbar_init(accFull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == WG_W_MMA) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(tmem_slot)), "r"(WG_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int ntl = a.tiles > (int)blockIdx.x ? (a.tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp == WG_W_LOAD) {
    if (lane == 0) {
      tma_prefetch_desc(&tmXh); tma_prefetch_desc(&tmXl); tma_prefetch_desc(&tmGh); tma_prefetch_desc(&tmGl);
      uint32_t s = 0, ph = 1;
      for (int i = 0; i < ntl; ++i) {
        const int t = blockIdx.x + i * gridDim.x;
        const int tyi = t / a.tiles_x, txi = t - tyi * a.tiles_x;
        mbar_wait(empty(s), ph);
        mbar_arrive_expect_tx(full(s), 2u * a.x_box + 2u * WG_G_PART);
        const uint32_t dst = smem_u32(stage0 + s * a.stage);
        tma_load_3d(dst, &tmXh, 0, txi * 8 - a.pw, tyi * 16 - a.ph, full(s));
        tma_load_3d(dst + a.x_part, &tmXl, 0, txi * 8 - a.pw, tyi * 16 - a.ph, full(s));
        tma_load_3d(dst + 2 * a.x_part, &tmGh, 0, txi * 8, tyi * 16, full(s));
        tma_load_3d(dst + 2 * a.x_part + WG_G_PART, &tmGl, 0, txi * 8, tyi * 16, full(s));
        if (++s == WG_STAGES) { s = 0; ph ^= 1; }
      }
    }
    __syncwarp();
  } else if (warp == WG_W_MMA) {
    // M = 128 (two taps x 64 input channels), N = ceil16(cout), K = 16 pixels; A and B MN-major, bf16, fp32 accumulate
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | (((uint32_t)a.N >> 3) << 17) | (8u << 24);
    const uint32_t PW = a.PW, ntaps = a.kh * a.kw;
    uint32_t s = 0, ph = 0;
    for (int i = 0; i < ntl; ++i) {
      mbar_wait(full(s), ph);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t xh = smem_u32(stage0 + s * a.stage), xl = xh + a.x_part;
        const uint32_t gh = xh + 2 * a.x_part, gl = gh + WG_G_PART;
        for (uint32_t j = 0; j < 8; ++j) {                       // K step = image rows 2j, 2j + 1 of the tile
          const uint64_t dGh = wg_desc(gh + j * 2048u, 0u, 1024u), dGl = wg_desc(gl + j * 2048u, 0u, 1024u);
          for (uint32_t q = 0; q < (uint32_t)a.npairs; ++q) {
            const uint32_t t0 = 2 * q, t1 = min(2 * q + 1, ntaps - 1);     // an odd tap count repeats the last tap (ignored)
            const uint32_t o0 = ((t0 / a.kw) * PW + (t0 % a.kw)) * 128u, o1 = ((t1 / a.kw) * PW + (t1 % a.kw)) * 128u;
            const uint32_t off = (2 * j * PW) * 128u + o0;
            const uint64_t dXh = wg_desc(xh + off, o1 - o0, PW * 128u), dXl = wg_desc(xl + off, o1 - o0, PW * 128u);
            const uint32_t d = tmem_base + q * 64u;
            const uint32_t first = (i == 0 && j == 0) ? 0u : 1u;
            tc_mma_bf16(d, dXh, dGh, idesc, first);
            tc_mma_bf16(d, dXl, dGh, idesc, 1u);
            tc_mma_bf16(d, dXh, dGl, idesc, 1u);
          }
        }
        tc_commit(empty(s));
        if (i + 1 == ntl) tc_commit(accFull);
      }
      __syncwarp();
      if (++s == WG_STAGES) { s = 0; ph ^= 1; }
    }
  } else if (ntl > 0) {
    // final flush: TMEM lane = (tap parity, input channel), columns = output channels of pair q
    mbar_wait(accFull, 0);
    tc_fence_after();
    const int row = warp * 32 + lane, half = row >> 6, ci = row & 63;
    const int ntaps = a.kh * a.kw;
    for (int q = 0; q < a.npairs; ++q) {
      const int tap = 2 * q + half;
      const bool ok = tap < ntaps && ci < a.cin;
      float* dst = a.dW + ((int64_t)tap * a.cin_p + ci) * a.w_cs;
      for (int c0 = 0; c0 < a.N; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(q * 64 + c0), v);
        if (ok) {
#pragma unroll
          for (int e = 0; e < 16; ++e)
            if (c0 + e < a.cout) atomicAdd(dst + c0 + e, __uint_as_float(v[e]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == WG_W_MMA) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(WG_TMEM_COLS) : "memory");
  }
}

// fp32 NHWC [M][in_rs] -> bf16 hi / lo [M][out_rs] (hi = bf16(x), lo = bf16(x - hi)), C channels (C % 4 == 0), and,
// when colsum != NULL, colsum[c] += sum over rows (the bias gradient of the layer whose output gradient this is):
// one pass over the tensor instead of a separate reduction.
__global__ void __launch_bounds__(256) split_colsum_kernel(const float* __restrict__ in, int in_rs, __nv_bfloat16* __restrict__ hi,
                                                          __nv_bfloat16* __restrict__ lo, int out_rs, int C4, int64_t M,
                                                          int rows_per_block, float* __restrict__ colsum) {
  const int cg = threadIdx.x % C4, r0 = threadIdx.x / C4, rstep = blockDim.x / C4;
  const int64_t m0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t m1 = r0 >= rstep ? 0 : ((m0 + rows_per_block < M) ? m0 + rows_per_block : M);     // idle threads: empty range
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  auto emit = [&](int64_t m, const float4 v) {
    const float f[4] = {v.x, v.y, v.z, v.w};
    unsigned short h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat16 hh = __float2bfloat16_rn(f[e]);
      h[e] = __bfloat16_as_ushort(hh);
      l[e] = __bfloat16_as_ushort(__float2bfloat16_rn(f[e] - __bfloat162float(hh)));
    }
    *reinterpret_cast<uint2*>(hi + m * out_rs + cg * 4) = make_uint2((uint32_t)h[0] | ((uint32_t)h[1] << 16), (uint32_t)h[2] | ((uint32_t)h[3] << 16));
    *reinterpret_cast<uint2*>(lo + m * out_rs + cg * 4) = make_uint2((uint32_t)l[0] | ((uint32_t)l[1] << 16), (uint32_t)l[2] | ((uint32_t)l[3] << 16));
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  };
  int64_t m = m0 + r0;
  for (; m + 3 * rstep < m1; m += 4 * rstep) {        // four independent rows in flight per thread
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) v[u] = __ldg(reinterpret_cast<const float4*>(in + (m + (int64_t)u * rstep) * in_rs + cg * 4));
#pragma unroll
    for (int u = 0; u < 4; ++u) emit(m + (int64_t)u * rstep, v[u]);
  }
  for (; m < m1; m += rstep) emit(m, __ldg(reinterpret_cast<const float4*>(in + m * in_rs + cg * 4)));
  if (colsum) {
    // block-level reduction first: thousands of blocks adding into the same <= 64 addresses serialise in L2
    __shared__ float4 red[256];
    red[threadIdx.x] = acc;
    __syncthreads();
    if (r0 == 0) {
      for (int r = 1; r < rstep; ++r) {
        const float4 t = red[r * C4 + cg];
        acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
      }
      if (acc.x != 0.f) atomicAdd(colsum + cg * 4, acc.x);
      if (acc.y != 0.f) atomicAdd(colsum + cg * 4 + 1, acc.y);
      if (acc.z != 0.f) atomicAdd(colsum + cg * 4 + 2, acc.z);
      if (acc.w != 0.f) atomicAdd(colsum + cg * 4 + 3, acc.w);
    }
  }
}

}  // namespace zt

/* fp32 [M][in_rs] -> bf16 hi / lo [M][out_rs] for C channels (C % 4 == 0, C <= 256); colsum (may be NULL, >= C floats):
 * colsum[c] += sum_m in[m][c] */
extern "C" int zt_split_bf16_hilo_colsum(const float* in, int in_rs, void* hi, void* lo, int out_rs, int C, int64_t M,
                                         float* colsum, void* stream) {
  using namespace zt;
  ZT_REQUIRE(in && hi && lo && C > 0 && C <= 256 && (C & 3) == 0 && (in_rs & 3) == 0 && (out_rs & 3) == 0 && in_rs >= C &&
             out_rs >= C && M > 0, "split_bf16_hilo_colsum: bad args");
  // ~8 rows per thread: enough blocks to fill the machine several times over, few enough atomics for the column sums
  const int rstep = 256 / (C / 4);
  const int rpb = 32 * rstep;
  split_colsum_kernel<<<(int)ceil_div64(M, rpb), 256, 0, (cudaStream_t)stream>>>(in, in_rs, (__nv_bfloat16*)hi, (__nv_bfloat16*)lo,
                                                                                 out_rs, C / 4, M, rpb, colsum);
  return check_launch("split_colsum_kernel");
}

/* dW[kh][kw][cin_p][w_cs] (fp32) += sum_p X[p + tap][ci] * G[p][co] on the tensor cores; X, G as bf16 hi/lo NHWC pairs
 * ([H*W][x_cs] / [H*W][g_cs], channel strides multiples of 8).  cin, cout <= 64; 1x1 or 3x3 (pad = k / 2). */
extern "C" int zt_conv2d_wgrad_tc(const void* Xhi, const void* Xlo, int x_cs, const void* Ghi, const void* Glo, int g_cs,
                                  float* dW, int H, int W, int kh, int kw, int cin, int cout, int cin_p, int w_cs,
                                  void* stream) {
  using namespace zt;
  ZT_REQUIRE(Xhi && Xlo && Ghi && Glo && dW && H > 0 && W > 0, "conv2d_wgrad_tc: bad args");
  ZT_REQUIRE(((kh == 3 && kw == 3) || (kh == 1 && kw == 1)), "conv2d_wgrad_tc: 1x1 or 3x3 only (got %dx%d)", kh, kw);
  ZT_REQUIRE(cin >= 1 && cin <= 64 && cout >= 1 && cout <= 64 && cin_p >= cin && w_cs >= cout, "conv2d_wgrad_tc: cin/cout must be <= 64");
  ZT_REQUIRE((x_cs & 7) == 0 && (g_cs & 7) == 0 && x_cs >= cin && g_cs >= cout, "conv2d_wgrad_tc: channel strides must be multiples of 8");
  ZT_REQUIRE((((uintptr_t)Xhi | (uintptr_t)Xlo | (uintptr_t)Ghi | (uintptr_t)Glo) & 15) == 0, "conv2d_wgrad_tc: operands must be 16-byte aligned");
  WgArgs a;
  memset(&a, 0, sizeof(a));
  a.dW = dW; a.H = H; a.W = W; a.kh = kh; a.kw = kw; a.ph = kh / 2; a.pw = kw / 2;
  a.cin = cin; a.cout = cout; a.cin_p = cin_p; a.w_cs = w_cs;
  a.PW = 8 + kw - 1;
  const int PH = 16 + kh - 1;
  a.tiles_x = ceil_div(W, 8);
  a.tiles = a.tiles_x * ceil_div(H, 16);
  a.x_box = (uint32_t)(a.PW * PH) * 128u;
  a.x_part = (uint32_t)ceil_div((int)a.x_box, 1024) * 1024u;
  a.stage = 2 * a.x_part + 2 * WG_G_PART;
  a.npairs = ceil_div(kh * kw, 2);
  a.N = ceil_div(cout, 16) * 16;
  CUtensorMap tm[4];
  const void* bases[4] = {Xhi, Xlo, Ghi, Glo};
  for (int i = 0; i < 4; ++i) {
    const bool isx = i < 2;
    const int c = isx ? cin : cout, cs = isx ? x_cs : g_cs;
    // the tensor map carries the REAL channel count: channels beyond it are zero filled by TMA whatever the padding holds
    const int dims[3] = {c, W, H};
    const int64_t strides[2] = {cs, (int64_t)W * cs};
    const int box[3] = {64, isx ? a.PW : 8, isx ? PH : 16};
    if (int e = tma_map_bf16(bases[i], 3, dims, strides, box, &tm[i])) return e;
  }
  const size_t smem = WG_BAR_BYTES + (size_t)WG_STAGES * a.stage;
  static bool attr_done = false;
  if (!attr_done) {
    ZT_CUDA(cudaFuncSetAttribute(wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_done = true;
  }
  const int gx = std::min(kNumSMs, a.tiles);
  wgrad_tc_kernel<<<gx, WG_THREADS, smem, (cudaStream_t)stream>>>(tm[0], tm[1], tm[2], tm[3], a);
  return check_launch("wgrad_tc_kernel");
}
