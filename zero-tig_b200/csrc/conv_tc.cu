// Stride-1 convolution as an implicit GEMM on the 5th-gen tensor cores (tcgen05 + TMEM).
//
// GEMM view: M = output pixels (a TH x 8 pixel patch per 128-row MMA tile), N = cout,
// K = taps * cin.  Activations stay fp32 NHWC in HBM (the layout every other kernel of the
// path uses); producer warps load the input patch INCLUDING its halo once per 32-channel
// group, split every value into bf16 hi + bf16 lo and store both in shared memory in the
// UMMA no-swizzle K-major canonical layout [k-chunk(8ch)][patch pixel][16 B].  Every filter
// tap is then just a *shifted view* of the same patch: the A-operand shared-memory
// descriptor starts at pixel (ky,kx) of the patch and uses SBO = patch-row pitch, so the
// 9 (or 5, 25, 49) taps re-use one copy of the data (no im2col, no re-reads).
// Weights are pre-packed on the host (bf16 hi/lo, canonical layout) and streamed per
// (channel group, tap) with cp.async.bulk + mbarrier complete_tx.
// One elected thread issues tcgen05.mma (kind::f16, bf16 x bf16 -> fp32 in TMEM):
//   SPLIT=1 : D += Ahi*Bhi                       (bf16 inputs, ~2^-8 relative)
//   SPLIT=3 : D += Ahi*Bhi + Alo*Bhi + Ahi*Blo   (~2^-16 relative: what RAFT's 0.01 px
//             EPE tolerance needs; TF32 alone is ~0.03 px - SURVEY F12)
// Epilogue: tcgen05.ld 32x32b -> bias/activation/GRU/residual epilogues -> fp32 NHWC.
//
// Replaces nn.Conv2d call sites with stride 1: model/model.py:24-28,40-44,74-81;
// model/RAFT/update.py:13-14,45-60,89-97,127-136; model/RAFT/extractor.py:46-55,167-191.
#include <cuda_bf16.h>
#include "common.cuh"

namespace zt {

constexpr int TC_TW = 8;          // tile width in pixels (= one 8-row core-matrix group)
constexpr int TC_TH = 16;         // tile height -> 128 GEMM rows
constexpr int TC_CG = 32;         // channels per A stage (4 chunks of 8)
constexpr int TC_NA = 2;          // A ring depth
constexpr int TC_THREADS = 192;   // warps 0-3 producer+epilogue, warp 4 MMA, warp 5 weight loader

struct TcGeom {
  int BN;            // N tile (multiple of 16, <= 256)
  int ngroups;       // ceil(cin / 32)
  int ntaps;         // kh*kw
  int PW, P;         // patch width, patch pixels
  int lboA;          // bytes between k-chunks of the A patch
  int tiles_x, tiles_y;
  int NB;            // weight ring depth
  int tmem_cols;     // power of two >= 32
};

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// shared-memory matrix descriptor, no swizzle, K-major: ((8,m),2):((16B,SBO),LBO)
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// split 4 fp32 into 4 bf16 hi (8 B) and 4 bf16 lo (8 B)
__device__ __forceinline__ void split4(const float4 v, uint2& hi, uint2& lo) {
  const __nv_bfloat16 h0 = __float2bfloat16_rn(v.x), h1 = __float2bfloat16_rn(v.y);
  const __nv_bfloat16 h2 = __float2bfloat16_rn(v.z), h3 = __float2bfloat16_rn(v.w);
  const __nv_bfloat16 l0 = __float2bfloat16_rn(v.x - __bfloat162float(h0));
  const __nv_bfloat16 l1 = __float2bfloat16_rn(v.y - __bfloat162float(h1));
  const __nv_bfloat16 l2 = __float2bfloat16_rn(v.z - __bfloat162float(h2));
  const __nv_bfloat16 l3 = __float2bfloat16_rn(v.w - __bfloat162float(h3));
  hi.x = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
  hi.y = (uint32_t)__bfloat16_as_ushort(h2) | ((uint32_t)__bfloat16_as_ushort(h3) << 16);
  lo.x = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
  lo.y = (uint32_t)__bfloat16_as_ushort(l2) | ((uint32_t)__bfloat16_as_ushort(l3) << 16);
}

// ---------------------------------------------------------------- epilogue (4 consecutive channels)
__device__ __forceinline__ void tc_epilogue4(const ZtConv& d, int64_t m, int nb, float x0, float x1, float x2,
                                             float x3) {
  float v[4] = {x0, x1, x2, x3};
  const int Ch = d.cout >> 1;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int n = nb + e;
    if (n >= d.cout) { v[e] = 0.f; continue; }
    float x = v[e];
    if (d.bias) x += __ldg(d.bias + n);
    switch (d.epi) {
      case ZT_EPI_GRU_ZR:
        x = 1.f / (1.f + expf(-x));
        if (n >= Ch) x *= d.aux0[m * d.aux0_cs + (n - Ch)];
        break;
      case ZT_EPI_GRU_Q: {
        const float z = d.aux1[m * d.aux1_cs + n];
        const float h = d.aux0[m * d.aux0_cs + n];
        x = (1.f - z) * h + z * tanhf(x);
      } break;
      case ZT_EPI_ADD_RELU:
        x = fmaxf(d.aux0[m * d.aux0_cs + n] + apply_act(x, d.act) * d.out_scale, 0.f);
        break;
      case ZT_EPI_ADD:
        x = d.aux0[m * d.aux0_cs + n] + apply_act(x, d.act) * d.out_scale;
        break;
      default:
        x = apply_act(x, d.act) * d.out_scale;
    }
    v[e] = x;
  }
  float* dst;
  int dn, dcs;
  if (d.epi == ZT_EPI_GRU_ZR && nb < Ch) { dst = d.aux1; dn = nb; dcs = d.aux1_cs; }
  else if (d.epi == ZT_EPI_GRU_ZR)       { dst = d.out; dn = nb - Ch; dcs = d.out_cs; }
  else                                   { dst = d.out; dn = nb; dcs = d.out_cs; }
  float* p = dst + m * dcs + dn;
  if (nb + 3 < d.cout && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (nb + e < d.cout) p[e] = v[e];
  }
}

// ---------------------------------------------------------------- kernel
template <int SPLIT>
__global__ void __launch_bounds__(TC_THREADS) conv_tc_kernel(const ZtConv d, const TcGeom g) {
  extern __shared__ __align__(128) uint8_t smem[];
  constexpr int NPART = SPLIT == 3 ? 2 : 1;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 240);
  const uint32_t a_part = 4u * g.lboA, a_stage = a_part * NPART;
  const uint32_t b_part = 4u * g.BN * 16u, b_stage = b_part * NPART;
  uint8_t* smA = smem + 256;
  uint8_t* smB = smA + TC_NA * a_stage;
  const uint32_t bar0 = smem_u32(bars);
  auto fullA = [&](int s) { return bar0 + 8u * s; };
  auto emptyA = [&](int s) { return bar0 + 8u * (TC_NA + s); };
  auto fullB = [&](int s) { return bar0 + 8u * (2 * TC_NA + s); };
  auto emptyB = [&](int s) { return bar0 + 8u * (2 * TC_NA + g.NB + s); };
  const uint32_t accFull = bar0 + 8u * (2 * TC_NA + 2 * g.NB);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tyi = blockIdx.x / g.tiles_x, txi = blockIdx.x - tyi * g.tiles_x;
  const int oy0 = tyi * TC_TH, ox0 = txi * TC_TW;
  const int n0 = blockIdx.y * g.BN;
  const int bimg = blockIdx.z;
  const int cin = d.c0 + d.c1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < TC_NA; ++s) { mbar_init(fullA(s), 128); mbar_init(emptyA(s), 1); }
    for (int s = 0; s < g.NB; ++s) { mbar_init(fullB(s), 1); mbar_init(emptyB(s), 1); }
    mbar_init(accFull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(tmem_slot)), "r"((uint32_t)g.tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    // ===================== producers: fp32 patch -> bf16 hi/lo canonical smem =====================
    const int tid = threadIdx.x;
    const int q = tid & 7, pl = tid >> 3;
    const uint32_t soff = (uint32_t)(q >> 1) * g.lboA + (uint32_t)(q & 1) * 8u;
    const int64_t img_pix = (int64_t)bimg * d.H * d.W;
    for (int gi = 0; gi < g.ngroups; ++gi) {
      const int s = gi & 1;
      mbar_wait(emptyA(s), ((gi >> 1) & 1) ^ 1);
      uint8_t* dst_hi = smA + s * a_stage;
      const int c = gi * TC_CG + q * 4;
      const bool cvalid = c < cin;
      const float* src = d.in0;
      int cs = d.in0_cs;
      if (cvalid) {
        if (c < d.c0) { src = d.in0 + c; }
        else { src = d.in1 + (c - d.c0); cs = d.in1_cs; }
      }
      int py = pl / g.PW, px = pl - py * g.PW;
      for (int p0 = pl; p0 < g.P; p0 += 64) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int p = p0 + 16 * u;
          const int iy = oy0 - d.ph + py, ix = ox0 - d.pw + px;
          v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (cvalid && p < g.P && iy >= 0 && iy < d.H && ix >= 0 && ix < d.W)
            v[u] = __ldg(reinterpret_cast<const float4*>(src + (img_pix + (int64_t)iy * d.W + ix) * cs));
          px += 16;
          while (px >= g.PW) { px -= g.PW; ++py; }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int p = p0 + 16 * u;
          if (p < g.P) {
            uint2 hi, lo;
            split4(v[u], hi, lo);
            *reinterpret_cast<uint2*>(dst_hi + soff + (uint32_t)p * 16u) = hi;
            if (SPLIT == 3) *reinterpret_cast<uint2*>(dst_hi + a_part + soff + (uint32_t)p * 16u) = lo;
          }
        }
      }
      fence_async_smem();
      mbar_arrive(fullA(s));
    }
    // ===================== epilogue: TMEM -> registers -> global =====================
    mbar_wait(accFull, 0);
    tc_fence_after();
    const int r = warp * 32 + lane;
    const int oy = oy0 + (r >> 3), ox = ox0 + (r & 7);
    const bool ok = oy < d.Ho && ox < d.Wo;
    const int64_t m = ((int64_t)bimg * d.Ho + oy) * d.Wo + ox;
    for (int col0 = 0; col0 < g.BN; col0 += 16) {
      if (n0 + col0 >= d.cout) break;
      uint32_t v[16];
      tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)col0, v);
      if (ok) {
#pragma unroll
        for (int e = 0; e < 16; e += 4)
          tc_epilogue4(d, m, n0 + col0 + e, __uint_as_float(v[e]), __uint_as_float(v[e + 1]),
                       __uint_as_float(v[e + 2]), __uint_as_float(v[e + 3]));
      }
    }
  } else if (warp == 4) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(g.BN >> 3) << 17) | (8u << 24);
      const uint32_t sboA = (uint32_t)g.PW * 16u, lboB = (uint32_t)g.BN * 16u;
      uint32_t it = 0, acc = 0;
      for (int gi = 0; gi < g.ngroups; ++gi) {
        const int s = gi & 1;
        mbar_wait(fullA(s), (gi >> 1) & 1);
        tc_fence_after();
        const int nks = (cin - gi * TC_CG) > 16 ? 2 : 1;
        const uint32_t a_hi = smem_u32(smA + s * a_stage), a_lo = a_hi + a_part;
        for (int tap = 0; tap < g.ntaps; ++tap, ++it) {
          const int ky = tap / d.kw, kx = tap - ky * d.kw;
          const uint32_t sb = it % g.NB;
          mbar_wait(fullB(sb), (it / g.NB) & 1);
          tc_fence_after();
          const uint32_t b_hi = smem_u32(smB + sb * b_stage), b_lo = b_hi + b_part;
          const uint32_t toff = (uint32_t)(ky * g.PW + kx) * 16u;
          for (int ks = 0; ks < nks; ++ks) {
            const uint64_t dah = make_desc(a_hi + toff + ks * 2 * g.lboA, g.lboA, sboA);
            const uint64_t dbh = make_desc(b_hi + ks * 2 * lboB, lboB, 128u);
            tc_mma_bf16(tmem_base, dah, dbh, idesc, acc);
            acc = 1;
            if (SPLIT == 3) {
              const uint64_t dal = make_desc(a_lo + toff + ks * 2 * g.lboA, g.lboA, sboA);
              const uint64_t dbl = make_desc(b_lo + ks * 2 * lboB, lboB, 128u);
              tc_mma_bf16(tmem_base, dal, dbh, idesc, 1);
              tc_mma_bf16(tmem_base, dah, dbl, idesc, 1);
            }
          }
          tc_commit(emptyB(sb));
        }
        tc_commit(emptyA(s));
      }
      tc_commit(accFull);
    }
    __syncwarp();
  } else {
    // ===================== weight loader (one thread, bulk async copies) =====================
    if (lane == 0) {
      const uint32_t nstage = (uint32_t)g.ngroups * g.ntaps;
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(d.weight) + (size_t)blockIdx.y * nstage * (2u * b_part);
      for (uint32_t it = 0; it < nstage; ++it) {
        const uint32_t sb = it % g.NB;
        mbar_wait(emptyB(sb), ((it / g.NB) & 1) ^ 1);
        mbar_arrive_expect_tx(fullB(sb), b_stage);
        bulk_g2s(smem_u32(smB + sb * b_stage), wsrc + (size_t)it * (2u * b_part), b_stage, fullB(sb));
      }
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)g.tmem_cols)
                 : "memory");
  }
}

}  // namespace zt

extern "C" int zt_conv2d_tc_nhwc(const ZtConv* dp, int BN, int split, void* stream) {
  using namespace zt;
  ZT_REQUIRE(dp != nullptr, "zt_conv2d_tc_nhwc: null descriptor");
  const ZtConv& d = *dp;
  const int cin = d.c0 + d.c1;
  ZT_REQUIRE(d.in0 && d.weight && d.out, "conv_tc: null pointer");
  ZT_REQUIRE(split == 1 || split == 3, "conv_tc: split must be 1 (bf16) or 3 (bf16 hi/lo, 3 products); got %d", split);
  ZT_REQUIRE(BN >= 16 && BN <= 256 && (BN & 15) == 0, "conv_tc: BN=%d must be a multiple of 16 in [16,256]", BN);
  ZT_REQUIRE(d.sh == 1 && d.sw == 1, "conv_tc: stride-1 convolutions only (got %dx%d)", d.sh, d.sw);
  ZT_REQUIRE((d.c0 & 3) == 0 && (d.c1 & 3) == 0 && cin > 0, "conv_tc: channel counts must be multiples of 4 (c0=%d c1=%d)", d.c0, d.c1);
  ZT_REQUIRE((d.in0_cs & 3) == 0 && (d.c1 == 0 || ((d.in1_cs & 3) == 0 && d.in1)), "conv_tc: input channel strides must be multiples of 4");
  ZT_REQUIRE(((uintptr_t)d.in0 & 15) == 0 && ((uintptr_t)d.weight & 15) == 0 && (d.c1 == 0 || ((uintptr_t)d.in1 & 15) == 0),
             "conv_tc: pointers must be 16-byte aligned");
  ZT_REQUIRE(d.B > 0 && d.H > 0 && d.W > 0 && d.cout > 0 && d.kh >= 1 && d.kw >= 1 && d.kh <= 7 && d.kw <= 7, "conv_tc: bad dims");
  ZT_REQUIRE(d.Ho == d.H + 2 * d.ph - d.kh + 1 && d.Wo == d.W + 2 * d.pw - d.kw + 1 && d.Ho > 0 && d.Wo > 0,
             "conv_tc: output dims inconsistent");
  if (d.epi == ZT_EPI_GRU_ZR) ZT_REQUIRE(d.aux0 && d.aux1 && (d.cout & 7) == 0, "conv_tc: GRU_ZR needs aux0/aux1 and cout%%8==0");
  if (d.epi == ZT_EPI_GRU_Q) ZT_REQUIRE(d.aux0 && d.aux1, "conv_tc: GRU_Q needs aux0/aux1");
  if (d.epi == ZT_EPI_ADD_RELU || d.epi == ZT_EPI_ADD) ZT_REQUIRE(d.aux0, "conv_tc: residual epilogue needs aux0");
  ZT_REQUIRE((int64_t)d.B * d.H * d.W < (1ll << 31) && d.B <= 65535, "conv_tc: too many pixels");

  TcGeom g;
  g.BN = BN;
  g.ngroups = ceil_div(cin, TC_CG);
  g.ntaps = d.kh * d.kw;
  g.PW = TC_TW + d.kw - 1;
  g.P = (TC_TH + d.kh - 1) * g.PW;
  g.lboA = ceil_div(g.P * 16, 128) * 128 + 32;   // +32 B: conflict-free producer stores
  g.tiles_x = ceil_div(d.Wo, TC_TW);
  g.tiles_y = ceil_div(d.Ho, TC_TH);
  g.tmem_cols = 32;
  while (g.tmem_cols < BN) g.tmem_cols <<= 1;
  const int npart = split == 3 ? 2 : 1;
  const size_t a_bytes = (size_t)TC_NA * 4 * g.lboA * npart;
  const size_t b_stage = (size_t)4 * BN * 16 * npart;
  // aim for <= ~110 KB per CTA so two CTAs share an SM (one's epilogue overlaps the other's MMAs)
  const size_t budget = 110 * 1024;
  int nb = a_bytes + 256 < budget ? (int)((budget - a_bytes - 256) / b_stage) : 2;
  g.NB = nb < 2 ? 2 : (nb > 8 ? 8 : nb);
  const size_t smem = 256 + a_bytes + (size_t)g.NB * b_stage;
  ZT_REQUIRE(smem <= 227 * 1024, "conv_tc: tile needs %zu B of shared memory", smem);

  static bool attr_done = false;
  if (!attr_done) {
    ZT_CUDA(cudaFuncSetAttribute(conv_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    ZT_CUDA(cudaFuncSetAttribute(conv_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    attr_done = true;
  }
  dim3 grid(g.tiles_x * g.tiles_y, ceil_div(d.cout, BN), d.B);
  cudaStream_t st = (cudaStream_t)stream;
  if (split == 3) conv_tc_kernel<3><<<grid, TC_THREADS, smem, st>>>(d, g);
  else conv_tc_kernel<1><<<grid, TC_THREADS, smem, st>>>(d, g);
  return check_launch("conv_tc_kernel");
}
