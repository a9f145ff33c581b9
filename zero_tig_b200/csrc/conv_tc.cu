// Stride-1 convolution as an implicit GEMM on the 5th-gen tensor cores (tcgen05 + TMEM).
//
// GEMM view: M = output pixels (a TH x 8 pixel patch per 128-row MMA tile), N = cout,
// K = taps * cin.  Activations stay fp32 NHWC in HBM (the layout every other kernel of the
// path uses); producer warps load the input patch INCLUDING its halo once per 32-channel
// group, split every value into bf16 hi + bf16 lo and store both in shared memory in the
// UMMA 128-byte-swizzle K-major layout: one 128 B row (64 channels) per patch pixel, 16 B
// chunk c of row p stored at chunk position c ^ (p & 7).  (The no-swizzle layout works too
// but is fetched at ~16 B/clk: measured ~210 clk per M128 K16 MMA.)  Every filter tap is
// just a *shifted view* of the same patch: the A-operand shared-memory descriptor starts at
// pixel (ky,kx) of the patch and uses SBO = patch-row pitch; the hardware swizzle is a
// function of the absolute shared-memory address, so any row offset stays consistent.  The
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
#include <cuda_fp16.h>
#include <stdlib.h>
#include <algorithm>
#include "common.cuh"
#include "tc_ptx.cuh"

namespace zt {

constexpr int TC_TW = 8;          // tile width in pixels (= one 8-row core-matrix group)
constexpr int TC_TH = 16;         // tile height -> 128 GEMM rows
constexpr int TC_CG = 64;         // channels per A stage: one 128 B swizzled row per pixel
// warps 0-7 epilogue (warp%4 = TMEM lane quadrant, warp/4 = odd/even 16-column groups),
// 8 MMA issuer, 9 weight loader, 10-17 producers (all 256 threads fill each A stage together)
constexpr int TC_THREADS = 576;
constexpr int TC_PB = 12;         // patch pixels (float4 loads) in flight per producer thread
constexpr int TC_W_MMA = 8, TC_W_LOAD = 9, TC_W_PROD = 10;
constexpr int TC_BAR_BYTES = 1024;
constexpr int TC_EPI_LD = 36;                               // staging row pitch in floats (32 + 4 pad)
constexpr int TC_EPI_BYTES = 8 * 32 * TC_EPI_LD * 4;        // 8 epilogue warps x 32 rows

struct TcGeom {
  int BN;            // N tile (multiple of 16, <= 256)
  int ngroups;       // ceil(cin / 64)
  int ntaps;         // kh*kw
  int PW, P;         // patch width, patch pixels
  int a_part;        // bytes of one (hi or lo) patch: P*128 rounded up to 1024
  int tiles_x, tiles_y;
  int NA;            // A (patch) ring depth
  int NB;            // weight ring depth (resident: number of weight stages = ngroups*ntaps)
  int resident;      // 1: all weights of this N tile stay in shared memory for every pixel tile
  int tmem_cols;     // power of two >= 32, holds two accumulators of BN columns
  int S;             // pixel tiles over all images = tiles_x*tiles_y*B
  int pdl;           // launched with programmatic stream serialization: prologue + constant weights run early
  int w_early;       // weights are constants (not produced by the preceding kernel): fetch them before the wait
  int a_off, epi_off;  // byte offsets of the A ring and of the epilogue staging tile (aliased onto the A ring when a CTA
                     // owns a single pixel tile: the ring is dead once the accumulator is complete)
  long long* trace;  // debug: per-role timestamps of CTA (0,0) (zt_debug_set_tc_trace), else NULL
  int trace_cap;
  int fp16;          // split mode 4: operands are fp16 hi/lo pairs of the SCALED values (activation * ZT_F16_ACT_SCALE,
                     // weight * per-layer power of two): 2^-22 per operand instead of the 2^-17 of the bf16 pair
  float acc_scale;   // accumulator * acc_scale = the unscaled convolution sum (1 unless fp16)
};

// debug timeline: record (event, index, globaltimer ns) for CTA (0,0); every role (warp) owns a
// region of trace_cap entries and a private counter, so tracing costs one fire-and-forget store
struct TcTracer {
  long long* base;
  int n, cap;
  __device__ __forceinline__ TcTracer(const TcGeom& g, int role)
      : base(nullptr), n(0), cap(g.trace_cap) {
    if (g.trace && blockIdx.x == 0 && blockIdx.y == 0) base = g.trace + (size_t)role * 2 * g.trace_cap;
  }
  __device__ __forceinline__ void operator()(int idx, int ev) {
    if (base && n < cap) {
      long long tns;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tns));
      base[2 * n] = ((long long)(ev + 1) << 32) | (unsigned)idx;
      base[2 * n + 1] = tns;
      ++n;
    }
  }
};

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

// the same for split mode 4: 4 fp32 (already multiplied by the activation scale) -> 4 fp16 hi + 4 fp16 lo
__device__ __forceinline__ void split4h(const float4 v, uint2& hi, uint2& lo) {
  const __half h0 = __float2half_rn(v.x), h1 = __float2half_rn(v.y), h2 = __float2half_rn(v.z), h3 = __float2half_rn(v.w);
  const __half l0 = __float2half_rn(v.x - __half2float(h0)), l1 = __float2half_rn(v.y - __half2float(h1));
  const __half l2 = __float2half_rn(v.z - __half2float(h2)), l3 = __float2half_rn(v.w - __half2float(h3));
  hi.x = (uint32_t)__half_as_ushort(h0) | ((uint32_t)__half_as_ushort(h1) << 16);
  hi.y = (uint32_t)__half_as_ushort(h2) | ((uint32_t)__half_as_ushort(h3) << 16);
  lo.x = (uint32_t)__half_as_ushort(l0) | ((uint32_t)__half_as_ushort(l1) << 16);
  lo.y = (uint32_t)__half_as_ushort(l2) | ((uint32_t)__half_as_ushort(l3) << 16);
}

// ---------------------------------------------------------------- epilogue
__device__ __forceinline__ float4 ld4_or_scalar(const float* p, int nvalid) {
  if (nvalid >= 4 && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) return *reinterpret_cast<const float4*>(p);
  float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
  if (nvalid > 0) r.x = p[0];
  if (nvalid > 1) r.y = p[1];
  if (nvalid > 2) r.z = p[2];
  if (nvalid > 3) r.w = p[3];
  return r;
}

template <int ACT>
__device__ __forceinline__ float act_t(float v, int act_rt) {
  if constexpr (ACT == ZT_ACT_NONE) return v;
  else if constexpr (ACT == ZT_ACT_RELU) return fmaxf(v, 0.f);
  else if constexpr (ACT == ZT_ACT_LRELU02) return v > 0.f ? v : 0.2f * v;
  else if constexpr (ACT == ZT_ACT_SIGMOID) return 1.f / (1.f + expf(-v));
  else if constexpr (ACT == ZT_ACT_TANH) return tanhf(v);
  else return apply_act(v, act_rt);
}

// EPI / ACT are compile-time for the combinations the hot path uses (the epilogue warps
// bound the kernel otherwise); EPI_T = -1 is the generic run-time fallback.
template <int EPI_T, int ACT_T>
__device__ __forceinline__ void tc_epilogue4(const ZtConv& d, int64_t m, int nb, int nend, float x0, float x1,
                                             float x2, float x3, const float4 bv, const float4 a0v,
                                             const float4 a1v) {
  const int epi = EPI_T >= 0 ? EPI_T : d.epi;
  float v[4] = {x0, x1, x2, x3};
  const float a0[4] = {a0v.x, a0v.y, a0v.z, a0v.w}, a1[4] = {a1v.x, a1v.y, a1v.z, a1v.w};
  const int Ch = d.cout >> 1;
  const int nvalid = min(4, nend - nb);   // nend: end of this CTA's channel range
  if (nvalid <= 0) return;
  const float b[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    float x = v[e] + b[e];
    if (epi == ZT_EPI_GRU_ZR) {
      x = __fdividef(1.f, 1.f + __expf(-x));            // ~1e-6 absolute: far inside the 0.01 px EPE gate
      if (nb >= Ch) x *= a0[e];
    } else if (epi == ZT_EPI_GRU_Q) {
      const float th = 1.f - __fdividef(2.f, 1.f + __expf(2.f * x));   // tanh(x), saturates cleanly to +-1
      x = (1.f - a1[e]) * a0[e] + a1[e] * th;
    } else if (epi == ZT_EPI_ADD_RELU) {
      x = fmaxf(a0[e] + act_t<ACT_T>(x, d.act) * d.out_scale, 0.f);
    } else if (epi == ZT_EPI_ADD) {
      x = a0[e] + act_t<ACT_T>(x, d.act) * d.out_scale;
    } else {
      x = act_t<ACT_T>(x, d.act) * d.out_scale;
    }
    v[e] = x;
  }
  float* dst;
  int dn, dcs;
  if (epi == ZT_EPI_GRU_ZR && nb < Ch) { dst = d.aux1; dn = nb; dcs = d.aux1_cs; }
  else if (epi == ZT_EPI_GRU_ZR)       { dst = d.out; dn = nb - Ch; dcs = d.out_cs; }
  else                                 { dst = d.out; dn = nb; dcs = d.out_cs; }
  float* p = dst + m * dcs + dn;
  if (nvalid == 4 && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (e < nvalid) p[e] = v[e];
  }
}

// ---------------------------------------------------------------- kernel
// Persistent, warp-specialised: each CTA (one per SM) walks pixel tiles t = blockIdx.x,
// +gridDim.x, ... of N tile blockIdx.y.  Two producer groups alternate A stages so two patch
// loads are in flight; accumulators are double-buffered in TMEM so the epilogue of tile i
// overlaps the MMAs of tile i+1; small weight sets stay resident in shared memory.
template <int SPLIT, int EPI_T, int ACT_T>
__global__ void __launch_bounds__(TC_THREADS, 1) conv_tc_kernel(const ZtConv d, const TcGeom g) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int NPART = SPLIT == 3 ? 2 : 1;
  if ((smem_u32(smem) & 1023u) != 0) __trap();     // swizzle atoms need 1024 B alignment
  // Programmatic dependent launch: let the next kernel of the stream start its prologue (barrier init, TMEM
  // allocation, constant weights) while this one runs; every warp that touches activations first executes
  // griddepcontrol.wait (= the preceding grid has completed and flushed).
  if (g.pdl) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + TC_BAR_BYTES - 16);
  const uint32_t a_part = (uint32_t)g.a_part, a_stage = a_part * NPART;
  const uint32_t b_part = (uint32_t)g.BN * 128u, b_stage = b_part * NPART;
  uint8_t* smA = smem + g.a_off;                       // 1024 B aligned (swizzle atoms)
  uint8_t* smB = smA + g.NA * a_stage;
  const uint32_t bar0 = smem_u32(smem);
  const int NA = g.NA, NB = g.NB;
  const int NBb = g.resident ? 1 : NB;      // number of weight barriers in use
  auto fullA = [&](uint32_t s) { return bar0 + 8u * s; };
  auto emptyA = [&](uint32_t s) { return bar0 + 8u * (NA + s); };
  auto fullB = [&](uint32_t s) { return bar0 + 8u * (2 * NA + s); };
  auto emptyB = [&](uint32_t s) { return bar0 + 8u * (2 * NA + NBb + s); };
  auto accFull = [&](uint32_t s) { return bar0 + 8u * (2 * NA + 2 * NBb + s); };
  auto accEmpty = [&](uint32_t s) { return bar0 + 8u * (2 * NA + 2 * NBb + 2 + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.y * g.BN;
  const int cin = d.c0 + d.c1;
  const int tiles_per_img = g.tiles_x * g.tiles_y;
  const int S = g.S, ngroups = g.ngroups, ntaps = g.ntaps;
  const uint32_t nstage = (uint32_t)ngroups * ntaps;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NA; ++s) { mbar_init(fullA(s), 8); mbar_init(emptyA(s), 1); }
    for (int s = 0; s < NBb; ++s) { mbar_init(fullB(s), 1); mbar_init(emptyB(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(accFull(s), 1); mbar_init(accEmpty(s), 256); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == TC_W_MMA) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(tmem_slot)), "r"((uint32_t)g.tmem_cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 8) {
    if (g.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
    // ===================== epilogue: TMEM -> registers -> smem transpose -> global ===========
    // tcgen05.ld hands thread i row i of the tile; a direct store would touch 32 lines per
    // instruction.  Each warp transposes its 32 rows x 32 columns through a padded staging
    // tile so that 8 lanes cover 128 contiguous bytes of one pixel (4 full lines per
    // instruction) for the residual/gate loads and the output stores alike.
    const int quad = warp & 3, half = warp >> 2;
    float* stage = reinterpret_cast<float*>(smem + g.epi_off) + warp * 32 * TC_EPI_LD;
    const int piece = lane & 7, rsub = lane >> 3;
    const int nend = min(d.cout, n0 + g.BN);        // this CTA owns output channels [n0, nend)
    const int nchunk = (nend - n0 + 31) >> 5;
    const int epi = EPI_T >= 0 ? EPI_T : d.epi;
    const int Ch = d.cout >> 1;
    uint32_t lt = 0;
    TcTracer trace(g, warp);
    for (int t = blockIdx.x; t < S; t += gridDim.x, ++lt) {
      const int bimg = t / tiles_per_img, tr = t - bimg * tiles_per_img;
      const int tyi = tr / g.tiles_x, txi = tr - tyi * g.tiles_x;
      const int oyb = tyi * TC_TH + quad * 4, oxb = txi * TC_TW + rsub;
      const int64_t mb = ((int64_t)bimg * d.Ho + oyb) * d.Wo + oxb;
      const uint32_t buf = lt & 1;
      bool waited = false;
      for (int ch = half; ch < nchunk; ch += 2) {
        const int nb = n0 + ch * 32 + piece * 4;          // this thread's 4 output channels
        const int nvalid = nend - nb;
        // residual / gate operands do not depend on the MMAs: fetch them before waiting
        // (the GRU epilogues need h, z and the bias plane: 96 live registers next to the accumulator would
        // spill, so those instances fetch their operands in two batches after the accumulator is staged)
        constexpr bool GRU = (EPI_T == ZT_EPI_GRU_ZR || EPI_T == ZT_EPI_GRU_Q);
        float4 a0[GRU ? 1 : 8], a1[GRU ? 1 : 8];
#pragma unroll
        for (int k = 0; k < (GRU ? 0 : 8); ++k) {
          a0[k] = make_float4(0.f, 0.f, 0.f, 0.f);
          a1[k] = make_float4(0.f, 0.f, 0.f, 0.f);
          const bool okk = (oyb + (k >> 1) < d.Ho) && (oxb + 4 * (k & 1) < d.Wo) && nvalid > 0;
          if (okk && epi != ZT_EPI_PLAIN) {
            const int64_t m = mb + (int64_t)(k >> 1) * d.Wo + 4 * (k & 1);
            if (epi == ZT_EPI_GRU_ZR) {
              if (nb >= Ch) a0[k] = ld4_or_scalar(d.aux0 + m * d.aux0_cs + (nb - Ch), nvalid);
            } else {
              a0[k] = ld4_or_scalar(d.aux0 + m * d.aux0_cs + nb, nvalid);
              if (epi == ZT_EPI_GRU_Q) a1[k] = ld4_or_scalar(d.aux1 + m * d.aux1_cs + nb, nvalid);
            }
          }
        }
        float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (d.bias && nvalid > 0) bv = ld4_or_scalar(d.bias + nb, nvalid);
        if (!waited) {
          if (lane == 0) trace(lt, 0);
          mbar_wait(accFull(buf), (lt >> 1) & 1);
          tc_fence_after();
          waited = true;
          if (lane == 0) trace(lt, 1);
        }
        // SPLIT=3: the accumulator is [Ahi*Bhi + Alo*Bhi | Ahi*Blo] (2*BN columns); add the halves
        const uint32_t accw = (SPLIT == 3 ? 2u : 1u) * (uint32_t)g.BN;
        const uint32_t taddr = tmem_base + buf * accw + ((uint32_t)(quad * 32) << 16) + (uint32_t)(ch * 32);
        uint32_t v[16];
        tmem_ld16(taddr, v);
        if (SPLIT == 3) {
          uint32_t v2[16];
          tmem_ld16(taddr + (uint32_t)g.BN, v2);
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = __float_as_uint((__uint_as_float(v[i]) + __uint_as_float(v2[i])) * g.acc_scale);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
          *reinterpret_cast<uint4*>(stage + lane * TC_EPI_LD + 4 * i) = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        if (ch * 32 + 16 < g.BN) {      // the second half of the chunk exists in TMEM (BN % 32 may be 16)
          tmem_ld16(taddr + 16, v);
          if (SPLIT == 3) {
            uint32_t v2[16];
            tmem_ld16(taddr + 16 + (uint32_t)g.BN, v2);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = __float_as_uint((__uint_as_float(v[i]) + __uint_as_float(v2[i])) * g.acc_scale);
          }
#pragma unroll
          for (int i = 0; i < 4; ++i)
            *reinterpret_cast<uint4*>(stage + lane * TC_EPI_LD + 16 + 4 * i) = make_uint4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
        __syncwarp();
        if constexpr (GRU) {
          // operands: h (aux0), z (aux1, GRU_Q only) and the optional per-frame bias plane (the
          // iteration-invariant share of the convolution); 4 pixels per batch, all loads of a batch in flight
#pragma unroll
          for (int kb = 0; kb < 8; kb += 4) {
            float4 h4[4], z4[4], b4[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int k = kb + j;
              const bool okk = (oyb + (k >> 1) < d.Ho) && (oxb + 4 * (k & 1) < d.Wo) && nvalid > 0;
              const int64_t m = mb + (int64_t)(k >> 1) * d.Wo + 4 * (k & 1);
              h4[j] = z4[j] = b4[j] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (okk) {
                if (EPI_T == ZT_EPI_GRU_ZR) {
                  if (nb >= Ch) h4[j] = ld4_or_scalar(d.aux0 + m * d.aux0_cs + (nb - Ch), nvalid);
                } else {
                  h4[j] = ld4_or_scalar(d.aux0 + m * d.aux0_cs + nb, nvalid);
                  z4[j] = ld4_or_scalar(d.aux1 + m * d.aux1_cs + nb, nvalid);
                }
                if (d.bplane) b4[j] = ld4_or_scalar(d.bplane + m * d.bplane_cs + nb, nvalid);
              }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int k = kb + j;
              const bool okk = (oyb + (k >> 1) < d.Ho) && (oxb + 4 * (k & 1) < d.Wo) && nvalid > 0;
              if (okk) {
                const float4 x = *reinterpret_cast<const float4*>(stage + (4 * k + rsub) * TC_EPI_LD + piece * 4);
                const int64_t m = mb + (int64_t)(k >> 1) * d.Wo + 4 * (k & 1);
                tc_epilogue4<EPI_T, ACT_T>(d, m, nb, nend, x.x + b4[j].x, x.y + b4[j].y, x.z + b4[j].z, x.w + b4[j].w,
                                           bv, h4[j], z4[j]);
              }
            }
          }
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const bool okk = (oyb + (k >> 1) < d.Ho) && (oxb + 4 * (k & 1) < d.Wo) && nvalid > 0;
            if (okk) {
              const float4 x = *reinterpret_cast<const float4*>(stage + (4 * k + rsub) * TC_EPI_LD + piece * 4);
              const int64_t m = mb + (int64_t)(k >> 1) * d.Wo + 4 * (k & 1);
              tc_epilogue4<EPI_T, ACT_T>(d, m, nb, nend, x.x, x.y, x.z, x.w, bv, a0[k], a1[k]);
            }
          }
        }
        __syncwarp();
      }
      if (!waited) { mbar_wait(accFull(buf), (lt >> 1) & 1); tc_fence_after(); }
      tc_fence_before();
      mbar_arrive(accEmpty(buf));
      if (lane == 0) trace(lt, 2);
    }
  } else if (warp == TC_W_MMA) {
    // ===================== MMA issuer (whole warp converged, one elected lane issues) ==========
    {
      const uint32_t BN = g.BN, PW = g.PW, kh = d.kh, kw = d.kw;
      // operand formats of kind::f16: 1 = bf16 (default), 0 = fp16 (split mode 4)
      const uint32_t fmt = g.fp16 ? 0u : ((1u << 7) | (1u << 10));
      const uint32_t idesc = (1u << 4) | fmt | ((BN >> 3) << 17) | (8u << 24);
      // SPLIT=3: the hi and lo weight blocks of a stage are contiguous rows of one swizzled matrix, so
      // Ahi * [Bhi | Blo] is ONE MMA with N = 2*BN (A is fetched once); Alo * Bhi is the second
      const uint32_t idesc2 = (1u << 4) | fmt | ((BN >> 2) << 17) | (8u << 24);
      const uint32_t accw = (SPLIT == 3 ? 2u : 1u) * BN;
      // descriptor templates: only the 14-bit start-address field (16 B units) changes per MMA;
      // a K step of 16 channels is +32 B inside the swizzled 128 B row
      const uint64_t descA0 = make_desc(0, PW * 128u), descB0 = make_desc(0, 1024u);
      const uint32_t apart16 = a_part >> 4;
      const uint32_t smA16 = smem_u32(smA) >> 4, smB16 = smem_u32(smB) >> 4;
      const uint32_t a_stage16 = a_stage >> 4, b_stage16 = b_stage >> 4;
      const bool resident = g.resident != 0;
      uint32_t lt = 0;
      TcTracer trace(g, TC_W_MMA);
      uint32_t sa = 0, pa = 0;          // A ring slot / parity
      uint32_t sbr = 0, pb = 0;         // B ring slot / parity (streamed mode)
      if (resident) { mbar_wait(fullB(0), 0); tc_fence_after(); }
      for (int t = blockIdx.x; t < S; t += gridDim.x, ++lt) {
        const uint32_t buf = lt & 1;
        if (lane == 0) trace(lt, 3);
        mbar_wait(accEmpty(buf), ((lt >> 1) & 1) ^ 1);
        tc_fence_after();
        if (lane == 0) trace(lt, 4);
        const uint32_t tmem_d = tmem_base + buf * accw;
        uint32_t acc = 0;
        int crem = cin;
        uint32_t bres16 = smB16;        // resident mode: walks the weight stages in order
        for (int gi = 0; gi < ngroups; ++gi, crem -= TC_CG) {
          if (lane == 0) trace(lt * ngroups + gi, 0);
          mbar_wait(fullA(sa), pa);
          tc_fence_after();
          if (lane == 0) trace(lt * ngroups + gi, 1);
          const int nks = crem >= TC_CG ? 4 : (crem + 15) >> 4;
          // One elected lane issues the whole stage (all taps) from inside a single elect region, so nothing but the
          // weight-stage waits separates the MMAs of consecutive taps (an elect / __syncwarp / descriptor rebuild per
          // tap left the tensor pipe idle ~400 clk per tap).  The ring state that outlives the region (sbr, pb, bres16,
          // acc) is advanced by ALL lanes after it, from uniform values: state mutated inside the region and carried
          // around the loop would be lane-private, and descriptors built from it are moved to uniform registers by
          // per-lane "waterfall" loops (~120 clk per MMA).
          const uint32_t ahi = (uint32_t)(descA0 >> 32), bhi = (uint32_t)(descB0 >> 32);
          const uint32_t rowlo0 = (uint32_t)descA0 + smA16 + sa * a_stage16;
          const uint32_t blo0 = (uint32_t)descB0;
          if (elect_one()) {
            uint32_t s_ = sbr, p_ = pb, bres_ = bres16, acc_ = acc, rowlo = rowlo0;
            for (uint32_t ky = 0; ky < kh; ++ky, rowlo += PW * 8u) {
              uint32_t alo = rowlo;
              for (uint32_t kx = 0; kx < kw; ++kx, alo += 8u) {
                uint32_t blo;
                if (resident) {
                  blo = blo0 + bres_;
                  bres_ += b_stage16;
                } else {
                  mbar_wait(fullB(s_), p_);
                  tc_fence_after();
                  blo = blo0 + smB16 + s_ * b_stage16;
                }
                if (nks == 4) {
#pragma unroll
                  for (uint32_t ks = 0; ks < 4; ++ks) {
                    if (SPLIT == 3) {
                      tc_mma_bf16_lh(tmem_d, alo + 2 * ks, ahi, blo + 2 * ks, bhi, idesc2, ks == 0 ? acc_ : 1u);
                      tc_mma_bf16_lh(tmem_d, alo + 2 * ks + apart16, ahi, blo + 2 * ks, bhi, idesc, 1);
                    } else {
                      tc_mma_bf16_lh(tmem_d, alo + 2 * ks, ahi, blo + 2 * ks, bhi, idesc, ks == 0 ? acc_ : 1u);
                    }
                  }
                } else {
                  for (int ks = 0; ks < nks; ++ks) {
                    if (SPLIT == 3) {
                      tc_mma_bf16_lh(tmem_d, alo + 2 * ks, ahi, blo + 2 * ks, bhi, idesc2, ks == 0 ? acc_ : 1u);
                      tc_mma_bf16_lh(tmem_d, alo + 2 * ks + apart16, ahi, blo + 2 * ks, bhi, idesc, 1);
                    } else {
                      tc_mma_bf16_lh(tmem_d, alo + 2 * ks, ahi, blo + 2 * ks, bhi, idesc, ks == 0 ? acc_ : 1u);
                    }
                  }
                }
                acc_ = 1;
                if (!resident) {
                  tc_commit(emptyB(s_));
                  if (++s_ == (uint32_t)NB) { s_ = 0; p_ ^= 1; }
                }
              }
            }
            tc_commit(emptyA(sa));
          }
          __syncwarp();
          // advance the carried state on every lane
          acc = 1;
          if (resident) {
            bres16 += (uint32_t)ntaps * b_stage16;
          } else {
            sbr += (uint32_t)ntaps;
            while (sbr >= (uint32_t)NB) { sbr -= (uint32_t)NB; pb ^= 1; }
          }
          if (lane == 0) trace(lt * ngroups + gi, 2);
          if (++sa == (uint32_t)NA) { sa = 0; pa ^= 1; }
        }
        if (elect_one()) tc_commit(accFull(buf));
        __syncwarp();
      }
    }
  } else if (warp == TC_W_LOAD) {
    // ===================== weight loader (one thread, bulk async copies) =====================
    if (lane == 0) {
      if (g.pdl && !g.w_early) asm volatile("griddepcontrol.wait;" ::: "memory");
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(d.weight) + (size_t)blockIdx.y * nstage * (2u * b_part);
      if (g.resident) {
        if ((int)blockIdx.x < S) {
          mbar_arrive_expect_tx(fullB(0), nstage * b_stage);
          for (uint32_t it = 0; it < nstage; ++it)
            bulk_g2s(smem_u32(smB + it * b_stage), wsrc + (size_t)it * (2u * b_part), b_stage, fullB(0));
        }
      } else {
        uint32_t sb = 0, pb = 1;
        for (int t = blockIdx.x; t < S; t += gridDim.x) {
          for (uint32_t it = 0; it < nstage; ++it) {
            mbar_wait(emptyB(sb), pb);
            mbar_arrive_expect_tx(fullB(sb), b_stage);
            bulk_g2s(smem_u32(smB + sb * b_stage), wsrc + (size_t)it * (2u * b_part), b_stage, fullB(sb));
            if (++sb == (uint32_t)NB) { sb = 0; pb ^= 1; }
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== producers: fp32 patch -> bf16 hi/lo canonical smem =====================
    // All 256 producer threads fill every A stage together (16 patch pixels per pass, up to 192 pixels =
    // a whole 3x3 / 1x5 / 5x1 patch in ONE batch of loads).  The global loads of a stage are issued BEFORE
    // waiting for its ring slot (they only need registers), so the HBM/L2 latency overlaps the
    // wait for the MMAs that free the slot.
    const int tid = threadIdx.x - TC_W_PROD * 32;                // 0..255
    const int q = tid & 15, pl = tid >> 4;          // 16 float4 pieces per pixel, 16 pixels per pass
    const int P = g.P, PW = g.PW;
    const int py0 = pl / PW, px0 = pl - py0 * PW;
    const int ntl = S > (int)blockIdx.x ? (S - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    const uint32_t nia = (uint32_t)ntl * ngroups;

    // per-stage source description
    const float* src = nullptr;
    int cs = 0, iy0 = 0, ix0 = 0;
    int64_t img_pix = 0;
    bool cvalid = false;
    auto setup = [&](uint32_t ia) {
      const int lt = ia / ngroups, gi = ia - lt * ngroups;
      const int t = blockIdx.x + lt * gridDim.x;
      const int bimg = t / tiles_per_img, tr = t - bimg * tiles_per_img;
      const int tyi = tr / g.tiles_x, txi = tr - tyi * g.tiles_x;
      iy0 = tyi * TC_TH - d.ph;
      ix0 = txi * TC_TW - d.pw;
      img_pix = (int64_t)bimg * d.H * d.W;
      const int c = gi * TC_CG + q * 4;
      cvalid = c < cin;
      src = d.in0; cs = d.in0_cs;
      if (cvalid) {
        if (c < d.c0) { src = d.in0 + c; }
        else { src = d.in1 + (c - d.c0); cs = d.in1_cs; }
      }
    };
    auto load_batch = [&](float4 (&v)[TC_PB], int p0, int& py, int& px) {
#pragma unroll
      for (int u = 0; u < TC_PB; ++u) {
        const int p = p0 + 16 * u;
        const int iy = iy0 + py, ix = ix0 + px;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (cvalid && p < P && iy >= 0 && iy < d.H && ix >= 0 && ix < d.W)
          v[u] = __ldg(reinterpret_cast<const float4*>(src + (img_pix + (int64_t)iy * d.W + ix) * cs));
        px += 16;
        while (px >= PW) { px -= PW; ++py; }
      }
    };
    auto store_batch = [&](const float4 (&v)[TC_PB], int p0, uint8_t* dst_hi) {
#pragma unroll
      for (int u = 0; u < TC_PB; ++u) {
        const int p = p0 + 16 * u;
        if (p < P) {
          uint2 hi, lo;
          if (SPLIT == 3 && g.fp16) {
            const float4 t = v[u];
            split4h(make_float4(t.x * ZT_F16_ACT_SCALE, t.y * ZT_F16_ACT_SCALE, t.z * ZT_F16_ACT_SCALE, t.w * ZT_F16_ACT_SCALE), hi, lo);
          } else {
            split4(v[u], hi, lo);
          }
          // 128B swizzle: 16 B chunk (q>>1) of row p lives at chunk position (q>>1) ^ (p & 7)
          const uint32_t off = (uint32_t)p * 128u + ((uint32_t)((q >> 1) ^ (p & 7)) << 4) + (uint32_t)(q & 1) * 8u;
          *reinterpret_cast<uint2*>(dst_hi + off) = hi;
          if (SPLIT == 3) *reinterpret_cast<uint2*>(dst_hi + a_part + off) = lo;
        }
      }
    };

    float4 v[TC_PB];
    int py = py0, px = px0;
    uint32_t ia = 0;
    TcTracer trace(g, TC_W_PROD);
    if (g.pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
    if (ia < nia) { setup(ia); load_batch(v, pl, py, px); }
    for (; ia < nia; ++ia) {
      const uint32_t sa = ia % (uint32_t)NA, pa = ((ia / (uint32_t)NA) & 1) ^ 1;
      if (tid == 0) trace(ia, 0);
      mbar_wait(emptyA(sa), pa);
      if (tid == 0) trace(ia, 1);
      uint8_t* dst_hi = smA + sa * a_stage;
      store_batch(v, pl, dst_hi);
      if (tid == 0) trace(ia, 2);
      for (int p0 = pl + 16 * TC_PB; p0 < P; p0 += 16 * TC_PB) {   // rest of a large patch (192 pixels per batch)
        load_batch(v, p0, py, px);
        store_batch(v, p0, dst_hi);
      }
      // one arrival per warp: 256 per-thread arrivals on one mbarrier word serialise (~1 us per stage)
      fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(fullA(sa));
      if (tid == 0) trace(ia, 3);
      if (ia + 1 < nia) {          // prefetch the next stage into registers
        setup(ia + 1);
        py = py0; px = px0;
        load_batch(v, pl, py, px);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == TC_W_MMA) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)g.tmem_cols)
                 : "memory");
  }
}

}  // namespace zt

namespace zt {
// Pack R rows of C fp32 values (row-major, stride rs) as the B operand of zt_conv2d_tc_nhwc for a
// 1x1 "convolution" whose weights change every frame (the all-pairs correlation, corr.py:52-60:
// weights = fmap2): [n_tile][64-ch group][hi|lo][row BN][128 B swizzled], rows beyond R zero.
__global__ void pack_tc_rows_kernel(const float* __restrict__ rows, int R, int C, int rs, int BN, int ngroups,
                                    uint8_t* __restrict__ out, int64_t total) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int chunk_all = (int)(i % (ngroups * 8));          // 8-channel chunk index over all groups
    const int64_t n = i / (ngroups * 8);                      // padded row index
    const int g = chunk_all >> 3, chunk = chunk_all & 7;
    const int nt = (int)(n / BN), r = (int)(n - (int64_t)nt * BN);
    const int c = chunk_all * 8;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    if (n < R && c < C) {
      a = __ldg(reinterpret_cast<const float4*>(rows + n * rs + c));
      b = __ldg(reinterpret_cast<const float4*>(rows + n * rs + c + 4));
    }
    uint2 h0, l0, h1, l1;
    split4(a, h0, l0);
    split4(b, h1, l1);
    const size_t part = (size_t)BN * 128;
    uint8_t* dst = out + ((size_t)(nt * ngroups + g) * 2) * part + (size_t)r * 128 + ((size_t)(chunk ^ (r & 7)) << 4);
    *reinterpret_cast<uint4*>(dst) = make_uint4(h0.x, h0.y, h1.x, h1.y);
    *reinterpret_cast<uint4*>(dst + part) = make_uint4(l0.x, l0.y, l1.x, l1.y);
  }
}
}  // namespace zt

namespace zt {
// nn.Conv2d weights [cout][cin][kh][kw] fp32 (live parameter storage) -> the hi/lo operand layout of conv_tc_kernel
// [n_tile][64-ch group][tap][hi|lo][row BN][128 B swizzled], on the device: the zero-shot training loop changes the
// weights every step, and re-packing them with host tensor ops cost ~20 small launches and a stream synchronise
// per layer.  element(n, c, tap) = w[n * sn + c * sc + tap'] with tap' = tap, or the mirrored tap when `flip`
// (data-gradient convolution: in/out channels swapped through sn / sc, taps mirrored).
__global__ void pack_tc_weights_kernel(const float* __restrict__ w, int N, int Cin, int ntaps, int sn, int sc, int flip,
                                       int BN, int ngroups, int fp16, float scale, uint8_t* __restrict__ out, int64_t total) {
  pdl_enter();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int chunk = (int)(i & 7);
    int64_t r = i >> 3;
    const int row = (int)(r % BN); r /= BN;
    const int tap = (int)(r % ntaps); r /= ntaps;
    const int g = (int)(r % ngroups);
    const int nt = (int)(r / ngroups);
    const int n = nt * BN + row;
    const int st = flip ? ntaps - 1 - tap : tap;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = g * 64 + chunk * 8 + e;
      v[e] = (n < N && c < Cin) ? __ldg(w + (int64_t)n * sn + (int64_t)c * sc + st) * scale : 0.f;
    }
    uint2 h0, l0, h1, l1;
    if (fp16) { split4h(make_float4(v[0], v[1], v[2], v[3]), h0, l0); split4h(make_float4(v[4], v[5], v[6], v[7]), h1, l1); }
    else { split4(make_float4(v[0], v[1], v[2], v[3]), h0, l0); split4(make_float4(v[4], v[5], v[6], v[7]), h1, l1); }
    const size_t part = (size_t)BN * 128;
    uint8_t* dst = out + ((size_t)((nt * ngroups + g) * ntaps + tap) * 2) * part + (size_t)row * 128 + ((size_t)(chunk ^ (row & 7)) << 4);
    *reinterpret_cast<uint4*>(dst) = make_uint4(h0.x, h0.y, h1.x, h1.y);
    *reinterpret_cast<uint4*>(dst + part) = make_uint4(l0.x, l0.y, l1.x, l1.y);
  }
}
}  // namespace zt

extern "C" int zt_pack_tc_weights_f32(const float* w, int cout, int cin, int kh, int kw, int cin_pad, int dgrad, int BN,
                                      int fp16, float scale, void* out, void* stream) {
  using namespace zt;
  ZT_REQUIRE(w && out && cout > 0 && cin > 0 && kh >= 1 && kw >= 1 && kh <= 7 && kw <= 7, "pack_tc_weights: bad args");
  ZT_REQUIRE(BN >= 16 && BN <= 256 && (BN & 15) == 0 && ((uintptr_t)out & 15) == 0, "pack_tc_weights: bad BN %d / alignment", BN);
  ZT_REQUIRE(scale > 0.f, "pack_tc_weights: scale must be positive");
  // forward: rows = cout, K channels = cin; data gradient: rows = cin, K channels = cout, taps mirrored
  const int N = dgrad ? cin : cout, C = dgrad ? cout : cin, taps = kh * kw;
  const int sn = dgrad ? taps : cin * taps, sc = dgrad ? cin * taps : taps;
  const int cp = cin_pad > 0 ? cin_pad : (C + 3) / 4 * 4;
  ZT_REQUIRE(cp >= C, "pack_tc_weights: cin_pad %d < %d", cp, C);
  const int ngroups = ceil_div(cp, TC_CG), nt = ceil_div(N, BN);
  const int64_t total = (int64_t)nt * ngroups * taps * BN * 8;
  const int blocks = (int)std::min<int64_t>(ceil_div64(total, 256), 4 * kNumSMs);
  ZT_CUDA(launch_pdl(pack_tc_weights_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, w, N, C, taps, sn, sc, dgrad ? 1 : 0,
                     BN, ngroups, fp16 ? 1 : 0, scale, (uint8_t*)out, total));
  return check_launch("pack_tc_weights_kernel");
}

extern "C" int zt_pack_tc_rows_f32(const float* rows, int R, int C, int rs, int BN, void* out, void* stream) {
  using namespace zt;
  ZT_REQUIRE(rows && out && R > 0 && C > 0 && (C & 7) == 0 && (rs & 3) == 0 && rs >= C, "pack_tc_rows: bad args");
  ZT_REQUIRE(BN >= 16 && BN <= 256 && (BN & 15) == 0, "pack_tc_rows: bad BN %d", BN);
  ZT_REQUIRE(((uintptr_t)rows & 15) == 0 && ((uintptr_t)out & 15) == 0, "pack_tc_rows: pointers must be 16-byte aligned");
  const int ngroups = ceil_div(C, TC_CG);
  const int64_t total = (int64_t)ceil_div(R, BN) * BN * ngroups * 8;
  const int blocks = (int)std::min<int64_t>(ceil_div64(total, 256), 8 * kNumSMs);
  pack_tc_rows_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(rows, R, C, rs, BN, ngroups, (uint8_t*)out, total);
  return check_launch("pack_tc_rows_kernel");
}

static long long* g_tc_trace = nullptr;
static int g_tc_trace_cap = 0;
// debug: device buffer [1 + 2*cap] int64 (counter, then (tag, ns) pairs) or NULL to disable
extern "C" int zt_debug_set_tc_trace(void* buf, int cap) {
  g_tc_trace = (long long*)buf;
  g_tc_trace_cap = cap;
  return ZT_OK;
}

extern "C" int zt_conv2d_tc_nhwc(const ZtConv* dp, int BN, int split, void* stream) {
  using namespace zt;
  ZT_REQUIRE(dp != nullptr, "zt_conv2d_tc_nhwc: null descriptor");
  const ZtConv& d = *dp;
  const int cin = d.c0 + d.c1;
  ZT_REQUIRE(d.in0 && d.weight && d.out, "conv_tc: null pointer");
  // bit 8 of `split`: the weight operand was produced by the preceding kernel of the stream (all-pairs
  // correlation: weights = fmap2 packed per frame), so it must not be fetched before that kernel completes
  const bool w_dynamic = (split & 0x100) != 0;
  split &= 0xff;
  ZT_REQUIRE(split == 1 || split == 3 || split == 4,
             "conv_tc: split must be 1 (bf16), 3 (bf16 hi/lo, 3 products) or 4 (scaled fp16 hi/lo, 3 products); got %d", split);
  const bool fp16 = split == 4;
  if (fp16) split = 3;
  ZT_REQUIRE(!fp16 || d.acc_scale > 0.f, "conv_tc: split 4 needs acc_scale = 1 / (activation scale * weight scale)");
  ZT_REQUIRE(BN >= 16 && BN <= 256 && (BN & 15) == 0, "conv_tc: BN=%d must be a multiple of 16 in [16,256]", BN);
  ZT_REQUIRE(d.sh == 1 && d.sw == 1, "conv_tc: stride-1 convolutions only (got %dx%d)", d.sh, d.sw);
  ZT_REQUIRE((d.c0 & 3) == 0 && (d.c1 & 3) == 0 && cin > 0, "conv_tc: channel counts must be multiples of 4 (c0=%d c1=%d)", d.c0, d.c1);
  ZT_REQUIRE((d.in0_cs & 3) == 0 && (d.c1 == 0 || ((d.in1_cs & 3) == 0 && d.in1)), "conv_tc: input channel strides must be multiples of 4");
  ZT_REQUIRE(((uintptr_t)d.in0 & 15) == 0 && ((uintptr_t)d.weight & 15) == 0 && (d.c1 == 0 || ((uintptr_t)d.in1 & 15) == 0),
             "conv_tc: pointers must be 16-byte aligned");
  ZT_REQUIRE(d.B > 0 && d.H > 0 && d.W > 0 && d.cout > 0 && d.kh >= 1 && d.kw >= 1 && d.kh <= 7 && d.kw <= 7, "conv_tc: bad dims");
  // Ho/Wo may be cropped below the full correlation size (even kernels from the stride-2 -> stride-1
  // space-to-depth rewrite produce one extra row/column that the reference convolution does not have)
  ZT_REQUIRE(d.Ho <= d.H + 2 * d.ph - d.kh + 1 && d.Wo <= d.W + 2 * d.pw - d.kw + 1 && d.Ho > 0 && d.Wo > 0,
             "conv_tc: output dims inconsistent");
  if (d.epi == ZT_EPI_GRU_ZR) ZT_REQUIRE(d.aux0 && d.aux1 && (d.cout & 7) == 0, "conv_tc: GRU_ZR needs aux0/aux1 and cout%%8==0");
  ZT_REQUIRE(!d.bplane || d.epi == ZT_EPI_GRU_ZR || d.epi == ZT_EPI_GRU_Q, "conv_tc: the bias plane is wired for the GRU epilogues only");
  if (d.epi == ZT_EPI_GRU_Q) ZT_REQUIRE(d.aux0 && d.aux1, "conv_tc: GRU_Q needs aux0/aux1");
  if (d.epi == ZT_EPI_ADD_RELU || d.epi == ZT_EPI_ADD) ZT_REQUIRE(d.aux0, "conv_tc: residual epilogue needs aux0");
  ZT_REQUIRE((int64_t)d.B * d.H * d.W < (1ll << 31), "conv_tc: too many pixels");

  TcGeom g;
  g.BN = BN;
  g.ngroups = ceil_div(cin, TC_CG);
  g.ntaps = d.kh * d.kw;
  g.PW = TC_TW + d.kw - 1;
  g.P = (TC_TH + d.kh - 1) * g.PW;
  g.a_part = ceil_div(g.P * 128, 1024) * 1024;
  g.tiles_x = ceil_div(d.Wo, TC_TW);
  g.tiles_y = ceil_div(d.Ho, TC_TH);
  g.S = g.tiles_x * g.tiles_y * d.B;
  g.trace = g_tc_trace;
  g.trace_cap = g_tc_trace_cap;
  static const bool pdl_on = getenv("ZT_NO_PDL") == nullptr;
  g.pdl = pdl_on ? 1 : 0;
  g.w_early = w_dynamic ? 0 : 1;
  g.fp16 = fp16 ? 1 : 0;
  g.acc_scale = fp16 ? d.acc_scale : 1.f;
  g.tmem_cols = 32;
  ZT_REQUIRE(split == 1 || BN <= 128, "conv_tc: the hi/lo split keeps 2*BN accumulator columns, BN=%d > 128", BN);
  while (g.tmem_cols < (split == 3 ? 4 : 2) * BN) g.tmem_cols <<= 1;
  const int npart = split == 3 ? 2 : 1;
  const size_t a_stage = (size_t)g.a_part * npart;
  const size_t b_stage = (size_t)BN * 128 * npart;
  const size_t nstage = (size_t)g.ngroups * g.ntaps;
  const int n_tiles = ceil_div(d.cout, BN);
  int gx = kNumSMs / n_tiles;
  if (gx < 1) gx = 1;
  if (gx > g.S) gx = g.S;
  // A CTA that owns a single pixel tile (the RAFT update block: 29 tiles) never refills its A ring after the last
  // MMA, so the epilogue staging tile can live on top of it; the 36 KB go to the weight ring, whose depth bounds
  // those kernels (each CTA streams all weights of its N tile once: bytes in flight / L2 round trip).
  const bool alias = g.S <= gx && 2 * a_stage >= (size_t)TC_EPI_BYTES;
  const size_t budget = 226 * 1024 - TC_BAR_BYTES - (alias ? 0 : TC_EPI_BYTES);
  g.a_off = TC_BAR_BYTES + (alias ? 0 : TC_EPI_BYTES);
  g.epi_off = TC_BAR_BYTES;
  // resident weights pay off when a CTA walks more than one pixel tile
  g.resident = (nstage * b_stage + 2 * a_stage <= budget && g.S > gx) ? 1 : 0;
  if (g.resident) {
    g.NB = (int)nstage;
    g.NA = (int)((budget - nstage * b_stage) / a_stage);
  } else {
    g.NA = (int)(budget / 2 / a_stage);
    if (g.NA < 2) g.NA = 2;
    if (alias && g.NA > 2) g.NA = 2;                               // single tile: the weight ring needs the room
    if ((size_t)g.NA * a_stage + 2 * b_stage > budget) g.NA = 1;   // huge patches (7x7 hi+lo): no overlap
  }
  if (g.NA > 8) g.NA = 8;
  if (!g.resident) {
    ZT_REQUIRE((size_t)g.NA * a_stage + b_stage <= budget, "conv_tc: tile does not fit shared memory");
    g.NB = (int)((budget - (size_t)g.NA * a_stage) / b_stage);
    if (g.NB > 32) g.NB = 32;
    if ((size_t)g.NB > nstage && nstage >= 2) g.NB = (int)nstage;
  }
  ZT_REQUIRE(!alias || (size_t)g.NA * a_stage >= (size_t)TC_EPI_BYTES, "conv_tc: staging alias does not fit the A ring");
  const size_t smem = (size_t)g.a_off + (size_t)g.NA * a_stage + (size_t)g.NB * b_stage;
  ZT_REQUIRE(smem <= 227 * 1024, "conv_tc: tile needs %zu B of shared memory", smem);

  dim3 grid(gx, n_tiles, 1);
  cudaStream_t st = (cudaStream_t)stream;
#define ZT_TC_LAUNCH(SP, EP, AC)                                                                          \
  do {                                                                                                    \
    static bool attr_done = false;                                                                        \
    if (!attr_done) {                                                                                     \
      ZT_CUDA(cudaFuncSetAttribute(conv_tc_kernel<SP, EP, AC>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                   227 * 1024));                                                          \
      attr_done = true;                                                                                   \
    }                                                                                                     \
    if (g.pdl) {                                                                                          \
      cudaLaunchConfig_t cfg = {};                                                                        \
      cfg.gridDim = grid; cfg.blockDim = dim3(TC_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;  \
      cudaLaunchAttribute at[1];                                                                          \
      at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                      \
      at[0].val.programmaticStreamSerializationAllowed = 1;                                               \
      cfg.attrs = at; cfg.numAttrs = 1;                                                                   \
      ZT_CUDA(cudaLaunchKernelEx(&cfg, conv_tc_kernel<SP, EP, AC>, d, g));                                \
    } else {                                                                                              \
      conv_tc_kernel<SP, EP, AC><<<grid, TC_THREADS, smem, st>>>(d, g);                                   \
    }                                                                                                     \
  } while (0)
#define ZT_TC_CASE(EP, AC)                                  \
  if (d.epi == (EP) && ((AC) < 0 || d.act == (AC))) {       \
    if (split == 3) ZT_TC_LAUNCH(3, EP, AC);                \
    else ZT_TC_LAUNCH(1, EP, AC);                           \
    return check_launch("conv_tc_kernel");                  \
  }
  ZT_TC_CASE(ZT_EPI_PLAIN, ZT_ACT_RELU)
  ZT_TC_CASE(ZT_EPI_PLAIN, ZT_ACT_NONE)
  ZT_TC_CASE(ZT_EPI_PLAIN, ZT_ACT_LRELU02)
  ZT_TC_CASE(ZT_EPI_PLAIN, ZT_ACT_SIGMOID)
  ZT_TC_CASE(ZT_EPI_PLAIN, ZT_ACT_TANH)
  ZT_TC_CASE(ZT_EPI_GRU_ZR, -1)
  ZT_TC_CASE(ZT_EPI_GRU_Q, -1)
  ZT_TC_CASE(ZT_EPI_ADD_RELU, ZT_ACT_RELU)
  ZT_TC_CASE(ZT_EPI_ADD, ZT_ACT_RELU)
  if (split == 3) ZT_TC_LAUNCH(3, -1, -1);
  else ZT_TC_LAUNCH(1, -1, -1);
  return check_launch("conv_tc_kernel");
}
