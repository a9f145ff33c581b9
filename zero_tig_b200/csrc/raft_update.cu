// RAFT update block (model/RAFT/update.py:127-136, the loop of model/RAFT/raft.py:111-120) as ONE persistent
// cooperative kernel for all iterations of a frame pair.
//
// Round 1 ran an iteration as a chain of ten dependent launches (lookup, convc1, convc2, conv, 2 x (z|r, q), flow
// head) of ~20 us each on 60-120 CTAs: each launch paid a dependent-launch wait, a TMEM allocation, a first-stage
// fill by CUDA-core producer warps that converted the fp32 activations to bf16 hi/lo, and a drain.  Here
//   * the layers are PHASES of one kernel separated by a grid barrier; barriers, TMEM and the tensor maps are set
//     up once per frame; a dedicated thread streams the (constant) weights of the CTA's next work item through a
//     ring while the current phase is still in its epilogue / at the barrier;
//   * every activation that feeds a convolution is STORED as a pair of bf16 tensors (hi, lo with hi + lo = the
//     fp32 value to 2^-17: the same bytes as fp32), so the A operand arrives by TMA (cp.async.bulk.tensor.3d,
//     SWIZZLE_128B, zero fill outside the image = the convolution's padding) - no producer warps, no conversions;
//     the patch (16 x 8 pixels + halo) lands in the layout in which every filter tap is a shifted view;
//   * work items are (pixel tile, output-channel chunk) pairs: 120 CTAs work in every tensor-core phase at 1080p
//     (45 x 80 positions = 30 pixel tiles x 4 chunks);
//   * the CUDA-core pieces are phases too: flow head conv2 + coordinate update + the 4-level 9x9 correlation
//     lookup of the NEXT iteration run together (one warp per position), convf1 (7x7, 2 -> 128) runs on the
//     epilogue warps beside the convc1 MMAs.
// Arithmetic is the round-1 path's: bf16 hi/lo split with three products (Ahi*Bhi + Ahi*Blo + Alo*Bhi), fp32
// accumulation in TMEM in the same K order, the same gate formulas; the hidden state keeps an fp32 master copy.
//
// Phases per iteration:  [convc1 | convf1]  [convc2 | convf2]  conv  zr1  q1  zr2  q2  fh1  [fh2 + update + lookup]
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <string.h>
#include <algorithm>
#include "common.cuh"
#include "tc_ptx.cuh"
#include "tma_host.cuh"

namespace zt {

constexpr int RU_THREADS = 352;            // warps 0-7 epilogue / CUDA-core work, 8 MMA issuer, 9 A loader, 10 weight loader
constexpr int RU_W_MMA = 8, RU_W_LDA = 9, RU_W_LDB = 10;
constexpr int RU_NLAYER = 9, RU_MAXG = 6, RU_NMAPS = 18;
constexpr int RU_BAR_BYTES = 1024;
constexpr uint32_t RU_A_STAGE = 49152;     // 2 parts x 24 KB (largest patch: 1x5 -> 16 x 12 pixels x 128 B)
constexpr uint32_t RU_B_SLOT = 16384;      // BN 64: 2 parts x 64 rows x 128 B
constexpr int RU_NA = 2, RU_NB = 4;
constexpr uint32_t RU_F1W_BYTES = 49 * 2 * 128 * 4;     // convf1 weights [7][7][2][128] fp32, resident in shared memory
constexpr int RU_R = 2;                                 // positions a warp works on at once in the CUDA-core phases
constexpr uint32_t RU_TMEM_COLS = 256;     // two accumulators of 2 * BN <= 128 columns

enum { RU_EPI_RELU_HL = 0, RU_EPI_RELU_F32 = 1, RU_EPI_ZR = 2, RU_EPI_Q = 3 };
enum { L_CONVC1 = 0, L_CONVF2, L_CONVC2, L_CONV, L_ZR1, L_Q1, L_ZR2, L_Q2, L_FH1 };

struct RuLayer {
  short kh, kw, ph, pw, PW, ngroups, nks_last, BN, nchunks, cout, epi;
  short flow_tail;                         // 1: output channels cout, cout + 1 are the current flow (cat[out, flow], update.py:96)
  short map[RU_MAXG];                      // tensor map (hi; lo = +1) of each 64-channel group
  short cofs[RU_MAXG];                     // channel offset of the group inside that tensor
  uint32_t a_part;                         // bytes between the hi and the lo patch (1024-aligned)
  uint32_t box_bytes;                      // bytes one TMA box writes
  const uint8_t* w;                        // [chunk][group][tap][hi|lo][BN][128 B swizzled]
  const float* bias;
  __nv_bfloat16 *out_hi, *out_lo;          // RELU_HL / ZR (r*h) / Q (h): [N][out_cs] at channel out_off
  float* out_f32;                          // RELU_F32
  const float* bplane;                     // ZR / Q: per-frame context share [N][cout]
  int out_cs, out_off;
  float acc_scale;                         // accumulator * acc_scale = convolution sum (1 unless fp16 operands)
};

struct alignas(64) RuParams {
  CUtensorMap maps[RU_NMAPS];
  RuLayer L[RU_NLAYER];
  int h8, w8, N, tiles_x, tiles, iters;
  float *h, *z, *fhid, *coords1, *flow4;
  __nv_bfloat16 *corr_hi, *corr_lo, *f1_hi, *f1_lo, *hm_hi, *hm_lo;
  const void* lvl[4];                      // correlation pyramid planes, fp32 or (lvl_bf16) bf16
  int lh[4], lw[4];
  int lvl_bf16;
  int fp16;                                // operands are scaled fp16 hi/lo pairs (split mode 4 of conv_tc) instead of bf16 pairs
  const float *wf1, *bf1, *wfh2, *bfh2;
  unsigned int* counter;                   // grid barrier (zeroed by the host before the launch)
  long long* trace;                        // debug timeline, see ru_trace (or NULL)
  int trace_cta;
};

static_assert(sizeof(RuParams) <= 4096, "kernel parameter block too large");

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned ld_acquire_cta_s(uint32_t saddr) {
  unsigned v;
  asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_cta_s(uint32_t saddr, unsigned v) {
  asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// TMEM load of 16 columns without the wait, and one wait covering four of them (the registers are operands of the
// wait so that no use can be scheduled above it)
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
#define ZT_R16(a) "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]), \
                  "+r"(a[8]), "+r"(a[9]), "+r"(a[10]), "+r"(a[11]), "+r"(a[12]), "+r"(a[13]), "+r"(a[14]), "+r"(a[15])
__device__ __forceinline__ void tmem_ld_wait2(uint32_t (&a)[16], uint32_t (&b)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;" : ZT_R16(a), ZT_R16(b) :: "memory");
}

// 16-bit operand halves travel as raw bit patterns (bf16, or fp16 of the scaled value when f16)
struct Half16 { unsigned short u; };
__device__ __forceinline__ void split1(float v, Half16& hi, Half16& lo, bool f16) {
  if (f16) {
    const float s = v * ZT_F16_ACT_SCALE;
    const __half h = __float2half_rn(s);
    hi.u = __half_as_ushort(h);
    lo.u = __half_as_ushort(__float2half_rn(s - __half2float(h)));
  } else {
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    hi.u = __bfloat16_as_ushort(h);
    lo.u = __bfloat16_as_ushort(__float2bfloat16_rn(v - __bfloat162float(h)));
  }
}
__device__ __forceinline__ uint32_t pack2(Half16 a, Half16 b) { return (uint32_t)a.u | ((uint32_t)b.u << 16); }

// debug timeline (zt_debug_set_raft_update_trace): CTA `trace_cta` stores %globaltimer into trace[8 * phase + slot];
// phase 0 = init, 1 + 9 * it + ph afterwards (ph 8 = the CUDA-core update / lookup phase).  Slots: 0 barrier passed
// (phase end), 1 A loader released, 2 first A stage landed, 3 last MMA issued, 4 accumulator complete (epilogue
// woke up), 5 epilogue stores done, 6 arrived at the grid barrier, 7 unused.
__device__ __forceinline__ void ru_trace(const long long* tr_base, int cta, int phase, int slot) {
  if (tr_base && (int)blockIdx.x == cta) {
    long long tns;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tns));
    const_cast<long long*>(tr_base)[8 * phase + slot] = tns;
  }
}

// the two layers (or one, second = -1) whose work items make up tensor-core phase ph of an iteration
__device__ __forceinline__ void phase_layers(int ph, int& la, int& lb) {
  lb = -1;
  switch (ph) {
    case 0: la = L_CONVC1; break;
    case 1: la = L_CONVC2; lb = L_CONVF2; break;
    case 2: la = L_CONV; break;
    case 3: la = L_ZR1; break;
    case 4: la = L_Q1; break;
    case 5: la = L_ZR2; break;
    case 6: la = L_Q2; break;
    default: la = L_FH1; break;
  }
}
struct RuItem { int layer, tile, chunk; };
__device__ __forceinline__ int phase_nitems(const RuParams& p, int ph) {
  int la, lb;
  phase_layers(ph, la, lb);
  return p.tiles * (p.L[la].nchunks + (lb >= 0 ? p.L[lb].nchunks : 0));
}
__device__ __forceinline__ RuItem phase_item(const RuParams& p, int ph, int idx) {
  int la, lb;
  phase_layers(ph, la, lb);
  const int na = p.tiles * p.L[la].nchunks;
  RuItem it;
  if (idx < na) { it.layer = la; it.tile = idx / p.L[la].nchunks; it.chunk = idx - it.tile * p.L[la].nchunks; }
  else { idx -= na; it.layer = lb; it.tile = idx / p.L[lb].nchunks; it.chunk = idx - it.tile * p.L[lb].nchunks; }
  return it;
}

// ------------------------------------------------------------------------------------------------------------
// CUDA-core pieces (epilogue warps; one warp per position)
// ------------------------------------------------------------------------------------------------------------
// BasicMotionEncoder.convf1 + ReLU (update.py:84,92): 7x7, 2 -> 128 on the current flow; lane = 4 output channels.
// Same arithmetic, in the same order, as convf1_kernel (raft_small.cu); the result is stored as 16-bit hi / lo for
// convf2's TMA.  A CTA has only 8 warps for this, so latency is hidden by ILP instead of occupancy: the weights
// (49 KB) sit in shared memory for the whole kernel, a warp works on R positions at once, the 49 neighbour flows
// of each are fetched with two loads per lane (all in flight together) and broadcast by shuffles.
template <int R>
__device__ __forceinline__ void ru_convf1(const RuParams& p, const float* __restrict__ wf1s, int pos0, int stride, int pend, int lane) {
  const int W = p.w8, H = p.h8;
  float2 fl[R][2];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int pos = pos0 + r * stride;
    const int y = pos / W, x = pos - y * W;
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int idx = lane + 32 * hh;
      const int ky = idx / 7, kx = idx - ky * 7;
      const int iy = y + ky - 3, ix = x + kx - 3;
      fl[r][hh] = make_float2(0.f, 0.f);      // taps outside the image contribute fma(0, w, acc) = acc
      if (pos < pend && idx < 49 && iy >= 0 && iy < H && ix >= 0 && ix < W)
        fl[r][hh] = *reinterpret_cast<const float2*>(p.flow4 + (int64_t)(iy * W + ix) * 4);
    }
  }
  float4 acc[R];
  const float4 bias = __ldg(reinterpret_cast<const float4*>(p.bf1 + lane * 4));
#pragma unroll
  for (int r = 0; r < R; ++r) acc[r] = bias;
#pragma unroll 7
  for (int t = 0; t < 49; ++t) {
    const float4 w0 = *reinterpret_cast<const float4*>(wf1s + t * 256 + lane * 4);
    const float4 w1 = *reinterpret_cast<const float4*>(wf1s + t * 256 + 128 + lane * 4);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float2 src = t < 32 ? fl[r][0] : fl[r][1];
      const float fx = __shfl_sync(0xffffffffu, src.x, t & 31), fy = __shfl_sync(0xffffffffu, src.y, t & 31);
      acc[r].x = fmaf(fx, w0.x, acc[r].x); acc[r].y = fmaf(fx, w0.y, acc[r].y);
      acc[r].z = fmaf(fx, w0.z, acc[r].z); acc[r].w = fmaf(fx, w0.w, acc[r].w);
      acc[r].x = fmaf(fy, w1.x, acc[r].x); acc[r].y = fmaf(fy, w1.y, acc[r].y);
      acc[r].z = fmaf(fy, w1.z, acc[r].z); acc[r].w = fmaf(fy, w1.w, acc[r].w);
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int pos = pos0 + r * stride;
    if (pos >= pend) continue;
    const float v[4] = {fmaxf(acc[r].x, 0.f), fmaxf(acc[r].y, 0.f), fmaxf(acc[r].z, 0.f), fmaxf(acc[r].w, 0.f)};
    Half16 hi[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) split1(v[e], hi[e], lo[e], p.fp16 != 0);
    *reinterpret_cast<uint2*>(p.f1_hi + (int64_t)pos * 128 + lane * 4) = make_uint2(pack2(hi[0], hi[1]), pack2(hi[2], hi[3]));
    *reinterpret_cast<uint2*>(p.f1_lo + (int64_t)pos * 128 + lane * 4) = make_uint2(pack2(lo[0], lo[1]), pack2(lo[2], lo[3]));
  }
}

// Flow head conv2 (update.py:13-14: 3x3, 256 -> 2) + coords1 += delta + flow = coords1 - coords0 (raft.py:117-118),
// then the 4-level 9x9 lookup at the new coordinates (corr.py:29-50; channel order Q1) for the next iteration.
// mode 0: first call of a frame (coords1 = pixel grid, flow = 0, hidden state split into hm), mode 1: update.
// A warp works on R positions at once (pos0 + r * stride) so that the loads of all of them are in flight together.
template <int R>
__device__ __forceinline__ void ru_update_lookup(const RuParams& p, int pos0, int stride, int pend, int lane, int mode, bool lookup,
                                                 float (*win)[4][10][12], float (*frac)[4][2], int trphase = -1) {
  const int W = p.w8, H = p.h8;
  float cx[R], cy[R];
  int px[R], py[R];
  bool ok[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int pos = pos0 + r * stride;
    ok[r] = pos < pend;
    py[r] = pos / W;
    px[r] = pos - py[r] * W;
  }
  if (mode == 0) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      cx[r] = (float)px[r]; cy[r] = (float)py[r];
      if (!ok[r]) continue;
      const int pos = pos0 + r * stride;
      const float4 hv = *reinterpret_cast<const float4*>(p.h + (int64_t)pos * 128 + lane * 4);
      const float v[4] = {hv.x, hv.y, hv.z, hv.w};
      Half16 hi[4], lo[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) split1(v[e], hi[e], lo[e], p.fp16 != 0);
      *reinterpret_cast<uint2*>(p.hm_hi + (int64_t)pos * 256 + lane * 4) = make_uint2(pack2(hi[0], hi[1]), pack2(hi[2], hi[3]));
      *reinterpret_cast<uint2*>(p.hm_lo + (int64_t)pos * 256 + lane * 4) = make_uint2(pack2(lo[0], lo[1]), pack2(lo[2], lo[3]));
    }
  } else {
    float a0[R], a1[R];
    float2 c_old[R];                   // coords1 of the previous iteration: fetched while the flow head runs
#pragma unroll
    for (int r = 0; r < R; ++r) {
      a0[r] = a1[r] = 0.f;
      c_old[r] = *reinterpret_cast<const float2*>(p.coords1 + 2 * (ok[r] ? pos0 + r * stride : 0));
    }
    // Same accumulation order per position as conv3x3_n2_kernel: taps outer, this lane's 8 channels inner.  The
    // loads of one filter row (3 taps x R positions x 2 float4) are issued together, unconditionally, from clamped
    // addresses; taps outside the image are multiplied by zero weights' worth of nothing: fma(0, w, a) == a.
#pragma unroll
    for (int ty = 0; ty < 3; ++ty) {
      float4 v[R][3][2];
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int tx = 0; tx < 3; ++tx) {
          const int iy = py[r] + ty - 1, ix = px[r] + tx - 1;
          const bool in = ok[r] && iy >= 0 && iy < H && ix >= 0 && ix < W;
          const float* ip = p.fhid + (int64_t)(in ? iy * W + ix : 0) * 256 + lane * 4;
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            const float4 t4 = *reinterpret_cast<const float4*>(ip + 128 * k);
            v[r][tx][k] = in ? t4 : make_float4(0.f, 0.f, 0.f, 0.f);
          }
        }
#pragma unroll
      for (int tx = 0; tx < 3; ++tx) {
        const int t = ty * 3 + tx;
        float4 wa[2], wb[2];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const float* wp = p.wfh2 + (int64_t)t * 512 + (lane * 4 + 128 * k) * 2;
          wa[k] = __ldg(reinterpret_cast<const float4*>(wp));
          wb[k] = __ldg(reinterpret_cast<const float4*>(wp + 4));
        }
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            const float4 x4 = v[r][tx][k];
            a0[r] = fmaf(x4.x, wa[k].x, a0[r]); a1[r] = fmaf(x4.x, wa[k].y, a1[r]);
            a0[r] = fmaf(x4.y, wa[k].z, a0[r]); a1[r] = fmaf(x4.y, wa[k].w, a1[r]);
            a0[r] = fmaf(x4.z, wb[k].x, a0[r]); a1[r] = fmaf(x4.z, wb[k].y, a1[r]);
            a0[r] = fmaf(x4.w, wb[k].z, a0[r]); a1[r] = fmaf(x4.w, wb[k].w, a1[r]);
          }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        a0[r] += __shfl_xor_sync(0xffffffffu, a0[r], o);
        a1[r] += __shfl_xor_sync(0xffffffffu, a1[r], o);
      }
    }
    const float b0 = __ldg(p.bfh2), b1 = __ldg(p.bfh2 + 1);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      cx[r] = c_old[r].x + (a0[r] + b0);
      cy[r] = c_old[r].y + (a1[r] + b1);
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (lane == 0 && ok[r]) {
      const int pos = pos0 + r * stride;
      p.coords1[2 * pos] = cx[r];
      p.coords1[2 * pos + 1] = cy[r];
      const float fx = cx[r] - (float)px[r], fy = cy[r] - (float)py[r];
      // (the motion encoder's conv epilogue copies it into channels 254, 255 of hm: cat[conv out (126), flow (2)])
      *reinterpret_cast<float2*>(p.flow4 + (int64_t)pos * 4) = make_float2(fx, fy);
    }
  }
  if (trphase >= 0 && lane == 0) ru_trace(p.trace, p.trace_cta, trphase, 1);      // flow head + coordinate update done
  if (!lookup) return;
  // Level by level, so that the plane size, base pointer and scale are warp-uniform: the 10x10 window that covers
  // the 9x9 bilinear taps is fetched with 4 loads per lane (elements lane, lane + 32, ...), all levels and all R
  // positions in flight together, then staged in shared memory with zeros outside the plane (grid_sample
  // padding_mode='zeros'); the 81 outputs of a level are then 3 per lane.
  float wv[R][4][4];
  float fxl[R][4], fyl[R][4];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int pos = ok[r] ? pos0 + r * stride : 0;
#pragma unroll
    for (int l = 0; l < 4; ++l) {
      const float sc = 1.f / (float)(1 << l);
      const float xs = cx[r] * sc, ys = cy[r] * sc;        // coords / 2**l  (exact: power-of-two divide)
      const float fx0 = floorf(xs), fy0 = floorf(ys);
      fxl[r][l] = xs - fx0;
      fyl[r][l] = ys - fy0;
      const int bx = (int)fx0 - 4, by = (int)fy0 - 4;
      const int hl = p.lh[l], wl = p.lw[l];
      const int64_t plane = (int64_t)pos * hl * wl;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = lane + 32 * j;
        const int wy = i / 10, wx = i - wy * 10;
        const int ix = bx + wx, iy = by + wy;
        const bool in = ok[r] && i < 100 && ix >= 0 && ix < wl && iy >= 0 && iy < hl;
        const int64_t e = plane + (in ? iy * wl + ix : 0);
        float v;
        if (p.lvl_bf16) v = __bfloat162float(__ushort_as_bfloat16(__ldg(reinterpret_cast<const unsigned short*>(p.lvl[l]) + e)));
        else v = __ldg(reinterpret_cast<const float*>(p.lvl[l]) + e);
        wv[r][l][j] = in ? v : 0.f;
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int l = 0; l < 4; ++l)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int i = lane + 32 * j;
        const int wy = i / 10, wx = i - wy * 10;
        if (i < 100) win[r][l][wy][wx] = wv[r][l][j];
      }
  __syncwarp();
  if (trphase >= 0 && lane == 0) ru_trace(p.trace, p.trace_cta, trphase, 2);      // windows staged (loads landed)
#pragma unroll
  for (int r = 0; r < R; ++r) {
    if (!ok[r]) continue;
    const int pos = pos0 + r * stride;
    unsigned short* oh = reinterpret_cast<unsigned short*>(p.corr_hi + (int64_t)pos * 384);
    unsigned short* ol = reinterpret_cast<unsigned short*>(p.corr_lo + (int64_t)pos * 384);
#pragma unroll
    for (int l = 0; l < 4; ++l) {
      const float fx = fxl[r][l], fy = fyl[r][l];
      const float w00 = (1.f - fx) * (1.f - fy), w01 = fx * (1.f - fy), w10 = (1.f - fx) * fy, w11 = fx * fy;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int i = lane + 32 * j;
        if (i < 81) {
          const int ax = i / 9, by = i - ax * 9;   // channel = l*81 + a*9 + b : x-offset a, y-offset b (Q1)
          const float v00 = win[r][l][by][ax], v01 = win[r][l][by][ax + 1];
          const float v10 = win[r][l][by + 1][ax], v11 = win[r][l][by + 1][ax + 1];
          // grid_sample bilinear: nw*(1-fx)(1-fy) + ne*fx(1-fy) + sw*(1-fx)fy + se*fx*fy
          const float o = v00 * w00 + v01 * w01 + v10 * w10 + v11 * w11;
          Half16 hi, lo;
          split1(o, hi, lo, p.fp16 != 0);
          oh[l * 81 + i] = hi.u;
          ol[l * 81 + i] = lo.u;
        }
      }
    }
  }
  __syncwarp();
  if (trphase >= 0 && lane == 0) ru_trace(p.trace, p.trace_cta, trphase, 3);      // first round of this warp complete
}

// ------------------------------------------------------------------------------------------------------------
// epilogue of one work item: this thread owns TMEM lane (= pixel) m and BN/2 consecutive output channels
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ru_store_hl16(__nv_bfloat16* oh, __nv_bfloat16* ol, const float (&y)[16], int nvalid, bool f16) {
  Half16 hi[16], lo[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) split1(y[e], hi[e], lo[e], f16);
  if (nvalid >= 16) {
    *reinterpret_cast<uint4*>(oh) = make_uint4(pack2(hi[0], hi[1]), pack2(hi[2], hi[3]), pack2(hi[4], hi[5]), pack2(hi[6], hi[7]));
    *reinterpret_cast<uint4*>(oh + 8) = make_uint4(pack2(hi[8], hi[9]), pack2(hi[10], hi[11]), pack2(hi[12], hi[13]), pack2(hi[14], hi[15]));
    *reinterpret_cast<uint4*>(ol) = make_uint4(pack2(lo[0], lo[1]), pack2(lo[2], lo[3]), pack2(lo[4], lo[5]), pack2(lo[6], lo[7]));
    *reinterpret_cast<uint4*>(ol + 8) = make_uint4(pack2(lo[8], lo[9]), pack2(lo[10], lo[11]), pack2(lo[12], lo[13]), pack2(lo[14], lo[15]));
  } else {
#pragma unroll
    for (int e = 0; e < 16; ++e)
      if (e < nvalid) { reinterpret_cast<unsigned short*>(oh)[e] = hi[e].u; reinterpret_cast<unsigned short*>(ol)[e] = lo[e].u; }
  }
}

// Operands of the epilogue that do not depend on the accumulator (bias / per-frame GRU plane in b, hidden state in
// h, z gate in z, 16 channels from n on): fetched BEFORE waiting for the MMAs so that their L2 latency is hidden.
// (h and z are fetched separately, ru_prefetch_hz: only the first 16-channel block of an item gets them early)
__device__ __forceinline__ void ru_prefetch_b(const RuParams& p, const RuLayer& L, int64_t pix, int n, float (&b)[16]) {
#pragma unroll
  for (int e = 0; e < 16; ++e) b[e] = 0.f;
  if (L.epi == RU_EPI_RELU_HL || L.epi == RU_EPI_RELU_F32) {
    const int nvalid = L.cout - n;
#pragma unroll
    for (int e = 0; e < 16; ++e)
      if (L.bias && e < nvalid) b[e] = __ldg(L.bias + n + e);
    return;
  }
  const float4* bp = reinterpret_cast<const float4*>(L.bplane + pix * L.cout + n);
#pragma unroll
  for (int j = 0; j < 4; ++j) { const float4 t = bp[j]; b[4 * j] = t.x; b[4 * j + 1] = t.y; b[4 * j + 2] = t.z; b[4 * j + 3] = t.w; }
}
__device__ __forceinline__ void ru_prefetch_hz(const RuParams& p, const RuLayer& L, int64_t pix, int n, float (&h)[16], float (&z)[16]) {
#pragma unroll
  for (int e = 0; e < 16; ++e) h[e] = z[e] = 0.f;
  if (L.epi == RU_EPI_RELU_HL || L.epi == RU_EPI_RELU_F32) {
    if (L.flow_tail && L.cout - n == 14) {  // motion features: the flow itself rides in the last two channels of the block
      const float2 f = *reinterpret_cast<const float2*>(p.flow4 + pix * 4);
      h[14] = f.x; h[15] = f.y;
    }
    return;
  }
  if (L.epi == RU_EPI_ZR && n < 128) return;
  const int c = L.epi == RU_EPI_ZR ? n - 128 : n;
  const float4* hp = reinterpret_cast<const float4*>(p.h + pix * 128 + c);
#pragma unroll
  for (int j = 0; j < 4; ++j) { const float4 t = hp[j]; h[4 * j] = t.x; h[4 * j + 1] = t.y; h[4 * j + 2] = t.z; h[4 * j + 3] = t.w; }
  if (L.epi == RU_EPI_Q) {
    const float4* zp = reinterpret_cast<const float4*>(p.z + pix * 128 + n);
#pragma unroll
    for (int j = 0; j < 4; ++j) { const float4 t = zp[j]; z[4 * j] = t.x; z[4 * j + 1] = t.y; z[4 * j + 2] = t.z; z[4 * j + 3] = t.w; }
  }
}

__device__ __forceinline__ void ru_epilogue16(const RuParams& p, const RuLayer& L, int64_t pix, int n, const float (&x)[16],
                                              const float (&b)[16], const float (&h)[16], const float (&z)[16]) {
  // n = first of 16 consecutive output channels of layer L held in x (accumulator sums)
  if (L.epi == RU_EPI_RELU_HL || L.epi == RU_EPI_RELU_F32) {
    const int nvalid = L.cout - n;
    if (nvalid <= 0) return;
    float y[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) y[e] = fmaxf(x[e] + b[e], 0.f);
    int nstore = nvalid;
    if (L.flow_tail && nvalid == 14) { y[14] = h[14]; y[15] = h[15]; nstore = 16; }
    if (L.epi == RU_EPI_RELU_F32) {
      float* o = L.out_f32 + pix * L.out_cs + L.out_off + n;
#pragma unroll
      for (int j = 0; j < 4; ++j) *reinterpret_cast<float4*>(o + 4 * j) = make_float4(y[4 * j], y[4 * j + 1], y[4 * j + 2], y[4 * j + 3]);
    } else {
      ru_store_hl16(L.out_hi + pix * L.out_cs + L.out_off + n, L.out_lo + pix * L.out_cs + L.out_off + n, y, nstore, p.fp16 != 0);
    }
    return;
  }
  if (L.epi == RU_EPI_ZR) {
    float s[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) s[e] = __fdividef(1.f, 1.f + __expf(-(x[e] + b[e])));
    if (n < 128) {            // z gate (update.py:50,56)
      float* o = p.z + pix * 128 + n;
#pragma unroll
      for (int j = 0; j < 4; ++j) *reinterpret_cast<float4*>(o + 4 * j) = make_float4(s[4 * j], s[4 * j + 1], s[4 * j + 2], s[4 * j + 3]);
    } else {                  // r gate times the hidden state: the q convolution's input (update.py:51-52)
      const int c = n - 128;
      float y[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) y[e] = s[e] * h[e];
      ru_store_hl16(L.out_hi + pix * L.out_cs + c, L.out_lo + pix * L.out_cs + c, y, 16, p.fp16 != 0);
    }
    return;
  }
  // RU_EPI_Q: h = (1 - z) * h + z * tanh(q)  (update.py:52-53)
  float* hp = p.h + pix * 128 + n;
  float y[16];
#pragma unroll
  for (int e = 0; e < 16; ++e) {
    const float th = 1.f - __fdividef(2.f, 1.f + __expf(2.f * (x[e] + b[e])));
    y[e] = (1.f - z[e]) * h[e] + z[e] * th;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) *reinterpret_cast<float4*>(hp + 4 * j) = make_float4(y[4 * j], y[4 * j + 1], y[4 * j + 2], y[4 * j + 3]);
  ru_store_hl16(L.out_hi + pix * L.out_cs + n, L.out_lo + pix * L.out_cs + n, y, 16, p.fp16 != 0);
}

// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RU_THREADS, 1) raft_update_kernel(const __grid_constant__ RuParams p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const uint32_t bar0 = smem_u32(smem);
  auto fullA = [&](uint32_t s) { return bar0 + 8u * s; };
  auto emptyA = [&](uint32_t s) { return bar0 + 8u * (RU_NA + s); };
  auto fullB = [&](uint32_t s) { return bar0 + 8u * (2 * RU_NA + s); };
  auto emptyB = [&](uint32_t s) { return bar0 + 8u * (2 * RU_NA + RU_NB + s); };
  auto accFull = [&](uint32_t s) { return bar0 + 8u * (2 * RU_NA + 2 * RU_NB + s); };
  auto accEmpty = [&](uint32_t s) { return bar0 + 8u * (2 * RU_NA + 2 * RU_NB + 2 + s); };
  const uint32_t phase_done = bar0 + 512;                        // number of grid barriers this CTA has passed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 520);
  uint8_t* smA = smem + RU_BAR_BYTES;
  uint8_t* smB = smA + RU_NA * RU_A_STAGE;
  float* wf1s = reinterpret_cast<float*>(smB + RU_NB * RU_B_SLOT);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = gridDim.x;
  if (threadIdx.x == 0) {
    for (int s = 0; s < RU_NA; ++s) { mbar_init(fullA(s), 1); mbar_init(emptyA(s), 1); }
    for (int s = 0; s < RU_NB; ++s) { mbar_init(fullB(s), 1); mbar_init(emptyB(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(accFull(s), 1); mbar_init(accEmpty(s), 256); }
    *reinterpret_cast<volatile uint32_t*>(smem + 512) = 0u;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = threadIdx.x; i < (int)(RU_F1W_BYTES / 16); i += RU_THREADS)
    reinterpret_cast<float4*>(wf1s)[i] = __ldg(reinterpret_cast<const float4*>(p.wf1) + i);
  if (warp == RU_W_MMA) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(tmem_slot)), "r"(RU_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int iters = p.iters;

  if (warp < 8) {
    // ============ epilogue / CUDA-core warps ============
    unsigned nbar = 0;
    auto grid_barrier = [&]() {
      ++nbar;
      epi_bar();                               // every epilogue thread's stores are ordered before thread 0's fence
      if (threadIdx.x == 0) {
        ru_trace(p.trace, p.trace_cta, (int)nbar - 1, 5);
        __threadfence();                       // (cumulative) release of the CTA's writes at GPU scope
        fence_proxy_async_all();
        atomicAdd(p.counter, 1u);
        ru_trace(p.trace, p.trace_cta, (int)nbar - 1, 6);
        const unsigned target = nbar * (unsigned)G;
        while (ld_acquire_gpu(p.counter) < target) { }
        __threadfence();
        st_release_cta_s(phase_done, nbar);
        ru_trace(p.trace, p.trace_cta, (int)nbar - 1, 0);
      }
      epi_bar();
    };
    if (threadIdx.x == 0) ru_trace(p.trace, p.trace_cta, 0, 7);          // kernel start
    // lookup windows (RU_R positions per warp) alias A stage 0: the activation loader is parked at the grid barrier
    float (*win)[4][10][12] = reinterpret_cast<float (*)[4][10][12]>(smA + warp * 4096);
    float (*frac)[4][2] = reinterpret_cast<float (*)[4][2]>(smA + warp * 4096 + RU_R * 1920);
    // CUDA-core phases: a CTA owns a CONTIGUOUS range of positions (their correlation planes, hidden-state rows and
    // flow neighbourhoods then share pages and cache lines), one warp per position, RU_R positions at a time
    const int ppc = (p.N + G - 1) / G;
    const int pbeg = blockIdx.x * ppc, pend = min(p.N, pbeg + ppc);
    const int quad = warp & 3, half = warp >> 2;
    const int m = quad * 32 + lane;
    uint32_t lt = 0;

    for (int pos = pbeg + warp; pos < pend; pos += RU_R * 8)
      ru_update_lookup<RU_R>(p, pos, 8, pend, lane, 0, iters > 0, win, frac);
    grid_barrier();
    for (int it = 0; it < iters; ++it) {
      for (int ph = 0; ph < 8; ++ph) {
        if (ph == 0)
          for (int pos = pbeg + warp; pos < pend; pos += 4 * 8) ru_convf1<4>(p, wf1s, pos, 8, pend, lane);
        const int nitems = phase_nitems(p, ph);
        for (int idx = blockIdx.x; idx < nitems; idx += G, ++lt) {
          const RuItem item = phase_item(p, ph, idx);
          const RuLayer& L = p.L[item.layer];
          const int tyi = item.tile / p.tiles_x, txi = item.tile - tyi * p.tiles_x;
          const int oy = tyi * 16 + (m >> 3), ox = txi * 8 + (m & 7);
          const bool ok = oy < p.h8 && ox < p.w8;
          const int64_t pix = (int64_t)oy * p.w8 + ox;
          const uint32_t buf = lt & 1;
          const int BN = L.BN, cols = BN >> 1;
          const int n0 = item.chunk * BN + half * cols;
          float pb[2][16], phh[16], pz[16];
          if (ok) {
            ru_prefetch_b(p, L, pix, n0, pb[0]);
            ru_prefetch_hz(p, L, pix, n0, phh, pz);
            if (cols > 16) ru_prefetch_b(p, L, pix, n0 + 16, pb[1]);
          }
          mbar_wait(accFull(buf), (lt >> 1) & 1);
          tc_fence_after();
          if (threadIdx.x == 0) ru_trace(p.trace, p.trace_cta, 1 + 9 * it + ph, 4);
          float x[32];
          const uint32_t taddr = tmem_base + buf * 128u + ((uint32_t)(quad * 32) << 16) + (uint32_t)(half * cols);
          {
            // columns [BN, 2BN) hold the Ahi*Blo products (see the MMA warp); two loads in flight per wait
            uint32_t v[16], v2[16];
            tmem_ld16_nowait(taddr, v);
            tmem_ld16_nowait(taddr + (uint32_t)BN, v2);
            tmem_ld_wait2(v, v2);
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = (__uint_as_float(v[i]) + __uint_as_float(v2[i])) * L.acc_scale;
            if (cols > 16) {
              tmem_ld16_nowait(taddr + 16, v);
              tmem_ld16_nowait(taddr + 16 + (uint32_t)BN, v2);
              tmem_ld_wait2(v, v2);
#pragma unroll
              for (int i = 0; i < 16; ++i) x[16 + i] = (__uint_as_float(v[i]) + __uint_as_float(v2[i])) * L.acc_scale;
            }
          }
          tc_fence_before();
          mbar_arrive(accEmpty(buf));                      // the accumulator is in registers
          if (ok) {
            float x0[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) x0[i] = x[i];
            ru_epilogue16(p, L, pix, n0, x0, pb[0], phh, pz);
            if (cols > 16) {
#pragma unroll
              for (int i = 0; i < 16; ++i) x0[i] = x[16 + i];
              ru_prefetch_hz(p, L, pix, n0 + 16, phh, pz);       // second block: fetched late (registers)
              ru_epilogue16(p, L, pix, n0 + 16, x0, pb[1], phh, pz);
            }
          }
        }
        grid_barrier();
      }
      const bool last = it + 1 == iters;
      for (int pos = pbeg + warp, rnd = 0; pos < pend; pos += RU_R * 8, ++rnd)
        ru_update_lookup<RU_R>(p, pos, 8, pend, lane, 1, !last, win, frac, (warp == 0 && rnd == 0) ? 9 + 9 * it : -1);
      if (threadIdx.x == 0) ru_trace(p.trace, p.trace_cta, 9 + 9 * it, 4);              // warp 0 finished all its rounds
      if (!last) grid_barrier();
    }
  } else if (warp == RU_W_MMA) {
    // ============ MMA issuer (whole warp converged, one elected lane issues) ============
    uint32_t lt = 0, sa = 0, pa = 0, sb = 0, pb = 0;
    const uint32_t smA16 = smem_u32(smA) >> 4, smB16 = smem_u32(smB) >> 4;
    for (int it = 0; it < iters; ++it) {
      for (int ph = 0; ph < 8; ++ph) {
        const int nitems = phase_nitems(p, ph);
        for (int idx = blockIdx.x; idx < nitems; idx += G, ++lt) {
          const RuItem item = phase_item(p, ph, idx);
          const RuLayer& L = p.L[item.layer];
          const uint32_t BN = L.BN, PW = L.PW, kh = L.kh, kw = L.kw;
          const uint32_t fmt = p.fp16 ? 0u : ((1u << 7) | (1u << 10));      // kind::f16 operand formats: fp16 / bf16
          const uint32_t idesc = (1u << 4) | fmt | ((BN >> 3) << 17) | (8u << 24);
          const uint32_t idesc2 = (1u << 4) | fmt | ((BN >> 2) << 17) | (8u << 24);
          const uint64_t descA0 = make_desc(0, PW * 128u), descB0 = make_desc(0, 1024u);
          const uint32_t ahi = (uint32_t)(descA0 >> 32), bhi = (uint32_t)(descB0 >> 32);
          const uint32_t apart16 = L.a_part >> 4;
          const uint32_t buf = lt & 1;
          mbar_wait(accEmpty(buf), ((lt >> 1) & 1) ^ 1);
          tc_fence_after();
          const uint32_t tmem_d = tmem_base + buf * 128u;
          uint32_t acc = 0;
          const int ngroups = L.ngroups, ntaps = kh * kw;
          for (int gi = 0; gi < ngroups; ++gi) {
            mbar_wait(fullA(sa), pa);
            tc_fence_after();
            if (gi == 0 && lane == 0) ru_trace(p.trace, p.trace_cta, 1 + 9 * it + ph, 2);
            const int nks = (gi + 1 == ngroups) ? L.nks_last : 4;
            const uint32_t rowlo0 = (uint32_t)descA0 + smA16 + sa * (RU_A_STAGE >> 4);
            const uint32_t blo0 = (uint32_t)descB0 + smB16;
            if (elect_one()) {
              uint32_t s_ = sb, p_ = pb, acc_ = acc, rowlo = rowlo0;
              for (uint32_t ky = 0; ky < kh; ++ky, rowlo += PW * 8u) {
                uint32_t alo = rowlo;
                for (uint32_t kx = 0; kx < kw; ++kx, alo += 8u) {
                  mbar_wait(fullB(s_), p_);
                  tc_fence_after();
                  const uint32_t blo = blo0 + s_ * (RU_B_SLOT >> 4);
                  for (int ks = 0; ks < nks; ++ks) {
                    // Ahi * [Bhi | Blo] as ONE MMA with N = 2*BN (hi and lo weight rows are contiguous), then Alo * Bhi
                    tc_mma_bf16_lh(tmem_d, alo + 2 * ks, ahi, blo + 2 * ks, bhi, idesc2, ks == 0 ? acc_ : 1u);
                    tc_mma_bf16_lh(tmem_d, alo + 2 * ks + apart16, ahi, blo + 2 * ks, bhi, idesc, 1);
                  }
                  acc_ = 1;
                  tc_commit(emptyB(s_));
                  if (++s_ == (uint32_t)RU_NB) { s_ = 0; p_ ^= 1; }
                }
              }
              tc_commit(emptyA(sa));
            }
            __syncwarp();
            acc = 1;
            sb += (uint32_t)ntaps;
            while (sb >= (uint32_t)RU_NB) { sb -= (uint32_t)RU_NB; pb ^= 1; }
            if (++sa == (uint32_t)RU_NA) { sa = 0; pa ^= 1; }
          }
          if (elect_one()) tc_commit(accFull(buf));
          __syncwarp();
          if (lane == 0) ru_trace(p.trace, p.trace_cta, 1 + 9 * it + ph, 3);
        }
      }
    }
  } else if (warp == RU_W_LDA) {
    // ============ activation loader: one TMA box per (64-channel group, hi / lo) ============
    if (lane == 0) {
      for (int i = 0; i < RU_NMAPS; ++i) tma_prefetch_desc(&p.maps[i]);
      uint32_t sa = 0, pa = 1;
      for (int it = 0; it < iters; ++it) {
        for (int ph = 0; ph < 8; ++ph) {
          const int nitems = phase_nitems(p, ph);
          if ((int)blockIdx.x >= nitems) continue;
          // activations of this phase were written by other CTAs before grid barrier number 1 + 9*it + ph
          const unsigned need = 1u + 9u * (unsigned)it + (unsigned)ph;
          while (ld_acquire_cta_s(phase_done) < need) { }
          fence_proxy_async_all();
          ru_trace(p.trace, p.trace_cta, 1 + 9 * it + ph, 1);
          for (int idx = blockIdx.x; idx < nitems; idx += G) {
            const RuItem item = phase_item(p, ph, idx);
            const RuLayer& L = p.L[item.layer];
            const int tyi = item.tile / p.tiles_x, txi = item.tile - tyi * p.tiles_x;
            const int x0 = txi * 8 - L.pw, y0 = tyi * 16 - L.ph;
            for (int gi = 0; gi < L.ngroups; ++gi) {
              mbar_wait(emptyA(sa), pa);
              mbar_arrive_expect_tx(fullA(sa), 2u * L.box_bytes);
              const uint32_t dst = smem_u32(smA + sa * RU_A_STAGE);
              tma_load_3d(dst, &p.maps[L.map[gi]], L.cofs[gi], x0, y0, fullA(sa));
              tma_load_3d(dst + L.a_part, &p.maps[L.map[gi] + 1], L.cofs[gi], x0, y0, fullA(sa));
              if (++sa == (uint32_t)RU_NA) { sa = 0; pa ^= 1; }
            }
          }
        }
      }
    }
    __syncwarp();
  } else {
    // ============ weight loader: constants, runs ahead of the phases through the ring ============
    if (lane == 0) {
      uint32_t sb = 0, pb = 1;
      for (int it = 0; it < iters; ++it) {
        for (int ph = 0; ph < 8; ++ph) {
          const int nitems = phase_nitems(p, ph);
          for (int idx = blockIdx.x; idx < nitems; idx += G) {
            const RuItem item = phase_item(p, ph, idx);
            const RuLayer& L = p.L[item.layer];
            const uint32_t ntaps = (uint32_t)L.kh * L.kw;
            const uint32_t b_stage = 2u * (uint32_t)L.BN * 128u;
            const uint8_t* wsrc = L.w + (size_t)item.chunk * L.ngroups * ntaps * b_stage;
            const uint32_t nst = (uint32_t)L.ngroups * ntaps;
            for (uint32_t s = 0; s < nst; ++s) {
              mbar_wait(emptyB(sb), pb);
              mbar_arrive_expect_tx(fullB(sb), b_stage);
              bulk_g2s(smem_u32(smB + sb * RU_B_SLOT), wsrc + (size_t)s * b_stage, b_stage, fullB(sb));
              if (++sb == (uint32_t)RU_NB) { sb = 0; pb ^= 1; }
            }
          }
        }
      }
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == RU_W_MMA) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(RU_TMEM_COLS) : "memory");
  }
}

// fp32 [N][C] (row stride rs) -> bf16 hi / lo [N][C] (row stride os): the hidden state entering the update loop
__global__ void __launch_bounds__(256) split_hilo_kernel(const float* __restrict__ in, int rs, __nv_bfloat16* __restrict__ hi,
                                                         __nv_bfloat16* __restrict__ lo, int os, int C4, int64_t total) {
  pdl_enter();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / C4;
    const int c = (int)(i - n * C4) * 4;
    const float4 v = __ldg(reinterpret_cast<const float4*>(in + n * rs + c));
    const float f[4] = {v.x, v.y, v.z, v.w};
    Half16 h[4], l[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) split1(f[e], h[e], l[e], false);
    *reinterpret_cast<uint2*>(hi + n * os + c) = make_uint2(pack2(h[0], h[1]), pack2(h[2], h[3]));
    *reinterpret_cast<uint2*>(lo + n * os + c) = make_uint2(pack2(l[0], l[1]), pack2(l[2], l[3]));
  }
}

}  // namespace zt

extern "C" int zt_split_bf16_hilo(const float* in, int in_rs, void* hi, void* lo, int out_rs, int C, int64_t N, void* stream) {
  using namespace zt;
  ZT_REQUIRE(in && hi && lo && C > 0 && (C & 3) == 0 && (in_rs & 3) == 0 && (out_rs & 3) == 0 && N > 0, "split_bf16_hilo: bad args");
  const int64_t total = N * (C / 4);
  const int blocks = (int)std::min<int64_t>(ceil_div64(total, 256), 8 * kNumSMs);
  ZT_CUDA(launch_pdl(split_hilo_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, in, in_rs, (__nv_bfloat16*)hi,
                     (__nv_bfloat16*)lo, out_rs, C / 4, total));
  return check_launch("split_hilo_kernel");
}

static long long* g_ru_trace = nullptr;
static int g_ru_trace_cta = 0;
/* debug: device buffer of 8 * (2 + 9 * iters) int64 (zeroed by the caller); CTA `cta` of the following zt_raft_update
 * launches stores %globaltimer (ns) per phase and role, see ru_trace (NULL disables) */
extern "C" int zt_debug_set_raft_update_trace(void* buf, int cta) {
  g_ru_trace = (long long*)buf;
  g_ru_trace_cta = cta;
  return ZT_OK;
}

extern "C" int zt_raft_update(const ZtRaftUpdate* dp, void* stream) {
  using namespace zt;
  ZT_REQUIRE(dp != nullptr, "raft_update: null descriptor");
  const ZtRaftUpdate& d = *dp;
  ZT_REQUIRE(d.h8 >= 2 && d.w8 >= 2 && d.iters >= 0 && d.iters <= 64, "raft_update: bad grid %dx%d / iters %d", d.h8, d.w8, d.iters);
  ZT_REQUIRE(d.corr_hi && d.corr_lo && d.c1_hi && d.c1_lo && d.f1_hi && d.f1_lo && d.c2f_hi && d.c2f_lo && d.hm_hi &&
             d.hm_lo && d.rh_hi && d.rh_lo && d.h && d.z && d.fhid && d.coords1 && d.flow4 && d.counter,
             "raft_update: null buffer");
  ZT_REQUIRE(d.bp_zr1 && d.bp_q1 && d.bp_zr2 && d.bp_q2 && d.wf1 && d.bf1 && d.wfh2 && d.bfh2, "raft_update: null constant");
  for (int i = 0; i < 4; ++i) ZT_REQUIRE(d.lvl[i] && d.lh[i] >= 1 && d.lw[i] >= 1, "raft_update: bad pyramid level %d", i);
  for (int i = 0; i < RU_NLAYER; ++i)
    ZT_REQUIRE(d.w[i] && (d.bn[i] == 32 || d.bn[i] == 64), "raft_update: layer %d needs packed weights and BN 32 or 64", i);
  const int N = d.h8 * d.w8;
  RuParams p;
  memset(&p, 0, sizeof(p));
  // ---- tensor maps: [C][w8][h8] bf16, box (64, 8 + kw - 1, 16 + kh - 1) ----
  struct MapDef { const void *hi, *lo; int C, kh, kw; };
  const MapDef defs[9] = {{d.corr_hi, d.corr_lo, 384, 1, 1}, {d.f1_hi, d.f1_lo, 128, 3, 3}, {d.c1_hi, d.c1_lo, 256, 3, 3},
                          {d.c2f_hi, d.c2f_lo, 256, 3, 3},  {d.hm_hi, d.hm_lo, 256, 1, 5},  {d.rh_hi, d.rh_lo, 128, 1, 5},
                          {d.hm_hi, d.hm_lo, 256, 5, 1},    {d.rh_hi, d.rh_lo, 128, 5, 1},  {d.hm_hi, d.hm_lo, 256, 3, 3}};
  for (int i = 0; i < 9; ++i) {
    const int dims[3] = {defs[i].C, d.w8, d.h8};
    const int64_t strides[2] = {defs[i].C, (int64_t)d.w8 * defs[i].C};
    const int box[3] = {64, 8 + defs[i].kw - 1, 16 + defs[i].kh - 1};
    ZT_REQUIRE(((uintptr_t)defs[i].hi & 15) == 0 && ((uintptr_t)defs[i].lo & 15) == 0, "raft_update: activation buffers must be 16-byte aligned");
    if (int e = tma_map_bf16(defs[i].hi, 3, dims, strides, box, &p.maps[2 * i])) return e;
    if (int e = tma_map_bf16(defs[i].lo, 3, dims, strides, box, &p.maps[2 * i + 1])) return e;
  }
  // ---- layers ----
  auto layer = [&](int li, int kh, int kw, int cin, int cout, int epi) -> RuLayer& {
    RuLayer& L = p.L[li];
    L.kh = (short)kh; L.kw = (short)kw; L.ph = (short)(kh / 2); L.pw = (short)(kw / 2);
    L.PW = (short)(8 + kw - 1);
    L.ngroups = (short)ceil_div(cin, 64);
    const int rem = cin - (L.ngroups - 1) * 64;
    L.nks_last = (short)ceil_div(rem, 16);
    L.BN = (short)d.bn[li];
    L.nchunks = (short)ceil_div(cout, d.bn[li]);
    L.cout = (short)cout;
    L.epi = (short)epi;
    const int P = L.PW * (16 + kh - 1);
    L.a_part = (uint32_t)ceil_div(P * 128, 1024) * 1024u;
    L.box_bytes = (uint32_t)P * 128u;
    L.w = (const uint8_t*)d.w[li];
    L.bias = d.b[li];
    L.acc_scale = d.fp16 ? d.acc_scale[li] : 1.f;
    return L;
  };
  auto src = [&](RuLayer& L, int g0, int ng, int mapdef, int c0) {
    for (int g = 0; g < ng; ++g) { L.map[g0 + g] = (short)(2 * mapdef); L.cofs[g0 + g] = (short)(c0 + 64 * g); }
  };
  {
    RuLayer& L = layer(L_CONVC1, 1, 1, 324, 256, RU_EPI_RELU_HL);
    src(L, 0, 6, 0, 0);
    L.out_hi = (__nv_bfloat16*)d.c1_hi; L.out_lo = (__nv_bfloat16*)d.c1_lo; L.out_cs = 256; L.out_off = 0;
  }
  {
    RuLayer& L = layer(L_CONVF2, 3, 3, 128, 64, RU_EPI_RELU_HL);
    src(L, 0, 2, 1, 0);
    L.out_hi = (__nv_bfloat16*)d.c2f_hi; L.out_lo = (__nv_bfloat16*)d.c2f_lo; L.out_cs = 256; L.out_off = 192;
  }
  {
    RuLayer& L = layer(L_CONVC2, 3, 3, 256, 192, RU_EPI_RELU_HL);
    src(L, 0, 4, 2, 0);
    L.out_hi = (__nv_bfloat16*)d.c2f_hi; L.out_lo = (__nv_bfloat16*)d.c2f_lo; L.out_cs = 256; L.out_off = 0;
  }
  {
    RuLayer& L = layer(L_CONV, 3, 3, 256, 126, RU_EPI_RELU_HL);
    src(L, 0, 4, 3, 0);
    L.flow_tail = 1;
    ZT_REQUIRE(d.bn[L_CONV] == 32 || d.bn[L_CONV] == 64, "raft_update: conv needs BN 32 or 64");
    L.out_hi = (__nv_bfloat16*)d.hm_hi; L.out_lo = (__nv_bfloat16*)d.hm_lo; L.out_cs = 256; L.out_off = 128;
  }
  for (int s = 0; s < 2; ++s) {
    const int kh = s == 0 ? 1 : 5, kw = s == 0 ? 5 : 1;
    RuLayer& Z = layer(s == 0 ? L_ZR1 : L_ZR2, kh, kw, 256, 256, RU_EPI_ZR);
    src(Z, 0, 4, s == 0 ? 4 : 6, 0);
    Z.out_hi = (__nv_bfloat16*)d.rh_hi; Z.out_lo = (__nv_bfloat16*)d.rh_lo; Z.out_cs = 128; Z.out_off = 0;
    Z.bplane = s == 0 ? d.bp_zr1 : d.bp_zr2;
    RuLayer& Q = layer(s == 0 ? L_Q1 : L_Q2, kh, kw, 256, 128, RU_EPI_Q);
    src(Q, 0, 2, s == 0 ? 5 : 7, 0);          // r * h
    src(Q, 2, 2, s == 0 ? 4 : 6, 128);        // motion features (channels 128..255 of hm)
    Q.out_hi = (__nv_bfloat16*)d.hm_hi; Q.out_lo = (__nv_bfloat16*)d.hm_lo; Q.out_cs = 256; Q.out_off = 0;
    Q.bplane = s == 0 ? d.bp_q1 : d.bp_q2;
  }
  {
    RuLayer& L = layer(L_FH1, 3, 3, 128, 256, RU_EPI_RELU_F32);
    src(L, 0, 2, 8, 0);
    L.out_f32 = d.fhid; L.out_cs = 256; L.out_off = 0;
  }
  for (int i = 0; i < RU_NLAYER; ++i) {
    ZT_REQUIRE(2u * p.L[i].a_part <= RU_A_STAGE && 2u * p.L[i].BN * 128u <= RU_B_SLOT, "raft_update: layer %d does not fit a stage", i);
    ZT_REQUIRE(((uintptr_t)p.L[i].w & 15) == 0, "raft_update: weights must be 16-byte aligned");
  }
  p.h8 = d.h8; p.w8 = d.w8; p.N = N; p.iters = d.iters;
  p.tiles_x = ceil_div(d.w8, 8);
  p.tiles = p.tiles_x * ceil_div(d.h8, 16);
  p.h = d.h; p.z = d.z; p.fhid = d.fhid; p.coords1 = d.coords1; p.flow4 = d.flow4;
  p.corr_hi = (__nv_bfloat16*)d.corr_hi; p.corr_lo = (__nv_bfloat16*)d.corr_lo;
  p.f1_hi = (__nv_bfloat16*)d.f1_hi; p.f1_lo = (__nv_bfloat16*)d.f1_lo;
  p.hm_hi = (__nv_bfloat16*)d.hm_hi; p.hm_lo = (__nv_bfloat16*)d.hm_lo;
  for (int i = 0; i < 4; ++i) { p.lvl[i] = d.lvl[i]; p.lh[i] = d.lh[i]; p.lw[i] = d.lw[i]; }
  p.lvl_bf16 = d.lvl_bf16;
  p.fp16 = d.fp16 ? 1 : 0;
  if (d.fp16)
    for (int i = 0; i < RU_NLAYER; ++i) ZT_REQUIRE(d.acc_scale[i] > 0.f, "raft_update: fp16 operands need acc_scale[%d] > 0", i);
  p.wf1 = d.wf1; p.bf1 = d.bf1; p.wfh2 = d.wfh2; p.bfh2 = d.bfh2;
  p.counter = d.counter;
  p.trace = g_ru_trace;
  p.trace_cta = g_ru_trace_cta;

  int max_items = 1;
  for (int ph = 0; ph < 8; ++ph) {
    int la, lb;
    switch (ph) {
      case 0: la = L_CONVC1; lb = -1; break;
      case 1: la = L_CONVC2; lb = L_CONVF2; break;
      case 2: la = L_CONV; lb = -1; break;
      case 3: la = L_ZR1; lb = -1; break;
      case 4: la = L_Q1; lb = -1; break;
      case 5: la = L_ZR2; lb = -1; break;
      case 6: la = L_Q2; lb = -1; break;
      default: la = L_FH1; lb = -1; break;
    }
    max_items = std::max(max_items, p.tiles * (p.L[la].nchunks + (lb >= 0 ? p.L[lb].nchunks : 0)));
  }
  const size_t smem = RU_BAR_BYTES + (size_t)RU_NA * RU_A_STAGE + (size_t)RU_NB * RU_B_SLOT + RU_F1W_BYTES;
  static int max_coresident = -1;
  if (max_coresident < 0) {
    ZT_CUDA(cudaFuncSetAttribute(raft_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0, dev = 0, sms = 0;
    ZT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, raft_update_kernel, RU_THREADS, smem));
    ZT_CUDA(cudaGetDevice(&dev));
    ZT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    ZT_REQUIRE(per_sm >= 1, "raft_update: the kernel does not fit an SM (%zu B of shared memory)", smem);
    max_coresident = per_sm * sms;
  }
  // every resident CTA helps in the CUDA-core phases (one warp per position), also those without a tensor-core item
  int G = max_coresident;
  if (d.max_ctas > 0) G = std::min(G, d.max_ctas);
  G = std::min(G, std::max(max_items, ceil_div(N, 8)));
  cudaStream_t st = (cudaStream_t)stream;
  ZT_CUDA(cudaMemsetAsync(d.counter, 0, sizeof(unsigned int), st));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(G); cfg.blockDim = dim3(RU_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;           // the grid barrier needs every CTA resident
  at[0].val.cooperative = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  ZT_CUDA(cudaLaunchKernelEx(&cfg, raft_update_kernel, p));
  return check_launch("raft_update_kernel");
}
