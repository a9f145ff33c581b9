// update_cache pre-processing (resize, *255, uint8 truncation, histogram equalisation, RAFT
// input packing) and the Retinex elementwise / layout kernels of Finetunemodel.forward.
// Reference: model/model.py:312-340,346-360; torchvision _functional_tensor._scale_channel;
// model/RAFT/raft.py:80-83,132-138; predict.py:57-61.
#include <algorithm>
#include "common.cuh"

namespace zt {

struct AxisTap2 { int i0, i1; float l0, l1; };
__device__ __forceinline__ AxisTap2 src_index2(float scale, int dst, int n_in) {
  float s = scale * ((float)dst + 0.5f) - 0.5f;
  s = s < 0.f ? 0.f : s;
  AxisTap2 t;
  t.i0 = min((int)s, n_in - 1);
  t.i1 = t.i0 + (t.i0 < n_in - 1 ? 1 : 0);
  t.l1 = fminf(fmaxf(s - (float)t.i0, 0.f), 1.f);
  t.l0 = 1.f - t.l1;
  return t;
}

__device__ __forceinline__ float raft_norm(float v255) { return 2.f * (v255 / 255.0f) - 1.0f; }  // raft.py:82-83

// mode 0: one thread per PADDED pixel, writes NHWC4.  mode 1: one thread per low-res pixel,
// writes u8 planes + per-block shared histograms flushed with atomics.
__global__ void __launch_bounds__(256) resize255_kernel(const float* __restrict__ img, int H, int W, int ht,
                                                        int wd, int mode, float* __restrict__ out_f32, int hp,
                                                        int wp, uint8_t* __restrict__ out_u8,
                                                        int* __restrict__ hist) {
  __shared__ int sh[3 * 256];
  if (mode == 1) {
    for (int i = threadIdx.x; i < 768; i += blockDim.x) sh[i] = 0;
    __syncthreads();
  }
  const float sy = (float)H / (float)ht, sx = (float)W / (float)wd;
  const int oh = mode == 0 ? hp : ht, ow = mode == 0 ? wp : wd;
  const int pt = (hp - ht) / 2, pl = (wp - wd) / 2;   // raft.py:136 [pw//2, ., ph//2, .]
  const int total = oh * ow;
  const int64_t plane = (int64_t)H * W;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    int y = i / ow, x = i - y * ow;
    if (mode == 0) {
      y = min(max(y - pt, 0), ht - 1);   // replicate padding
      x = min(max(x - pl, 0), wd - 1);
    }
    const AxisTap2 ty = src_index2(sy, y, H), tx = src_index2(sx, x, W);
    float v[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float* p = img + c * plane;
      const float v00 = __ldg(p + (int64_t)ty.i0 * W + tx.i0), v01 = __ldg(p + (int64_t)ty.i0 * W + tx.i1);
      const float v10 = __ldg(p + (int64_t)ty.i1 * W + tx.i0), v11 = __ldg(p + (int64_t)ty.i1 * W + tx.i1);
      v[c] = (ty.l0 * (tx.l0 * v00 + tx.l1 * v01) + ty.l1 * (tx.l0 * v10 + tx.l1 * v11)) * 255.f;
    }
    if (mode == 0) {
      *reinterpret_cast<float4*>(out_f32 + (int64_t)i * 4) =
          make_float4(raft_norm(v[0]), raft_norm(v[1]), raft_norm(v[2]), 0.f);
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int q = (int)fminf(fmaxf(v[c], 0.f), 255.f);   // .to(torch.uint8): truncation
        out_u8[(int64_t)c * total + i] = (uint8_t)q;
        atomicAdd(&sh[c * 256 + q], 1);
      }
    }
  }
  if (mode == 1) {
    __syncthreads();
    for (int i = threadIdx.x; i < 768; i += blockDim.x)
      if (sh[i]) atomicAdd(hist + i, sh[i]);
  }
}

// torchvision _scale_channel, integer exact.  One block per channel, one thread per bin: block-wide total / last
// non-zero bin, warp-scan prefix sum, one 64-bit division per thread.
__global__ void __launch_bounds__(256) equalize_lut_kernel(const int* __restrict__ hist, uint8_t* __restrict__ lut) {
  pdl_enter();
  const int c = blockIdx.x, i = threadIdx.x, lane = i & 31, w = i >> 5;
  __shared__ long long wsum[8];
  __shared__ int wlast[8];
  const int hi = hist[c * 256 + i];
  // inclusive prefix sum of the histogram
  long long incl = hi;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const long long t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  int last = hi != 0 ? i : -1;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
  if (lane == 31) wsum[w] = incl;
  if (lane == 0) wlast[w] = last;
  __syncthreads();
  long long before = 0, total = 0;
  int last_nz = -1;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (k < w) before += wsum[k];
    total += wsum[k];
    last_nz = max(last_nz, wlast[k]);
  }
  const long long cum = before + incl;                                    // sum(h[0..i])
  const long long step = last_nz >= 0 ? (total - hist[c * 256 + last_nz]) / 255 : 0;   // sum(nonzero_hist[:-1]) // 255
  uint8_t* l = lut + c * 256;
  if (step == 0) {
    l[i] = (uint8_t)i;                                                     // identity
    return;
  }
  if (i == 0) l[0] = 0;
  if (i < 255) {
    const long long v = (cum + step / 2) / step;                           // shifted right by one bin
    l[i + 1] = (uint8_t)(v > 255 ? 255 : v);
  }
}

__global__ void __launch_bounds__(256) equalize_apply_pack_kernel(const uint8_t* __restrict__ q,
                                                                  const uint8_t* __restrict__ lut, int ht, int wd,
                                                                  int hp, int wp, float* __restrict__ out_f32,
                                                                  uint8_t* __restrict__ eq_u8) {
  __shared__ uint8_t sl[768];
  for (int i = threadIdx.x; i < 768; i += blockDim.x) sl[i] = lut[i];
  __syncthreads();
  const int pt = (hp - ht) / 2, pl = (wp - wd) / 2;
  const int total = hp * wp, lowtotal = ht * wd;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int py = i / wp, px = i - py * wp;
    const int y = min(max(py - pt, 0), ht - 1), x = min(max(px - pl, 0), wd - 1);
    const int s = y * wd + x;
    float v[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const uint8_t e = sl[c * 256 + q[c * lowtotal + s]];
      v[c] = (float)e;
      if (eq_u8 && py - pt == y && px - pl == x) eq_u8[c * lowtotal + s] = e;
    }
    *reinterpret_cast<float4*>(out_f32 + (int64_t)i * 4) =
        make_float4(raft_norm(v[0]), raft_norm(v[1]), raft_norm(v[2]), 0.f);
  }
}

// ---------------- layout / Retinex elementwise ----------------
__global__ void __launch_bounds__(256) pack_nhwc_kernel(const float* __restrict__ s0, const float* __restrict__ s1,
                                                        const float* __restrict__ s2, const float* __restrict__ s3,
                                                        int c_each, int nsrc, float add, float* __restrict__ out,
                                                        int cs, int64_t HW) {
  const float* src[4] = {s0, s1, s2, s3};
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += (int64_t)gridDim.x * blockDim.x) {
    float* o = out + p * cs;
    int ch = 0;
    for (int s = 0; s < nsrc; ++s)
      for (int c = 0; c < c_each; ++c) o[ch++] = __ldg(src[s] + c * HW + p) + add;
    for (; ch < cs; ++ch) o[ch] = 0.f;
  }
}

__global__ void __launch_bounds__(256) retinex_l2_kernel(const float* __restrict__ x, const float* __restrict__ d,
                                                         int d_cs, float* __restrict__ L2, int64_t HW, float eps) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c)
      L2[c * HW + p] = clampf((x[c * HW + p] + eps) - d[p * d_cs + c], eps, 1.f);
  }
}

__global__ void __launch_bounds__(256) retinex_h2_kernel(const float* __restrict__ x, const float* __restrict__ s,
                                                         int s_cs, float* __restrict__ s2, float* __restrict__ H2,
                                                         int64_t HW, float eps) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float sv = clampf(s[p * s_cs + c], 0.0001f, 1.f);
      s2[c * HW + p] = sv;
      H2[c * HW + p] = clampf((x[c * HW + p] + eps) / sv, eps, 1.f);
    }
  }
}

__global__ void __launch_bounds__(256) retinex_h5_kernel(const float* __restrict__ H2, const float* __restrict__ s2,
                                                         const float* __restrict__ d, int d_cs,
                                                         float* __restrict__ H3, float* __restrict__ s3, int64_t HW,
                                                         float eps) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      H3[c * HW + p] = clampf(H2[c * HW + p] - d[p * d_cs + c], eps, 1.f);
      s3[c * HW + p] = clampf(s2[c * HW + p] - d[p * d_cs + 3 + c], eps, 1.f);
    }
  }
}

__global__ void __launch_bounds__(256) quantize_u8_hwc_kernel(const float* __restrict__ x, uint8_t* __restrict__ out,
                                                              int C, int64_t HW) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += (int64_t)gridDim.x * blockDim.x)
    for (int c = 0; c < C; ++c) {
      const float v = fminf(fmaxf(x[c * HW + p] * 255.f, 0.f), 255.f);   // np.clip(x*255,0,255).astype(uint8)
      out[p * C + c] = (uint8_t)(int)v;
    }
}

// three channels, four pixels per thread: float4 plane loads, 12 contiguous output bytes (three 32-bit stores)
__global__ void __launch_bounds__(256) quantize_u8_hwc3_x4_kernel(const float* __restrict__ x, uint32_t* __restrict__ out,
                                                                  int64_t HW) {
  const int64_t n4 = HW >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 r = __ldg(reinterpret_cast<const float4*>(x) + i);
    const float4 g = __ldg(reinterpret_cast<const float4*>(x + HW) + i);
    const float4 b = __ldg(reinterpret_cast<const float4*>(x + 2 * HW) + i);
    auto q = [](float v) { return (uint32_t)(int)fminf(fmaxf(v * 255.f, 0.f), 255.f); };
    const uint32_t by[12] = {q(r.x), q(g.x), q(b.x), q(r.y), q(g.y), q(b.y), q(r.z), q(g.z), q(b.z), q(r.w), q(g.w), q(b.w)};
    uint32_t* o = out + i * 3;
#pragma unroll
    for (int k = 0; k < 3; ++k) o[k] = by[4 * k] | (by[4 * k + 1] << 8) | (by[4 * k + 2] << 16) | (by[4 * k + 3] << 24);
  }
}

// ToTensor (dataloader/multi_read_data.py:86-87,131): uint8 HWC -> float NCHW, value / 255 (IEEE division, like torch)
__global__ void __launch_bounds__(256) u8_hwc_to_f32_kernel(const uint8_t* __restrict__ in, float* __restrict__ out, int C,
                                                            int64_t HW) {
  for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < HW; p += (int64_t)gridDim.x * blockDim.x)
    for (int c = 0; c < C; ++c) out[c * HW + p] = __fdiv_rn((float)in[p * C + c], 255.f);
}
__global__ void __launch_bounds__(256) u8_hwc3_to_f32_x4_kernel(const uint32_t* __restrict__ in, float* __restrict__ out,
                                                                int64_t HW) {
  const int64_t n4 = HW >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t w0 = __ldg(in + i * 3), w1 = __ldg(in + i * 3 + 1), w2 = __ldg(in + i * 3 + 2);
    auto f = [](uint32_t v) { return __fdiv_rn((float)(v & 255u), 255.f); };
    // bytes: r0 g0 b0 r1 | g1 b1 r2 g2 | b2 r3 g3 b3
    reinterpret_cast<float4*>(out)[i] = make_float4(f(w0), f(w0 >> 24), f(w1 >> 16), f(w2 >> 8));
    reinterpret_cast<float4*>(out + HW)[i] = make_float4(f(w0 >> 8), f(w1), f(w1 >> 24), f(w2 >> 16));
    reinterpret_cast<float4*>(out + 2 * HW)[i] = make_float4(f(w0 >> 16), f(w1 >> 8), f(w2), f(w2 >> 24));
  }
}

static inline int ew_blocks(int64_t n) { return (int)std::min<int64_t>(ceil_div64(n, 256), 16 * kNumSMs); }

}  // namespace zt

extern "C" int zt_resize255_f32(const float* img, int H, int W, int ht, int wd, int mode, float* out_f32, int hp,
                                int wp, uint8_t* out_u8, int* hist, void* stream) {
  using namespace zt;
  ZT_REQUIRE(img && H > 0 && W > 0 && ht > 0 && wd > 0, "resize255: bad args");
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 0) {
    ZT_REQUIRE(out_f32 && hp >= ht && wp >= wd, "resize255: mode 0 needs out_f32 and padded dims");
    resize255_kernel<<<ew_blocks((int64_t)hp * wp), 256, 0, st>>>(img, H, W, ht, wd, 0, out_f32, hp, wp, nullptr, nullptr);
  } else {
    ZT_REQUIRE(out_u8 && hist, "resize255: mode 1 needs out_u8 and hist");
    ZT_CUDA(cudaMemsetAsync(hist, 0, sizeof(int) * 768, st));
    const int blocks = min(ew_blocks((int64_t)ht * wd), 2 * kNumSMs);
    resize255_kernel<<<blocks, 256, 0, st>>>(img, H, W, ht, wd, 1, nullptr, ht, wd, out_u8, hist);
  }
  return check_launch("resize255_kernel");
}

extern "C" int zt_equalize_lut(const int* hist, uint8_t* lut, void* stream) {
  using namespace zt;
  ZT_REQUIRE(hist && lut, "equalize_lut: null pointer");
  equalize_lut_kernel<<<3, 256, 0, (cudaStream_t)stream>>>(hist, lut);
  return check_launch("equalize_lut_kernel");
}

extern "C" int zt_equalize_apply_pack(const uint8_t* q, const uint8_t* lut, int ht, int wd, int hp, int wp,
                                      float* out_f32, uint8_t* eq_u8, void* stream) {
  using namespace zt;
  ZT_REQUIRE(q && lut && out_f32 && ht > 0 && wd > 0 && hp >= ht && wp >= wd, "equalize_apply_pack: bad args");
  equalize_apply_pack_kernel<<<ew_blocks((int64_t)hp * wp), 256, 0, (cudaStream_t)stream>>>(q, lut, ht, wd, hp, wp,
                                                                                            out_f32, eq_u8);
  return check_launch("equalize_apply_pack_kernel");
}

extern "C" int zt_pack_nhwc_f32(const float* s0, const float* s1, const float* s2, const float* s3, int c_each,
                                int nsrc, float add, float* out, int cs, int64_t HW, void* stream) {
  using namespace zt;
  ZT_REQUIRE(s0 && out && nsrc >= 1 && nsrc <= 4 && c_each >= 1 && c_each * nsrc <= cs && HW > 0, "pack_nhwc: bad args");
  ZT_REQUIRE((nsrc < 2 || s1) && (nsrc < 3 || s2) && (nsrc < 4 || s3), "pack_nhwc: missing source");
  pack_nhwc_kernel<<<ew_blocks(HW), 256, 0, (cudaStream_t)stream>>>(s0, s1, s2, s3, c_each, nsrc, add, out, cs, HW);
  return check_launch("pack_nhwc_kernel");
}

extern "C" int zt_retinex_l2_f32(const float* x, const float* d, int d_cs, float* L2, int64_t HW, float eps,
                                 void* stream) {
  using namespace zt;
  ZT_REQUIRE(x && d && L2 && d_cs >= 3 && HW > 0, "retinex_l2: bad args");
  retinex_l2_kernel<<<ew_blocks(HW), 256, 0, (cudaStream_t)stream>>>(x, d, d_cs, L2, HW, eps);
  return check_launch("retinex_l2_kernel");
}

extern "C" int zt_retinex_h2_f32(const float* x, const float* s, int s_cs, float* s2, float* H2, int64_t HW,
                                 float eps, void* stream) {
  using namespace zt;
  ZT_REQUIRE(x && s && s2 && H2 && s_cs >= 3 && HW > 0, "retinex_h2: bad args");
  retinex_h2_kernel<<<ew_blocks(HW), 256, 0, (cudaStream_t)stream>>>(x, s, s_cs, s2, H2, HW, eps);
  return check_launch("retinex_h2_kernel");
}

extern "C" int zt_retinex_h5_f32(const float* H2, const float* s2, const float* d, int d_cs, float* H3, float* s3,
                                 int64_t HW, float eps, void* stream) {
  using namespace zt;
  ZT_REQUIRE(H2 && s2 && d && H3 && s3 && d_cs >= 6 && HW > 0, "retinex_h5: bad args");
  retinex_h5_kernel<<<ew_blocks(HW), 256, 0, (cudaStream_t)stream>>>(H2, s2, d, d_cs, H3, s3, HW, eps);
  return check_launch("retinex_h5_kernel");
}

extern "C" int zt_quantize_u8_hwc(const float* x, uint8_t* out, int C, int64_t HW, void* stream) {
  using namespace zt;
  ZT_REQUIRE(x && out && C >= 1 && HW > 0, "quantize_u8: bad args");
  if (C == 3 && (HW & 3) == 0 && (((uintptr_t)x | (uintptr_t)out) & 15) == 0) {
    quantize_u8_hwc3_x4_kernel<<<ew_blocks(HW >> 2), 256, 0, (cudaStream_t)stream>>>(x, (uint32_t*)out, HW);
    return check_launch("quantize_u8_hwc3_x4_kernel");
  }
  quantize_u8_hwc_kernel<<<ew_blocks(HW), 256, 0, (cudaStream_t)stream>>>(x, out, C, HW);
  return check_launch("quantize_u8_hwc_kernel");
}

extern "C" int zt_u8_hwc_to_f32(const uint8_t* in, float* out, int C, int64_t HW, void* stream) {
  using namespace zt;
  ZT_REQUIRE(in && out && C >= 1 && HW > 0, "u8_hwc_to_f32: bad args");
  if (C == 3 && (HW & 3) == 0 && (((uintptr_t)in | (uintptr_t)out) & 15) == 0) {
    u8_hwc3_to_f32_x4_kernel<<<ew_blocks(HW >> 2), 256, 0, (cudaStream_t)stream>>>((const uint32_t*)in, out, HW);
    return check_launch("u8_hwc3_to_f32_x4_kernel");
  }
  u8_hwc_to_f32_kernel<<<ew_blocks(HW), 256, 0, (cudaStream_t)stream>>>(in, out, C, HW);
  return check_launch("u8_hwc_to_f32_kernel");
}
