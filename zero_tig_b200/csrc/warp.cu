// Fused backward warp of the two recurrent state planes by one flow field.
// Reference: warp_tensor utils/utils.py:203-230, called twice with the same flow
// (model/model.py:374-375).  One pass computes, per output pixel, the align_corners=False
// bilinear upsample of the low-res sampling map, the [-1,1] normalisation by (W-1)/2 and
// grid_sample's align_corners=False un-normalisation (quirks Q2, Q4, Q11 reproduced), then
// gathers both images with the same coordinates.  HBM-bound: algorithmic bytes per pair =
// flow + 2 images read + 2 images written.
#include <stdlib.h>
#include <algorithm>
#include "common.cuh"

namespace zt {

struct AxisTap { int i0, i1; float l0, l1; };

__device__ __forceinline__ AxisTap src_index(float scale, int dst, int n_in) {
  float s = scale * ((float)dst + 0.5f) - 0.5f;   // area_pixel_compute_source_index
  s = s < 0.f ? 0.f : s;
  AxisTap t;
  t.i0 = min((int)s, n_in - 1);
  t.i1 = t.i0 + (t.i0 < n_in - 1 ? 1 : 0);
  t.l1 = fminf(fmaxf(s - (float)t.i0, 0.f), 1.f);
  t.l0 = 1.f - t.l1;
  return t;
}

template <int VEC>
__global__ void __launch_bounds__(256) warp_pair_kernel(const float* __restrict__ flow, int hf, int wf,
                                                        const float* __restrict__ img_a,
                                                        const float* __restrict__ img_b,
                                                        float* __restrict__ out_a, float* __restrict__ out_b,
                                                        int C, int H, int W) {
  const int wq = (W + VEC - 1) / VEC;
  const int64_t total = (int64_t)H * wq;
  const float h_scale = (float)((double)H / (double)hf);   // applied to x (Q2)
  const float w_scale = (float)((double)W / (double)wf);   // applied to y (Q2)
  const float sy = (float)hf / (float)H, sx = (float)wf / (float)W;
  const float nx = (float)((double)(W - 1) / 2.0), ny = (float)((double)(H - 1) / 2.0);
  const int64_t plane = (int64_t)H * W;
  const float* fxp = flow;
  const float* fyp = flow + (int64_t)hf * wf;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int y = (int)(i / wq);
    const int x0 = (int)(i - (int64_t)y * wq) * VEC;
    const AxisTap ty = src_index(sy, y, hf);
    float ra[3][VEC], rb[3][VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
      const int x = x0 + v;
#pragma unroll
      for (int c = 0; c < 3; ++c) { ra[c][v] = 0.f; rb[c][v] = 0.f; }
      if (x >= W) continue;
      const AxisTap tx = src_index(sx, x, wf);
      const int o00 = ty.i0 * wf + tx.i0, o01 = ty.i0 * wf + tx.i1;
      const int o10 = ty.i1 * wf + tx.i0, o11 = ty.i1 * wf + tx.i1;
      // low-res maps: (grid - flow) * scale
      const float mx00 = ((float)tx.i0 - __ldg(fxp + o00)) * h_scale, mx01 = ((float)tx.i1 - __ldg(fxp + o01)) * h_scale;
      const float mx10 = ((float)tx.i0 - __ldg(fxp + o10)) * h_scale, mx11 = ((float)tx.i1 - __ldg(fxp + o11)) * h_scale;
      const float my00 = ((float)ty.i0 - __ldg(fyp + o00)) * w_scale, my01 = ((float)ty.i0 - __ldg(fyp + o01)) * w_scale;
      const float my10 = ((float)ty.i1 - __ldg(fyp + o10)) * w_scale, my11 = ((float)ty.i1 - __ldg(fyp + o11)) * w_scale;
      const float mx = ty.l0 * (tx.l0 * mx00 + tx.l1 * mx01) + ty.l1 * (tx.l0 * mx10 + tx.l1 * mx11);
      const float my = ty.l0 * (tx.l0 * my00 + tx.l1 * my01) + ty.l1 * (tx.l0 * my10 + tx.l1 * my11);
      const float gx = mx / nx - 1.f, gy = my / ny - 1.f;
      const float ix = ((gx + 1.f) * (float)W - 1.f) / 2.f;   // grid_sampler_unnormalize, align_corners=False
      const float iy = ((gy + 1.f) * (float)H - 1.f) / 2.f;
      const float fx0 = floorf(ix), fy0 = floorf(iy);
      const float wx1 = ix - fx0, wy1 = iy - fy0;
      const float wx0 = (fx0 + 1.f) - ix, wy0 = (fy0 + 1.f) - iy;
      // keep the int conversion defined for wild coordinates
      const int xi = (int)fminf(fmaxf(fx0, -2.f), (float)W + 1.f);
      const int yi = (int)fminf(fmaxf(fy0, -2.f), (float)H + 1.f);
      const bool x0ok = xi >= 0 && xi < W, x1ok = xi + 1 >= 0 && xi + 1 < W;
      const bool y0ok = yi >= 0 && yi < H, y1ok = yi + 1 >= 0 && yi + 1 < H;
      const float w00 = wx0 * wy0, w01 = wx1 * wy0, w10 = wx0 * wy1, w11 = wx1 * wy1;
      const int64_t b00 = (int64_t)yi * W + xi;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        if (c >= C) break;
        const float* pa = img_a + c * plane + b00;
        float acc = 0.f;
        if (y0ok && x0ok) acc += __ldg(pa) * w00;
        if (y0ok && x1ok) acc += __ldg(pa + 1) * w01;
        if (y1ok && x0ok) acc += __ldg(pa + W) * w10;
        if (y1ok && x1ok) acc += __ldg(pa + W + 1) * w11;
        ra[c][v] = acc;
        if (img_b) {
          const float* pb = img_b + c * plane + b00;
          float acb = 0.f;
          if (y0ok && x0ok) acb += __ldg(pb) * w00;
          if (y0ok && x1ok) acb += __ldg(pb + 1) * w01;
          if (y1ok && x0ok) acb += __ldg(pb + W) * w10;
          if (y1ok && x1ok) acb += __ldg(pb + W + 1) * w11;
          rb[c][v] = acb;
        }
      }
    }
    const int64_t o = (int64_t)y * W + x0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      if (c >= C) break;
      if (VEC == 4) {
        *reinterpret_cast<float4*>(out_a + c * plane + o) = make_float4(ra[c][0], ra[c][1], ra[c][2], ra[c][3]);
        if (img_b) *reinterpret_cast<float4*>(out_b + c * plane + o) = make_float4(rb[c][0], rb[c][1], rb[c][2], rb[c][3]);
      } else {
        out_a[c * plane + o] = ra[c][0];
        if (img_b) out_b[c * plane + o] = rb[c][0];
      }
    }
  }
}

// Three-channel pair, R consecutive output rows of one column per thread (2-D grid, no index divisions).
// The low-res sampling map is piecewise bilinear: consecutive output rows mostly share their two flow rows, so the
// horizontally interpolated map values of those rows (hx*, hy*) are kept in registers and re-fetched only when the
// row pair changes (warp-uniform test: a warp owns 32 consecutive columns of the same rows).  Same arithmetic, in
// the same order, as warp_pair_kernel; taps inside the image (all but a border fringe) take a predicate-free path.
struct WarpScales { float h_scale, w_scale, sy, sx, nx, ny; };   // the same expressions as in warp_pair_kernel

template <int R>
__global__ void __launch_bounds__(128) warp_pair3_rows_kernel(const float* __restrict__ flow, int hf, int wf,
                                                              const float* __restrict__ img_a,
                                                              const float* __restrict__ img_b,
                                                              float* __restrict__ out_a, float* __restrict__ out_b,
                                                              int H, int W, WarpScales k) {
  const int x = blockIdx.x * 128 + threadIdx.x;
  const int y_begin = blockIdx.y * R;
  if (x >= W) return;
  const float h_scale = k.h_scale, w_scale = k.w_scale, sy = k.sy, sx = k.sx, nx = k.nx, ny = k.ny;
  const int plane = H * W;
  const float* fxp = flow;
  const float* fyp = flow + (int64_t)hf * wf;
  const AxisTap tx = src_index(sx, x, wf);
  const float gx0 = (float)tx.i0, gx1 = (float)tx.i1;
  int ci0 = -1, ci1 = -1;
  float hx0 = 0.f, hx1 = 0.f, hy0 = 0.f, hy1 = 0.f;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int y = y_begin + r;
    if (y >= H) break;
    const AxisTap ty = src_index(sy, y, hf);
    if (ty.i0 != ci0 || ty.i1 != ci1) {
      ci0 = ty.i0; ci1 = ty.i1;
      const int o00 = ty.i0 * wf + tx.i0, o01 = ty.i0 * wf + tx.i1;
      const int o10 = ty.i1 * wf + tx.i0, o11 = ty.i1 * wf + tx.i1;
      const float mx00 = (gx0 - __ldg(fxp + o00)) * h_scale, mx01 = (gx1 - __ldg(fxp + o01)) * h_scale;
      const float mx10 = (gx0 - __ldg(fxp + o10)) * h_scale, mx11 = (gx1 - __ldg(fxp + o11)) * h_scale;
      const float my00 = ((float)ty.i0 - __ldg(fyp + o00)) * w_scale, my01 = ((float)ty.i0 - __ldg(fyp + o01)) * w_scale;
      const float my10 = ((float)ty.i1 - __ldg(fyp + o10)) * w_scale, my11 = ((float)ty.i1 - __ldg(fyp + o11)) * w_scale;
      hx0 = tx.l0 * mx00 + tx.l1 * mx01; hx1 = tx.l0 * mx10 + tx.l1 * mx11;
      hy0 = tx.l0 * my00 + tx.l1 * my01; hy1 = tx.l0 * my10 + tx.l1 * my11;
    }
    const float mx = ty.l0 * hx0 + ty.l1 * hx1;
    const float my = ty.l0 * hy0 + ty.l1 * hy1;
    const float gx = mx / nx - 1.f, gy = my / ny - 1.f;
    const float ix = ((gx + 1.f) * (float)W - 1.f) / 2.f;   // grid_sampler_unnormalize, align_corners=False
    const float iy = ((gy + 1.f) * (float)H - 1.f) / 2.f;
    const float fx0 = floorf(ix), fy0 = floorf(iy);
    const float wx1 = ix - fx0, wy1 = iy - fy0;
    const float wx0 = (fx0 + 1.f) - ix, wy0 = (fy0 + 1.f) - iy;
    const int xi = (int)fminf(fmaxf(fx0, -2.f), (float)W + 1.f);
    const int yi = (int)fminf(fmaxf(fy0, -2.f), (float)H + 1.f);
    const float w00 = wx0 * wy0, w01 = wx1 * wy0, w10 = wx0 * wy1, w11 = wx1 * wy1;
    const int b00 = yi * W + xi;
    const int o = y * W + x;
    if (xi >= 0 && xi + 1 < W && yi >= 0 && yi + 1 < H) {
      float va[3][4], vb[3][4];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float* pa = img_a + c * plane + b00;
        const float* pb = img_b + c * plane + b00;
        va[c][0] = __ldg(pa); va[c][1] = __ldg(pa + 1); va[c][2] = __ldg(pa + W); va[c][3] = __ldg(pa + W + 1);
        vb[c][0] = __ldg(pb); vb[c][1] = __ldg(pb + 1); vb[c][2] = __ldg(pb + W); vb[c][3] = __ldg(pb + W + 1);
      }
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        out_a[c * plane + o] = (((0.f + va[c][0] * w00) + va[c][1] * w01) + va[c][2] * w10) + va[c][3] * w11;
        out_b[c * plane + o] = (((0.f + vb[c][0] * w00) + vb[c][1] * w01) + vb[c][2] * w10) + vb[c][3] * w11;
      }
    } else {
      const bool x0ok = xi >= 0 && xi < W, x1ok = xi + 1 >= 0 && xi + 1 < W;
      const bool y0ok = yi >= 0 && yi < H, y1ok = yi + 1 >= 0 && yi + 1 < H;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float* pa = img_a + c * plane + b00;
        const float* pb = img_b + c * plane + b00;
        float acc = 0.f, acb = 0.f;
        if (y0ok && x0ok) { acc += __ldg(pa) * w00; acb += __ldg(pb) * w00; }
        if (y0ok && x1ok) { acc += __ldg(pa + 1) * w01; acb += __ldg(pb + 1) * w01; }
        if (y1ok && x0ok) { acc += __ldg(pa + W) * w10; acb += __ldg(pb + W) * w10; }
        if (y1ok && x1ok) { acc += __ldg(pa + W + 1) * w11; acb += __ldg(pb + W + 1) * w11; }
        out_a[c * plane + o] = acc;
        out_b[c * plane + o] = acb;
      }
    }
  }
}

}  // namespace zt

extern "C" int zt_warp_pair_f32(const float* flow, int hf, int wf, const float* img_a, const float* img_b,
                                float* out_a, float* out_b, int C, int H, int W, void* stream) {
  using namespace zt;
  ZT_REQUIRE(flow && img_a && out_a && (!img_b || out_b), "warp_pair: null pointer");
  ZT_REQUIRE(C >= 1 && C <= 3 && H > 1 && W > 1 && hf > 0 && wf > 0, "warp_pair: bad dims (C=%d H=%d W=%d)", C, H, W);
  cudaStream_t st = (cudaStream_t)stream;
  // 4 pixels per thread with 128-bit stores is the faster mapping on real (smooth) flow: 58 us vs 65 us in the
  // 1080p frame loop; on a random +-9 px flow the one-pixel mapping wins (77 us vs 105 us, tools/warp_bench.py,
  // L1-sector bound: profiles/r01_ncu_warp.txt).  ZT_WARP_SCALAR=1 selects the one-pixel mapping.
  static const bool force_scalar = getenv("ZT_WARP_SCALAR") != nullptr;
  static const int cap = getenv("ZT_WARP_CAP") ? atoi(getenv("ZT_WARP_CAP")) : 64;
  // Default for the 3-channel pair (the only shape the frame loop uses): the row-strip kernel above.
  static const int rows_mode = getenv("ZT_WARP_ROWS") ? atoi(getenv("ZT_WARP_ROWS")) : 3;
  if (rows_mode > 0 && !force_scalar && C == 3 && img_b && (int64_t)H * W * 3 < (1ll << 31)) {
    WarpScales k;
    k.h_scale = (float)((double)H / (double)hf);   // applied to x (Q2)
    k.w_scale = (float)((double)W / (double)wf);   // applied to y (Q2)
    k.sy = (float)hf / (float)H; k.sx = (float)wf / (float)W;
    k.nx = (float)((double)(W - 1) / 2.0); k.ny = (float)((double)(H - 1) / 2.0);
    if (rows_mode >= 6) {
      dim3 grid((unsigned)ceil_div64(W, 128), (unsigned)ceil_div64(H, 6));
      warp_pair3_rows_kernel<6><<<grid, 128, 0, st>>>(flow, hf, wf, img_a, img_b, out_a, out_b, H, W, k);
    } else {
      dim3 grid((unsigned)ceil_div64(W, 128), (unsigned)ceil_div64(H, 3));
      warp_pair3_rows_kernel<3><<<grid, 128, 0, st>>>(flow, hf, wf, img_a, img_b, out_a, out_b, H, W, k);
    }
    return check_launch("warp_pair3_rows_kernel");
  }
  const bool vec = !force_scalar && (W & 3) == 0 && ((uintptr_t)out_a & 15) == 0 && (!out_b || ((uintptr_t)out_b & 15) == 0);
  if (vec) {
    const int64_t total = (int64_t)H * (W / 4);
    const int blocks = (int)std::min<int64_t>(ceil_div64(total, 256), 16 * kNumSMs);
    warp_pair_kernel<4><<<blocks, 256, 0, st>>>(flow, hf, wf, img_a, img_b, out_a, out_b, C, H, W);
  } else {
    const int64_t total = (int64_t)H * W;
    const int blocks = (int)std::min<int64_t>(ceil_div64(total, 256), (int64_t)cap * kNumSMs);
    warp_pair_kernel<1><<<blocks, 256, 0, st>>>(flow, hf, wf, img_a, img_b, out_a, out_b, C, H, W);
  }
  return check_launch("warp_pair_kernel");
}
