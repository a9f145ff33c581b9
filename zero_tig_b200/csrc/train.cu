// Training-path kernels: pair_downsampler, Retinex elementwise variants, TextureDifference,
// separable 21-tap blur, and the reductions of LossFunction (forward).
// Reference: utils/utils.py:15-24,41-79; model/model.py:144-203; loss.py:23-152,173-311.
// All tensors NCHW fp32, batch 1 (train.py:16).  Reductions: warp shuffle -> block -> one fp64
// atomicAdd per block into a small accumulator array (deterministic enough for 1e-3 relative).
#include <algorithm>
#include "common.cuh"

namespace zt {

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// block-wide sum of `v`, one atomicAdd(double) by thread 0
__device__ __forceinline__ void block_accumulate(double v, double* acc) {
  __shared__ double red[32];
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  if (w == 0) {
    v = l < (int)((blockDim.x + 31) >> 5) ? red[l] : 0.0;
    v = warp_sum(v);
    if (l == 0 && v != 0.0) atomicAdd(acc, v);
  }
}
__device__ __forceinline__ int reflect_idx(int i, int n) {   // F.pad(mode='reflect')
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return i;
}
inline int grid_for(int64_t total, int threads = 256) {
  return (int)std::min<int64_t>(ceil_div64(total, threads), 32 * kNumSMs);
}

// ---------------- pair_downsampler (utils.py:15-24) ----------------
__global__ void pair_down_kernel(const float* __restrict__ in, float* __restrict__ o1, float* __restrict__ o2,
                                 int C, int H, int W, float add) {
  const int ho = H >> 1, wo = W >> 1;
  const int64_t total = (int64_t)C * ho * wo;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % wo);
    const int64_t r = i / wo;
    const int y = (int)(r % ho), c = (int)(r / ho);
    const float* p = in + ((int64_t)c * H + 2 * y) * W + 2 * x;
    const float a = p[0] + add, b = p[1] + add, cc = p[W] + add, d = p[W + 1] + add;
    // conv2d with [[0,.5],[.5,0]] / [[.5,0],[0,.5]]: products summed in tap order (0*a + .5*b + .5*c + 0*d)
    o1[i] = 0.5f * b + 0.5f * cc;
    o2[i] = 0.5f * a + 0.5f * d;
  }
}

// out[c][p] = clamp?( (a[c][p] + add) - d[p][c] )   d is NHWC with channel stride d_cs
__global__ void sub_clamp_kernel(const float* __restrict__ a, float add, const float* __restrict__ d, int d_cs,
                                 float* __restrict__ out, int C, int64_t HW, float lo, float hi, int do_clamp) {
  const int64_t total = (int64_t)C * HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i / HW);
    const int64_t p = i - (int64_t)c * HW;
    float v = (a[i] + add) - d[p * d_cs + c];
    if (do_clamp) v = clampf(v, lo, hi);
    out[i] = v;
  }
}
// out = clamp((a + add) / b, lo, hi)
__global__ void div_clamp_kernel(const float* __restrict__ a, float add, const float* __restrict__ b,
                                 float* __restrict__ out, int64_t n, float lo, float hi) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = clampf((a[i] + add) / b[i], lo, hi);
}

// ---------------- TextureDifference (loss.py:106-136) ----------------
__device__ __forceinline__ float gray_at(const float* im, int64_t plane, int y, int x, int W) {
  const int64_t o = (int64_t)y * W + x;
  return 0.144f * im[o] + 0.5870f * im[plane + o] + 0.299f * im[2 * plane + o];
}
__global__ void texture_diff_kernel(const float* __restrict__ i1, const float* __restrict__ i2, float* __restrict__ out,
                                    int H, int W, float Cc, float thr) {
  const int64_t plane = (int64_t)H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < plane; i += (int64_t)gridDim.x * blockDim.x) {
    const int y = (int)(i / W), x = (int)(i - (int64_t)y * W);
    float g1[25], g2[25], m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) {
      const int yy = reflect_idx(y + k / 5 - 2, H), xx = reflect_idx(x + k % 5 - 2, W);
      g1[k] = gray_at(i1, plane, yy, xx, W);
      g2[k] = gray_at(i2, plane, yy, xx, W);
      m1 += g1[k];
      m2 += g2[k];
    }
    m1 /= 25.f;
    m2 /= 25.f;
    float v1 = 0.f, v2 = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) {
      v1 += (g1[k] - m1) * (g1[k] - m1);
      v2 += (g2[k] - m2) * (g2[k] - m2);
    }
    const float s1 = sqrtf(v1 / 25.f + 1e-9f), s2 = sqrtf(v2 / 25.f + 1e-9f);
    const float dd = 2.f * s1 * s2 / (s1 * s1 + s2 * s2 + Cc);
    out[i] = dd > thr ? 1.f : 0.f;
  }
}

// ---------------- separable 21-tap blur with reflect padding (utils.py:52-58) ----------------
struct BlurTaps { float k[21]; };
// dir 0: along x, dir 1: along y
__global__ void blur21_kernel(const float* __restrict__ in, float* __restrict__ out, int C, int H, int W, int dir,
                              const BlurTaps t) {
  const int64_t total = (int64_t)C * H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % W);
    const int64_t r = i / W;
    const int y = (int)(r % H);
    const float* base = in + (r - y) * W;   // start of the plane
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 21; ++k) {
      const int yy = dir ? reflect_idx(y + k - 10, H) : y;
      const int xx = dir ? x : reflect_idx(x + k - 10, W);
      acc += t.k[k] * base[(int64_t)yy * W + xx];
    }
    out[i] = acc;
  }
}


// ---------------- LossFunction reductions (loss.py:23-78) ----------------
// per-channel sums of a 3-plane image -> acc[0..2]
__global__ void chan_sum_kernel(const float* __restrict__ x, int64_t HW, double* __restrict__ acc) {
  const int c = blockIdx.y;
  double s = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += (int64_t)gridDim.x * blockDim.x)
    s += x[(int64_t)c * HW + i];
  block_accumulate(s, acc + c);
}

// illumination targets (loss.py:26-47): acc[0] += (s2 - bright)^2, acc[1] += (norm_ll - adj)^2
__global__ void illum_kernel(const float* __restrict__ L2, const float* __restrict__ s2,
                             const double* __restrict__ chansum, int is_WB, int64_t HW, double* __restrict__ acc) {
  const int c = blockIdx.y;
  const float eps = 1e-9f;
  float ef;
  if (is_WB) {
    ef = 0.3f / ((float)(chansum[c] / (double)HW) + eps);
  } else {
    const float ym = (float)((chansum[2] * 0.299 + chansum[1] * 0.587 + chansum[0] * 0.144) / (double)HW);
    ef = 0.5f / (ym + eps);
  }
  ef = clampf(ef, 1.f, 25.f);
  const float ratio = powf(0.7f, -ef) / ef;
  double a0 = 0.0, a1 = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += (int64_t)gridDim.x * blockDim.x) {
    const float l = L2[(int64_t)c * HW + i], s = s2[(int64_t)c * HW + i];
    const float nl = clampf(l / s, eps, 0.8f);
    const float bright = clampf(powf(l * ef, ef) * ratio, eps, 1.f);
    const float adj = clampf(l * ef, eps, 1.f);
    a0 += (double)((s - bright) * (s - bright));
    a1 += (double)((nl - adj) * (nl - adj));
  }
  block_accumulate(a0, acc);
  block_accumulate(a1, acc + 1);
}

// SmoothLoss.rgb2yCbCr on FLAT triples of the NCHW buffer (quirk Q8, loss.py:178-190)
__global__ void ycbcr_flat_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t ntriples) {
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < ntriples; t += (int64_t)gridDim.x * blockDim.x) {
    const float a = in[3 * t], b = in[3 * t + 1], c = in[3 * t + 2];
    out[3 * t + 0] = a * 0.257f + b * 0.564f + c * 0.098f + 16.0f / 255.0f;
    out[3 * t + 1] = a * -0.148f + b * -0.291f + c * 0.439f + 128.0f / 255.0f;
    out[3 * t + 2] = a * 0.439f + b * -0.368f + c * -0.071f + 128.0f / 255.0f;
  }
}

__constant__ int c_smooth_dy[12] = {1, 0, 1, 1, 2, 0, 2, 2, 1, 1, 2, 2};
__constant__ int c_smooth_dx[12] = {0, 1, 1, -1, 0, 2, 1, -1, 2, -2, 2, -2};
// acc[o] += exp(-sum_c dY^2/200) * sum_c |d s2| over all pixel pairs (p, p + offset o)
__global__ void smooth_kernel(const float* __restrict__ y, const float* __restrict__ s2, int H, int W,
                              double* __restrict__ acc) {
  const int64_t plane = (int64_t)H * W;
  double part[12];
#pragma unroll
  for (int o = 0; o < 12; ++o) part[o] = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < plane; i += (int64_t)gridDim.x * blockDim.x) {
    const int py = (int)(i / W), px = (int)(i - (int64_t)py * W);
    const float y0 = y[i], y1 = y[plane + i], y2 = y[2 * plane + i];
    const float s0 = s2[i], s1 = s2[plane + i], s2v = s2[2 * plane + i];
#pragma unroll
    for (int o = 0; o < 12; ++o) {
      const int qy = py + c_smooth_dy[o], qx = px + c_smooth_dx[o];
      if (qy < H && qx >= 0 && qx < W) {
        const int64_t j = (int64_t)qy * W + qx;
        const float d0 = y0 - y[j], d1 = y1 - y[plane + j], d2 = y2 - y[2 * plane + j];
        const float wgt = expf((d0 * d0 + d1 * d1 + d2 * d2) * (-1.0f / 200.0f));
        const float g = fabsf(s0 - s2[j]) + fabsf(s1 - s2[plane + j]) + fabsf(s2v - s2[2 * plane + j]);
        part[o] += (double)(wgt * g);
      }
    }
  }
#pragma unroll
  for (int o = 0; o < 12; ++o) block_accumulate(part[o], acc + o);
}

// L_TV (loss.py:139-152): acc[0] += (x[y+1]-x[y])^2, acc[1] += (x[x+1]-x[x])^2
__global__ void tv_kernel(const float* __restrict__ x, int C, int H, int W, double* __restrict__ acc) {
  const int64_t total = (int64_t)C * H * W;
  double h = 0.0, w = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int px = (int)(i % W);
    const int py = (int)((i / W) % H);
    const float v = x[i];
    if (py + 1 < H) { const float d = x[i + W] - v; h += (double)(d * d); }
    if (px + 1 < W) { const float d = x[i + 1] - v; w += (double)(d * d); }
  }
  block_accumulate(h, acc);
  block_accumulate(w, acc + 1);
}

__global__ void mse_sum_kernel(const float* __restrict__ a, const float* __restrict__ b, float add_a, int64_t n,
                               double* __restrict__ acc) {
  double s = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float d = (a[i] + add_a) - b[i];
    s += (double)(d * d);
  }
  block_accumulate(s, acc);
}

// Loss_inter (loss.py:68-73): 5x5 reflect local means, weighted_diff1/2 (quirk Q7), two MSE sums
__global__ void inter_kernel(const float* __restrict__ d1, const float* __restrict__ d2, const float* __restrict__ diff,
                             int C, int H, int W, double* __restrict__ acc) {
  const int64_t plane = (int64_t)H * W, total = (int64_t)C * plane;
  double a0 = 0.0, a1 = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i % plane;
    const int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    const float* b1 = d1 + (i - p);
    const float* b2 = d2 + (i - p);
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) {
      const int64_t q = (int64_t)reflect_idx(y + k / 5 - 2, H) * W + reflect_idx(x + k % 5 - 2, W);
      m1 += b1[q];
      m2 += b2[q];
    }
    m1 /= 25.f;
    m2 /= 25.f;
    const float dd = diff[p], v1 = d1[i], v2 = d2[i];
    const float w1 = (1.f - dd) * m1 + v1 * dd, w2 = (1.f - dd) * m2 + v1 * dd;
    a0 += (double)((v1 - w1) * (v1 - w1));
    a1 += (double)((v2 - w2) * (v2 - w2));
  }
  block_accumulate(a0, acc);
  block_accumulate(a1, acc + 1);
}

// 5x5 box mean with ZERO padding, dividing by 25 everywhere (AvgPool2d count_include_pad, Q9) of z = a - b
__global__ void box5_zero_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out,
                                 int C, int H, int W) {
  const int64_t plane = (int64_t)H * W, total = (int64_t)C * plane;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i % plane;
    const int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    const int64_t base = i - p;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) {
      const int yy = y + k / 5 - 2, xx = x + k % 5 - 2;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
        const int64_t q = base + (int64_t)yy * W + xx;
        s += b ? a[q] - b[q] : a[q];
      }
    }
    out[i] = s / 25.f;
  }
}
// Loss_Var (loss.py:75-77): var_k[p] = mean_{q in 5x5, zero pad}(z_k[q] - avg_k[q])^2, acc += (var1 - var2)^2
// with z1 = H2, z2 = H3 - H2
__global__ void var_mse_kernel(const float* __restrict__ H2, const float* __restrict__ H3, const float* __restrict__ avg1,
                               const float* __restrict__ avg2, int C, int H, int W, double* __restrict__ acc) {
  const int64_t plane = (int64_t)H * W, total = (int64_t)C * plane;
  double s = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i % plane;
    const int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    const int64_t base = i - p;
    float v1 = 0.f, v2 = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) {
      const int yy = y + k / 5 - 2, xx = x + k % 5 - 2;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
        const int64_t q = base + (int64_t)yy * W + xx;
        const float h2 = H2[q];
        const float e1 = h2 - avg1[q], e2 = (H3[q] - h2) - avg2[q];
        v1 += e1 * e1;
        v2 += e2 * e2;
      }
    }
    const float dv = v1 / 25.f - v2 / 25.f;
    s += (double)(dv * dv);
  }
  block_accumulate(s, acc);
}

}  // namespace zt

#define ZT_ST (cudaStream_t) stream
extern "C" int zt_pair_downsample_f32(const float* in, float* out1, float* out2, int C, int H, int W, float add,
                                      void* stream) {
  using namespace zt;
  ZT_REQUIRE(in && out1 && out2 && C > 0 && H >= 2 && W >= 2, "pair_downsample: bad args");
  const int64_t total = (int64_t)C * (H / 2) * (W / 2);
  pair_down_kernel<<<grid_for(total), 256, 0, ZT_ST>>>(in, out1, out2, C, H, W, add);
  return check_launch("pair_down_kernel");
}
extern "C" int zt_sub_clamp_f32(const float* a, float add, const float* d, int d_cs, float* out, int C, int64_t HW,
                                float lo, float hi, int do_clamp, void* stream) {
  using namespace zt;
  ZT_REQUIRE(a && d && out && C > 0 && HW > 0 && d_cs >= C, "sub_clamp: bad args");
  sub_clamp_kernel<<<grid_for((int64_t)C * HW), 256, 0, ZT_ST>>>(a, add, d, d_cs, out, C, HW, lo, hi, do_clamp);
  return check_launch("sub_clamp_kernel");
}
extern "C" int zt_div_clamp_f32(const float* a, float add, const float* b, float* out, int64_t n, float lo, float hi,
                                void* stream) {
  using namespace zt;
  ZT_REQUIRE(a && b && out && n > 0, "div_clamp: bad args");
  div_clamp_kernel<<<grid_for(n), 256, 0, ZT_ST>>>(a, add, b, out, n, lo, hi);
  return check_launch("div_clamp_kernel");
}
extern "C" int zt_texture_diff_f32(const float* img1, const float* img2, float* out, int H, int W, void* stream) {
  using namespace zt;
  ZT_REQUIRE(img1 && img2 && out && H >= 3 && W >= 3, "texture_diff: needs H,W >= 3 (reflect pad 2)");
  texture_diff_kernel<<<grid_for((int64_t)H * W, 128), 128, 0, ZT_ST>>>(img1, img2, out, H, W, 1e-5f, 0.975f);
  return check_launch("texture_diff_kernel");
}
/* taps: 21 host floats (sqrt(k)/sum sqrt(k), utils.py:29-39 with nsig=1); tmp: C*H*W scratch */
extern "C" int zt_blur21_f32(const float* in, float* tmp, float* out, const float* taps, int C, int H, int W,
                             void* stream) {
  using namespace zt;
  ZT_REQUIRE(in && tmp && out && taps && C > 0 && H > 10 && W > 10, "blur21: needs H,W > 10 (reflect pad 10)");
  BlurTaps t;
  for (int i = 0; i < 21; ++i) t.k[i] = taps[i];
  const int64_t total = (int64_t)C * H * W;
  blur21_kernel<<<grid_for(total), 256, 0, ZT_ST>>>(in, tmp, C, H, W, 0, t);
  if (int e = check_launch("blur21_kernel")) return e;
  blur21_kernel<<<grid_for(total), 256, 0, ZT_ST>>>(tmp, out, C, H, W, 1, t);
  return check_launch("blur21_kernel");
}
/* acc: device double[>=2] pre-zeroed; chansum: device double[3] scratch (zeroed inside) */
extern "C" int zt_loss_illum_f32(const float* L2, const float* s2, int is_WB, int64_t HW, double* chansum, double* acc,
                                 void* stream) {
  using namespace zt;
  ZT_REQUIRE(L2 && s2 && chansum && acc && HW > 0, "loss_illum: bad args");
  ZT_CUDA(cudaMemsetAsync(chansum, 0, 3 * sizeof(double), ZT_ST));
  const int gx = std::min(grid_for(HW), 4 * kNumSMs);
  chan_sum_kernel<<<dim3(gx, 3), 256, 0, ZT_ST>>>(L2, HW, chansum);
  if (int e = check_launch("chan_sum_kernel")) return e;
  illum_kernel<<<dim3(gx, 3), 256, 0, ZT_ST>>>(L2, s2, chansum, is_WB, HW, acc);
  return check_launch("illum_kernel");
}
/* SmoothLoss: ycc = C*H*W scratch; acc: double[12] pre-zeroed (one raw sum per unique offset) */
extern "C" int zt_loss_smooth_f32(const float* L2, const float* s2, float* ycc, int H, int W, double* acc,
                                  void* stream) {
  using namespace zt;
  ZT_REQUIRE(L2 && s2 && ycc && acc && H > 2 && W > 2, "loss_smooth: bad args");
  const int64_t plane = (int64_t)H * W;
  ycbcr_flat_kernel<<<grid_for(plane), 256, 0, ZT_ST>>>(L2, ycc, plane);
  if (int e = check_launch("ycbcr_flat_kernel")) return e;
  smooth_kernel<<<std::min(grid_for(plane), 8 * kNumSMs), 256, 0, ZT_ST>>>(ycc, s2, H, W, acc);
  return check_launch("smooth_kernel");
}
extern "C" int zt_loss_tv_f32(const float* x, int C, int H, int W, double* acc, void* stream) {
  using namespace zt;
  ZT_REQUIRE(x && acc && C > 0 && H > 1 && W > 1, "loss_tv: bad args");
  tv_kernel<<<std::min(grid_for((int64_t)C * H * W), 8 * kNumSMs), 256, 0, ZT_ST>>>(x, C, H, W, acc);
  return check_launch("tv_kernel");
}
/* acc[0] += sum((a + add_a - b)^2) */
extern "C" int zt_mse_sum_f32(const float* a, const float* b, float add_a, int64_t n, double* acc, void* stream) {
  using namespace zt;
  ZT_REQUIRE(a && b && acc && n > 0, "mse_sum: bad args");
  mse_sum_kernel<<<std::min(grid_for(n), 8 * kNumSMs), 256, 0, ZT_ST>>>(a, b, add_a, n, acc);
  return check_launch("mse_sum_kernel");
}
extern "C" int zt_loss_inter_f32(const float* d1, const float* d2, const float* diff, int C, int H, int W, double* acc,
                                 void* stream) {
  using namespace zt;
  ZT_REQUIRE(d1 && d2 && diff && acc && C > 0 && H >= 3 && W >= 3, "loss_inter: bad args");
  inter_kernel<<<std::min(grid_for((int64_t)C * H * W), 8 * kNumSMs), 256, 0, ZT_ST>>>(d1, d2, diff, C, H, W, acc);
  return check_launch("inter_kernel");
}
/* Loss_Var: avg1, avg2 = C*H*W scratch each; acc[0] += sum((var(H2) - var(H3-H2))^2) */
extern "C" int zt_loss_var_f32(const float* H2, const float* H3, float* avg1, float* avg2, int C, int H, int W,
                               double* acc, void* stream) {
  using namespace zt;
  ZT_REQUIRE(H2 && H3 && avg1 && avg2 && acc && C > 0 && H > 0 && W > 0, "loss_var: bad args");
  const int64_t total = (int64_t)C * H * W;
  box5_zero_kernel<<<grid_for(total), 256, 0, ZT_ST>>>(H2, nullptr, avg1, C, H, W);
  if (int e = check_launch("box5_zero_kernel")) return e;
  box5_zero_kernel<<<grid_for(total), 256, 0, ZT_ST>>>(H3, H2, avg2, C, H, W);
  if (int e = check_launch("box5_zero_kernel")) return e;
  var_mse_kernel<<<std::min(grid_for(total), 8 * kNumSMs), 256, 0, ZT_ST>>>(H2, H3, avg1, avg2, C, H, W, acc);
  return check_launch("var_mse_kernel");
}
