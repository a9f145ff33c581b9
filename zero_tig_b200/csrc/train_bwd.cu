// Backward kernels of the zero-shot training step (train.py:126-131): adjoints of the loss terms
// (loss.py:23-78,139-152,173-311), of the Retinex elementwise glue (model/model.py:144-203), of
// pair_downsampler / LocalMean / blur / local variance (utils/utils.py:15-79), train-mode BatchNorm
// backward, convolution weight gradients, and the fused global-norm clip + Adam update
// (train.py:98,130-131).  Gradients w.r.t. frame-level tensors are NCHW fp32 planes that the
// kernels ACCUMULATE into (the host zeroes them once per step); gradients inside a network are
// NHWC fp32 like the activations.  Convolution data gradients reuse the forward tensor-core
// convolution with spatially flipped, transposed weights (host side).
#include <math.h>
#include <algorithm>
#include "common.cuh"

namespace zt {

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_sum_f(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
inline int grid_bw(int64_t total, int threads = 256) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div64(total, threads), 16 * kNumSMs));
}
__device__ __forceinline__ float sgnf(float v) { return v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f); }

// pre-images of index j under F.pad(mode='reflect') with `pad` border: j itself, -j, 2(n-1)-j
__device__ __forceinline__ int reflect_pre(int j, int n, int pad, int* q) {
  int k = 0;
  q[k++] = j;
  if (j >= 1 && j <= pad) q[k++] = -j;
  if (j <= n - 2 && j >= n - 1 - pad) q[k++] = 2 * (n - 1) - j;
  return k;
}

// ------------------------------------------------------------------------------------------
// loss adjoints
// ------------------------------------------------------------------------------------------
// d = coef * ((a + add) - b);  ga += d;  gb -= d     (MSE: coef = weight * 2 / n)
__global__ void mse_bwd_kernel(const float* __restrict__ a, const float* __restrict__ b, float add, int64_t n,
                               float coef, float* __restrict__ ga, float* __restrict__ gb) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float d = coef * ((a[i] + add) - b[i]);
    if (ga) ga[i] += d;
    if (gb) gb[i] -= d;
  }
}

// loss.py:26-47: g_s2 += c0 * (s - bright) + c1 * (nl - adj) * d nl/d s   (targets are constants)
__global__ void illum_bwd_kernel(const float* __restrict__ L2, const float* __restrict__ s2,
                                 const double* __restrict__ chansum, int is_WB, int64_t HW, float c0, float c1,
                                 float* __restrict__ g_s2) {
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
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < HW; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t k = (int64_t)c * HW + i;
    const float l = L2[k], s = s2[k];
    const float q = l / s;
    const float nl = clampf(q, eps, 0.8f);
    const float bright = clampf(powf(l * ef, ef) * ratio, eps, 1.f);
    const float adj = clampf(l * ef, eps, 1.f);
    float g = c0 * (s - bright);
    if (q >= eps && q <= 0.8f) g += c1 * (nl - adj) * (-q / s);
    g_s2[k] += g;
  }
}

__constant__ int c_bw_dy[12] = {1, 0, 1, 1, 2, 0, 2, 2, 1, 1, 2, 2};
__constant__ int c_bw_dx[12] = {0, 1, 1, -1, 0, 2, 1, -1, 2, -2, 2, -2};
struct SmoothCoef { float k[12]; };
// SmoothLoss adjoint w.r.t. the illumination map (the luma weights come from detached L2):
// for every offset o: term = k_o * sum_p w_o(p) * sum_c |s[c,p] - s[c,p+o]|
__global__ void smooth_bwd_kernel(const float* __restrict__ y, const float* __restrict__ s2, int H, int W,
                                  const SmoothCoef kc, float* __restrict__ g_s2) {
  const int64_t plane = (int64_t)H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < plane; i += (int64_t)gridDim.x * blockDim.x) {
    const int py = (int)(i / W), px = (int)(i - (int64_t)py * W);
    const float y0 = y[i], y1 = y[plane + i], y2 = y[2 * plane + i];
    const float s0 = s2[i], s1 = s2[plane + i], s2v = s2[2 * plane + i];
    float g0 = 0.f, g1 = 0.f, g2 = 0.f;
#pragma unroll
    for (int o = 0; o < 12; ++o) {
#pragma unroll
      for (int dir = 0; dir < 2; ++dir) {
        // dir 0: this pixel is the first of the pair (p, p+o); dir 1: it is the second (p-o, p)
        const int qy = dir ? py - c_bw_dy[o] : py + c_bw_dy[o];
        const int qx = dir ? px - c_bw_dx[o] : px + c_bw_dx[o];
        if (qy >= 0 && qy < H && qx >= 0 && qx < W) {
          const int64_t j = (int64_t)qy * W + qx;
          const float d0 = y0 - y[j], d1 = y1 - y[plane + j], d2 = y2 - y[2 * plane + j];
          const float wgt = kc.k[o] * expf((d0 * d0 + d1 * d1 + d2 * d2) * (-1.0f / 200.0f));
          g0 += wgt * sgnf(s0 - s2[j]);
          g1 += wgt * sgnf(s1 - s2[plane + j]);
          g2 += wgt * sgnf(s2v - s2[2 * plane + j]);
        }
      }
    }
    g_s2[i] += g0;
    g_s2[plane + i] += g1;
    g_s2[2 * plane + i] += g2;
  }
}

// L_TV adjoint: g += ch * d/dx sum (x[y+1]-x[y])^2 + cw * d/dx sum (x[x+1]-x[x])^2
__global__ void tv_bwd_kernel(const float* __restrict__ x, int C, int H, int W, float ch, float cw,
                              float* __restrict__ g) {
  const int64_t total = (int64_t)C * H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int px = (int)(i % W);
    const int py = (int)((i / W) % H);
    const float v = x[i];
    float a = 0.f, b = 0.f;
    if (py > 0) a += v - x[i - W];
    if (py + 1 < H) a -= x[i + W] - v;
    if (px > 0) b += v - x[i - 1];
    if (px + 1 < W) b -= x[i + 1] - v;
    g[i] += 2.f * (ch * a + cw * b);
  }
}

// Loss_inter adjoint, pass A (loss.py:66-73):  e1 = (1-D)(v1 - m1), e2 = v2 - (1-D) m2 - D v1
//   u1 = (1-D) * 2c e1, u2 = (1-D) * 2c e2;  g_d1 += u1 - D * 2c e2;  g_d2 += 2c e2
__global__ void inter_bwd_a_kernel(const float* __restrict__ d1, const float* __restrict__ d2,
                                   const float* __restrict__ diff, int C, int H, int W, float coef,
                                   float* __restrict__ u1, float* __restrict__ u2, float* __restrict__ g1,
                                   float* __restrict__ g2) {
  const int64_t plane = (int64_t)H * W, total = (int64_t)C * plane;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i % plane;
    const int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    const float* b1 = d1 + (i - p);
    const float* b2 = d2 + (i - p);
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) {
      int yy = y + k / 5 - 2, xx = x + k % 5 - 2;
      yy = yy < 0 ? -yy : (yy >= H ? 2 * H - 2 - yy : yy);
      xx = xx < 0 ? -xx : (xx >= W ? 2 * W - 2 - xx : xx);
      const int64_t q = (int64_t)yy * W + xx;
      m1 += b1[q];
      m2 += b2[q];
    }
    m1 /= 25.f;
    m2 /= 25.f;
    const float dd = diff[p], v1 = d1[i], v2 = d2[i];
    const float e1 = v1 - ((1.f - dd) * m1 + v1 * dd), e2 = v2 - ((1.f - dd) * m2 + v1 * dd);
    const float ge1 = coef * e1, ge2 = coef * e2;
    const float a1 = (1.f - dd) * ge1, a2 = (1.f - dd) * ge2;
    u1[i] = a1;
    u2[i] = a2;
    g1[i] += a1 - dd * ge2;
    g2[i] += ge2;
  }
}
// pass B: g -= LocalMean^T(u)  (adjoint of the 5x5 mean over a reflect-padded image)
__global__ void localmean_adj_sub_kernel(const float* __restrict__ u, int C, int H, int W, float* __restrict__ g) {
  const int64_t plane = (int64_t)H * W, total = (int64_t)C * plane;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i % plane;
    const int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    const float* b = u + (i - p);
    int qy[3], qx[3];
    const int ny = reflect_pre(y, H, 2, qy), nx = reflect_pre(x, W, 2, qx);
    float s = 0.f;
    for (int a = 0; a < ny; ++a)
      for (int c = 0; c < nx; ++c)
#pragma unroll
        for (int k = 0; k < 25; ++k) {
          // padded position (qy,qx) is read by the windows centred at (qy-dy, qx-dx)
          const int yy = qy[a] - (k / 5 - 2), xx = qx[c] - (k % 5 - 2);
          if (yy >= 0 && yy < H && xx >= 0 && xx < W) s += b[(int64_t)yy * W + xx];
        }
    g[i] -= s / 25.f;
  }
}

// Loss_Var (loss.py:75-77) adjoint.  v_k = box5((z_k - box5 z_k)^2), loss = c * sum (v1 - v2)^2,
// z1 = H2, z2 = H3 - H2; box5 = zero-padded 5x5 mean / 25 (self-adjoint).
// step 1: dv[p] = coef * (v1 - v2)
__global__ void var_dv_kernel(const float* __restrict__ H2, const float* __restrict__ H3, const float* __restrict__ avg1,
                              const float* __restrict__ avg2, int C, int H, int W, float coef, float* __restrict__ dv) {
  const int64_t plane = (int64_t)H * W, total = (int64_t)C * plane;
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
    dv[i] = coef * (v1 / 25.f - v2 / 25.f);
  }
}
// step 2: gr1 = 2 r1 box5(dv), gr2 = -2 r2 box5(dv)   (r_k = z_k - avg_k)
__global__ void var_gr_kernel(const float* __restrict__ H2, const float* __restrict__ H3, const float* __restrict__ avg1,
                              const float* __restrict__ avg2, const float* __restrict__ dv, int C, int H, int W,
                              float* __restrict__ gr1, float* __restrict__ gr2) {
  const int64_t plane = (int64_t)H * W, total = (int64_t)C * plane;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i % plane;
    const int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    const int64_t base = i - p;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) {
      const int yy = y + k / 5 - 2, xx = x + k % 5 - 2;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) s += dv[base + (int64_t)yy * W + xx];
    }
    s /= 25.f;
    const float h2 = H2[i];
    gr1[i] = 2.f * (h2 - avg1[i]) * s;
    gr2[i] = -2.f * ((H3[i] - h2) - avg2[i]) * s;
  }
}
// step 3: gz_k = gr_k - box5(gr_k);  g_H2 += gz1 - gz2;  g_H3 += gz2
__global__ void var_gz_kernel(const float* __restrict__ gr1, const float* __restrict__ gr2, int C, int H, int W,
                              float* __restrict__ g_H2, float* __restrict__ g_H3) {
  const int64_t plane = (int64_t)H * W, total = (int64_t)C * plane;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i % plane;
    const int y = (int)(p / W), x = (int)(p - (int64_t)y * W);
    const int64_t base = i - p;
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < 25; ++k) {
      const int yy = y + k / 5 - 2, xx = x + k % 5 - 2;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
        const int64_t q = base + (int64_t)yy * W + xx;
        s1 += gr1[q];
        s2 += gr2[q];
      }
    }
    const float z1 = gr1[i] - s1 / 25.f, z2 = gr2[i] - s2 / 25.f;
    g_H2[i] += z1 - z2;
    g_H3[i] += z2;
  }
}

// adjoint of one pass of the reflect-padded 21-tap filter (utils.py:52-58); dir 0 along x, 1 along y.
// accumulate != 0: out += result
struct BlurTapsB { float k[21]; };
__global__ void blur21_adj_kernel(const float* __restrict__ in, float* __restrict__ out, int C, int H, int W, int dir,
                                  const BlurTapsB t, int accumulate) {
  const int64_t total = (int64_t)C * H * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % W);
    const int64_t r = i / W;
    const int y = (int)(r % H);
    const float* base = in + (r - y) * W;
    const int n = dir ? H : W, j = dir ? y : x;
    int q[3];
    const int nq = reflect_pre(j, n, 10, q);
    float acc = 0.f;
    for (int a = 0; a < nq; ++a)
#pragma unroll
      for (int k = 0; k < 21; ++k) {
        const int src = q[a] - (k - 10);   // output position whose tap k reads padded index q[a]
        if (src >= 0 && src < n) acc += t.k[k] * (dir ? base[(int64_t)src * W + x] : base[(int64_t)y * W + src]);
      }
    out[i] = accumulate ? out[i] + acc : acc;
  }
}

// ------------------------------------------------------------------------------------------
// Retinex glue adjoints
// ------------------------------------------------------------------------------------------
// pair_downsampler^T: gx[2y][2x+1], gx[2y+1][2x] += g1/2;  gx[2y][2x], gx[2y+1][2x+1] += g2/2
__global__ void pair_down_bwd_kernel(const float* __restrict__ g1, const float* __restrict__ g2, int C, int H, int W,
                                     float* __restrict__ gx) {
  const int ho = H >> 1, wo = W >> 1;
  const int64_t total = (int64_t)C * ho * wo;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % wo);
    const int64_t r = i / wo;
    const int y = (int)(r % ho), c = (int)(r / ho);
    float* p = gx + ((int64_t)c * H + 2 * y) * W + 2 * x;
    const float a = 0.5f * g1[i], b = 0.5f * g2[i];
    p[0] += b;
    p[1] += a;
    p[W] += a;
    p[W + 1] += b;
  }
}
// out = clamp((a + add) / b, lo, hi):  g_b += g_out * [lo <= q <= hi] * (-(a + add) / b^2)
__global__ void div_clamp_bwd_kernel(const float* __restrict__ a, float add, const float* __restrict__ b,
                                     const float* __restrict__ g_out, int64_t n, float lo, float hi,
                                     float* __restrict__ g_b) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float bv = b[i], q = (a[i] + add) / bv;
    if (q >= lo && q <= hi) g_b[i] += g_out[i] * (-q / bv);
  }
}
// NHWC gradient of a network output from NCHW plane gradients:
//   out[p][k*c_each + c] = sign * g_k[c][p] * [lo < o_k[c][p] < hi]   (mask only when o_k given)
// (torch.clamp passes the gradient where the un-clamped value lies inside the bounds; an output
// strictly inside them is exactly that case)
__global__ void mask_pack_kernel(const float* __restrict__ g0, const float* __restrict__ o0,
                                 const float* __restrict__ g1, const float* __restrict__ o1, int c_each, int nsrc,
                                 float sign, float lo, float hi, float* __restrict__ out, int cs, int64_t HW) {
  const int cpk = c_each * nsrc;
  const int64_t total = HW * cs;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % cs);
    const int64_t p = i / cs;
    float v = 0.f;
    if (c < cpk) {
      const int k = c / c_each, cc = c - k * c_each;
      const float* g = k ? g1 : g0;
      const float* o = k ? o1 : o0;
      const int64_t q = (int64_t)cc * HW + p;
      v = sign * g[q];
      if (o) {
        const float ov = o[q];
        if (!(ov > lo && ov < hi)) v = 0.f;
      }
    }
    out[i] = v;
  }
}
// dst_k[c][p] += g[p][c_off + k*c_each + c]
__global__ void unpack_add_kernel(const float* __restrict__ g, int cs, int c_off, int c_each, int nd,
                                  float* __restrict__ d0, float* __restrict__ d1, int64_t HW) {
  const int64_t total = (int64_t)nd * c_each * HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t p = i % HW;
    const int kc = (int)(i / HW);
    const int k = kc / c_each, c = kc - k * c_each;
    float* d = k ? d1 : d0;
    d[(int64_t)c * HW + p] += g[p * cs + c_off + kc];
  }
}
// g[m][c] *= act'(y[m][c])   (derivative expressed through the activation OUTPUT y)
__global__ void act_bwd_kernel(float* __restrict__ g, int g_cs, const float* __restrict__ y, int y_cs, int C, int act,
                               int64_t M) {
  const int c4n = C >> 2;
  const int64_t total = M * c4n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c4n);
    const int64_t m = i / c4n;
    float4 gv = *reinterpret_cast<float4*>(g + m * g_cs + cg * 4);
    const float4 yv = __ldg(reinterpret_cast<const float4*>(y + m * y_cs + cg * 4));
    float ga[4] = {gv.x, gv.y, gv.z, gv.w};
    const float ya[4] = {yv.x, yv.y, yv.z, yv.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float d = 1.f;
      if (act == ZT_ACT_RELU) d = ya[e] > 0.f ? 1.f : 0.f;
      else if (act == ZT_ACT_LRELU02) d = ya[e] > 0.f ? 1.f : 0.2f;
      else if (act == ZT_ACT_SIGMOID) d = ya[e] * (1.f - ya[e]);
      else if (act == ZT_ACT_TANH) d = 1.f - ya[e] * ya[e];
      ga[e] *= d;
    }
    *reinterpret_cast<float4*>(g + m * g_cs + cg * 4) = make_float4(ga[0], ga[1], ga[2], ga[3]);
  }
}

// ------------------------------------------------------------------------------------------
// train-mode BatchNorm + ReLU + residual backward:  out = res + relu(gamma * xhat + beta)
// ------------------------------------------------------------------------------------------
// pass 1: per channel sums of dy and dy * xhat (dy = dout * [y > 0]) -> red[c*2], red[c*2+1] (fp64)
__global__ void bn_bwd_reduce_kernel(const float* __restrict__ x, int x_cs, const float* __restrict__ dout, int d_cs,
                                     const double* __restrict__ stats, const float* __restrict__ gamma,
                                     const float* __restrict__ beta, int HW, int C, float eps, int ppb,
                                     double* __restrict__ red) {
  // thread t owns channel t % C; rows strided by blockDim.x / C
  const int c = threadIdx.x % C, r0 = threadIdx.x / C, rstep = blockDim.x / C;
  const double mean = stats[c * 2] / HW;
  double var = stats[c * 2 + 1] / HW - mean * mean;
  var = var < 0 ? 0 : var;
  const float rstd = (float)(1.0 / sqrt(var + (double)eps)), mu = (float)mean;
  const float ga = gamma[c], be = beta[c];
  const int p0 = blockIdx.x * ppb, p1 = r0 >= rstep ? 0 : min(HW, p0 + ppb);
  double s0 = 0.0, s1 = 0.0;
  for (int p = p0 + r0; p < p1; p += rstep) {
    const float xh = (x[(int64_t)p * x_cs + c] - mu) * rstd;
    const float y = xh * ga + be;
    const float dy = y > 0.f ? dout[(int64_t)p * d_cs + c] : 0.f;
    s0 += (double)dy;
    s1 += (double)(dy * xh);
  }
  // block-level reduction first: hundreds of blocks adding fp64 values into the same 2*C addresses serialise in L2
  __shared__ double rs[256][2];
  rs[threadIdx.x][0] = s0;
  rs[threadIdx.x][1] = s1;
  __syncthreads();
  if (r0 == 0) {
    for (int r = 1; r < rstep; ++r) { s0 += rs[r * C + c][0]; s1 += rs[r * C + c][1]; }
    atomicAdd(red + c * 2, s0);
    atomicAdd(red + c * 2 + 1, s1);
  }
}
// pass 2: dx = gamma * rstd * (dy - mean(dy) - xhat * mean(dy * xhat))
__global__ void bn_bwd_apply_kernel(const float* __restrict__ x, int x_cs, const float* __restrict__ dout, int d_cs,
                                    const double* __restrict__ stats, const double* __restrict__ red,
                                    const float* __restrict__ gamma, const float* __restrict__ beta, int HW, int C,
                                    float eps, float* __restrict__ dx, int dx_cs, int64_t total) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int64_t p = i / C;
    const double mean = stats[c * 2] / HW;
    double var = stats[c * 2 + 1] / HW - mean * mean;
    var = var < 0 ? 0 : var;
    const float rstd = (float)(1.0 / sqrt(var + (double)eps));
    const float xh = (x[p * x_cs + c] - (float)mean) * rstd;
    const float ga = gamma[c];
    const float y = xh * ga + beta[c];
    const float dy = y > 0.f ? dout[p * d_cs + c] : 0.f;
    const float m0 = (float)(red[c * 2] / HW), m1 = (float)(red[c * 2 + 1] / HW);
    dx[p * dx_cs + c] = ga * rstd * (dy - m0 - xh * m1);
  }
}
// Vectorised forms of the two passes for C % 4 == 0 (every BatchNorm of the model: C = 64).  The scalar kernels above
// evaluate the fp64 mean / variance / 1/sqrt chain per ELEMENT (the apply pass spent its time in fp64 division and
// square roots: 550 us for a 960x540x64 tensor that moves 670 MB); here the per-channel constants are computed once
// per thread (reduce) or once per block into shared memory (apply) - the same fp64 expressions, so the same fp32
// constants and the same per-element fp32 arithmetic - and the tensor is streamed as float4.
__global__ void __launch_bounds__(256) bn_bwd_reduce4_kernel(const float* __restrict__ x, int x_cs,
                                                             const float* __restrict__ dout, int d_cs,
                                                             const double* __restrict__ stats,
                                                             const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, int HW, int C, float eps,
                                                             int ppb, double* __restrict__ red) {
  const int c4n = C >> 2, lanes = 256 / c4n;
  const int cg = threadIdx.x % c4n, pl = threadIdx.x / c4n;
  float mu[4], rstd[4], ga[4], be[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int c = cg * 4 + e;
    const double mean = stats[c * 2] / HW;
    double var = stats[c * 2 + 1] / HW - mean * mean;
    var = var < 0 ? 0 : var;
    rstd[e] = (float)(1.0 / sqrt(var + (double)eps));
    mu[e] = (float)mean;
    ga[e] = gamma[c];
    be[e] = beta[c];
  }
  const int p0 = blockIdx.x * ppb, p1 = pl < lanes ? min(HW, p0 + ppb) : 0;
  double s0[4] = {0, 0, 0, 0}, s1[4] = {0, 0, 0, 0};
  int p = p0 + pl;
  for (; p + lanes < p1; p += 2 * lanes) {                       // two pixels in flight per thread
    const float4 xa = __ldg(reinterpret_cast<const float4*>(x + (int64_t)p * x_cs + cg * 4));
    const float4 da = __ldg(reinterpret_cast<const float4*>(dout + (int64_t)p * d_cs + cg * 4));
    const float4 xb = __ldg(reinterpret_cast<const float4*>(x + (int64_t)(p + lanes) * x_cs + cg * 4));
    const float4 db = __ldg(reinterpret_cast<const float4*>(dout + (int64_t)(p + lanes) * d_cs + cg * 4));
    const float xv[2][4] = {{xa.x, xa.y, xa.z, xa.w}, {xb.x, xb.y, xb.z, xb.w}};
    const float dv[2][4] = {{da.x, da.y, da.z, da.w}, {db.x, db.y, db.z, db.w}};
#pragma unroll
    for (int u = 0; u < 2; ++u)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float xh = (xv[u][e] - mu[e]) * rstd[e];
        const float y = xh * ga[e] + be[e];
        const float dy = y > 0.f ? dv[u][e] : 0.f;
        s0[e] += (double)dy;
        s1[e] += (double)(dy * xh);
      }
  }
  for (; p < p1; p += lanes) {
    const float4 xa = __ldg(reinterpret_cast<const float4*>(x + (int64_t)p * x_cs + cg * 4));
    const float4 da = __ldg(reinterpret_cast<const float4*>(dout + (int64_t)p * d_cs + cg * 4));
    const float xv[4] = {xa.x, xa.y, xa.z, xa.w}, dv[4] = {da.x, da.y, da.z, da.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float xh = (xv[e] - mu[e]) * rstd[e];
      const float y = xh * ga[e] + be[e];
      const float dy = y > 0.f ? dv[e] : 0.f;
      s0[e] += (double)dy;
      s1[e] += (double)(dy * xh);
    }
  }
  __shared__ double rs[256 * 8];                                 // [thread][s0 x4 | s1 x4]
#pragma unroll
  for (int e = 0; e < 4; ++e) { rs[threadIdx.x * 8 + e] = s0[e]; rs[threadIdx.x * 8 + 4 + e] = s1[e]; }
  __syncthreads();
  for (int j = threadIdx.x; j < c4n * 8; j += 256) {
    double acc = 0;
    for (int l = 0; l < lanes; ++l) acc += rs[l * c4n * 8 + j];
    const int g4 = j >> 3, e = j & 7;
    atomicAdd(red + (g4 * 4 + (e & 3)) * 2 + (e >> 2), acc);
  }
}
__global__ void __launch_bounds__(256) bn_bwd_apply4_kernel(const float* __restrict__ x, int x_cs,
                                                            const float* __restrict__ dout, int d_cs,
                                                            const double* __restrict__ stats,
                                                            const double* __restrict__ red,
                                                            const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, int HW, int C, float eps,
                                                            float* __restrict__ dx, int dx_cs, int64_t total4) {
  __shared__ float cst[256][6];                                  // mu, rstd, gamma, beta, mean(dy), mean(dy * xhat)
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const double mean = stats[c * 2] / HW;
    double var = stats[c * 2 + 1] / HW - mean * mean;
    var = var < 0 ? 0 : var;
    cst[c][0] = (float)mean;
    cst[c][1] = (float)(1.0 / sqrt(var + (double)eps));
    cst[c][2] = gamma[c];
    cst[c][3] = beta[c];
    cst[c][4] = (float)(red[c * 2] / HW);
    cst[c][5] = (float)(red[c * 2 + 1] / HW);
  }
  __syncthreads();
  const int c4n = C >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c4n);
    const int64_t p = i / c4n;
    const float4 xa = __ldg(reinterpret_cast<const float4*>(x + p * x_cs + cg * 4));
    const float4 da = __ldg(reinterpret_cast<const float4*>(dout + p * d_cs + cg * 4));
    const float xv[4] = {xa.x, xa.y, xa.z, xa.w}, dv[4] = {da.x, da.y, da.z, da.w};
    float r[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float* k = cst[cg * 4 + e];
      const float xh = (xv[e] - k[0]) * k[1];
      const float y = xh * k[2] + k[3];
      const float dy = y > 0.f ? dv[e] : 0.f;
      r[e] = k[2] * k[1] * (dy - k[4] - xh * k[5]);
    }
    *reinterpret_cast<float4*>(dx + p * dx_cs + cg * 4) = make_float4(r[0], r[1], r[2], r[3]);
  }
}
// dgamma[c] += sum dy xhat; dbeta[c] += sum dy
__global__ void bn_bwd_param_kernel(const double* __restrict__ red, int C, float* __restrict__ dgamma,
                                    float* __restrict__ dbeta) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  dgamma[c] += (float)red[c * 2 + 1];
  dbeta[c] += (float)red[c * 2];
}

// ------------------------------------------------------------------------------------------
// convolution weight / bias gradients (NHWC fp32 activations, stride 1)
//   dW[ky][kx][ci][co] += sum_p X[p + (ky - ph, kx - pw)][ci] * G[p][co]
// One CTA owns one filter row ky and walks (row, 64-pixel segment) tiles: the input segment (+ kw-1
// halo) and the gradient segment are staged in shared memory and every thread keeps a
// [KW][4 ci][4 co] register tile.  Thin layers (few ci or co tiles) split the pixels of a segment
// over PS thread groups.  Partial sums leave through fp32 atomics once per CTA.
// ------------------------------------------------------------------------------------------
constexpr int WG_PX = 64;
template <int KW>
__global__ void __launch_bounds__(256) wgrad_kernel(const float* __restrict__ X, int x_cs, const float* __restrict__ G,
                                                    int g_cs, float* __restrict__ dW, int H, int W, int ph, int pw,
                                                    int cin, int cout, int w_cs, int TI, int TO, int PS) {
  extern __shared__ float sm[];
  const int ci_n = TI * 4, co_n = TO * 4;           // staged channel counts (cin, cout rounded up to 4)
  float* Xs = sm;                                   // [WG_PX + KW - 1][ci_n]
  float* Gs = sm + (WG_PX + KW - 1) * ci_n;         // [WG_PX][co_n]
  const int ky = blockIdx.y;
  const int tid = threadIdx.x;
  const int nthr = TI * TO * PS;
  const int to = tid % TO, ti = (tid / TO) % TI, ps = tid / (TO * TI);
  const bool active = tid < nthr;
  float acc[KW][4][4];
#pragma unroll
  for (int k = 0; k < KW; ++k)
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) acc[k][a][b] = 0.f;
  const int segs = ceil_div(W, WG_PX);
  const int tiles = H * segs;
  for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int y = t / segs, x0 = (t - y * segs) * WG_PX;
    const int yi = y + ky - ph;
    if (yi < 0 || yi >= H) continue;                // whole input row is padding: contributes nothing
    __syncthreads();
    // stage G[y][x0 .. x0+PX)[co] and X[yi][x0-pw .. x0+PX-1+pw][ci]
    const int c4g = co_n >> 2, c4x = ci_n >> 2;
    for (int i = tid; i < WG_PX * c4g; i += blockDim.x) {
      const int p = i / c4g, cg = i - p * c4g;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (x0 + p < W) v = __ldg(reinterpret_cast<const float4*>(G + ((int64_t)y * W + x0 + p) * g_cs + cg * 4));
      *reinterpret_cast<float4*>(Gs + p * co_n + cg * 4) = v;
    }
    for (int i = tid; i < (WG_PX + KW - 1) * c4x; i += blockDim.x) {
      const int p = i / c4x, cg = i - p * c4x;
      const int xi = x0 + p - pw;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (xi >= 0 && xi < W) v = __ldg(reinterpret_cast<const float4*>(X + ((int64_t)yi * W + xi) * x_cs + cg * 4));
      *reinterpret_cast<float4*>(Xs + p * ci_n + cg * 4) = v;
    }
    __syncthreads();
    if (active) {
#pragma unroll 4
      for (int p = ps; p < WG_PX; p += PS) {
        const float4 b = *reinterpret_cast<const float4*>(Gs + p * co_n + to * 4);
        const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < KW; ++k) {
          const float4 a = *reinterpret_cast<const float4*>(Xs + (p + k) * ci_n + ti * 4);
          const float aa[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
          for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int v = 0; v < 4; ++v) acc[k][u][v] = fmaf(aa[u], bb[v], acc[k][u][v]);
        }
      }
    }
  }
  if (active) {
#pragma unroll
    for (int k = 0; k < KW; ++k)
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int ci = ti * 4 + u;
        if (ci >= cin) continue;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const int co = to * 4 + v;
          if (co < cout && acc[k][u][v] != 0.f)
            atomicAdd(dW + (((int64_t)ky * KW + k) * cin + ci) * w_cs + co, acc[k][u][v]);
        }
      }
  }
}
// out[c] += sum_m g[m][c]
__global__ void colsum_kernel(const float* __restrict__ g, int cs, int C, int64_t M, int rows_per_block,
                              float* __restrict__ out) {
  const int c = threadIdx.x % C, r0 = threadIdx.x / C, rstep = blockDim.x / C;
  if (r0 >= rstep) return;
  const int64_t m0 = (int64_t)blockIdx.x * rows_per_block, m1 = (m0 + rows_per_block < M) ? m0 + rows_per_block : M;
  float s = 0.f;
  for (int64_t m = m0 + r0; m < m1; m += rstep) s += g[m * cs + c];
  if (s != 0.f) atomicAdd(out + c, s);
}

// ------------------------------------------------------------------------------------------
// fused global-norm clip + Adam (train.py:98,130-131: clip_grad_norm_(5), Adam lr 1e-4,
// betas (0.9, 0.999), eps 1e-8, weight_decay 3e-4 added to the gradient - not AdamW)
// ------------------------------------------------------------------------------------------
__global__ void sumsq_kernel(const float* __restrict__ g, int64_t n, double* __restrict__ acc) {
  double s = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    s += (double)g[i] * (double)g[i];
  s = warp_sum_d(s);
  if ((threadIdx.x & 31) == 0 && s != 0.0) atomicAdd(acc, s);
}
__global__ void clip_adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                                 float* __restrict__ v, int64_t n, const double* __restrict__ sumsq, float grad_scale,
                                 float max_norm, float lr, float b1, float b2, float eps, float wd, float bc1,
                                 float bc2_sqrt) {
  // clip_grad_norm_: coef = max_norm / (norm + 1e-6), clamped to 1; applied in place to the grads
  const float norm = (float)sqrt(*sumsq) * grad_scale;
  float coef = max_norm > 0.f ? max_norm / (norm + 1e-6f) : 1.f;
  coef = fminf(coef, 1.f) * grad_scale;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gc = g[i] * coef;
    g[i] = gc;
    const float pv = p[i];
    const float gd = gc + wd * pv;
    const float mi = b1 * m[i] + (1.f - b1) * gd;
    const float vi = b2 * v[i] + (1.f - b2) * gd * gd;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = pv - (lr / bc1) * (mi / denom);
  }
}

}  // namespace zt

#define ZT_ST (cudaStream_t) stream
extern "C" int zt_mse_bwd_f32(const float* a, const float* b, float add_a, int64_t n, float coef, float* ga, float* gb,
                              void* stream) {
  using namespace zt;
  ZT_REQUIRE(a && b && n > 0 && (ga || gb), "mse_bwd: bad args");
  mse_bwd_kernel<<<grid_bw(n), 256, 0, ZT_ST>>>(a, b, add_a, n, coef, ga, gb);
  return check_launch("mse_bwd_kernel");
}
/* chansum: the per-channel sums of L2 left by zt_loss_illum_f32 */
extern "C" int zt_loss_illum_bwd_f32(const float* L2, const float* s2, int is_WB, int64_t HW, const double* chansum,
                                     float c0, float c1, float* g_s2, void* stream) {
  using namespace zt;
  ZT_REQUIRE(L2 && s2 && chansum && g_s2 && HW > 0, "loss_illum_bwd: bad args");
  illum_bwd_kernel<<<dim3(std::min(grid_bw(HW), 4 * kNumSMs), 3), 256, 0, ZT_ST>>>(L2, s2, chansum, is_WB, HW, c0, c1,
                                                                                 g_s2);
  return check_launch("illum_bwd_kernel");
}
/* ycc: the flat-triple YCbCr planes left by zt_loss_smooth_f32; coef[12] = weight * 2 / count per offset (host) */
extern "C" int zt_loss_smooth_bwd_f32(const float* ycc, const float* s2, int H, int W, const float* coef, float* g_s2,
                                      void* stream) {
  using namespace zt;
  ZT_REQUIRE(ycc && s2 && coef && g_s2 && H > 2 && W > 2, "loss_smooth_bwd: bad args");
  SmoothCoef kc;
  for (int i = 0; i < 12; ++i) kc.k[i] = coef[i];
  smooth_bwd_kernel<<<grid_bw((int64_t)H * W, 128), 128, 0, ZT_ST>>>(ycc, s2, H, W, kc, g_s2);
  return check_launch("smooth_bwd_kernel");
}
extern "C" int zt_loss_tv_bwd_f32(const float* x, int C, int H, int W, float ch, float cw, float* g, void* stream) {
  using namespace zt;
  ZT_REQUIRE(x && g && C > 0 && H > 1 && W > 1, "loss_tv_bwd: bad args");
  tv_bwd_kernel<<<grid_bw((int64_t)C * H * W), 256, 0, ZT_ST>>>(x, C, H, W, ch, cw, g);
  return check_launch("tv_bwd_kernel");
}
/* u1, u2: C*H*W scratch; coef = weight * 2 / n */
extern "C" int zt_loss_inter_bwd_f32(const float* d1, const float* d2, const float* diff, int C, int H, int W,
                                     float coef, float* u1, float* u2, float* g_d1, float* g_d2, void* stream) {
  using namespace zt;
  ZT_REQUIRE(d1 && d2 && diff && u1 && u2 && g_d1 && g_d2 && C > 0 && H >= 5 && W >= 5, "loss_inter_bwd: bad args");
  const int64_t total = (int64_t)C * H * W;
  inter_bwd_a_kernel<<<grid_bw(total), 256, 0, ZT_ST>>>(d1, d2, diff, C, H, W, coef, u1, u2, g_d1, g_d2);
  if (int e = check_launch("inter_bwd_a_kernel")) return e;
  localmean_adj_sub_kernel<<<grid_bw(total), 256, 0, ZT_ST>>>(u1, C, H, W, g_d1);
  if (int e = check_launch("localmean_adj_sub_kernel")) return e;
  localmean_adj_sub_kernel<<<grid_bw(total), 256, 0, ZT_ST>>>(u2, C, H, W, g_d2);
  return check_launch("localmean_adj_sub_kernel");
}
/* avg1, avg2: the box means left by zt_loss_var_f32; t0,t1,t2: C*H*W scratch; coef = weight * 2 / n */
extern "C" int zt_loss_var_bwd_f32(const float* H2, const float* H3, const float* avg1, const float* avg2, int C, int H,
                                   int W, float coef, float* t0, float* t1, float* t2, float* g_H2, float* g_H3,
                                   void* stream) {
  using namespace zt;
  ZT_REQUIRE(H2 && H3 && avg1 && avg2 && t0 && t1 && t2 && g_H2 && g_H3 && C > 0 && H > 0 && W > 0,
             "loss_var_bwd: bad args");
  const int64_t total = (int64_t)C * H * W;
  var_dv_kernel<<<grid_bw(total), 256, 0, ZT_ST>>>(H2, H3, avg1, avg2, C, H, W, coef, t0);
  if (int e = check_launch("var_dv_kernel")) return e;
  var_gr_kernel<<<grid_bw(total), 256, 0, ZT_ST>>>(H2, H3, avg1, avg2, t0, C, H, W, t1, t2);
  if (int e = check_launch("var_gr_kernel")) return e;
  var_gz_kernel<<<grid_bw(total), 256, 0, ZT_ST>>>(t1, t2, C, H, W, g_H2, g_H3);
  return check_launch("var_gz_kernel");
}
/* g_out += blur^T(g_in)  (tmp: C*H*W scratch) */
extern "C" int zt_blur21_bwd_f32(const float* g_in, float* tmp, float* g_out, const float* taps, int C, int H, int W,
                                 void* stream) {
  using namespace zt;
  ZT_REQUIRE(g_in && tmp && g_out && taps && C > 0 && H > 10 && W > 10, "blur21_bwd: needs H,W > 10");
  BlurTapsB t;
  for (int i = 0; i < 21; ++i) t.k[i] = taps[i];
  const int64_t total = (int64_t)C * H * W;
  blur21_adj_kernel<<<grid_bw(total), 256, 0, ZT_ST>>>(g_in, tmp, C, H, W, 1, t, 0);
  if (int e = check_launch("blur21_adj_kernel")) return e;
  blur21_adj_kernel<<<grid_bw(total), 256, 0, ZT_ST>>>(tmp, g_out, C, H, W, 0, t, 1);
  return check_launch("blur21_adj_kernel");
}
extern "C" int zt_pair_downsample_bwd_f32(const float* g1, const float* g2, int C, int H, int W, float* gx,
                                          void* stream) {
  using namespace zt;
  ZT_REQUIRE(g1 && g2 && gx && C > 0 && H >= 2 && W >= 2, "pair_downsample_bwd: bad args");
  pair_down_bwd_kernel<<<grid_bw((int64_t)C * (H / 2) * (W / 2)), 256, 0, ZT_ST>>>(g1, g2, C, H, W, gx);
  return check_launch("pair_down_bwd_kernel");
}
extern "C" int zt_div_clamp_bwd_f32(const float* a, float add, const float* b, const float* g_out, int64_t n, float lo,
                                    float hi, float* g_b, void* stream) {
  using namespace zt;
  ZT_REQUIRE(a && b && g_out && g_b && n > 0, "div_clamp_bwd: bad args");
  div_clamp_bwd_kernel<<<grid_bw(n), 256, 0, ZT_ST>>>(a, add, b, g_out, n, lo, hi, g_b);
  return check_launch("div_clamp_bwd_kernel");
}
extern "C" int zt_mask_pack_nhwc_f32(const float* g0, const float* o0, const float* g1, const float* o1, int c_each,
                                     int nsrc, float sign, float lo, float hi, float* out, int cs, int64_t HW,
                                     void* stream) {
  using namespace zt;
  ZT_REQUIRE(g0 && out && nsrc >= 1 && nsrc <= 2 && (nsrc == 1 || g1) && c_each * nsrc <= cs && HW > 0,
             "mask_pack: bad args");
  mask_pack_kernel<<<grid_bw(HW * cs), 256, 0, ZT_ST>>>(g0, o0, g1, o1, c_each, nsrc, sign, lo, hi, out, cs, HW);
  return check_launch("mask_pack_kernel");
}
extern "C" int zt_unpack_add_nchw_f32(const float* g, int cs, int c_off, int c_each, int nd, float* d0, float* d1,
                                      int64_t HW, void* stream) {
  using namespace zt;
  ZT_REQUIRE(g && d0 && nd >= 1 && nd <= 2 && (nd == 1 || d1) && c_off + nd * c_each <= cs && HW > 0,
             "unpack_add: bad args");
  unpack_add_kernel<<<grid_bw((int64_t)nd * c_each * HW), 256, 0, ZT_ST>>>(g, cs, c_off, c_each, nd, d0, d1, HW);
  return check_launch("unpack_add_kernel");
}
extern "C" int zt_act_bwd_nhwc_f32(float* g, int g_cs, const float* y, int y_cs, int C, int act, int64_t M,
                                   void* stream) {
  using namespace zt;
  ZT_REQUIRE(g && y && (C & 3) == 0 && (g_cs & 3) == 0 && (y_cs & 3) == 0 && M > 0, "act_bwd: bad args");
  act_bwd_kernel<<<grid_bw(M * (C >> 2)), 256, 0, ZT_ST>>>(g, g_cs, y, y_cs, C, act, M);
  return check_launch("act_bwd_kernel");
}
/* stats: the (sum, sumsq) pairs of the forward call; red: double[2*C] scratch; dgamma/dbeta accumulate */
extern "C" int zt_bn_train_bwd_nhwc_f32(const float* x, int x_cs, const float* dout, int d_cs, const double* stats,
                                        double* red, const float* gamma, const float* beta, int HW, int C, float eps,
                                        float* dx, int dx_cs, float* dgamma, float* dbeta, void* stream) {
  using namespace zt;
  ZT_REQUIRE(x && dout && stats && red && gamma && beta && dx && dgamma && dbeta, "bn_train_bwd: null pointer");
  ZT_REQUIRE(C > 0 && C <= 256 && HW > 0, "bn_train_bwd: C must be in 1..256");
  ZT_CUDA(cudaMemsetAsync(red, 0, sizeof(double) * 2 * C, ZT_ST));
  const int threads = (256 / C) * C;
  const int blocks = std::max(1, std::min(ceil_div(HW, 64), 4 * kNumSMs));
  const int ppb = ceil_div(HW, blocks);
  const int64_t total = (int64_t)HW * C;
  const bool vec4 = (C & 3) == 0 && ((x_cs | d_cs | dx_cs) & 3) == 0 &&
                    (((uintptr_t)x | (uintptr_t)dout | (uintptr_t)dx) & 15) == 0;
  if (vec4) {
    bn_bwd_reduce4_kernel<<<ceil_div(HW, ppb), 256, 0, ZT_ST>>>(x, x_cs, dout, d_cs, stats, gamma, beta, HW, C, eps, ppb,
                                                                red);
    if (int e = check_launch("bn_bwd_reduce4_kernel")) return e;
    bn_bwd_apply4_kernel<<<grid_bw(total / 4), 256, 0, ZT_ST>>>(x, x_cs, dout, d_cs, stats, red, gamma, beta, HW, C, eps,
                                                                dx, dx_cs, total / 4);
    if (int e = check_launch("bn_bwd_apply4_kernel")) return e;
  } else {
    bn_bwd_reduce_kernel<<<ceil_div(HW, ppb), threads, 0, ZT_ST>>>(x, x_cs, dout, d_cs, stats, gamma, beta, HW, C, eps,
                                                                    ppb, red);
    if (int e = check_launch("bn_bwd_reduce_kernel")) return e;
    bn_bwd_apply_kernel<<<grid_bw(total), 256, 0, ZT_ST>>>(x, x_cs, dout, d_cs, stats, red, gamma, beta, HW, C, eps, dx,
                                                           dx_cs, total);
    if (int e = check_launch("bn_bwd_apply_kernel")) return e;
  }
  bn_bwd_param_kernel<<<ceil_div(C, 128), 128, 0, ZT_ST>>>(red, C, dgamma, dbeta);
  return check_launch("bn_bwd_param_kernel");
}
/* dW: [kh][kw][cin][w_cs] fp32, accumulated into (zero it once per step); dbias[cout] accumulated (may be null).
   X: [H*W][x_cs] input of the layer, G: [H*W][g_cs] gradient of its pre-activation output. */
extern "C" int zt_conv2d_wgrad_nhwc_f32(const float* X, int x_cs, const float* G, int g_cs, float* dW, float* dbias,
                                        int H, int W, int kh, int kw, int ph, int pw, int cin, int cout, int w_cs,
                                        void* stream) {
  using namespace zt;
  ZT_REQUIRE(X && G && dW, "conv2d_wgrad: null pointer");
  ZT_REQUIRE((kw == 1 || kw == 3) && kh >= 1 && kh <= 7, "conv2d_wgrad: kw must be 1 or 3 (got %dx%d)", kh, kw);
  const int TI = ceil_div(cin, 4), TO = ceil_div(cout, 4);
  ZT_REQUIRE(TI * TO <= 256 && TI * 4 <= x_cs && TO * 4 <= g_cs && (x_cs & 3) == 0 && (g_cs & 3) == 0,
             "conv2d_wgrad: cin/cout tile (%d x %d) unsupported", cin, cout);
  int PS = 256 / (TI * TO);
  PS = std::min(PS, WG_PX);
  const size_t smem = sizeof(float) * ((size_t)(WG_PX + kw - 1) * TI * 4 + (size_t)WG_PX * TO * 4);
  const int tiles = H * ceil_div(W, WG_PX);
  const int gx = std::max(1, std::min(tiles, (2 * kNumSMs) / kh));
  if (kw == 3)
    wgrad_kernel<3><<<dim3(gx, kh), 256, smem, ZT_ST>>>(X, x_cs, G, g_cs, dW, H, W, ph, pw, cin, cout, w_cs, TI, TO, PS);
  else
    wgrad_kernel<1><<<dim3(gx, kh), 256, smem, ZT_ST>>>(X, x_cs, G, g_cs, dW, H, W, ph, pw, cin, cout, w_cs, TI, TO, PS);
  if (int e = check_launch("wgrad_kernel")) return e;
  if (dbias) {
    ZT_REQUIRE(cout <= 256, "conv2d_wgrad: bias gradient needs cout <= 256");
    const int64_t M = (int64_t)H * W;
    const int threads = (256 / cout) * cout;
    const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div64(M, 256), 4 * kNumSMs));
    const int rpb = (int)ceil_div64(M, blocks);
    colsum_kernel<<<(int)ceil_div64(M, rpb), threads, 0, ZT_ST>>>(G, g_cs, cout, M, rpb, dbias);
    return check_launch("colsum_kernel");
  }
  return ZT_OK;
}
/* One optimiser step over flat fp32 buffers.  sumsq: device double[1] scratch.  grad_scale multiplies the
   gradients first (1/world after an all-reduce sum).  step >= 1.  Gradients are left clipped in place. */
extern "C" int zt_clip_adam_f32(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double* sumsq,
                                float grad_scale, float max_norm, float lr, float beta1, float beta2, float eps,
                                float weight_decay, int step, void* stream) {
  using namespace zt;
  ZT_REQUIRE(param && grad && exp_avg && exp_avg_sq && sumsq && n > 0 && step >= 1, "clip_adam: bad args");
  ZT_CUDA(cudaMemsetAsync(sumsq, 0, sizeof(double), ZT_ST));
  const int blocks = std::min(grid_bw(n), 2 * kNumSMs);
  sumsq_kernel<<<blocks, 256, 0, ZT_ST>>>(grad, n, sumsq);
  if (int e = check_launch("sumsq_kernel")) return e;
  const float bc1 = (float)(1.0 - pow((double)beta1, (double)step));
  const float bc2s = (float)sqrt(1.0 - pow((double)beta2, (double)step));
  clip_adam_kernel<<<blocks, 256, 0, ZT_ST>>>(param, grad, exp_avg, exp_avg_sq, n, sumsq, grad_scale, max_norm, lr,
                                              beta1, beta2, eps, weight_decay, bc1, bc2s);
  return check_launch("clip_adam_kernel");
}
