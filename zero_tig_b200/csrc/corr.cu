// RAFT correlation pyramid helpers, 4-level 9x9 bilinear lookup, convex upsampling.
// Reference: model/RAFT/corr.py:13-60, utils/utils.py:285-299, model/RAFT/raft.py:64-75,111-120.
#include <cuda_bf16.h>
#include <algorithm>
#include "common.cuh"

namespace zt {

// ---------------- transpose (fmap2 [N][256] -> [256][N]) ----------------
__global__ void transpose_kernel(const float* __restrict__ in, int R, int C, int in_rs,
                                 float* __restrict__ out, int out_rs) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int r = r0 + j, c = c0 + threadIdx.x;
    tile[j][threadIdx.x] = (r < R && c < C) ? in[(int64_t)r * in_rs + c] : 0.f;
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int c = c0 + j, r = r0 + threadIdx.x;
    if (c < C && r < out_rs) out[(int64_t)c * out_rs + r] = (r < R) ? tile[threadIdx.x][j] : 0.f;
  }
}

// ---------------- 2x2 average pooling over many small planes ----------------
__global__ void avgpool2_kernel(const float* __restrict__ in, float* __restrict__ out, int64_t total,
                                int h, int w, int ho, int wo) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int x = (int)(i % wo);
    const int64_t r = i / wo;
    const int y = (int)(r % ho);
    const int64_t pl = r / ho;
    const float* p = in + (pl * h + 2 * y) * w + 2 * x;
    // same summation order as ATen avg_pool2d: row-major accumulate then divide by 4
    out[i] = (p[0] + p[1] + p[w] + p[w + 1]) * 0.25f;
  }
}

// ---------------- 4-level lookup ----------------
struct LookupArgs {
  const void* lvl[4];    // fp32 or bf16 planes (template parameter of the kernel)
  int h[4], w[4];
};
__device__ __forceinline__ float lk_get(const float* p, int64_t i) { return __ldg(p + i); }
__device__ __forceinline__ float lk_get(const __nv_bfloat16* p, int64_t i) {
  return __bfloat162float(__ushort_as_bfloat16(__ldg(reinterpret_cast<const unsigned short*>(p) + i)));
}

// One WARP per query position (4 queries per 128-thread CTA, no block-wide barrier): per level the 10x10
// window of the position's own correlation plane that covers the 9x9 bilinear taps is staged in shared memory
// with zeros outside the plane (grid_sample padding_mode='zeros') - 12-13 independent loads in flight per lane -
// then 324 outputs are written contiguously (channel-fastest NHWC, the layout convc1 consumes).
template <typename T>
__global__ void __launch_bounds__(128) corr_lookup_kernel(const LookupArgs a, const float* __restrict__ coords,
                                                          float* __restrict__ out, int out_cs, int N) {
  __shared__ float win_s[4][4][10][12];
  __shared__ float frac_s[4][4][2];
  pdl_enter();
  const int wq = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * 4 + wq;
  if (n >= N) return;
  float (*win)[10][12] = win_s[wq];
  float (*frac)[2] = frac_s[wq];
  const float cx = coords[2 * n], cy = coords[2 * n + 1];
#pragma unroll
  for (int k = 0; k < 13; ++k) {
    const int i = lane + 32 * k;
    if (i < 400) {
      const int l = i / 100, r = i - l * 100;
      const int wy = r / 10, wx = r - wy * 10;
      const float s = 1.f / (float)(1 << l);
      const float x = cx * s, y = cy * s;      // coords / 2**i  (exact: power-of-two divide)
      const float fx0 = floorf(x), fy0 = floorf(y);
      const int ix = (int)fx0 - 4 + wx, iy = (int)fy0 - 4 + wy;
      // selects instead of dynamic indexing: keeps the kernel parameters out of local memory
      const int hl = l == 0 ? a.h[0] : l == 1 ? a.h[1] : l == 2 ? a.h[2] : a.h[3];
      const int wl = l == 0 ? a.w[0] : l == 1 ? a.w[1] : l == 2 ? a.w[2] : a.w[3];
      const T* pl = reinterpret_cast<const T*>(l == 0 ? a.lvl[0] : l == 1 ? a.lvl[1] : l == 2 ? a.lvl[2] : a.lvl[3]);
      float v = 0.f;
      if (ix >= 0 && ix < wl && iy >= 0 && iy < hl) v = lk_get(pl, ((int64_t)n * hl + iy) * wl + ix);
      win[l][wy][wx] = v;
      if (r == 0) { frac[l][0] = x - fx0; frac[l][1] = y - fy0; }
    }
  }
  __syncwarp();
  float* o = out + (int64_t)n * out_cs;
#pragma unroll
  for (int k = 0; k < 11; ++k) {
    const int i = lane + 32 * k;
    if (i < 324) {
      const int l = i / 81, r = i - l * 81;
      const int ax = r / 9, by = r - ax * 9;   // channel = l*81 + a*9 + b : x-offset a, y-offset b (Q1)
      const float fx = frac[l][0], fy = frac[l][1];
      const float v00 = win[l][by][ax], v01 = win[l][by][ax + 1];
      const float v10 = win[l][by + 1][ax], v11 = win[l][by + 1][ax + 1];
      // grid_sample bilinear: nw*(1-fx)(1-fy) + ne*fx(1-fy) + sw*(1-fx)fy + se*fx*fy
      o[i] = v00 * ((1.f - fx) * (1.f - fy)) + v01 * (fx * (1.f - fy)) + v10 * ((1.f - fx) * fy) + v11 * (fx * fy);
    }
  }
}

// ---------------- convex upsampling (last iteration only) ----------------
// grid: one CTA of 64 threads per coarse pixel; thread = (i,j) sub-pixel
__global__ void __launch_bounds__(64) convex_upsample_kernel(const float* __restrict__ flow, int flow_cs,
                                                             const float* __restrict__ mask, int mask_cs,
                                                             float* __restrict__ out, int h8, int w8) {
  const int n = blockIdx.x;
  const int y = n / w8, x = n - y * w8;
  const int t = threadIdx.x;
  float m[9], mx = -INFINITY;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    m[k] = __ldg(mask + (int64_t)n * mask_cs + k * 64 + t);
    mx = fmaxf(mx, m[k]);
  }
  float sum = 0.f;
#pragma unroll
  for (int k = 0; k < 9; ++k) { m[k] = expf(m[k] - mx); sum += m[k]; }
  float ux = 0.f, uy = 0.f;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    const int yy = y + k / 3 - 1, xx = x + k % 3 - 1;   // F.unfold 3x3, padding 1 (zeros)
    float fx = 0.f, fy = 0.f;
    if (yy >= 0 && yy < h8 && xx >= 0 && xx < w8) {
      fx = 8.f * flow[(int64_t)flow_cs * (yy * w8 + xx)];
      fy = 8.f * flow[(int64_t)flow_cs * (yy * w8 + xx) + 1];
    }
    const float wgt = m[k] / sum;
    ux += wgt * fx;
    uy += wgt * fy;
  }
  const int i = t >> 3, j = t & 7;
  const int64_t H = 8 * h8, Wd = 8 * w8;
  const int64_t o = (int64_t)(8 * y + i) * Wd + 8 * x + j;
  out[o] = ux;
  out[H * Wd + o] = uy;
}

// coords1 += delta ; flow = coords1 - coords0 (coords0 = pixel grid) -> written to 1 or 2 buffers
__global__ void flow_update_kernel(float* __restrict__ coords1, const float* __restrict__ delta, int d_cs,
                                   float* __restrict__ f1, int f1_cs, float* __restrict__ f2, int f2_cs,
                                   int N, int w8) {
  pdl_enter();
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const int y = n / w8, x = n - y * w8;
  float c0 = coords1[2 * n], c1 = coords1[2 * n + 1];
  if (delta) { c0 += delta[(int64_t)n * d_cs]; c1 += delta[(int64_t)n * d_cs + 1]; }
  coords1[2 * n] = c0;
  coords1[2 * n + 1] = c1;
  const float fx = c0 - (float)x, fy = c1 - (float)y;
  if (f1) { f1[(int64_t)n * f1_cs] = fx; f1[(int64_t)n * f1_cs + 1] = fy; }
  if (f2) { f2[(int64_t)n * f2_cs] = fx; f2[(int64_t)n * f2_cs + 1] = fy; }
}

}  // namespace zt

extern "C" int zt_transpose_f32(const float* in, int R, int C, int in_rs, float* out, int out_rs, void* stream) {
  using namespace zt;
  ZT_REQUIRE(in && out && R > 0 && C > 0 && in_rs >= C && out_rs >= R, "transpose: bad args");
  transpose_kernel<<<dim3(ceil_div(C, 32), ceil_div(out_rs, 32)), dim3(32, 8), 0, (cudaStream_t)stream>>>(
      in, R, C, in_rs, out, out_rs);
  return check_launch("transpose_kernel");
}

extern "C" int zt_avgpool2_f32(const float* in, float* out, int64_t planes, int h, int w, void* stream) {
  using namespace zt;
  ZT_REQUIRE(in && out && planes > 0 && h >= 2 && w >= 2, "avgpool2: bad args (h=%d w=%d)", h, w);
  const int ho = h / 2, wo = w / 2;
  const int64_t total = planes * ho * wo;
  const int blocks = (int)std::min<int64_t>(ceil_div64(total, 256), 16 * kNumSMs);
  avgpool2_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(in, out, total, h, w, ho, wo);
  return check_launch("avgpool2_kernel");
}

static int corr_lookup_any(const void* l0, const void* l1, const void* l2, const void* l3, bool lvl_bf16,
                           const int* hs, const int* ws_, const float* coords, float* out, int out_cs, int N, void* stream) {
  using namespace zt;
  ZT_REQUIRE(l0 && l1 && l2 && l3 && hs && ws_ && coords && out && N > 0 && out_cs >= 324, "corr_lookup: bad args");
  LookupArgs a;
  a.lvl[0] = l0; a.lvl[1] = l1; a.lvl[2] = l2; a.lvl[3] = l3;
  for (int i = 0; i < 4; ++i) {
    a.h[i] = hs[i]; a.w[i] = ws_[i];
    // the reference divides by (W-1), (H-1): a 1-pixel level is NaN there; refuse.
    ZT_REQUIRE(hs[i] >= 2 && ws_[i] >= 2, "corr_lookup: pyramid level %d is %dx%d; the reference yields NaN below 2x2", i, hs[i], ws_[i]);
  }
  if (lvl_bf16)
    ZT_CUDA(launch_pdl(corr_lookup_kernel<__nv_bfloat16>, dim3(ceil_div(N, 4)), dim3(128), 0, (cudaStream_t)stream, a, coords, out, out_cs, N));
  else
    ZT_CUDA(launch_pdl(corr_lookup_kernel<float>, dim3(ceil_div(N, 4)), dim3(128), 0, (cudaStream_t)stream, a, coords, out, out_cs, N));
  return check_launch("corr_lookup_kernel");
}

extern "C" int zt_corr_lookup_f32(const float* l0, const float* l1, const float* l2, const float* l3,
                                  const int* hs, const int* ws_, const float* coords, float* out, int out_cs,
                                  int N, void* stream) {
  return corr_lookup_any(l0, l1, l2, l3, false, hs, ws_, coords, out, out_cs, N, stream);
}
/* the same lookup over a bf16-storage pyramid (zt_corr_pyramid_bf16); interpolation and output stay fp32 */
extern "C" int zt_corr_lookup_bf16(const void* l0, const void* l1, const void* l2, const void* l3,
                                   const int* hs, const int* ws_, const float* coords, float* out, int out_cs,
                                   int N, void* stream) {
  return corr_lookup_any(l0, l1, l2, l3, true, hs, ws_, coords, out, out_cs, N, stream);
}

extern "C" int zt_convex_upsample_f32(const float* flow, int flow_cs, const float* mask, int mask_cs, float* out,
                                      int h8, int w8, void* stream) {
  using namespace zt;
  ZT_REQUIRE(flow && mask && out && h8 > 0 && w8 > 0 && mask_cs >= 576 && flow_cs >= 2, "convex_upsample: bad args");
  convex_upsample_kernel<<<h8 * w8, 64, 0, (cudaStream_t)stream>>>(flow, flow_cs, mask, mask_cs, out, h8, w8);
  return check_launch("convex_upsample_kernel");
}

extern "C" int zt_flow_update_f32(float* coords1, const float* delta, int d_cs, float* flow_out, int f_cs,
                                  float* flow_out2, int f2_cs, int h8, int w8, void* stream) {
  using namespace zt;
  ZT_REQUIRE(coords1 && h8 > 0 && w8 > 0, "flow_update: bad args");
  const int N = h8 * w8;
  ZT_CUDA(launch_pdl(flow_update_kernel, dim3(ceil_div(N, 128)), dim3(128), 0, (cudaStream_t)stream, coords1, delta, d_cs,
                     flow_out, f_cs, flow_out2, f2_cs, N, w8));
  return check_launch("flow_update_kernel");
}
