// Small RAFT helpers around the tensor-core convolution:
//  * space_to_depth2: [B][H][W][C] -> [B][H/2][W/2][4C] so that the encoders' stride-2 convolutions
//    (model/RAFT/extractor.py:131,44-48 with stride 2) run as stride-1 tensor-core convolutions with
//    rearranged weights (7x7 s2 -> 4x4, 3x3 s2 -> 2x2 over 4C channels; 1x1 s2 -> 1x1 on block (0,0));
//  * convf1: BasicMotionEncoder.convf1 (update.py:84,92: 7x7, 2 -> 128, ReLU).  K = 98: a tensor-core
//    tile would spend 49 taps on 2 useful channels, so this one stays on the CUDA cores in fp32;
//  * flow head conv2 (update.py:13-14: 3x3, 256 -> 2): N = 2, one warp per position, fp32.
#include <algorithm>
#include "common.cuh"

namespace zt {

// out[b][y][x][(by*2+bx)*C + c] = in[b][2y+by][2x+bx][c]
__global__ void space_to_depth2_kernel(const float* __restrict__ in, int in_cs, int C, int B, int H, int W,
                                       float* __restrict__ out) {
  pdl_enter();
  const int c4n = C >> 2, ho = H >> 1, wo = W >> 1;
  const int64_t total = (int64_t)B * ho * wo * 4 * c4n;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c4n);
    int64_t r = i / c4n;
    const int blk = (int)(r & 3);
    r >>= 2;
    const int x = (int)(r % wo);
    r /= wo;
    const int y = (int)(r % ho), b = (int)(r / ho);
    const int64_t src = ((int64_t)b * H + 2 * y + (blk >> 1)) * W + 2 * x + (blk & 1);
    const float4 v = __ldg(reinterpret_cast<const float4*>(in + src * in_cs + cg * 4));
    *reinterpret_cast<float4*>(out + (((int64_t)b * ho + y) * wo + x) * (4 * C) + blk * C + cg * 4) = v;
  }
}

// one warp per output position, lane = 4 output channels; weights [7][7][2][128] (tap-major), flow [N][f_cs]
__global__ void __launch_bounds__(256) convf1_kernel(const float* __restrict__ flow, int f_cs, const float* __restrict__ w,
                                                     const float* __restrict__ bias, float* __restrict__ out, int out_cs,
                                                     int H, int W) {
  pdl_enter();
  const int lane = threadIdx.x & 31;
  const int pos = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (pos >= H * W) return;
  const int y = pos / W, x = pos - y * W;
  float4 acc = __ldg(reinterpret_cast<const float4*>(bias + lane * 4));
#pragma unroll 1
  for (int ky = 0; ky < 7; ++ky) {
    const int iy = y + ky - 3;
    if (iy < 0 || iy >= H) continue;
#pragma unroll
    for (int kx = 0; kx < 7; ++kx) {
      const int ix = x + kx - 3;
      if (ix < 0 || ix >= W) continue;
      const float2 f = __ldg(reinterpret_cast<const float2*>(flow + (int64_t)(iy * W + ix) * f_cs));
      const float* wp = w + (ky * 7 + kx) * 256 + lane * 4;
      const float4 w0 = __ldg(reinterpret_cast<const float4*>(wp));
      const float4 w1 = __ldg(reinterpret_cast<const float4*>(wp + 128));
      acc.x = fmaf(f.x, w0.x, acc.x); acc.y = fmaf(f.x, w0.y, acc.y);
      acc.z = fmaf(f.x, w0.z, acc.z); acc.w = fmaf(f.x, w0.w, acc.w);
      acc.x = fmaf(f.y, w1.x, acc.x); acc.y = fmaf(f.y, w1.y, acc.y);
      acc.z = fmaf(f.y, w1.z, acc.z); acc.w = fmaf(f.y, w1.w, acc.w);
    }
  }
  acc.x = fmaxf(acc.x, 0.f); acc.y = fmaxf(acc.y, 0.f); acc.z = fmaxf(acc.z, 0.f); acc.w = fmaxf(acc.w, 0.f);
  *reinterpret_cast<float4*>(out + (int64_t)pos * out_cs + lane * 4) = acc;
}

// one warp per position: out[pos][0..1] = bias + sum_{tap, c} in[pos+tap][c] * w[tap][c][0..1];  w = [9][C][2]
// coords1 != NULL fuses the flow update of the iteration (raft.py:117-118: coords1 += delta; flow = coords1 - coords0
// written to up to two NHWC buffers) into the flow head, so delta never goes through memory.
__global__ void __launch_bounds__(256) conv3x3_n2_kernel(const float* __restrict__ in, int in_cs, int C,
                                                         const float* __restrict__ w, const float* __restrict__ bias,
                                                         float* __restrict__ out, int out_cs, int H, int W,
                                                         float* __restrict__ coords1, float* __restrict__ f1, int f1_cs,
                                                         float* __restrict__ f2, int f2_cs) {
  pdl_enter();
  const int lane = threadIdx.x & 31;
  const int pos = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (pos >= H * W) return;
  const int y = pos / W, x = pos - y * W;
  float a0 = 0.f, a1 = 0.f;
  for (int t = 0; t < 9; ++t) {
    const int iy = y + t / 3 - 1, ix = x + t % 3 - 1;
    if (iy < 0 || iy >= H || ix < 0 || ix >= W) continue;
    const float* ip = in + (int64_t)(iy * W + ix) * in_cs;
    const float* wp = w + (int64_t)t * C * 2;
    for (int c = lane * 4; c < C; c += 128) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(ip + c));
      const float4 wa = __ldg(reinterpret_cast<const float4*>(wp + c * 2));
      const float4 wb = __ldg(reinterpret_cast<const float4*>(wp + c * 2 + 4));
      a0 = fmaf(v.x, wa.x, a0); a1 = fmaf(v.x, wa.y, a1);
      a0 = fmaf(v.y, wa.z, a0); a1 = fmaf(v.y, wa.w, a1);
      a0 = fmaf(v.z, wb.x, a0); a1 = fmaf(v.z, wb.y, a1);
      a0 = fmaf(v.w, wb.z, a0); a1 = fmaf(v.w, wb.w, a1);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, o);
    a1 += __shfl_xor_sync(0xffffffffu, a1, o);
  }
  if (lane == 0) {
    const float d0 = a0 + bias[0], d1 = a1 + bias[1];
    if (out) {
      out[(int64_t)pos * out_cs] = d0;
      out[(int64_t)pos * out_cs + 1] = d1;
    }
    if (coords1) {
      const float c0 = coords1[2 * pos] + d0, c1 = coords1[2 * pos + 1] + d1;
      coords1[2 * pos] = c0;
      coords1[2 * pos + 1] = c1;
      const float fx = c0 - (float)x, fy = c1 - (float)y;
      if (f1) { f1[(int64_t)pos * f1_cs] = fx; f1[(int64_t)pos * f1_cs + 1] = fy; }
      if (f2) { f2[(int64_t)pos * f2_cs] = fx; f2[(int64_t)pos * f2_cs + 1] = fy; }
    }
  }
}

}  // namespace zt

extern "C" int zt_space_to_depth2_f32(const float* in, int in_cs, int C, int B, int H, int W, float* out,
                                      void* stream) {
  using namespace zt;
  ZT_REQUIRE(in && out && C > 0 && (C & 3) == 0 && (in_cs & 3) == 0 && in_cs >= C && B > 0, "space_to_depth2: bad args");
  ZT_REQUIRE(H >= 2 && W >= 2 && (H & 1) == 0 && (W & 1) == 0, "space_to_depth2: H, W must be even (got %dx%d)", H, W);
  const int64_t total = (int64_t)B * (H / 2) * (W / 2) * C;
  const int blocks = (int)std::min<int64_t>(ceil_div64(total, 256), 16 * kNumSMs);
  ZT_CUDA(launch_pdl(space_to_depth2_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, in, in_cs, C, B, H, W, out));
  return check_launch("space_to_depth2_kernel");
}
/* flow: [H*W][f_cs] (2 channels used); weight: fp32 [7][7][2][128]; out: [H*W][out_cs] (128 channels, ReLU) */
extern "C" int zt_raft_convf1_f32(const float* flow, int f_cs, const float* weight, const float* bias, float* out,
                                  int out_cs, int H, int W, void* stream) {
  using namespace zt;
  ZT_REQUIRE(flow && weight && bias && out && f_cs >= 2 && (f_cs & 1) == 0 && out_cs >= 128 && (out_cs & 3) == 0 &&
             H > 0 && W > 0, "raft_convf1: bad args");
  ZT_CUDA(launch_pdl(convf1_kernel, dim3(ceil_div(H * W, 8)), dim3(256), 0, (cudaStream_t)stream, flow, f_cs, weight, bias,
                     out, out_cs, H, W));
  return check_launch("convf1_kernel");
}
/* in: [H*W][in_cs] with C channels (C % 4 == 0); weight: fp32 [9][C][2]; out[pos][0..1] (stride out_cs), no activation */
extern "C" int zt_conv3x3_n2_f32(const float* in, int in_cs, int C, const float* weight, const float* bias, float* out,
                                 int out_cs, int H, int W, void* stream) {
  using namespace zt;
  ZT_REQUIRE(in && weight && bias && out && C > 0 && (C & 3) == 0 && (in_cs & 3) == 0 && in_cs >= C && out_cs >= 2 &&
             H > 0 && W > 0, "conv3x3_n2: bad args");
  ZT_CUDA(launch_pdl(conv3x3_n2_kernel, dim3(ceil_div(H * W, 8)), dim3(256), 0, (cudaStream_t)stream, in, in_cs, C, weight,
                     bias, out, out_cs, H, W, (float*)nullptr, (float*)nullptr, 0, (float*)nullptr, 0));
  return check_launch("conv3x3_n2_kernel");
}
/* FlowHead.conv2 + the flow update of one RAFT iteration (update.py:13-14, raft.py:117-118) in one launch:
 * delta = conv3x3(in); coords1 [H*W][2] += delta; flow = coords1 - pixel grid -> f1 / f2 (NHWC, 2 channels, may be NULL) */
extern "C" int zt_flow_head2_update_f32(const float* in, int in_cs, int C, const float* weight, const float* bias,
                                        float* coords1, float* f1, int f1_cs, float* f2, int f2_cs, int H, int W,
                                        void* stream) {
  using namespace zt;
  ZT_REQUIRE(in && weight && bias && coords1 && C > 0 && (C & 3) == 0 && (in_cs & 3) == 0 && in_cs >= C && H > 0 && W > 0 &&
             (!f1 || f1_cs >= 2) && (!f2 || f2_cs >= 2), "flow_head2_update: bad args");
  ZT_CUDA(launch_pdl(conv3x3_n2_kernel, dim3(ceil_div(H * W, 8)), dim3(256), 0, (cudaStream_t)stream, in, in_cs, C, weight,
                     bias, (float*)nullptr, 0, H, W, coords1, f1, f1_cs, f2, f2_cs));
  return check_launch("conv3x3_n2_kernel");
}
