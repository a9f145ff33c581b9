// InstanceNorm2d (affine=False, track_running_stats=False) on NHWC fp32.
// Reference: nn.InstanceNorm2d in model/RAFT/extractor.py:27-31,129 (fnet only).
// Two kernels: per-(image,channel) sum / sum-of-squares accumulated in fp64 (one pass,
// no cancellation problem), then a fused normalise [+relu] [+residual] [+relu] pass.
#include <algorithm>
#include <mutex>
#include "common.cuh"

namespace zt {

// grid: (slabs, B); block: 256 threads; thread -> channel group of 4 (float4), strided pixels
__global__ void __launch_bounds__(256) instnorm_stats_kernel(const float* __restrict__ x, int x_cs,
                                                             double* __restrict__ stats, int HW, int C,
                                                             int pix_per_block) {
  pdl_enter();
  const int c4n = C >> 2;                 // float4 groups per pixel
  const int lanes = 256 / c4n;            // pixels processed concurrently
  const int cg = threadIdx.x % c4n;
  const int pl = threadIdx.x / c4n;
  const int b = blockIdx.y;
  const int p0 = blockIdx.x * pix_per_block;
  const int p1 = min(p0 + pix_per_block, HW);
  double s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0};
  if (pl < lanes) {
    const float* base = x + (int64_t)b * HW * x_cs + cg * 4;
    int p = p0 + pl;
    // four independent loads in flight per thread; the sums stay in fp64 (no cancellation problem later)
    for (; p + 3 * lanes < p1; p += 4 * lanes) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldg(reinterpret_cast<const float4*>(base + (int64_t)(p + u * lanes) * x_cs));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        s[0] += v[u].x; q[0] += (double)v[u].x * v[u].x;
        s[1] += v[u].y; q[1] += (double)v[u].y * v[u].y;
        s[2] += v[u].z; q[2] += (double)v[u].z * v[u].z;
        s[3] += v[u].w; q[3] += (double)v[u].w * v[u].w;
      }
    }
    for (; p < p1; p += lanes) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(base + (int64_t)p * x_cs));
      s[0] += v.x; q[0] += (double)v.x * v.x;
      s[1] += v.y; q[1] += (double)v.y * v.y;
      s[2] += v.z; q[2] += (double)v.z * v.z;
      s[3] += v.w; q[3] += (double)v.w * v.w;
    }
  }
  // red[pixel lane][channel group][sum x4 | sumsq x4]; one thread per (channel group, entry) folds the pixel lanes
  __shared__ double red[256 * 8];
#pragma unroll
  for (int e = 0; e < 4; ++e) { red[threadIdx.x * 8 + e] = s[e]; red[threadIdx.x * 8 + 4 + e] = q[e]; }
  __syncthreads();
  for (int j = threadIdx.x; j < c4n * 8; j += 256) {
    double acc = 0;
    for (int l = 0; l < lanes; ++l) acc += red[l * c4n * 8 + j];
    const int g4 = j >> 3, e = j & 7;
    // stats layout: [b][c][sum, sumsq]
    atomicAdd(stats + ((int64_t)b * C + g4 * 4 + (e & 3)) * 2 + (e >> 2), acc);
  }
}

// mean / 1/std of every (image, channel) once per block into shared memory, then a streaming normalise
__global__ void __launch_bounds__(256) instnorm_apply_kernel(const float* __restrict__ x, int x_cs,
                                                             float* __restrict__ out, int out_cs,
                                                             const float* __restrict__ res, int res_cs,
                                                             const double* __restrict__ stats, int HW, int C, int BC,
                                                             float eps, int relu_inner, int relu_outer,
                                                             int64_t total4) {
  extern __shared__ float2 mr[];          // [b*C + c] = (mean, rstd)
  pdl_enter();
  for (int i = threadIdx.x; i < BC; i += blockDim.x) {
    const double mean = stats[2 * i] / HW;
    double var = stats[2 * i + 1] / HW - mean * mean;   // biased variance
    var = var < 0 ? 0 : var;
    mr[i] = make_float2((float)mean, (float)(1.0 / sqrt(var + (double)eps)));
  }
  __syncthreads();
  const int c4n = C >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c4n);
    const int64_t pix = i / c4n;          // b*HW + p
    const int b = (int)(pix / HW);
    const float2* st = mr + b * C + cg * 4;
    float4 v = __ldg(reinterpret_cast<const float4*>(x + pix * x_cs + cg * 4));
    float r[4] = {v.x, v.y, v.z, v.w};
    float4 rv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (res) rv = __ldg(reinterpret_cast<const float4*>(res + pix * res_cs + cg * 4));
    const float rr[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float y = (r[e] - st[e].x) * st[e].y;
      if (relu_inner) y = fmaxf(y, 0.f);
      if (res) y += rr[e];
      if (relu_outer) y = fmaxf(y, 0.f);
      r[e] = y;
    }
    *reinterpret_cast<float4*>(out + pix * out_cs + cg * 4) = make_float4(r[0], r[1], r[2], r[3]);
  }
}

// BatchNorm2d in TRAIN mode for batch 1 (Enhancer block, model/model.py:60-67,74-81 after
// train.py:116): batch statistics over H*W, affine, ReLU, residual add: out = res + relu(bn(x)).
__global__ void __launch_bounds__(256) bn_train_apply_kernel(const float* __restrict__ x, int x_cs,
                                                             float* __restrict__ out, int out_cs,
                                                             const float* __restrict__ res, int res_cs,
                                                             const double* __restrict__ stats,
                                                             const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, int HW, int C, float eps,
                                                             int64_t total4) {
  // per-channel constants once per block (the fp64 mean / variance / 1/sqrt chain used to run per element)
  extern __shared__ float4 bn_cst[];      // [c] = (mean, rstd, gamma, beta)
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    const double mean = stats[c * 2] / HW;
    double var = stats[c * 2 + 1] / HW - mean * mean;
    var = var < 0 ? 0 : var;
    bn_cst[c] = make_float4((float)mean, (float)(1.0 / sqrt(var + (double)eps)), __ldg(gamma + c), __ldg(beta + c));
  }
  __syncthreads();
  const int c4n = C >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c4n);
    const int64_t pix = i / c4n;
    const float4 v = __ldg(reinterpret_cast<const float4*>(x + pix * x_cs + cg * 4));
    float r[4] = {v.x, v.y, v.z, v.w};
    float4 rv = make_float4(0.f, 0.f, 0.f, 0.f);
    if (res) rv = __ldg(reinterpret_cast<const float4*>(res + pix * res_cs + cg * 4));
    const float rr[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float4 k = bn_cst[cg * 4 + e];
      float y = (r[e] - k.x) * k.y * k.z + k.w;
      y = fmaxf(y, 0.f);
      r[e] = res ? rr[e] + y : y;
    }
    *reinterpret_cast<float4*>(out + pix * out_cs + cg * 4) = make_float4(r[0], r[1], r[2], r[3]);
  }
}
// running_mean / running_var update: momentum m, unbiased variance (torch BatchNorm semantics)
__global__ void bn_update_running_kernel(const double* __restrict__ stats, float* __restrict__ rmean,
                                         float* __restrict__ rvar, int HW, int C, float momentum) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const double mean = stats[c * 2] / HW;
  double var = stats[c * 2 + 1] / HW - mean * mean;
  var = var < 0 ? 0 : var;
  const double unb = HW > 1 ? var * (double)HW / (double)(HW - 1) : var;
  rmean[c] = (1.f - momentum) * rmean[c] + momentum * (float)mean;
  rvar[c] = (1.f - momentum) * rvar[c] + momentum * (float)unb;
}

}  // namespace zt

extern "C" int zt_instnorm_nhwc_f32(const float* x, int x_cs, float* out, int out_cs, const float* res,
                                    int res_cs, double* stats, int B, int HW, int C, float eps,
                                    int relu_inner, int relu_outer, void* stream) {
  using namespace zt;
  ZT_REQUIRE(x && out && stats, "instnorm: null pointer");
  ZT_REQUIRE((C & 3) == 0 && C <= 1024 && (x_cs & 3) == 0 && (out_cs & 3) == 0 && (!res || (res_cs & 3) == 0),
             "instnorm: C and strides must be multiples of 4 (C=%d)", C);
  ZT_REQUIRE(B > 0 && HW > 0 && (int64_t)B * C <= 6144, "instnorm: bad dims (B*C must be <= 6144)");
  cudaStream_t st = (cudaStream_t)stream;
  ZT_CUDA(cudaMemsetAsync(stats, 0, sizeof(double) * 2 * B * C, st));
  const int slabs = max(1, min(ceil_div(HW, 64), (4 * kNumSMs) / B));
  const int ppb = ceil_div(HW, slabs);
  instnorm_stats_kernel<<<dim3(ceil_div(HW, ppb), B), 256, 0, st>>>(x, x_cs, stats, HW, C, ppb);
  if (int e = check_launch("instnorm_stats_kernel")) return e;
  const int64_t total4 = (int64_t)B * HW * (C >> 2);
  const int blocks = (int)std::min<int64_t>(ceil_div64(total4, 256), 8 * kNumSMs);
  ZT_CUDA(launch_pdl(instnorm_apply_kernel, dim3(blocks), dim3(256), (size_t)B * C * sizeof(float2), st, x, x_cs, out, out_cs,
                     res, res_cs, stats, HW, C, B * C, eps, relu_inner, relu_outer, total4));
  return check_launch("instnorm_apply_kernel");
}

// ---- allocation-free variant ---------------------------------------------------------------------------------------
// Same two kernels, but the fp64 statistics live in a file-scope device array instead of a caller-provided buffer,
// for callers that cannot allocate (CUDA graph capture, tight serving loops).  No counterpart in the reference.
// Every GPU that loads the library holds the 16 KB until exit.  All calls on a device share the array: it is cleared
// at the start of every call, and two calls that overlap on the GPU (different streams) race on it.
namespace zt {
static __device__ __align__(256) double g_instnorm_scratch[2 * ZT_INSTNORM_NOALLOC_MAX_BC];

// In host code `g_instnorm_scratch` is only a host-side stand-in for the symbol, not a device address: resolve it
// once per device (with that device current) and cache the pointer.
static int instnorm_scratch_ptr(double** out) {
  constexpr int kMaxDev = 64;
  static std::mutex mu;
  static double* ptrs[kMaxDev] = {};
  int dev = -1;
  ZT_CUDA(cudaGetDevice(&dev));
  ZT_REQUIRE(dev >= 0 && dev < kMaxDev, "instnorm_noalloc: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    double* p = nullptr;
    ZT_CUDA(cudaGetSymbolAddress((void**)&p, g_instnorm_scratch));
    ZT_REQUIRE(p != nullptr, "instnorm_noalloc: cudaGetSymbolAddress returned a null pointer");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return ZT_OK;
}
}  // namespace zt

extern "C" int zt_instnorm_nhwc_f32_noalloc(const float* x, int x_cs, float* out, int out_cs, const float* res,
                                            int res_cs, int B, int HW, int C, float eps, int relu_inner,
                                            int relu_outer, void* stream) {
  using namespace zt;
  ZT_REQUIRE(x && out, "instnorm_noalloc: null pointer");
  ZT_REQUIRE((C & 3) == 0 && C > 0 && (x_cs & 3) == 0 && (out_cs & 3) == 0 && (!res || (res_cs & 3) == 0),
             "instnorm_noalloc: C and strides must be multiples of 4 (C=%d)", C);
  ZT_REQUIRE(B > 0 && HW > 0, "instnorm_noalloc: bad dims");
  ZT_REQUIRE((int64_t)B * C <= ZT_INSTNORM_NOALLOC_MAX_BC,
             "instnorm_noalloc: B*C = %lld exceeds the static scratch (ZT_INSTNORM_NOALLOC_MAX_BC = %d); use "
             "zt_instnorm_nhwc_f32 with a caller-provided stats buffer", (long long)B * C, ZT_INSTNORM_NOALLOC_MAX_BC);
  double* stats = nullptr;
  if (int e = instnorm_scratch_ptr(&stats)) return e;
  cudaStream_t st = (cudaStream_t)stream;
  ZT_CUDA(cudaMemsetAsync(stats, 0, sizeof(double) * 2 * B * C, st));     // the array keeps the last call's sums
  const int slabs = max(1, min(ceil_div(HW, 64), (4 * kNumSMs) / B));
  const int ppb = ceil_div(HW, slabs);
  instnorm_stats_kernel<<<dim3(ceil_div(HW, ppb), B), 256, 0, st>>>(x, x_cs, stats, HW, C, ppb);
  if (int e = check_launch("instnorm_stats_kernel")) return e;
  const int64_t total4 = (int64_t)B * HW * (C >> 2);
  const int blocks = (int)std::min<int64_t>(ceil_div64(total4, 256), 8 * kNumSMs);
  ZT_CUDA(launch_pdl(instnorm_apply_kernel, dim3(blocks), dim3(256), (size_t)B * C * sizeof(float2), st, x, x_cs, out,
                     out_cs, res, res_cs, (const double*)stats, HW, C, B * C, eps, relu_inner, relu_outer, total4));
  return check_launch("instnorm_apply_kernel");
}

extern "C" int zt_bn_train_nhwc_f32(const float* x, int x_cs, float* out, int out_cs, const float* res, int res_cs,
                                    double* stats, const float* gamma, const float* beta, float* running_mean,
                                    float* running_var, int HW, int C, float eps, float momentum, void* stream) {
  using namespace zt;
  ZT_REQUIRE(x && out && stats && gamma && beta, "bn_train: null pointer");
  ZT_REQUIRE((C & 3) == 0 && C <= 1024 && (x_cs & 3) == 0 && (out_cs & 3) == 0 && (!res || (res_cs & 3) == 0),
             "bn_train: C and strides must be multiples of 4 (C=%d)", C);
  ZT_REQUIRE(HW > 0, "bn_train: bad dims");
  cudaStream_t st = (cudaStream_t)stream;
  ZT_CUDA(cudaMemsetAsync(stats, 0, sizeof(double) * 2 * C, st));
  const int slabs = max(1, min(ceil_div(HW, 64), 4 * kNumSMs));
  const int ppb = ceil_div(HW, slabs);
  instnorm_stats_kernel<<<dim3(ceil_div(HW, ppb), 1), 256, 0, st>>>(x, x_cs, stats, HW, C, ppb);
  if (int e = check_launch("instnorm_stats_kernel")) return e;
  const int64_t total4 = (int64_t)HW * (C >> 2);
  const int blocks = (int)std::min<int64_t>(ceil_div64(total4, 256), 8 * kNumSMs);
  bn_train_apply_kernel<<<blocks, 256, (size_t)C * sizeof(float4), st>>>(x, x_cs, out, out_cs, res, res_cs, stats, gamma,
                                                                         beta, HW, C, eps, total4);
  if (int e = check_launch("bn_train_apply_kernel")) return e;
  if (running_mean && running_var) {
    bn_update_running_kernel<<<ceil_div(C, 128), 128, 0, st>>>(stats, running_mean, running_var, HW, C, momentum);
    return check_launch("bn_update_running_kernel");
  }
  return ZT_OK;
}
