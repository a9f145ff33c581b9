// Generic implicit-GEMM convolution on NHWC fp32, CUDA-core (FFMA) path.
//
// Used where the 0.01 px flow tolerance needs full fp32 products (RAFT encoders,
// correlation, update block) - see DESIGN.md "precision plan".  GEMM view:
//   M = B*Ho*Wo output pixels, N = cout, K = kh*kw*cin  (k = tap*cin + c)
// Tile BM x BN x 16 with 256 threads, (TM x TN) register tile per thread, smem
// double buffering with register-staged global prefetch.
//
// Replaces nn.Conv2d call sites: model/RAFT/extractor.py:46-55,167-191,
// model/RAFT/update.py:13-14,45-60,89-97,127-136, model/model.py:24-28,40-44,74-81.
#include "common.cuh"

namespace zt {

constexpr int BK = 16;

template <int TM, int TN>
__global__ void __launch_bounds__(256) conv_f32_kernel(const ZtConv d, const int M, const int cin) {
  constexpr int BM = 16 * TM;
  constexpr int BN = 16 * TN;
  constexpr int A_LD = (BM * 4) / 256;              // float4 loads of A per thread per chunk
  constexpr int B_F4 = BN >= 4 ? (BK * BN) / 4 : 0;  // float4s in a B tile
  constexpr int B_LD = (B_F4 + 255) / 256;
  static_assert(A_LD >= 1, "tile too small");

  __shared__ __align__(16) float As[2][BK][BM];
  __shared__ __align__(16) float Bs[2][BK][BN];

  const int t = threadIdx.x;
  const int tx = t & 15, ty = t >> 4;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int cin4 = cin >> 2;
  const int K4 = d.kh * d.kw * cin4;      // number of float4 groups along K
  const int Ktot = K4 * 4;
  const int nchunks = (K4 + 3) >> 2;
  const int HoWo = d.Ho * d.Wo;

  // ---- per-thread A slots: fixed pixel, fixed k-quad ----
  int a_m[A_LD], a_kq[A_LD], a_iy0[A_LD], a_ix0[A_LD];
  int64_t a_boff[A_LD];
  bool a_ok[A_LD];
#pragma unroll
  for (int i = 0; i < A_LD; ++i) {
    const int f = t + i * 256;
    a_m[i] = f % BM;
    a_kq[i] = f / BM;
    const int m = m0 + a_m[i];
    a_ok[i] = m < M;
    const int mm = a_ok[i] ? m : 0;
    const int b = mm / HoWo;
    const int r = mm - b * HoWo;
    const int oy = r / d.Wo, ox = r - oy * d.Wo;
    a_iy0[i] = oy * d.sh - d.ph;
    a_ix0[i] = ox * d.sw - d.pw;
    a_boff[i] = (int64_t)b * d.H * d.W;
  }

  float4 a_reg[A_LD];
  float4 b_reg[B_LD > 0 ? B_LD : 1];

  auto load_global = [&](int kc) {
#pragma unroll
    for (int i = 0; i < A_LD; ++i) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      const int g = kc * 4 + a_kq[i];
      if (a_ok[i] && g < K4) {
        const int tap = g / cin4;
        const int c = (g - tap * cin4) << 2;
        const int ky = tap / d.kw, kx = tap - ky * d.kw;
        const int iy = a_iy0[i] + ky, ix = a_ix0[i] + kx;
        if (iy >= 0 && iy < d.H && ix >= 0 && ix < d.W) {
          const int64_t pix = a_boff[i] + (int64_t)iy * d.W + ix;
          const float* p = (c < d.c0) ? d.in0 + pix * d.in0_cs + c : d.in1 + pix * d.in1_cs + (c - d.c0);
          v = __ldg(reinterpret_cast<const float4*>(p));
        }
      }
      a_reg[i] = v;
    }
    if (BN >= 4) {
#pragma unroll
      for (int i = 0; i < B_LD; ++i) {
        const int f = t + i * 256;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (f < B_F4) {
          const int row = f / (BN / 4), col = (f % (BN / 4)) * 4;
          const int k = kc * BK + row;
          if (k < Ktot && n0 + col < d.w_cs)
            v = __ldg(reinterpret_cast<const float4*>(d.weight + (int64_t)k * d.w_cs + n0 + col));
        }
        b_reg[i] = v;
      }
    }
  };
  auto store_smem = [&](int buf) {
#pragma unroll
    for (int i = 0; i < A_LD; ++i) {
      const int kb = a_kq[i] * 4;
      As[buf][kb + 0][a_m[i]] = a_reg[i].x;
      As[buf][kb + 1][a_m[i]] = a_reg[i].y;
      As[buf][kb + 2][a_m[i]] = a_reg[i].z;
      As[buf][kb + 3][a_m[i]] = a_reg[i].w;
    }
#pragma unroll
    for (int i = 0; i < B_LD; ++i) {
      const int f = t + i * 256;
      if (f < B_F4) {
        const int row = f / (BN / 4), col = (f % (BN / 4)) * 4;
        *reinterpret_cast<float4*>(&Bs[buf][row][col]) = b_reg[i];
      }
    }
  };

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  load_global(0);
  store_smem(0);
  __syncthreads();

  for (int kc = 0; kc < nchunks; ++kc) {
    const int buf = kc & 1;
    if (kc + 1 < nchunks) load_global(kc + 1);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[TM], b[TN];
#pragma unroll
      for (int i = 0; i < TM / 4; ++i) {
        const float4 v = *reinterpret_cast<const float4*>(&As[buf][k][i * 64 + ty * 4]);
        a[i * 4 + 0] = v.x; a[i * 4 + 1] = v.y; a[i * 4 + 2] = v.z; a[i * 4 + 3] = v.w;
      }
      if constexpr (TN == 1) {
        b[0] = Bs[buf][k][tx];
      } else {
#pragma unroll
        for (int j = 0; j < TN / 4; ++j) {
          const float4 v = *reinterpret_cast<const float4*>(&Bs[buf][k][j * 64 + tx * 4]);
          b[j * 4 + 0] = v.x; b[j * 4 + 1] = v.y; b[j * 4 + 2] = v.z; b[j * 4 + 3] = v.w;
        }
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kc + 1 < nchunks) {
      store_smem(buf ^ 1);
      __syncthreads();
    }
  }

  // ---- epilogue ----
  constexpr int NQ = TN >= 4 ? TN / 4 : 1;   // n-groups per thread
  constexpr int NE = TN >= 4 ? 4 : 1;        // elements per group
  const bool vec_ok = NE == 4 && (d.out_cs & 3) == 0 && (d.cout & 3) == 0 &&
                      (d.epi != ZT_EPI_GRU_ZR || (d.aux1_cs & 3) == 0);
  const int Ch = d.cout >> 1;  // GRU_ZR split
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int m = m0 + (i >> 2) * 64 + ty * 4 + (i & 3);
    if (m >= M) continue;
#pragma unroll
    for (int jq = 0; jq < NQ; ++jq) {
      const int nb = NE == 4 ? n0 + jq * 64 + tx * 4 : n0 + tx;
      if (nb >= d.cout) continue;
      float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int e = 0; e < NE; ++e) {
        const int n = nb + e;
        float x = acc[i][jq * NE + e];
        if (n < d.cout) {
          if (d.bias) x += __ldg(d.bias + n);
          switch (d.epi) {
            case ZT_EPI_GRU_ZR: {
              x = 1.f / (1.f + expf(-x));
              if (n >= Ch) x *= d.aux0[(int64_t)m * d.aux0_cs + (n - Ch)];
            } break;
            case ZT_EPI_GRU_Q: {
              const float z = d.aux1[(int64_t)m * d.aux1_cs + n];
              const float h = d.aux0[(int64_t)m * d.aux0_cs + n];
              x = (1.f - z) * h + z * tanhf(x);
            } break;
            case ZT_EPI_ADD_RELU:
              x = fmaxf(d.aux0[(int64_t)m * d.aux0_cs + n] + apply_act(x, d.act) * d.out_scale, 0.f);
              break;
            case ZT_EPI_ADD:
              x = d.aux0[(int64_t)m * d.aux0_cs + n] + apply_act(x, d.act) * d.out_scale;
              break;
            default:
              x = apply_act(x, d.act) * d.out_scale;
          }
        }
        v[e] = x;
      }
      float* dst;
      int dst_n, dst_cs;
      if (d.epi == ZT_EPI_GRU_ZR && nb < Ch) { dst = d.aux1; dst_n = nb; dst_cs = d.aux1_cs; }
      else if (d.epi == ZT_EPI_GRU_ZR)       { dst = d.out; dst_n = nb - Ch; dst_cs = d.out_cs; }
      else                                   { dst = d.out; dst_n = nb; dst_cs = d.out_cs; }
      float* p = dst + (int64_t)m * dst_cs + dst_n;
      if (vec_ok) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
      } else {
#pragma unroll
        for (int e = 0; e < NE; ++e)
          if (nb + e < d.cout) p[e] = v[e];
      }
    }
  }
}

template <int TM, int TN>
static int launch_conv(const ZtConv& d, int M, int cin, cudaStream_t st) {
  dim3 grid(ceil_div(M, 16 * TM), ceil_div(d.cout, 16 * TN));
  conv_f32_kernel<TM, TN><<<grid, 256, 0, st>>>(d, M, cin);
  return check_launch("conv_f32_kernel");
}

}  // namespace zt

extern "C" int zt_conv2d_nhwc_f32(const ZtConv* dp, void* stream) {
  using namespace zt;
  ZT_REQUIRE(dp != nullptr, "zt_conv2d_nhwc_f32: null descriptor");
  const ZtConv& d = *dp;
  const int cin = d.c0 + d.c1;
  ZT_REQUIRE(d.in0 && d.weight && d.out, "conv: null pointer");
  ZT_REQUIRE(d.bplane == nullptr, "conv: the per-pixel bias plane is a tensor-core-path feature");
  ZT_REQUIRE((d.c0 & 3) == 0 && (d.c1 & 3) == 0 && cin > 0, "conv: channel counts must be multiples of 4 (c0=%d c1=%d)", d.c0, d.c1);
  ZT_REQUIRE((d.in0_cs & 3) == 0 && (d.c1 == 0 || ((d.in1_cs & 3) == 0 && d.in1)), "conv: input channel strides must be multiples of 4");
  ZT_REQUIRE((d.w_cs & 3) == 0 && d.w_cs >= d.cout, "conv: weight row stride %d must be a multiple of 4 and >= cout %d", d.w_cs, d.cout);
  ZT_REQUIRE(((uintptr_t)d.in0 & 15) == 0 && ((uintptr_t)d.weight & 15) == 0 && (d.c1 == 0 || ((uintptr_t)d.in1 & 15) == 0), "conv: pointers must be 16-byte aligned");
  ZT_REQUIRE(d.B > 0 && d.H > 0 && d.W > 0 && d.Ho > 0 && d.Wo > 0 && d.cout > 0, "conv: bad dims");
  ZT_REQUIRE(d.Ho == (d.H + 2 * d.ph - d.kh) / d.sh + 1 && d.Wo == (d.W + 2 * d.pw - d.kw) / d.sw + 1, "conv: output dims inconsistent");
  if (d.epi == ZT_EPI_GRU_ZR) ZT_REQUIRE(d.aux0 && d.aux1 && (d.cout & 7) == 0, "conv: GRU_ZR needs aux0/aux1 and cout%%8==0");
  if (d.epi == ZT_EPI_GRU_Q) ZT_REQUIRE(d.aux0 && d.aux1, "conv: GRU_Q needs aux0/aux1");
  if (d.epi == ZT_EPI_ADD_RELU || d.epi == ZT_EPI_ADD) ZT_REQUIRE(d.aux0, "conv: residual epilogue needs aux0");
  const int64_t M64 = (int64_t)d.B * d.Ho * d.Wo;
  ZT_REQUIRE(M64 < (1ll << 31), "conv: too many output pixels");
  const int M = (int)M64;
  cudaStream_t st = (cudaStream_t)stream;
  // tile selection: wide N tiles for wide layers; 64-row tiles when 128-row tiles would not
  // give every SM at least ~2 CTAs (RAFT's 1/8-resolution grids).
  const bool small_m = (int64_t)ceil_div(M, 128) * ceil_div(d.cout, d.cout > 64 ? 128 : 64) < 2 * kNumSMs;
  if (d.cout <= 16) return small_m ? launch_conv<4, 1>(d, M, cin, st) : launch_conv<8, 1>(d, M, cin, st);
  if (d.cout <= 64 || small_m) return small_m ? launch_conv<4, 4>(d, M, cin, st) : launch_conv<8, 4>(d, M, cin, st);
  return launch_conv<8, 8>(d, M, cin, st);
}
