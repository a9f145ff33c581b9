/*
 * zerotig.h - C ABI of libzerotig.so, the B200 (sm_100a) kernels behind the Zero-TIG
 * per-frame-pair hot path.  Plain pointers and sizes only: every pointer is a DEVICE
 * pointer owned by the caller (torch tensors' data_ptr() in the Python host),
 * `stream` is a cudaStream_t passed as void*.  Every function returns 0 on success or
 * a non-zero code; zt_last_error() returns the message of the last failure on the
 * calling thread.  There is no CPU fallback: without a CUDA device every compute
 * entry point fails.
 *
 * The reference (L-Forster/Zero-TIG) has no native code and therefore no FFI of its
 * own: each entry point below names the reference Python function
 * (file:line under the upstream tree) whose torch ops it replaces.  The host-side
 * binding is ctypes (zero_tig_b200/_lib.py); INTEGRATION.md shows the stub.
 *
 * Internal activation layout: NHWC fp32 ("pixel-major", channel stride `cs` >= C,
 * cs % 4 == 0, base pointers 16-byte aligned).  Frame-level tensors that cross the
 * reference's Python interface stay NCHW fp32 like the reference's.
 */
#ifndef ZEROTIG_H_
#define ZEROTIG_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ZT_OK 0
#define ZT_ERR_ARG 1
#define ZT_ERR_CUDA 2

/* activation codes for ZtConv.act */
#define ZT_ACT_NONE 0
#define ZT_ACT_RELU 1
#define ZT_ACT_LRELU02 2
#define ZT_ACT_SIGMOID 3
#define ZT_ACT_TANH 4

/* epilogue codes for ZtConv.epi */
#define ZT_EPI_PLAIN 0        /* out = scale * act(conv + bias)                                   */
#define ZT_EPI_GRU_ZR 1       /* cout = 2*Ch: n<Ch  -> aux1[n]   = sigmoid(v)      (z gate)       */
                              /*              n>=Ch -> out[n-Ch] = sigmoid(v)*aux0[n-Ch] (r*h)    */
#define ZT_EPI_GRU_Q 2        /* q=tanh(v); out[n] = (1-z)*h + z*q with h=aux0[n], z=aux1[n]      */
#define ZT_EPI_ADD_RELU 3     /* out = relu(aux0[n] + act(v))   (ResidualBlock tail)              */
#define ZT_EPI_ADD 4          /* out = aux0[n] + act(v)         (Enhancer block)                  */

const char* zt_last_error(void);
int zt_version(void);
/* number of kernel launches issued through this library since load (bench's gpu_launches) */
int64_t zt_launch_count(void);
/* add the kernel count of a replayed CUDA graph (captured launches are counted once, at capture time) */
int zt_add_launches(int64_t n);

/* ---- generic implicit-GEMM convolution, fp32 CUDA-core path --------------------------------
 * Replaces every nn.Conv2d on the RAFT side (model/RAFT/extractor.py:9-10,44,131,140;
 * update.py:9-10,36-42,82-86,122-125) and, in the fp32 build, the enhancement nets
 * (model/model.py:20-22,36-38,54-71).  Input channels may come from two sources
 * (virtual torch.cat([in0, in1], 1), update.py:47,50,130).  Weight layout
 * [kh][kw][cin][w_cs] with w_cs = cout rounded up to 4 (zero padded). */
typedef struct ZtConv {
  const float* in0; int in0_cs; int c0;
  const float* in1; int in1_cs; int c1;
  const float* weight; int w_cs;
  const float* bias;               /* [cout] or NULL */
  float* out; int out_cs;
  int B, H, W, Ho, Wo, cout;
  int kh, kw, sh, sw, ph, pw;
  int act; float out_scale; int epi;
  const float* aux0; int aux0_cs;
  float* aux1; int aux1_cs;
  const float* bplane; int bplane_cs;   /* optional per-pixel bias [M][bplane_cs] added before the epilogue (tensor-core
                                          path only): the iteration-invariant part of a GRU convolution, i.e. the
                                          convolution of the context features `inp` (update.py:47,50), computed once per frame */
  float acc_scale;                     /* tensor-core split mode 4 only: 1 / (ZT_F16_ACT_SCALE * weight scale) */
} ZtConv;
int zt_conv2d_nhwc_f32(const ZtConv* d, void* stream);

/* Same convolution on the tcgen05 tensor cores (stride 1 only): `weight` points at bf16 hi/lo
 * weights packed by the host as [n_tile][64-ch group][tap][hi|lo][row BN][64 ch], every
 * 128-byte row stored with the UMMA 128B swizzle (16 B chunk c of row n at position c^(n&7)),
 * zero padded; w_cs is ignored.  split = 1: bf16 x bf16; split = 3: hi*hi + lo*hi + hi*lo over bf16 pairs
 * (about 2^-17 relative per operand, what RAFT's 0.01 px tolerance needs); split = 4: the same three products
 * over fp16 hi/lo pairs of the SCALED operands - activations * ZT_F16_ACT_SCALE (applied by the kernel), weights *
 * a power of two chosen by the packer so that the low halves stay normal fp16 numbers - about 2^-22 per operand,
 * i.e. fp32-grade; the accumulator is multiplied by desc.acc_scale = 1 / (both scales) before the epilogue.  An
 * activation above 65504 / ZT_F16_ACT_SCALE would overflow to inf (loud, not silent).  BN = N tile, multiple of 16
 * in [16,256].  Same call sites as zt_conv2d_nhwc_f32. */
#define ZT_F16_ACT_SCALE 16.0f
int zt_conv2d_tc_nhwc(const ZtConv* d, int BN, int split, void* stream);

/* nn.Conv2d weights [cout][cin][kh][kw] fp32 (the live parameter, train.py:98) -> the packed weight operand of
 * zt_conv2d_tc_nhwc, on the device (no host tensor ops, no synchronisation): bf16 hi/lo pairs, or (fp16 != 0) fp16
 * pairs of weight * scale (split mode 4).  dgrad != 0 packs the weights of the DATA-GRADIENT convolution instead
 * (in/out channels swapped, taps mirrored: dX = conv(dY, W')).  cin_pad: padded K channel count (0 = round up to 4).
 * out: ceil(rows/BN)*BN * ceil(cin_pad/64)*64 * kh*kw * 4 bytes.  Launch the convolution with split | 0x100 (weights
 * written by the preceding kernel). */
int zt_pack_tc_weights_f32(const float* w, int cout, int cin, int kh, int kw, int cin_pad, int dgrad, int BN, int fp16,
                           float scale, void* out, void* stream);

/* Pack R rows x C fp32 (row stride rs) as the weight operand of zt_conv2d_tc_nhwc for a 1x1
 * convolution with cout = R whose weights are produced on the device every frame: the all-pairs
 * correlation corr.py:52-60 (weights = fmap2).  out: ceil(R/BN)*BN * ceil(C/64)*64 * 4 bytes. */
int zt_pack_tc_rows_f32(const float* rows, int R, int C, int rs, int BN, void* out, void* stream);

/* debug only: per-role timeline of CTA (0,0) of the next zt_conv2d_tc_nhwc launches; buf is a
 * device int64 array [1 + 2*cap] (zeroed by the caller), NULL disables. */
int zt_debug_set_tc_trace(void* buf, int cap);

/* ---- InstanceNorm2d (affine=False, eps) on NHWC fp32; extractor.py:27-31,129 -------------
 * stats: per (image, channel) sums in fp64 scratch `stats` [B][C][2] (zeroed inside).
 * apply: y = norm(x); if relu_inner y=relu(y); if res: y = res + y; if relu_outer y=relu(y) */
int zt_instnorm_nhwc_f32(const float* x, int x_cs, float* out, int out_cs,
                         const float* res, int res_cs, double* stats,
                         int B, int HW, int C, float eps, int relu_inner, int relu_outer,
                         void* stream);
/* Allocation-free variant of zt_instnorm_nhwc_f32.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments minus `stats`: the fp64 statistics live in a file-scope static __device__ array of
 * 2 * ZT_INSTNORM_NOALLOC_MAX_BC doubles (16 KB, held by every GPU that loads the library until exit); the call
 * allocates nothing.  B * C > ZT_INSTNORM_NOALLOC_MAX_BC is rejected with ZT_ERR_ARG.  All calls on a device share
 * the array (cleared at the start of each call): calls that overlap on the GPU (different streams) race. */
#define ZT_INSTNORM_NOALLOC_MAX_BC 1024
int zt_instnorm_nhwc_f32_noalloc(const float* x, int x_cs, float* out, int out_cs,
                                 const float* res, int res_cs,
                                 int B, int HW, int C, float eps, int relu_inner, int relu_outer,
                                 void* stream);

/* ---- RAFT correlation: corr.py:13-27,52-60 ---------------------------------------------- */
/* out[c][r] = in[r][c]; out row stride out_rs (>= R) */
int zt_transpose_f32(const float* in, int R, int C, int in_rs, float* out, int out_rs, void* stream);
/* F.avg_pool2d(x,2,stride=2) over `planes` planes of h x w -> (h/2) x (w/2)  (corr.py:26) */
int zt_avgpool2_f32(const float* in, float* out, int64_t planes, int h, int w, void* stream);
/* CorrBlock.__init__ (corr.py:13-27) + CorrBlock.corr (corr.py:52-60) in one pass: fmap [2*h0*w0][fm_rs]
 * (rows 0..N-1 = fmap1, N..2N-1 = fmap2, 256 channels) -> level l [N][hl][wl], hl = h0 >> l (floor),
 * level 0 = <fmap1[n], fmap2[m]> / 16, levels 1..3 = successive 2x2 means.  tcgen05 GEMM (bf16 hi/lo
 * split, fp32 accumulate) fed by TMA; all four levels are written from the same accumulator.
 * scratch: zt_corr_pyramid_scratch_bytes(h0, w0) bytes, 128-byte aligned. */
int64_t zt_corr_pyramid_scratch_bytes(int h0, int w0);
int zt_corr_pyramid_f32(const float* fmap, int fm_rs, int h0, int w0, float* l0, float* l1, float* l2,
                        float* l3, void* scratch, int64_t scratch_bytes, void* stream);
/* bf16-STORAGE variant (BASELINE.json configs[4], "fp32 vs bf16 corr"): the same GEMM (fp32-grade products, fp32
 * accumulation); the four levels are rounded to bf16 on the way out, halving the bytes of this output-bound kernel */
int zt_corr_pyramid_bf16(const float* fmap, int fm_rs, int h0, int w0, void* l0, void* l1, void* l2,
                         void* l3, void* scratch, int64_t scratch_bytes, void* stream);
/* CorrBlock.__call__ (corr.py:29-50) + bilinear_sampler (utils/utils.py:285-299):
 * coords [N][2] (x,y), pyramid level l is [N][hl][wl]; out [N][out_cs], channel
 * l*81 + a*9 + b samples level l at (x/2^l + a-4, y/2^l + b-4)  (x-offset major). */
int zt_corr_lookup_f32(const float* l0, const float* l1, const float* l2, const float* l3,
                       const int* hs, const int* ws_, /* host arrays[4] */
                       const float* coords, float* out, int out_cs, int N, void* stream);
/* the same lookup over a bf16-storage pyramid (zt_corr_pyramid_bf16); interpolation and output stay fp32 */
int zt_corr_lookup_bf16(const void* l0, const void* l1, const void* l2, const void* l3,
                       const int* hs, const int* ws_, /* host arrays[4] */
                       const float* coords, float* out, int out_cs, int N, void* stream);
/* RAFT.upsample_flow (raft.py:64-75): flow [N][flow_cs] NHWC (first 2 channels), mask [N][576]
 * NHWC -> out NCHW [2][8h][8w] */
int zt_convex_upsample_f32(const float* flow, int flow_cs, const float* mask, int mask_cs, float* out,
                           int h8, int w8, void* stream);
/* coords1 += delta (raft.py:120) and flow = coords1 - coords0 (raft.py:115) written into a
 * channel slice; delta [N][d_cs], coords [N][2], flow_out [N][f_cs] (first 2 channels) */
int zt_flow_update_f32(float* coords1, const float* delta, int d_cs, float* flow_out, int f_cs,
                       float* flow_out2, int f2_cs, int h8, int w8, void* stream);

/* ---- update_cache pre-processing: model/model.py:346-360 -------------------------------- */
/* F.interpolate(bilinear, align_corners=False) of an NCHW 3-plane image to (ht,wd), *255.
 * mode 0: write RAFT image1: replicate-pad to (hp,wp) (raft.py:132-138), 2*(v/255)-1,
 *         NHWC cs=4 (4th channel 0) into out_f32.
 * mode 1: truncate to uint8 into out_u8 [3][ht][wd] and accumulate hist [3][256] (int32,
 *         zeroed inside). */
int zt_resize255_f32(const float* img, int H, int W, int ht, int wd, int mode,
                     float* out_f32, int hp, int wp, uint8_t* out_u8, int* hist, void* stream);
/* torchvision equalize LUT (_functional_tensor._scale_channel): hist [3][256] -> lut [3][256] u8 */
int zt_equalize_lut(const int* hist, uint8_t* lut, void* stream);
/* apply LUT, then RAFT image2 packing as mode 0 above; optionally also writes the equalised
 * u8 image (eq_u8 may be NULL) */
int zt_equalize_apply_pack(const uint8_t* q, const uint8_t* lut, int ht, int wd, int hp, int wp,
                           float* out_f32, uint8_t* eq_u8, void* stream);

/* ---- warp: utils/utils.py:203-230, called twice with the same flow (model.py:374-375) ---
 * flow NCHW [2][hf][wf]; img_a,img_b NCHW [C][H][W]; out_a,out_b same.  img_b/out_b may be
 * NULL (single warp). */
int zt_warp_pair_f32(const float* flow, int hf, int wf, const float* img_a, const float* img_b,
                     float* out_a, float* out_b, int C, int H, int W, void* stream);

/* ---- layout + Retinex elementwise: model/model.py:313-335 -------------------------------- */
/* concat up to 4 NCHW sources (c channels each) into NHWC [HW][cs]; channels beyond the
 * concatenation are zero; `add` is added to every source value (input + 1e-4, model.py:314) */
int zt_pack_nhwc_f32(const float* s0, const float* s1, const float* s2, const float* s3,
                     int c_each, int nsrc, float add, float* out, int cs, int64_t HW, void* stream);
/* L2 = clamp((x+eps) - d, eps, 1): x NCHW [3], d NHWC (cs), L2 NCHW   (model.py:315-316) */
int zt_retinex_l2_f32(const float* x, const float* d, int d_cs, float* L2, int64_t HW, float eps,
                      void* stream);
/* s2 = clamp(s, 1e-4, 1) (model.py:79); H2 = clamp((x+eps)/s2, eps, 1) (model.py:327-328);
 * s NHWC (cs) already sigmoid'ed; outputs NCHW */
int zt_retinex_h2_f32(const float* x, const float* s, int s_cs, float* s2, float* H2, int64_t HW,
                      float eps, void* stream);
/* H3 = clamp(H2 - d[0:3], eps, 1), s3 = clamp(s2 - d[3:6], eps, 1)  (model.py:334-337) */
int zt_retinex_h5_f32(const float* H2, const float* s2, const float* d, int d_cs, float* H3,
                      float* s3, int64_t HW, float eps, void* stream);
/* save_images quantisation, predict.py:57-61: clip(x*255,0,255) -> uint8 (truncate), HWC */
int zt_quantize_u8_hwc(const float* x, uint8_t* out, int C, int64_t HW, void* stream);
/* transforms.ToTensor on the device, dataloader/multi_read_data.py:86-87,131 (uint8 HWC image -> float NCHW, value/255
 * with IEEE division: bit-identical to torch's): lets a caller upload the 8-bit frame (1/4 of the bytes) */
int zt_u8_hwc_to_f32(const uint8_t* in, float* out, int C, int64_t HW, void* stream);

/* ---- Retinex networks at full resolution: TMA-fed tcgen05 convolution over bf16 NHWC activations ----
 * Replaces the nn.Conv2d layers of Denoise_1 / Enhancer / Denoise_2 (model/model.py:24-28,40-44,74-81) in
 * inference: stride 1, 3x3 (pad 1) or 1x1, cin <= 64 (multiple of 8), cout <= 64, batch 1.
 * in: bf16 [H*W][cin]; weight: bf16 [tap][BN=ceil16(cout)][64 ch], 128-byte swizzled rows (host packs);
 * out: bf16 [H*W][out_cs] (out_f32 = 0; optional bf16 residual: out = res + act(conv + bias)) or
 * fp32 [H*W][out_cs] with out_cs 4 or 8 (out_f32 = 1: the 3- and 6-channel heads). act = ZT_ACT_*. */
typedef struct ZtConvBf16 {
  const void* in; int cin;
  const void* weight; const float* bias;
  void* out; int out_cs; int out_f32;
  const void* res; int res_cs;
  int H; int W; int cout; int kh; int kw; int act;
  /* optional fused 1x1 head (the denoisers' last layer, model/model.py:27,43), head_n = 0: none.  Then out is fp32
   * [H*W][out_cs] (out_cs 4 or 8 >= head_n) = head_b + head_w . act(conv + bias): head_w fp32 [8][64] (row = head
   * channel, zero padded), head_b fp32 [8]; the cout-channel activation itself is not written. */
  const float* head_w; const float* head_b; int head_n;
} ZtConvBf16;
int zt_conv2d_bf16_tma(const ZtConvBf16* desc, void* stream);
/* NCHW fp32 planes (+add) -> bf16 NHWC [HW][cs], cs = 8 or 16, zero padded (virtual torch.cat, model.py:313,326,334) */
int zt_pack_nhwc_bf16(const float* s0, const float* s1, const float* s2, const float* s3, int c_each, int nsrc,
                      float add, void* out, int cs, int64_t HW, void* stream);

/* ---- RAFT helpers around the tensor-core convolution ----
 * [B][H][W][C] -> [B][H/2][W/2][4C] (channel = (by*2+bx)*C + c): lets the encoders' stride-2 convolutions
 * (model/RAFT/extractor.py:44-48,131) run as stride-1 tensor-core convolutions with rearranged weights */
int zt_space_to_depth2_f32(const float* in, int in_cs, int C, int B, int H, int W, float* out, void* stream);
/* BasicMotionEncoder.convf1 + ReLU (model/RAFT/update.py:84,92): 7x7, 2 -> 128; weight fp32 [7][7][2][128] */
int zt_raft_convf1_f32(const float* flow, int f_cs, const float* weight, const float* bias, float* out, int out_cs,
                       int H, int W, void* stream);
/* FlowHead.conv2 (model/RAFT/update.py:13-14): 3x3, C -> 2, no activation; weight fp32 [9][C][2] */
int zt_conv3x3_n2_f32(const float* in, int in_cs, int C, const float* weight, const float* bias, float* out,
                      int out_cs, int H, int W, void* stream);
/* FlowHead.conv2 fused with the per-iteration flow update (update.py:13-14 + raft.py:117-118): coords1 [H*W][2] +=
 * conv3x3(in); flow = coords1 - pixel grid is written to f1 / f2 (NHWC, first 2 channels; either may be NULL) */
int zt_flow_head2_update_f32(const float* in, int in_cs, int C, const float* weight, const float* bias,
                             float* coords1, float* f1, int f1_cs, float* f2, int f2_cs, int H, int W, void* stream);

/* Convolution weight gradient on the tensor cores (train.py:129 loss.backward(); the nn.Conv2d layers of
 * model/model.py:15-81): dW[kh][kw][cin_p][w_cs] (fp32, accumulated with atomics) += sum over pixels p of
 * X[p + tap][ci] * G[p][co].  X (layer input) and G (gradient of the layer output) are NHWC bf16 hi/lo pairs
 * ([H*W][x_cs], [H*W][g_cs]; zt_split_bf16_hilo; strides multiples of 8): TMA-fed tcgen05 with MN-major operands,
 * K = pixels, three-term hi/lo products, fp32 accumulation in TMEM across all tiles of a persistent CTA.
 * 1x1 or 3x3 (zero padding k/2, stride 1), cin <= 64, cout <= 64. */
/* the operand conversion of zt_conv2d_wgrad_tc: fp32 [M][in_rs] -> bf16 hi / lo [M][out_rs] for C channels (C % 4 == 0);
 * colsum (may be NULL): colsum[c] += sum_m in[m][c] = the bias gradient when `in` is the output gradient */
int zt_split_bf16_hilo_colsum(const float* in, int in_rs, void* hi, void* lo, int out_rs, int C, int64_t M, float* colsum,
                              void* stream);
int zt_conv2d_wgrad_tc(const void* Xhi, const void* Xlo, int x_cs, const void* Ghi, const void* Glo, int g_cs,
                       float* dW, int H, int W, int kh, int kw, int cin, int cout, int cin_p, int w_cs, void* stream);

/* ---- RAFT update block, all iterations of a frame pair in ONE persistent cooperative kernel ----
 * Replaces the loop body of RAFT.forward, model/RAFT/raft.py:111-120: CorrBlock.__call__ (corr.py:29-50),
 * BasicUpdateBlock.forward (update.py:127-136: BasicMotionEncoder :89-97, SepConvGRU :45-60, FlowHead :13-14) and
 * coords1 += delta_flow, for `iters` iterations (the convex-upsampling mask head runs once afterwards, Q15).
 * Every activation that feeds a convolution is a pair of 16-bit tensors [N][C] (hi, lo: bf16 with hi + lo = the fp32
 * value to 2^-17, or - fp16 != 0 - fp16 of the value * ZT_F16_ACT_SCALE, to 2^-22), N = h8*w8 positions, so TMA can
 * feed the tensor cores; products use the three-term hi/lo split.
 *   in : h [N][128] fp32 hidden state (tanh(cnet[:128]), raft.py:104), the four per-frame GRU planes
 *        bp_* = conv_{z|r,q}(context features) + bias (the iteration-invariant share of update.py:47-58),
 *        the 4-level correlation pyramid lvl[i] = [N][lh[i]*lw[i]] fp32
 *   out: flow4 [N][4] fp32 (flow in channels 0,1 = coords1 - coords0), coords1 [N][2], h (final hidden state)
 * Weights: w[i] = bf16 hi/lo packed as for zt_conv2d_tc_nhwc with N tile bn[i] (32 or 64), b[i] fp32 bias or NULL,
 * layer order convc1, convf2, convc2, conv, zr1, q1, zr2, q2, flow-head conv1; the GRU weights are the [h | motion]
 * input-channel slices (the context slice lives in bp_*).  wf1 / bf1: convf1 fp32 [7][7][2][128] / [128];
 * wfh2 / bfh2: flow-head conv2 fp32 [9][256][2] / [2].  corr_* [N][384] must have channels 324..383 zero.
 * counter: one device word for the grid barrier (the call zeroes it).  max_ctas: 0 = as many as fit. */
typedef struct ZtRaftUpdate {
  int h8; int w8; int iters; int max_ctas;
  void *corr_hi, *corr_lo;      /* [N][384] */
  void *c1_hi, *c1_lo;          /* [N][256] */
  void *f1_hi, *f1_lo;          /* [N][128] */
  void *c2f_hi, *c2f_lo;        /* [N][256] */
  void *hm_hi, *hm_lo;          /* [N][256]: hidden state | motion features */
  void *rh_hi, *rh_lo;          /* [N][128] */
  float *h, *z, *fhid, *coords1, *flow4;
  const float *bp_zr1, *bp_q1, *bp_zr2, *bp_q2;
  const void* lvl[4]; int lh[4]; int lw[4]; int lvl_bf16;   /* pyramid planes fp32, or bf16 when lvl_bf16 != 0 */
  const void* w[9]; const float* b[9]; int bn[9];
  int fp16; float acc_scale[9];  /* fp16 != 0: split mode 4 of zt_conv2d_tc_nhwc (scaled fp16 hi/lo operands, weights packed
                                    accordingly, acc_scale[i] = 1 / (ZT_F16_ACT_SCALE * weight scale of layer i)) */
  const float *wf1, *bf1, *wfh2, *bfh2;
  unsigned int* counter;
} ZtRaftUpdate;
int zt_raft_update(const ZtRaftUpdate* desc, void* stream);
/* debug only: device buffer of 8 * (2 + 9*iters) int64; CTA `cta` of the following zt_raft_update launches stores
 * %globaltimer (ns) per phase and role (see ru_trace in csrc/raft_update.cu); NULL disables */
int zt_debug_set_raft_update_trace(void* buf, int cta);
/* fp32 [N][C] (row stride in_rs) -> bf16 hi / lo [N][C] (row stride out_rs), hi = bf16(x), lo = bf16(x - hi) */
int zt_split_bf16_hilo(const float* in, int in_rs, void* hi, void* lo, int out_rs, int C, int64_t N, void* stream);

/* ---- training path: Network.forward (model/model.py:144-203) and loss.py -------------------
 * NCHW fp32, batch 1.  Reduction entry points ADD raw sums into fp64 accumulators the caller
 * zeroes; the host turns them into the 17 weighted terms of LossFunction.forward. */
/* BatchNorm2d in train mode fused with ReLU and the Enhancer residual (model.py:76-78):
 * out = res + relu(gamma*(x-mean)/sqrt(var+eps)+beta) with batch stats over HW; optionally updates
 * running_mean/var (momentum, unbiased var).  stats: fp64 scratch [C][2]. */
int zt_bn_train_nhwc_f32(const float* x, int x_cs, float* out, int out_cs, const float* res, int res_cs,
                         double* stats, const float* gamma, const float* beta, float* running_mean,
                         float* running_var, int HW, int C, float eps, float momentum, void* stream);
/* pair_downsampler utils/utils.py:15-24 on (in + add): out1 = anti-diagonal mean, out2 = diagonal mean */
int zt_pair_downsample_f32(const float* in, float* out1, float* out2, int C, int H, int W, float add, void* stream);
/* out[c][p] = (a[c][p] + add) - d[p][c]  (d NHWC), optionally clamped: L_pred / L2 (model.py:149-152) */
int zt_sub_clamp_f32(const float* a, float add, const float* d, int d_cs, float* out, int C, int64_t HW,
                     float lo, float hi, int do_clamp, void* stream);
/* out = clamp((a + add) / b, lo, hi): H2, H11, H12, H1 (model.py:170-177,198-199) */
int zt_div_clamp_f32(const float* a, float add, const float* b, float* out, int64_t n, float lo, float hi,
                     void* stream);
/* TextureDifference.forward loss.py:106-136: 3-plane images -> binary map [H][W] */
int zt_texture_diff_f32(const float* img1, const float* img2, float* out, int H, int W, void* stream);
/* utils.blur (utils.py:52-58): reflect pad 10 + 21x21 kernel, run separably; taps = 21 HOST floats */
int zt_blur21_f32(const float* in, float* tmp, float* out, const float* taps, int C, int H, int W, void* stream);
/* loss.py:26-47: acc[0] += sum (s2-bright)^2, acc[1] += sum (norm_ll-adj)^2; chansum = fp64[3] scratch */
int zt_loss_illum_f32(const float* L2, const float* s2, int is_WB, int64_t HW, double* chansum, double* acc,
                      void* stream);
/* SmoothLoss loss.py:173-311: acc[0..11] += raw sums of the 12 unique offsets; ycc = 3*H*W scratch */
int zt_loss_smooth_f32(const float* L2, const float* s2, float* ycc, int H, int W, double* acc, void* stream);
/* L_TV loss.py:139-152: acc[0] += vertical, acc[1] += horizontal squared differences */
int zt_loss_tv_f32(const float* x, int C, int H, int W, double* acc, void* stream);
/* acc[0] += sum((a + add_a - b)^2): every nn.MSELoss term of loss.py:46-77 */
int zt_mse_sum_f32(const float* a, const float* b, float add_a, int64_t n, double* acc, void* stream);
/* Loss_inter loss.py:68-73 (LocalMean 5x5 reflect, quirk Q7): acc[0], acc[1] */
int zt_loss_inter_f32(const float* d1, const float* d2, const float* diff, int C, int H, int W, double* acc,
                      void* stream);
/* Loss_Var loss.py:75-77 with calculate_local_variance utils.py:66-79: acc[0]; avg1/avg2 = C*H*W scratch */
int zt_loss_var_f32(const float* H2, const float* H3, float* avg1, float* avg2, int C, int H, int W, double* acc,
                    void* stream);

/* ---- training backward (train.py:126-131): adjoints of the loss terms, of the Retinex glue and of
 * the networks.  Frame-level gradients are NCHW fp32 planes the entry points ACCUMULATE into (the
 * host zeroes them once per step); network-internal gradients are NHWC like the activations.  The
 * data gradient of a convolution is the forward convolution (zt_conv2d_tc_nhwc) with spatially
 * flipped, transposed weights. */
/* nn.MSELoss backward: d = coef*((a+add_a)-b); ga += d (if non-null); gb -= d (if non-null) */
int zt_mse_bwd_f32(const float* a, const float* b, float add_a, int64_t n, float coef, float* ga, float* gb,
                   void* stream);
/* loss.py:40-47 adjoint w.r.t. s2: g_s2 += c0*(s2-bright) + c1*(norm_ll-adj)*d norm_ll/d s2 */
int zt_loss_illum_bwd_f32(const float* L2, const float* s2, int is_WB, int64_t HW, const double* chansum, float c0,
                          float c1, float* g_s2, void* stream);
/* SmoothLoss loss.py:193-311 adjoint w.r.t. the illumination map; coef = 12 HOST floats (weight*2/count) */
int zt_loss_smooth_bwd_f32(const float* ycc, const float* s2, int H, int W, const float* coef, float* g_s2,
                           void* stream);
/* L_TV loss.py:139-152 adjoint: ch, cw = weight*2/count of the vertical / horizontal sums */
int zt_loss_tv_bwd_f32(const float* x, int C, int H, int W, float ch, float cw, float* g, void* stream);
/* Loss_inter loss.py:66-73 adjoint incl. LocalMean^T (reflect 5x5); u1,u2 = C*H*W scratch */
int zt_loss_inter_bwd_f32(const float* d1, const float* d2, const float* diff, int C, int H, int W, float coef,
                          float* u1, float* u2, float* g_d1, float* g_d2, void* stream);
/* Loss_Var loss.py:75-77 adjoint through calculate_local_variance (utils.py:66-79); t0..t2 scratch */
int zt_loss_var_bwd_f32(const float* H2, const float* H3, const float* avg1, const float* avg2, int C, int H, int W,
                        float coef, float* t0, float* t1, float* t2, float* g_H2, float* g_H3, void* stream);
/* utils.blur adjoint (reflect-pad folding included): g_out += blur^T(g_in) */
int zt_blur21_bwd_f32(const float* g_in, float* tmp, float* g_out, const float* taps, int C, int H, int W,
                      void* stream);
/* pair_downsampler^T (utils.py:15-24): gx += scatter of g1 (anti-diagonal) and g2 (diagonal) */
int zt_pair_downsample_bwd_f32(const float* g1, const float* g2, int C, int H, int W, float* gx, void* stream);
/* clamp((a+add)/b, lo, hi) adjoint w.r.t. b (model.py:170-177) */
int zt_div_clamp_bwd_f32(const float* a, float add, const float* b, const float* g_out, int64_t n, float lo, float hi,
                         float* g_b, void* stream);
/* NHWC gradient of a network output from one or two NCHW gradient planes, sign applied and masked by the
 * clamp of the matching forward output (model.py:152,180-191; o* may be null = no clamp) */
int zt_mask_pack_nhwc_f32(const float* g0, const float* o0, const float* g1, const float* o1, int c_each, int nsrc,
                          float sign, float lo, float hi, float* out, int cs, int64_t HW, void* stream);
/* d_k[c][p] += g[p][c_off + k*c_each + c]: gradient of the non-detached torch.cat inputs of denoise_2 */
int zt_unpack_add_nchw_f32(const float* g, int cs, int c_off, int c_each, int nd, float* d0, float* d1, int64_t HW,
                           void* stream);
/* g *= act'(y) with the derivative written through the activation output y (ZT_ACT_*) */
int zt_act_bwd_nhwc_f32(float* g, int g_cs, const float* y, int y_cs, int C, int act, int64_t M, void* stream);
/* backward of zt_bn_train_nhwc_f32 (BatchNorm2d train mode + ReLU; the residual passes dout through) */
int zt_bn_train_bwd_nhwc_f32(const float* x, int x_cs, const float* dout, int d_cs, const double* stats, double* red,
                             const float* gamma, const float* beta, int HW, int C, float eps, float* dx, int dx_cs,
                             float* dgamma, float* dbeta, void* stream);
/* nn.Conv2d weight / bias gradient, stride 1: dW[kh][kw][cin][w_cs] += X^T * G per tap; dbias += colsum(G) */
int zt_conv2d_wgrad_nhwc_f32(const float* X, int x_cs, const float* G, int g_cs, float* dW, float* dbias, int H, int W,
                             int kh, int kw, int ph, int pw, int cin, int cout, int w_cs, void* stream);
/* clip_grad_norm_ + torch.optim.Adam step (train.py:98,130-131) over flat fp32 buffers, optional
 * pre-scale of the gradients (1/world after the NCCL all-reduce) */
int zt_clip_adam_f32(float* param, float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double* sumsq,
                     float grad_scale, float max_norm, float lr, float beta1, float beta2, float eps,
                     float weight_decay, int step, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ZEROTIG_H_ */
