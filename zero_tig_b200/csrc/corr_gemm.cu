// All-pairs correlation volume + 4-level pyramid in ONE GEMM (CorrBlock.__init__ / CorrBlock.corr,
// model/RAFT/corr.py:13-27,52-60):  corr[n][m] = <fmap1[n], fmap2[m]> / sqrt(256), then three 2x2
// average poolings over m (floor on odd sizes).
//
// Average pooling over m is linear, so pooled level l is the correlation of fmap1 with the 2^l x 2^l
// window means of fmap2:  level_l[n][m_l] = <fmap1[n], pool_l(fmap2)[m_l]> / 16.  corr_pack_kernel
// builds P = [fmap2 ; pool2 ; pool4 ; pool8] (each level padded to 128 rows) and Q = fmap1, both split
// to bf16 hi/lo pairs; corr_gemm_kernel then writes ALL FOUR levels straight from the accumulator and
// the volume is never re-read.
//
// GEMM on tcgen05 fed by TMA: the M side (TMEM lanes) is 128 rows of P = 128 consecutive positions of
// one level, the N side is 128 queries, K = 256.  fp32-grade products from the hi/lo pairs (3 of the 4
// partial products: Phi*[Qhi|Qlo] as one N=256 MMA + Plo*Qhi), fp32 accumulation.  An epilogue thread
// owns one position m and walks the queries of the tile, so a warp store is 32 consecutive floats of one
// correlation plane (128 B, fully coalesced).  The P tile (hi/lo, 128 KB) stays resident in shared memory
// while the CTA walks its (persistent, contiguous) share of the tile grid; Q streams through a 3-deep TMA ring; accumulators are
// double buffered in TMEM (2 x 256 columns) so the stores of tile t overlap the MMAs of tile t+1.
// HBM roofline: 4*N*sum_l(h_l*w_l) bytes written (+ 2 KB per fmap row read once by the pack kernel).
#include <cuda.h>
#include <cuda_bf16.h>
#include <algorithm>
#include "common.cuh"
#include "tc_ptx.cuh"
#include "tma_host.cuh"

namespace zt {

constexpr int CG_EPI_WARPS = 8;          // warp w: TMEM lane quadrant w&3, accumulator columns 64(w>>2)..+63
constexpr int CG_THREADS = 32 * (CG_EPI_WARPS + 2);   // + MMA warp + TMA loader warp
constexpr int CG_BAR = 1024;
constexpr int CG_A_BYTES = 2 * 4 * 16384;            // P tile: hi/lo x 4 k-blocks x [128 rows][128 B]
constexpr int CG_B_STAGE = 2 * 16384;                // Q stage: [hi rows 0..127 | lo rows 128..255] x 128 B
constexpr int CG_NB = 3;

struct CgArgs {
  void* lvl[4];          // fp32 or bf16 planes (template parameter of the kernel)
  int npos[4];           // h_l * w_l
  int tile0[5];          // first M tile of level l (tile0[4] = total M tiles)
  int N;                 // queries (= npos[0])
  int q_tiles;           // query tiles of 128
  float scale;
};

struct CgPack {
  int h[4], w[4];
  int row0[4];           // first P row of level l
  int Npad, Ptot;
  int64_t items[5];      // pack work items: [0, items[0]) Q rows; [items[l], items[l+1]) level l (level 0 starts at items[0])
};

// fmap [2N][fm_rs] fp32 -> Q hi/lo [Npad][256] bf16 (fmap1, zero rows beyond N) and
// P hi/lo [Ptot][256] bf16 (level l rows row0[l].. in row-major (y_l, x_l) order = window means of fmap2)
__global__ void __launch_bounds__(256) corr_pack_kernel(const float* __restrict__ fmap, int fm_rs, const CgPack g,
                                                        __nv_bfloat16* __restrict__ Qhi, __nv_bfloat16* __restrict__ Qlo,
                                                        __nv_bfloat16* __restrict__ Phi, __nv_bfloat16* __restrict__ Plo) {
  pdl_enter();
  const int N = g.h[0] * g.w[0];
  // Work items: one per (row, 4 channels) for Q and level 0; for pooled level l the 2^l x 2^l window is split over
  // 2^l adjacent lanes (lane = window row dy, each sums its 2^l values with all loads in flight) and folded with
  // xor-shuffles - a single thread walking the 8x8 window (64 dependent-issue loads) was the long pole of this kernel.
  // Segment sizes are multiples of 32 items, so a warp never straddles two levels.
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < g.items[4]; i += (int64_t)gridDim.x * blockDim.x) {
    const int l = (i >= g.items[1]) + (i >= g.items[2]) + (i >= g.items[3]);
    int64_t j = i;
    int row, dy = 0;
    if (i < g.items[0]) {                      // Q rows
      row = (int)(j >> 6);
    } else {
      j -= (l == 0 ? g.items[0] : l == 1 ? g.items[1] : l == 2 ? g.items[2] : g.items[3]);
      dy = (int)(j & ((1 << l) - 1));
      j >>= l;
      row = g.Npad + (l == 0 ? g.row0[0] : l == 1 ? g.row0[1] : l == 2 ? g.row0[2] : g.row0[3]) + (int)(j >> 6);
    }
    const int c4 = (int)(j & 63);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    __nv_bfloat16 *dh, *dl;
    if (row < g.Npad) {
      if (row < N) v = __ldg(reinterpret_cast<const float4*>(fmap + (int64_t)row * fm_rs + c4 * 4));
      dh = Qhi + (int64_t)row * 256 + c4 * 4;
      dl = Qlo + (int64_t)row * 256 + c4 * 4;
    } else {
      const int r = row - g.Npad;
      const int hl = l == 0 ? g.h[0] : l == 1 ? g.h[1] : l == 2 ? g.h[2] : g.h[3];
      const int wl = l == 0 ? g.w[0] : l == 1 ? g.w[1] : l == 2 ? g.w[2] : g.w[3];
      const int m = r - (l == 0 ? g.row0[0] : l == 1 ? g.row0[1] : l == 2 ? g.row0[2] : g.row0[3]);
      const int win = 1 << l;
      if (m < hl * wl) {                       // uniform over the 2^l lanes of a window
        const int y = m / wl, x = m - y * wl;
        const float* src = fmap + ((int64_t)N + (int64_t)(y * win + dy) * g.w[0] + x * win) * fm_rs + c4 * 4;
        // window mean; equals the successive 2x2 means of corr.py:26 up to fp32 rounding
        float4 t[8];
#pragma unroll
        for (int dx = 0; dx < 8; ++dx)
          t[dx] = dx < win ? __ldg(reinterpret_cast<const float4*>(src + (int64_t)dx * fm_rs)) : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 acc = t[0];
#pragma unroll
        for (int dx = 1; dx < 8; ++dx) { acc.x += t[dx].x; acc.y += t[dx].y; acc.z += t[dx].z; acc.w += t[dx].w; }
        for (int o = 1; o < win; o <<= 1) {    // fold the window rows (lanes dy ^ o)
          acc.x += __shfl_xor_sync(0xffffffffu, acc.x, o); acc.y += __shfl_xor_sync(0xffffffffu, acc.y, o);
          acc.z += __shfl_xor_sync(0xffffffffu, acc.z, o); acc.w += __shfl_xor_sync(0xffffffffu, acc.w, o);
        }
        const float inv = 1.f / (float)(win * win);
        v = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
      }
      dh = Phi + (int64_t)r * 256 + c4 * 4;
      dl = Plo + (int64_t)r * 256 + c4 * 4;
    }
    if (dy != 0) continue;
    const float f[4] = {v.x, v.y, v.z, v.w};
    __nv_bfloat16 h[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      h[e] = __float2bfloat16_rn(f[e]);
      lo[e] = __float2bfloat16_rn(f[e] - __bfloat162float(h[e]));
    }
    *reinterpret_cast<uint2*>(dh) = *reinterpret_cast<const uint2*>(h);
    *reinterpret_cast<uint2*>(dl) = *reinterpret_cast<const uint2*>(lo);
  }
}

// issue one 16-column TMEM load (no wait)
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
}
// wait for the outstanding TMEM loads; the registers are operands so that no use can be scheduled above the wait
__device__ __forceinline__ void tmem_ld_wait2(uint32_t (&a)[16], uint32_t (&b)[16]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]),
                 "+r"(a[8]), "+r"(a[9]), "+r"(a[10]), "+r"(a[11]), "+r"(a[12]), "+r"(a[13]), "+r"(a[14]), "+r"(a[15]),
                 "+r"(b[0]), "+r"(b[1]), "+r"(b[2]), "+r"(b[3]), "+r"(b[4]), "+r"(b[5]), "+r"(b[6]), "+r"(b[7]),
                 "+r"(b[8]), "+r"(b[9]), "+r"(b[10]), "+r"(b[11]), "+r"(b[12]), "+r"(b[13]), "+r"(b[14]), "+r"(b[15])
               :: "memory");
}
// 16 queries x 32 positions: one coalesced 128 B warp store per query
__device__ __forceinline__ void cg_put(float* p, float v) { *p = v; }
__device__ __forceinline__ void cg_put(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
template <typename OutT>
__device__ __forceinline__ void cg_store16(OutT* dst, int64_t npos, int nv, bool m_ok, float scale,
                                           const uint32_t (&hi)[16], const uint32_t (&lo)[16]) {
  if (m_ok) {
#pragma unroll
    for (int e = 0; e < 16; ++e)
      if (e < nv) cg_put(dst + (int64_t)e * npos, (__uint_as_float(hi[e]) + __uint_as_float(lo[e])) * scale);
  }
}

// Persistent: CTA b owns the flat work items [b*T/G, (b+1)*T/G) of the (M tile, query tile) grid, M tile major,
// so the resident P tile changes at most a few times per CTA; it is reloaded k-block by k-block as the last
// tile of the previous segment releases each k-block.
template <typename OutT>
__global__ void __launch_bounds__(CG_THREADS, 1) corr_gemm_kernel(const __grid_constant__ CUtensorMap tmPhi,
                                                                  const __grid_constant__ CUtensorMap tmPlo,
                                                                  const __grid_constant__ CUtensorMap tmQhi,
                                                                  const __grid_constant__ CUtensorMap tmQlo,
                                                                  const CgArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + CG_BAR - 16);
  uint8_t* smA = smem + CG_BAR;
  uint8_t* smB = smA + CG_A_BYTES;
  const uint32_t bar0 = smem_u32(smem);
  auto fullA = [&](uint32_t k) { return bar0 + 8u * k; };
  auto emptyA = [&](uint32_t k) { return bar0 + 8u * (4 + k); };
  auto fullB = [&](uint32_t s) { return bar0 + 8u * (8 + s); };
  auto emptyB = [&](uint32_t s) { return bar0 + 8u * (8 + CG_NB + s); };
  auto accFull = [&](uint32_t s) { return bar0 + 8u * (8 + 2 * CG_NB + s); };
  auto accEmpty = [&](uint32_t s) { return bar0 + 8u * (10 + 2 * CG_NB + s); };
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int k = 0; k < 4; ++k) { mbar_init(fullA(k), 1); mbar_init(emptyA(k), 1); }
    for (int s = 0; s < CG_NB; ++s) { mbar_init(fullB(s), 1); mbar_init(emptyB(s), 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(accFull(s), 1); mbar_init(accEmpty(s), 32 * CG_EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == CG_EPI_WARPS) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int64_t total = (int64_t)a.tile0[4] * a.q_tiles;
  const int item0 = (int)(total * blockIdx.x / gridDim.x), item1 = (int)(total * (blockIdx.x + 1) / gridDim.x);

  if (warp == CG_EPI_WARPS + 1) {
    // ---------------- TMA loader ----------------
    if (lane == 0) {
      tma_prefetch_desc(&tmPhi); tma_prefetch_desc(&tmPlo); tma_prefetch_desc(&tmQhi); tma_prefetch_desc(&tmQlo);
      asm volatile("griddepcontrol.wait;" ::: "memory");
      uint32_t it = 0;
      int cur_mt = -1, nseg = 0;
      for (int item = item0; item < item1; ++item) {
        const int mt = item / a.q_tiles, j = item - mt * a.q_tiles;
        const bool new_seg = mt != cur_mt;
        if (new_seg) { cur_mt = mt; ++nseg; }
        for (int kb = 0; kb < 4; ++kb, ++it) {
          if (new_seg) {
            if (nseg > 1) mbar_wait(emptyA(kb), nseg & 1);          // completion #(nseg-1) -> parity (nseg-2)&1
            mbar_arrive_expect_tx(fullA(kb), 2 * 16384);
            tma_load_2d(smem_u32(smA + kb * 16384), &tmPhi, kb * 64, mt * 128, fullA(kb));
            tma_load_2d(smem_u32(smA + (4 + kb) * 16384), &tmPlo, kb * 64, mt * 128, fullA(kb));
          }
          const uint32_t s = it % CG_NB, ph = (it / CG_NB) & 1;
          mbar_wait(emptyB(s), ph ^ 1);
          mbar_arrive_expect_tx(fullB(s), CG_B_STAGE);
          const uint32_t dst = smem_u32(smB + s * CG_B_STAGE);
          tma_load_2d(dst, &tmQhi, kb * 64, j * 128, fullB(s));
          tma_load_2d(dst + 16384, &tmQlo, kb * 64, j * 128, fullB(s));
        }
      }
    }
  } else if (warp == CG_EPI_WARPS) {
    // ---------------- MMA issue ----------------
    const uint32_t idesc128 = (1u << 4) | (1u << 7) | (1u << 10) | ((128u >> 3) << 17) | (8u << 24);
    const uint32_t idesc256 = (1u << 4) | (1u << 7) | (1u << 10) | ((256u >> 3) << 17) | (8u << 24);
    const uint64_t desc0 = make_desc(0, 1024u);
    const uint32_t dlo0 = (uint32_t)desc0, dhi = (uint32_t)(desc0 >> 32);
    const uint32_t a_hi0 = dlo0 + (smem_u32(smA) >> 4), a_lo0 = a_hi0 + ((4 * 16384) >> 4);
    const uint32_t b0 = dlo0 + (smem_u32(smB) >> 4);
    uint32_t it = 0;
    int cur_mt = -1, nseg = 0;
    for (int item = item0, lt = 0; item < item1; ++item, ++lt) {
      const int mt = item / a.q_tiles;
      const bool new_seg = mt != cur_mt;
      if (new_seg) { cur_mt = mt; ++nseg; }
      const bool last_of_seg = item + 1 < item1 && (item + 1) / a.q_tiles != mt;
      const uint32_t buf = lt & 1, bph = (lt >> 1) & 1;
      mbar_wait(accEmpty(buf), bph ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + buf * 256;
      for (int kb = 0; kb < 4; ++kb, ++it) {
        const uint32_t s = it % CG_NB, ph = (it / CG_NB) & 1;
        if (new_seg) mbar_wait(fullA(kb), (nseg - 1) & 1);
        mbar_wait(fullB(s), ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t ah = a_hi0 + kb * (16384 >> 4), al = a_lo0 + kb * (16384 >> 4);
          const uint32_t bs = b0 + s * (CG_B_STAGE >> 4);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) {
            // columns 0..127: Phi*Qhi (+ Plo*Qhi below), columns 128..255: Phi*Qlo
            tc_mma_bf16_lh(tmem_d, ah + 2 * ks, dhi, bs + 2 * ks, dhi, idesc256, (kb | ks) ? 1u : 0u);
            tc_mma_bf16_lh(tmem_d, al + 2 * ks, dhi, bs + 2 * ks, dhi, idesc128, 1u);
          }
          tc_commit(emptyB(s));
          if (last_of_seg) tc_commit(emptyA(kb));
          if (kb == 3) tc_commit(accFull(buf));
        }
        __syncwarp();
      }
    }
  } else {
    // ------- epilogue: warp w owns TMEM lanes 32(w&3).. and columns 64(w>>2)..; a thread owns one position m -------
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int quad = warp & 3, half = warp >> 2;
    const uint32_t lane_addr = ((uint32_t)(quad * 32) << 16) + (uint32_t)(half * 64);
    int cur_mt = -1;
    int64_t npos = 0;
    bool m_ok = false;
    OutT* plane0 = nullptr;
    for (int item = item0, lt = 0; item < item1; ++item, ++lt) {
      const int mt = item / a.q_tiles, j = item - mt * a.q_tiles;
      if (mt != cur_mt) {
        cur_mt = mt;
        const int l = (mt >= a.tile0[1]) + (mt >= a.tile0[2]) + (mt >= a.tile0[3]);
        // selects instead of dynamic indexing: keeps the kernel parameters out of local memory
        npos = l == 0 ? a.npos[0] : l == 1 ? a.npos[1] : l == 2 ? a.npos[2] : a.npos[3];
        const int t0l = l == 0 ? a.tile0[0] : l == 1 ? a.tile0[1] : l == 2 ? a.tile0[2] : a.tile0[3];
        OutT* const lbase = reinterpret_cast<OutT*>(l == 0 ? a.lvl[0] : l == 1 ? a.lvl[1] : l == 2 ? a.lvl[2] : a.lvl[3]);
        const int m = (mt - t0l) * 128 + quad * 32 + lane;
        m_ok = m < npos;
        plane0 = lbase + (m_ok ? m : 0);
      }
      const int n0 = j * 128 + half * 64;
      const uint32_t buf = lt & 1, bph = (lt >> 1) & 1;
      mbar_wait(accFull(buf), bph);
      tc_fence_after();
      const uint32_t t0 = tmem_base + lane_addr + buf * 256;
      OutT* dst = plane0 + (int64_t)n0 * npos;
      const int nv = a.N - n0;                 // valid queries from n0 on (warp-uniform)
      uint32_t h0[16], l0[16], h1[16], l1[16];
      tmem_ld16_issue(t0, h0);
      tmem_ld16_issue(t0 + 128, l0);
      tmem_ld_wait2(h0, l0);
      tmem_ld16_issue(t0 + 16, h1);
      tmem_ld16_issue(t0 + 128 + 16, l1);
      cg_store16(dst, npos, nv, m_ok, a.scale, h0, l0);
      tmem_ld_wait2(h1, l1);
      tmem_ld16_issue(t0 + 32, h0);
      tmem_ld16_issue(t0 + 128 + 32, l0);
      cg_store16(dst + 16 * npos, npos, nv - 16, m_ok, a.scale, h1, l1);
      tmem_ld_wait2(h0, l0);
      tmem_ld16_issue(t0 + 48, h1);
      tmem_ld16_issue(t0 + 128 + 48, l1);
      cg_store16(dst + 32 * npos, npos, nv - 32, m_ok, a.scale, h0, l0);
      tmem_ld_wait2(h1, l1);
      tc_fence_before();
      mbar_arrive(accEmpty(buf));              // all TMEM reads of this accumulator buffer are done
      cg_store16(dst + 48 * npos, npos, nv - 48, m_ok, a.scale, h1, l1);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == CG_EPI_WARPS) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
  }
}

}  // namespace zt

namespace {
struct CgPlan {
  zt::CgPack pk;
  int tile0[5];
  int q_tiles;
};
CgPlan cg_plan(int h0, int w0) {
  CgPlan p;
  p.pk.h[0] = h0; p.pk.w[0] = w0;
  for (int i = 1; i < 4; ++i) { p.pk.h[i] = p.pk.h[i - 1] / 2; p.pk.w[i] = p.pk.w[i - 1] / 2; }
  int row = 0;
  for (int i = 0; i < 4; ++i) {
    p.pk.row0[i] = row;
    p.tile0[i] = row / 128;
    row += zt::ceil_div(p.pk.h[i] * p.pk.w[i], 128) * 128;
  }
  p.tile0[4] = row / 128;
  p.pk.Ptot = row;
  p.pk.Npad = zt::ceil_div(h0 * w0, 128) * 128;
  p.q_tiles = p.pk.Npad / 128;
  p.pk.items[0] = (int64_t)p.pk.Npad * 64;
  for (int i = 0; i < 4; ++i) {
    const int rows = (i < 3 ? p.pk.row0[i + 1] : p.pk.Ptot) - p.pk.row0[i];
    p.pk.items[i + 1] = p.pk.items[i] + ((int64_t)rows * 64 << i);
  }
  return p;
}
}  // namespace

// scratch bytes needed by zt_corr_pyramid_f32 for an h0 x w0 feature grid
extern "C" int64_t zt_corr_pyramid_scratch_bytes(int h0, int w0) {
  if (h0 <= 0 || w0 <= 0 || (int64_t)h0 * w0 >= (1 << 24)) return 0;
  const CgPlan p = cg_plan(h0, w0);
  return (int64_t)(2 * p.pk.Npad + 2 * p.pk.Ptot) * 256 * 2;
}

static int corr_pyramid_any(const float* fmap, int fm_rs, int h0, int w0, void* l0, void* l1, void* l2, void* l3,
                            bool out_bf16, void* scratch, int64_t scratch_bytes, void* stream) {
  using namespace zt;
  ZT_REQUIRE(fmap && l0 && l1 && l2 && l3 && scratch && h0 >= 8 && w0 >= 8 && fm_rs >= 256 && (fm_rs & 3) == 0,
             "corr_pyramid: bad args (grid %dx%d, row stride %d)", h0, w0, fm_rs);
  ZT_REQUIRE((int64_t)h0 * w0 < (1 << 24), "corr_pyramid: grid %dx%d too large", h0, w0);
  ZT_REQUIRE(scratch_bytes >= zt_corr_pyramid_scratch_bytes(h0, w0) && ((uintptr_t)scratch & 127) == 0,
             "corr_pyramid: scratch too small or unaligned (%lld bytes given, %lld needed)", (long long)scratch_bytes,
             (long long)zt_corr_pyramid_scratch_bytes(h0, w0));
  cudaStream_t st = (cudaStream_t)stream;
  const CgPlan p = cg_plan(h0, w0);
  const int Npad = p.pk.Npad, Ptot = p.pk.Ptot;
  __nv_bfloat16* Qhi = reinterpret_cast<__nv_bfloat16*>(scratch);
  __nv_bfloat16* Qlo = Qhi + (int64_t)Npad * 256;
  __nv_bfloat16* Phi = Qlo + (int64_t)Npad * 256;
  __nv_bfloat16* Plo = Phi + (int64_t)Ptot * 256;

  const int64_t pack_items = p.pk.items[4];
  const int pack_blocks = (int)std::min<int64_t>(ceil_div64(pack_items, 256), 8 * kNumSMs);
  ZT_CUDA(launch_pdl(corr_pack_kernel, dim3(pack_blocks), dim3(256), 0, st, fmap, fm_rs, p.pk, Qhi, Qlo, Phi, Plo));
  if (int rc = check_launch("corr_pack_kernel")) return rc;

  CUtensorMap tm[4];
  const __nv_bfloat16* bases[4] = {Phi, Plo, Qhi, Qlo};
  for (int i = 0; i < 4; ++i) {
    const int dims[2] = {256, i < 2 ? Ptot : Npad};
    const int64_t strides[1] = {256};
    const int box[2] = {64, 128};
    if (int rc = tma_map_bf16(bases[i], 2, dims, strides, box, &tm[i])) return rc;
  }
  CgArgs a;
  a.lvl[0] = l0; a.lvl[1] = l1; a.lvl[2] = l2; a.lvl[3] = l3;
  for (int i = 0; i < 4; ++i) { a.npos[i] = p.pk.h[i] * p.pk.w[i]; a.tile0[i] = p.tile0[i]; }
  a.tile0[4] = p.tile0[4];
  a.N = h0 * w0; a.q_tiles = p.q_tiles; a.scale = 0.0625f;
  const int64_t items = (int64_t)p.tile0[4] * p.q_tiles;
  const int grid = (int)std::min<int64_t>(items, kNumSMs);            // persistent: one CTA per SM
  const size_t smem = CG_BAR + CG_A_BYTES + CG_NB * CG_B_STAGE;
  static bool attr_done = false;
  if (!attr_done) {
    ZT_CUDA(cudaFuncSetAttribute(corr_gemm_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    ZT_CUDA(cudaFuncSetAttribute(corr_gemm_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  if (out_bf16)
    ZT_CUDA(launch_pdl(corr_gemm_kernel<__nv_bfloat16>, dim3(grid), dim3(CG_THREADS), smem, st, tm[0], tm[1], tm[2], tm[3], a));
  else
    ZT_CUDA(launch_pdl(corr_gemm_kernel<float>, dim3(grid), dim3(CG_THREADS), smem, st, tm[0], tm[1], tm[2], tm[3], a));
  return check_launch("corr_gemm_kernel");
}

extern "C" int zt_corr_pyramid_f32(const float* fmap, int fm_rs, int h0, int w0, float* l0, float* l1, float* l2,
                                   float* l3, void* scratch, int64_t scratch_bytes, void* stream) {
  return corr_pyramid_any(fmap, fm_rs, h0, w0, l0, l1, l2, l3, false, scratch, scratch_bytes, stream);
}
/* bf16-storage variant (BASELINE.json configs[4] "fp32 vs bf16 corr"): same GEMM (fp32-grade products, fp32
 * accumulation), the four levels are rounded to bf16 on the way out - half the bytes of the output-bound kernel */
extern "C" int zt_corr_pyramid_bf16(const float* fmap, int fm_rs, int h0, int w0, void* l0, void* l1, void* l2,
                                    void* l3, void* scratch, int64_t scratch_bytes, void* stream) {
  return corr_pyramid_any(fmap, fm_rs, h0, w0, l0, l1, l2, l3, true, scratch, scratch_bytes, stream);
}
