"""configs[4]: RAFT correlation volume + pyramid, pyramid lookup and grid_sample warp micro-benchmark sweep
(frame sizes 256^2 .. 4K at of_scale 3, plus of_scale 1 for the small frames) against the HBM roofline.
Algorithmic bytes per SURVEY 8d: volume+pyramid = 4*N^2*(1+1/4+1/16+1/64) written + 2*N*256*4 read;
lookup = N*(324*4 + 4*100*4 + 8) per iteration; warp pair = flow + 2 planes read + 2 planes written.
fp32 arm = the reference's storage; bf16 arm = zt_corr_pyramid_bf16 / zt_corr_lookup_bf16 (levels rounded to bf16 on
the way out: 2*N^2*(1+1/4+1/16+1/64) bytes written).
    python tools/corr_warp_sweep.py > profiles/r02_corr_warp_sweep.txt
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from zero_tig_b200 import ops          # noqa: E402

PEAK = 6547.2
p = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(p):
    PEAK = json.load(open(p))["hbm_gbs"]


def timed(fn, iters=10):
    """GPU time per call: `iters` calls captured in one CUDA graph (no CPU enqueue time in the number), replayed 5x."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        fn()
        side.synchronize()
        with torch.cuda.graph(graph, stream=side):
            for _ in range(iters):
                fn()
    graph.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (5 * iters)       # us


def main():
    g = torch.Generator().manual_seed(0)
    print("# HBM peak %.1f GB/s (measured copy bandwidth); times are CUDA-event averages, us" % PEAK)
    print("# corr3 = pack + conv_tc GEMM + 3 avgpool kernels (ZT_CORR_FUSED=0); corr = zt_corr_pyramid_f32 (pack + one tcgen05 kernel)")
    print("%-22s %7s | %9s %9s %7s %6s | %9s %7s %6s | %9s %7s %6s | %9s %7s %6s | %9s" % (
        "frame (of_scale)", "N", "corr3 us", "corr us", "GB/s", "frac", "lookup us", "GB/s", "frac", "warp us", "GB/s", "frac",
        "corr16 us", "GB/s", "frac", "lookup16"))
    cases = [(256, 256, 1), (512, 512, 1), (540, 960, 3), (1080, 1920, 3), (1080, 1920, 1), (2160, 3840, 3)]
    for H, W, s in cases:
        ht, wd = H // s, W // s
        hp, wp = (ht + 7) // 8 * 8, (wd + 7) // 8 * 8
        h8, w8 = hp // 8, wp // 8
        N = h8 * w8
        hs, ws_ = [h8], [w8]
        for _ in range(3):
            hs.append(hs[-1] // 2)
            ws_.append(ws_[-1] // 2)
        if hs[3] < 2 or ws_[3] < 2:
            print("%-22s %7d | grid too small for the 4-level pyramid (reference yields NaN, SURVEY F8)" % ("%dx%d (%d)" % (H, W, s), N))
            continue
        fmap = torch.randn(2 * N, 256, generator=g).cuda()
        lv = [torch.empty(N, hs[i] * ws_[i], device="cuda") for i in range(4)]
        cw = ops.RowsAsWeights(N, 256, 128, "cuda")

        def corr():
            cw.pack(fmap[N:], 256)
            ops.conv2d(cw, fmap, 256, 256, 1, h8, w8, lv[0], N, scale=0.0625, tc=3, bn=128)
            for i in range(3):
                ops.avgpool2(lv[i], lv[i + 1], N, hs[i], ws_[i])
        t_c3 = timed(corr)
        scratch = torch.empty(ops.corr_pyramid_scratch_bytes(h8, w8) // 4, device="cuda")
        t_c = timed(lambda: ops.corr_pyramid(fmap, 256, h8, w8, lv, scratch))
        b_c = 4.0 * N * sum(hs[i] * ws_[i] for i in range(4)) + 2.0 * N * 256 * 4
        coords = (torch.stack(torch.meshgrid(torch.arange(h8), torch.arange(w8), indexing="ij")[::-1], -1).float().reshape(N, 2)
                  + torch.randn(N, 2, generator=g) * 2).cuda().contiguous()
        out = torch.empty(N, 324, device="cuda")
        t_l = timed(lambda: ops.corr_lookup(lv, hs, ws_, coords, out, 324, N), 20)
        b_l = N * (324 * 4 + 4 * 100 * 4 + 8.0)
        flow = (torch.randn(1, 2, hp, wp, generator=g)).cuda()
        a, b = torch.rand(1, 3, H, W, generator=g).cuda(), torch.rand(1, 3, H, W, generator=g).cuda()
        oa, ob = torch.empty_like(a), torch.empty_like(b)
        t_w = timed(lambda: ops.warp_pair(flow, hp, wp, a, b, oa, ob, 3, H, W))
        b_w = 2.0 * hp * wp * 4 + 4.0 * 3 * H * W * 4
        r = lambda byt, us: (byt / us / 1e3, byt / us / 1e3 / PEAK)
        # bf16-storage arm
        lb = [torch.empty(N, hs[i] * ws_[i], device="cuda", dtype=torch.bfloat16) for i in range(4)]
        t_c16 = timed(lambda: ops.corr_pyramid(fmap, 256, h8, w8, lb, scratch))
        b_c16 = 2.0 * N * sum(hs[i] * ws_[i] for i in range(4)) + 2.0 * N * 256 * 4
        t_l16 = timed(lambda: ops.corr_lookup(lb, hs, ws_, coords, out, 324, N), 20)
        print("%-22s %7d | %9.1f %9.1f %7.0f %6.3f | %9.1f %7.0f %6.3f | %9.1f %7.0f %6.3f | %9.1f %7.0f %6.3f | %9.1f" % (
            "%dx%d (%d)" % (H, W, s), N, t_c3, t_c, *r(b_c, t_c), t_l, *r(b_l, t_l), t_w, *r(b_w, t_w),
            t_c16, *r(b_c16, t_c16), t_l16), flush=True)
        del fmap, lv, lb, cw, scratch
        torch.cuda.empty_cache()
    print("# corr / lookup = fp32 storage (the reference's layout); corr16 / lookup16 = bf16 storage (same GEMM, levels rounded on "
          "the way out).  Timed with CUDA-graph replay of back-to-back calls: inputs of the small sizes are L2-resident.")


if __name__ == "__main__":
    main()
