"""Micro-benchmark of the conv kernels on the layer shapes of the 1080p hot path (CUDA events)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zero_tig_b200 import ops

CASES = {
    # name: cin, cout, kh, kw, pad, H, W, B, epi
    "enh_blk": (64, 64, 3, 3, (1, 1), 1080, 1920, 1, "add"),
    "d_c2": (48, 48, 3, 3, (1, 1), 1080, 1920, 1, "plain"),
    "e_in": (12, 64, 3, 3, (1, 1), 1080, 1920, 1, "plain"),
    "e_out": (64, 3, 3, 3, (1, 1), 1080, 1920, 1, "plain"),
    "d_c3": (48, 6, 1, 1, (0, 0), 1080, 1920, 1, "plain"),
    "gru_zr": (384, 256, 1, 5, (0, 2), 45, 80, 1, "plain"),
    "convc1": (324, 256, 1, 1, (0, 0), 45, 80, 1, "plain"),
    "fh1": (128, 256, 3, 3, (1, 1), 45, 80, 1, "plain"),
    "k1_512": (512, 64, 1, 1, (0, 0), 540, 960, 1, "plain"),
    "k1x3_256": (256, 64, 1, 3, (0, 1), 540, 960, 1, "plain"),
    "enc_l1": (64, 64, 3, 3, (1, 1), 180, 320, 2, "plain"),
    "gru_zr256": (256, 256, 1, 5, (0, 2), 45, 80, 1, "plain"),
    "gru_v256": (256, 256, 5, 1, (2, 0), 45, 80, 1, "plain"),
    "convc2": (256, 192, 3, 3, (1, 1), 45, 80, 1, "plain"),
}


def run(name, tc, bn=None, iters=10):
    cin, cout, kh, kw, pad, H, W, B, epi = CASES[name]
    cs = (cin + 3) // 4 * 4
    x = torch.randn(B * H * W, cs, device="cuda")
    w = torch.randn(cout, cin, kh, kw, device="cuda") * 0.05
    wt = ops.ConvW(w, torch.zeros(cout, device="cuda"))
    ocs = (cout + 3) // 4 * 4
    out = torch.empty(B * H * W, ocs, device="cuda")
    kw_ = dict(pad=pad, act="relu", tc=tc, bn=bn)
    if epi == "add":
        kw_.update(epi="add", aux0=x, aux0_cs=cs)
    for _ in range(3):
        ops.conv2d(wt, x, cs, wt.cin, B, H, W, out, ocs, **kw_)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.conv2d(wt, x, cs, wt.cin, B, H, W, out, ocs, **kw_)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    gf = 2.0 * B * H * W * cout * cin * kh * kw / 1e9
    print("%-8s tc=%d bn=%s  %8.3f ms  %8.1f TFLOP/s (useful)  io %.1f GB/s" % (
        name, tc, bn, ms, gf / ms, (x.numel() + out.numel() * (2 if epi == "add" else 1)) * 4 / ms / 1e6), flush=True)


if __name__ == "__main__":
    names = sys.argv[1].split(",") if len(sys.argv) > 1 else list(CASES)
    modes = [int(m) for m in sys.argv[2].split(",")] if len(sys.argv) > 2 else [3, 1, 0]
    for n in names:
        for tc in modes:
            run(n, tc)
