"""GPU micro-benchmark: the RAFT update-block convolutions at M = 3600 (45x80), CUDA events."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from zero_tig_b200 import ops

H, W = 45, 80
N = H * W
g = torch.Generator().manual_seed(0)


def bench(name, fn, iters=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    t_cpu = (time.perf_counter() - t0) / iters * 1e6
    torch.cuda.synchronize()
    print("%-34s %7.1f us/launch (gpu)  %6.1f us/launch (cpu enqueue)" % (name, e0.elapsed_time(e1) * 1e3 / iters, t_cpu), flush=True)


hx = torch.randn(N, 384, generator=g).cuda()
rh = torch.randn(N, 128, generator=g).cuda()
z = torch.empty(N, 128).cuda()
out = torch.empty(N, 384).cuda()
P = torch.randn(N, 256, generator=g).cuda()
for cin, kh, kw, pad in ((384, 1, 5, (0, 2)), (256, 1, 5, (0, 2)), (128, 1, 5, (0, 2)), (384, 5, 1, (2, 0)), (384, 1, 1, (0, 0)),
                         (64, 1, 1, (0, 0))):
    for cout, bn in ((256, 64), (256, 128), (128, 32), (128, 64)):
        w = ops.ConvW((torch.randn(cout, cin, kh, kw, generator=g) * 0.02).cuda(), None)
        o = torch.empty(N, cout).cuda()
        bench("plain cin=%d %dx%d cout=%d bn=%d" % (cin, kh, kw, cout, bn),
              lambda: ops.conv2d(w, hx, 384, cin, 1, H, W, o, cout, pad=pad, tc=3, bn=bn))
w = ops.ConvW((torch.randn(256, 256, 1, 5, generator=g) * 0.02).cuda(), None)
bench("gru_zr 128+128 bplane", lambda: ops.conv2d(w, hx, 384, 128, 1, H, W, rh, 128, in1=hx, in1_cs=384, c1=128, in1_off=256,
                                                   pad=(0, 2), epi="gru_zr", aux0=hx, aux0_cs=384, aux1=z, aux1_cs=128, tc=3,
                                                   bn=64, bplane=P, bplane_cs=256))
w2 = ops.ConvW((torch.randn(256, 256, 1, 5, generator=g) * 0.02).cuda(), None)
o256 = torch.empty(N, 256).cuda()
bench("plain two-source 128+128", lambda: ops.conv2d(w2, hx, 384, 128, 1, H, W, o256, 256, in1=hx, in1_cs=384, c1=128,
                                                      in1_off=256, pad=(0, 2), tc=3, bn=64))
bench("plain single 256 (cs 384)", lambda: ops.conv2d(w2, hx, 384, 256, 1, H, W, o256, 256, pad=(0, 2), tc=3, bn=64))
bench("gru_zr single 256 no bplane", lambda: ops.conv2d(w2, hx, 384, 256, 1, H, W, rh, 128, pad=(0, 2), epi="gru_zr", aux0=hx,
                                                         aux0_cs=384, aux1=z, aux1_cs=128, tc=3, bn=64))
bench("relu single 256", lambda: ops.conv2d(w2, hx, 384, 256, 1, H, W, o256, 256, pad=(0, 2), act="relu", tc=3, bn=64))
bench("sigmoid single 256", lambda: ops.conv2d(w2, hx, 384, 256, 1, H, W, o256, 256, pad=(0, 2), act="sigmoid", tc=3, bn=64))
wq = ops.ConvW((torch.randn(128, 256, 1, 5, generator=g) * 0.02).cuda(), None)
hq = torch.empty(N, 128).cuda()
bench("gru_q single 256 bn32", lambda: ops.conv2d(wq, hx, 384, 256, 1, H, W, hq, 128, pad=(0, 2), epi="gru_q", aux0=hx, aux0_cs=384,
                                                   aux1=z, aux1_cs=128, tc=3, bn=32))
bench("plain 256->128 bn32", lambda: ops.conv2d(wq, hx, 384, 256, 1, H, W, hq, 128, pad=(0, 2), tc=3, bn=32))
