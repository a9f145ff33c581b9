"""GPU tool: in-situ time of the RAFT pass as a function of the iteration count (CUDA graph replays, 1080p / of_scale 3):
the slope is the cost of one update iteration, the intercept the encoders + correlation pyramid + upsampling."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                           # noqa: E402

dev = torch.device("cuda")
model, sd = bench.build_model(dev)
eng = model.raft.engine()
hp, wp = 360, 640
x2 = torch.randn(2, hp, wp, 4, device=dev).clamp_(-1, 1)
res = {}
for iters in (0, 2, 6, 12):
    for _ in range(3):
        eng.forward_packed_static(x2, hp, wp, iters)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        eng.forward_packed_static(x2, hp, wp, iters)
    e1.record()
    torch.cuda.synchronize()
    res[iters] = e0.elapsed_time(e1) / 20
    print("iters %2d: %.3f ms per RAFT pass" % (iters, res[iters]), flush=True)
print("per iteration %.1f us, fixed part %.3f ms" % ((res[12] - res[2]) / 10 * 1e3, res[0]))
