"""One fused correlation-pyramid call at a given feature grid (for ncu captures): python tools/corr_one.py 90 160"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from zero_tig_b200 import ops          # noqa: E402

h8, w8 = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
N = h8 * w8
hs, ws = [h8], [w8]
for _ in range(3):
    hs.append(hs[-1] // 2)
    ws.append(ws[-1] // 2)
fmap = torch.randn(2 * N, 256, device="cuda")
lv = [torch.empty(N, hs[i] * ws[i], device="cuda") for i in range(4)]
scratch = torch.empty(ops.corr_pyramid_scratch_bytes(h8, w8) // 4, device="cuda")
for _ in range(reps):
    ops.corr_pyramid(fmap, 256, h8, w8, lv, scratch)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    ops.corr_pyramid(fmap, 256, h8, w8, lv, scratch)
e1.record()
torch.cuda.synchronize()
byt = 4.0 * N * sum(hs[i] * ws[i] for i in range(4))
t = e0.elapsed_time(e1) / reps
print("grid %dx%d N=%d: %.1f us, %.0f GB/s written" % (h8, w8, N, t * 1e3, byt / t / 1e6))
