"""GPU micro-benchmark: warp_pair at 1080p (flow 360x640), CUDA events, inputs larger than L2 rotate."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from zero_tig_b200 import ops
H, W, hf, wf = 1080, 1920, 360, 640
g = torch.Generator().manual_seed(0)
if os.environ.get("WARP_FLOW", "random") == "smooth":     # what RAFT returns on a real clip: ~1 px, slowly varying
    yy, xx = torch.meshgrid(torch.arange(hf, dtype=torch.float32), torch.arange(wf, dtype=torch.float32), indexing="ij")
    flow = torch.stack([0.7 + 0.5 * torch.sin(xx / 31.0) * torch.cos(yy / 47.0),
                        0.35 + 0.4 * torch.cos(xx / 53.0 + yy / 29.0)])[None].contiguous().cuda()
else:
    flow = (torch.randn(1, 2, hf, wf, generator=g) * 3).cuda()
sets = [(torch.rand(1, 3, H, W, generator=g).cuda(), torch.rand(1, 3, H, W, generator=g).cuda(),
         torch.empty(1, 3, H, W).cuda(), torch.empty(1, 3, H, W).cuda()) for _ in range(4)]   # 4 x 100 MB > L2
for _ in range(3):
    for a, b, oa, ob in sets:
        ops.warp_pair(flow, hf, wf, a, b, oa, ob, 3, H, W)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
n = 0
for _ in range(10):
    for a, b, oa, ob in sets:
        ops.warp_pair(flow, hf, wf, a, b, oa, ob, 3, H, W)
        n += 1
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / n
byt = 2 * hf * wf * 4 + 4 * 3 * H * W * 4
print("warp_pair %.1f us  %.0f GB/s (algorithmic %.1f MB)" % (us, byt / us / 1e3, byt / 1e6))
