"""GPU tool: cost of one dependent launch in a CUDA graph with programmatic dependent launch (the floor under every
kernel of the RAFT update loop): a chain of 200 trivial flow_update launches (29 CTAs) replayed as a graph."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from zero_tig_b200 import ops          # noqa: E402

h8, w8 = 45, 80
N = h8 * w8
coords = torch.zeros(N, 2, device="cuda")
flow = torch.zeros(N, 4, device="cuda")
hx = torch.zeros(N, 384, device="cuda")


def chain(n):
    for _ in range(n):
        ops.flow_update(coords, None, 0, flow, 4, hx, 384, 382, h8, w8)


chain(3)
torch.cuda.synchronize()
side = torch.cuda.Stream()
g = torch.cuda.CUDAGraph()
with torch.cuda.stream(side):
    with torch.cuda.graph(g, stream=side):
        chain(200)
g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    g.replay()
e1.record()
torch.cuda.synchronize()
print("dependent launch in a graph (PDL %s): %.2f us per kernel" % ("off" if os.environ.get("ZT_NO_PDL") else "on",
                                                                    e0.elapsed_time(e1) * 1e3 / 2000))
