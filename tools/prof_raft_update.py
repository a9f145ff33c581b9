"""GPU tool for ncu: ONE launch of the persistent RAFT update kernel at 1080p / of_scale 3 inside a
cudaProfilerStart/Stop window (ncu --profile-from-start off -k regex:raft_update)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                           # noqa: E402

hp, wp = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (360, 640)
dev = torch.device("cuda")
model, sd = bench.build_model(dev)
eng = model.raft.engine()
x2 = torch.randn(2, hp, wp, 4, device=dev).clamp_(-1, 1)
for _ in range(2):
    eng.forward_packed(x2, hp, wp, 12)
torch.cuda.synchronize()
torch.cuda.profiler.start()
eng.forward_packed(x2, hp, wp, 12)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
