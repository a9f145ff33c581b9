"""GPU tool for ncu: warm up, then run N steady-state frames inside a cudaProfilerStart/Stop window
(use `ncu --profile-from-start off ...`), so the launch list holds exactly the kernels of those frames.
    python tools/prof_steady.py [frames H W]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                           # noqa: E402
from zero_tig_b200 import synth                        # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
H, W = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1080, 1920)
dev = torch.device("cuda")
model, sd = bench.build_model(dev)
clip = synth.synthetic_clip_mt(4 + n, H, W, seed=0).to(dev)
for t in range(4):                       # new-sequence frame, eager frame, RAFT graph capture, first replay
    model.is_new_seq = (t == 0)
    model(clip[t:t + 1])
torch.cuda.synchronize()
torch.cuda.profiler.start()
for t in range(4, 4 + n):
    model.is_new_seq = False
    model(clip[t:t + 1])
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
