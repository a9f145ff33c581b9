"""GPU tool: a few 1080p frames through Finetunemodel (new-sequence frame + steady frames) for ncu."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                           # noqa: E402
from zero_tig_b200 import synth                        # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
H, W = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1080, 1920)
dev = torch.device("cuda")
model, sd = bench.build_model(dev)
clip = synth.synthetic_clip(n, H, W, seed=0).to(dev)
for t in range(n):
    model.is_new_seq = (t == 0)
    model(clip[t:t + 1])
torch.cuda.synchronize()
print("ok")
