"""GPU tool: the oracle port of the reference (plain torch functional ops = cuDNN / cuBLAS eager kernels) on the B200 itself,
TF32 off (the parity setting) and on (fastest eager) - BASELINE.md section 3 "GPU eager baseline (same box)".
    python tools/eager_gpu_baseline.py [frames=6] [device=cuda]
The reference itself does not travel to the GPU box; its restatement runs the same torch ops in the same order."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import zt_oracle as O                                   # noqa: E402
import bench                                            # noqa: E402
from zero_tig_b200 import synth                         # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 6
dev = torch.device(sys.argv[2] if len(sys.argv) > 2 else "cuda")
H, W = (1080, 1920) if dev.type == "cuda" else (390, 400)
sd = {k: v.to(dev) for k, v in bench.synthetic_weights().items()}
clip = synth.synthetic_clip_mt(n + 2, H, W, seed=0).to(dev)


def sync():
    if dev.type == "cuda":
        torch.cuda.synchronize()


for tf32 in (False, True):
    torch.backends.cudnn.allow_tf32 = tf32
    torch.backends.cuda.matmul.allow_tf32 = tf32
    with torch.no_grad():
        st = O.InferState()
        O.finetune_forward(clip[0:1], sd, st, True, 3)          # new-sequence frame
        O.finetune_forward(clip[1:2], sd, st, False, 3)         # warm steady frame (cuDNN autotune, allocations)
        sync()
        t0 = time.perf_counter()
        for t in range(2, 2 + n):
            O.finetune_forward(clip[t:t + 1], sd, st, False, 3)
        sync()
        dt = (time.perf_counter() - t0) / n
    print("oracle port in torch eager on %s, %dx%d, TF32 %s: %.1f ms per steady frame = %.2f frames/s"
          % (dev, W, H, "on" if tf32 else "off", dt * 1e3, 1.0 / dt))
