"""GPU tool (python tools/two_shards.py S K [H W]): S independent frame-range shards of a 1080p clip on ONE GPU, one model instance + one stream per
shard, frames enqueued round-robin from one host thread.  Prints frames/s against the single-shard loop."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                           # noqa: E402
from zero_tig_b200 import synth                        # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 2
K = int(sys.argv[2]) if len(sys.argv) > 2 else 30
H, W = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (1080, 1920)
NF = 12 if H <= 1080 else 6
dev = torch.device("cuda")
models, clips, streams = [], [], []
for s in range(S):
    m, _ = bench.build_model(dev)
    models.append(m)
    clips.append(synth.synthetic_clip(NF, H, W, seed=s, underwater=(H > 1080)).to(dev))
    streams.append(torch.cuda.Stream())
torch.cuda.synchronize()


def run(n, shards):
    for i in range(n):
        for s in shards:
            with torch.cuda.stream(streams[s]):
                models[s].is_new_seq = False
                models[s](clips[s][1 + i % (NF - 1):2 + i % (NF - 1)])


for s in range(S):
    with torch.cuda.stream(streams[s]):
        models[s].is_new_seq = True
        models[s](clips[s][0:1])
run(4, range(S))
torch.cuda.synchronize()
for shards in ([0], list(range(S))):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run(K, shards)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print("%dx%d shards on the GPU: %d  %.1f frames/s (%.2f ms per frame)" % (W, H, len(shards), K * len(shards) / dt, 1e3 * dt / (K * len(shards))))
