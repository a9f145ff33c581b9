"""GPU tool: CPU enqueue time vs GPU time of the steady-state 1080p frame loop."""
import os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from zero_tig_b200 import synth
dev = torch.device("cuda")
model, sd = bench.build_model(dev)
clip = synth.synthetic_clip(8, 1080, 1920, seed=0).to(dev)
model.is_new_seq = True
model(clip[0:1])
for i in range(5):
    model.is_new_seq = False
    model(clip[1 + i % 7:2 + i % 7])
torch.cuda.synchronize()
n = 30
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for i in range(n):
    model.is_new_seq = False
    model(clip[1 + i % 7:2 + i % 7])
e1.record()
t_cpu = time.perf_counter() - t0
torch.cuda.synchronize()
print("cpu enqueue %.2f ms/frame, gpu %.2f ms/frame" % (t_cpu / n * 1e3, e0.elapsed_time(e1) / n))
