"""GPU tool: wall clock of predict.py's per-frame loop (decode + resize -> model -> two PNGs) on a synthetic RLV tree,
(a) serially, the way predict.py does it, (b) through clip_io.ClipRunner (threads + FramePipeline).
    python tools/clip_io_bench.py [frames=32] [src_H=540] [src_W=960]
Source PNGs are src_H x src_W and are resized to 1920x1080 by the loader, as the RLV loader does."""
import os
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                           # noqa: E402
from zero_tig_b200 import clip_io, synth               # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
sh, sw = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (540, 960)
from PIL import Image                                  # noqa: E402

tmp = tempfile.mkdtemp(prefix="zt_clipio_")
d = os.path.join(tmp, "input", "S01", "low_light_10")
os.makedirs(d)
clip = synth.synthetic_clip_mt(n, sh, sw, seed=0)
paths = []
for t in range(n):
    p = os.path.join(d, "%05d.png" % (t + 1))
    Image.fromarray((clip[t].permute(1, 2, 0).numpy() * 255.0 + 0.5).astype(np.uint8)).save(p)
    paths.append(p)
dev = torch.device("cuda")
model, _ = bench.build_model(dev)
H, W = 1080, 1920


def serial(out, paths=paths):
    """predict.py:75-104 as written: one frame at a time on one thread."""
    last = paths[0]
    for p in paths:
        x = torch.from_numpy(clip_io.decode_frame(p)).permute(2, 0, 1).float().div(255)[None].to(dev)
        model.is_new_seq = clip_io.is_new_sequence(p, last)
        last = p
        H2, H3, _ = model(x)
        a, b = clip_io.output_paths(p, out)
        for t, f in ((H3, a), (H2, b)):
            im = np.clip(np.transpose(t[0].cpu().float().numpy(), (1, 2, 0)) * 255.0, 0, 255.0).astype("uint8")
            clip_io.encode_png(im, f)


serial(os.path.join(tmp, "warm"), paths[:3])           # warm-up: allocations, weight packing
t0 = time.perf_counter()
serial(os.path.join(tmp, "serial"))
t_serial = time.perf_counter() - t0
runner = clip_io.ClipRunner.for_model(model, os.path.join(tmp, "warm2"))
runner.run(paths[:4])                                  # warm-up: graph capture
runner.save_root = os.path.join(tmp, "piped")
t0 = time.perf_counter()
runner.run(paths)
t_pipe = time.perf_counter() - t0
same = all(open(os.path.join(tmp, "serial", "S01", "low_light_10", f), "rb").read() ==
           open(os.path.join(tmp, "piped", "S01", "low_light_10", f), "rb").read()
           for f in sorted(os.listdir(os.path.join(tmp, "serial", "S01", "low_light_10"))))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
x = torch.from_numpy(clip_io.decode_frame(paths[1])).permute(2, 0, 1).float().div(255)[None].to(dev)
model.is_new_seq = False
e0.record()
for _ in range(n):
    model(x)
e1.record()
torch.cuda.synchronize()
print("host cores %d, %d frames %dx%d -> 1920x1080" % (os.cpu_count(), n, sw, sh))
print("serial loop (predict.py order)        : %.2f s  = %.1f frames/s" % (t_serial, n / t_serial))
print("ClipRunner (%d decode + %d encode thr): %.2f s  = %.1f frames/s" % (runner.decode_threads, runner.encode_threads,
                                                                          t_pipe, n / t_pipe))
print("module alone (device-resident frames) : %.1f frames/s" % (n / (e0.elapsed_time(e1) / 1e3)))
print("PNG files byte-identical between the two loops: %s" % same)
