"""GPU tool: per-phase, per-role timeline of the persistent RAFT update kernel (csrc/raft_update.cu) at 1080p /
of_scale 3.  One CTA stores %globaltimer at: the grid-barrier exit (phase boundary), the release of its activation
loader, the arrival of its first A stage, its last MMA, the completion of its accumulator, the end of its epilogue
stores and its arrival at the barrier.
    python tools/raft_update_trace.py [cta] [hp wp]"""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                                           # noqa: E402
from zero_tig_b200 import _lib                         # noqa: E402

cta = int(sys.argv[1]) if len(sys.argv) > 1 else 0
hp, wp = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (360, 640)
dev = torch.device("cuda")
model, sd = bench.build_model(dev)
eng = model.raft.engine()
x2 = torch.randn(2, hp, wp, 4, device=dev).clamp_(-1, 1)
iters = 12
for _ in range(3):
    eng.forward_packed(x2, hp, wp, iters)
torch.cuda.synchronize()
nph = 1 + 9 * iters
buf = torch.zeros(8 * (nph + 1), dtype=torch.int64, device=dev)
_lib.lib().zt_debug_set_raft_update_trace(C.c_void_p(buf.data_ptr()), cta)
runs, acc = 5, None
for _ in range(runs):
    buf.zero_()
    eng.forward_packed(x2, hp, wp, iters)
    torch.cuda.synchronize()
    t = buf.cpu().double().reshape(-1, 8) / 1e3      # us
    start = torch.cat([t[0:1, 7], t[:-1, 0]])          # phase start = previous barrier exit (kernel start for phase 0)
    rel = t - start[:, None]
    rel[t == 0] = float("nan")
    acc = rel if acc is None else acc + rel
_lib.lib().zt_debug_set_raft_update_trace(C.c_void_p(0), 0)
rel = acc / runs
names = ["convc1|convf1", "convc2|convf2", "conv", "zr1", "q1", "zr2", "q2", "fh1", "fh2+update+lookup"]
print("CTA %d; microseconds after the phase start (previous grid-barrier exit), averaged over the 12 iterations" % cta)
print("%-20s %9s %9s %9s %9s %9s %9s %9s" % ("phase", "A-release", "A-landed", "last-MMA", "acc-done", "epi-done", "arrived", "barrier"))
print("%-20s %9s %9s %9s %9s %9.1f %9.1f %9.1f" % ("init", "-", "-", "-", "-", rel[0, 5], rel[0, 6], rel[0, 0]))
tot = 0.0
for k in range(9):
    rows = rel[1 + k:nph:9]
    m = torch.nanmean(rows, 0)
    f = lambda v: ("%9.1f" % v) if v == v else "%9s" % "-"      # noqa: E731
    print("%-20s %s %s %s %s %s %s %s" % (names[k], f(m[1].item()), f(m[2].item()), f(m[3].item()), f(m[4].item()),
                                         f(m[5].item()), f(m[6].item()), f(m[0].item())))
    tot += m[0].item() if m[0] == m[0] else 0.0
print("iteration (sum of phases): %.1f us" % tot)
print("(fh2+update+lookup row: A-release = flow head + update of warp 0's first round done, A-landed = its lookup windows "
      "staged, last-MMA = its first round complete, acc-done = all rounds of warp 0 complete)")
