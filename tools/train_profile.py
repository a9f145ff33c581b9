"""GPU tool: per-C-ABI-call CUDA-event timing of one zero-shot training step at 960x540 (configs[2])."""
import os
import sys
import collections

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from zero_tig_b200 import _lib, synth                      # noqa: E402
from zero_tig_b200.model.model import Network              # noqa: E402
from zero_tig_b200.optim import ClipAdam                   # noqa: E402


def main():
    H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (540, 960)
    dev = torch.device("cuda")
    sd = synth.synthetic_state_dict(Network(synth.make_args()).state_dict(), seed=1234)
    net = Network(synth.make_args(of_scale=3)).to(dev)
    net.load_state_dict(sd)
    net.train()
    opt = ClipAdam(net.trainable())
    clip = synth.synthetic_clip(4, H, W, seed=1).to(dev)

    def step(i):
        net.is_new_seq = (i == 0)
        opt.zero_grad()
        loss = net._loss(clip[i:i + 1])
        loss.backward()
        opt.step()
    for i in range(3):
        step(i)
    torch.cuda.synchronize()
    # wrap every ctypes entry point with events
    L = _lib.lib()
    rec = collections.defaultdict(list)
    orig = {}
    for name in _lib.SIGNATURES:
        if name in ("zt_last_error", "zt_version", "zt_launch_count", "zt_debug_set_tc_trace"):
            continue
        f = getattr(L, name)
        orig[name] = f

        def wrap(*a, _f=f, _n=name):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = _f(*a)
            e1.record()
            rec[_n].append((e0, e1))
            return r
        setattr(L, name, wrap)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    step(3)
    e1.record()
    torch.cuda.synchronize()
    for n, f in orig.items():
        setattr(L, n, f)
    total = e0.elapsed_time(e1)
    rows = sorted(((sum(a.elapsed_time(b) for a, b in v), len(v), n) for n, v in rec.items()), reverse=True)
    print("step %.2f ms (with event overhead); sum of calls %.2f ms" % (total, sum(r[0] for r in rows)))
    for ms, cnt, n in rows:
        print("%8.3f ms %4d calls  %s" % (ms, cnt, n))
    e0.record()
    for i in range(3):
        step(1 + i)
    e1.record()
    torch.cuda.synchronize()
    print("clean step: %.2f ms" % (e0.elapsed_time(e1) / 3))


if __name__ == "__main__":
    main()
