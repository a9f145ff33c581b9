"""Timeline of one CTA of the tcgen05 conv kernel (debug)."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from zero_tig_b200 import ops, _lib
from conv_bench import CASES
name = sys.argv[1]; tc = int(sys.argv[2]); ntile = int(sys.argv[3]) if len(sys.argv) > 3 else 6
cin, cout, kh, kw, pad, H, W, B, epi = CASES[name]
cs = (cin + 3) // 4 * 4
x = torch.randn(B * H * W, cs, device="cuda"); w = torch.randn(cout, cin, kh, kw, device="cuda") * 0.05
wt = ops.ConvW(w, torch.zeros(cout, device="cuda")); ocs = (cout + 3) // 4 * 4
out = torch.empty(B * H * W, ocs, device="cuda")
kw_ = dict(pad=pad, act="relu", tc=tc, bn=int(os.environ.get("BN", "64")))
if epi == "add": kw_.update(epi="add", aux0=x, aux0_cs=cs)
for _ in range(3): ops.conv2d(wt, x, cs, wt.cin, B, H, W, out, ocs, **kw_)
cap = 4000
NR = 12
buf = torch.zeros(NR * 2 * cap, dtype=torch.int64, device="cuda")
_lib.lib().zt_debug_set_tc_trace(C.c_void_p(buf.data_ptr()), cap)
ops.conv2d(wt, x, cs, wt.cin, B, H, W, out, ocs, **kw_)
torch.cuda.synchronize()
_lib.lib().zt_debug_set_tc_trace(None, 0)
b = buf.cpu().view(NR, cap, 2)
ev = []
for role in range(NR):
    for i in range(cap):
        tag, t = int(b[role, i, 0]), int(b[role, i, 1])
        if tag == 0: break
        ev.append((t, role, (tag >> 32) - 1, tag & 0xffffffff))
ev.sort()
t0 = ev[0][0]
names = {8: "MMA", 10: "PG0", 11: "PG1"}
evn = {"MMA": {0: "waitA", 1: "gotA", 2: "issued", 3: "waitAcc", 4: "gotAcc", 5: "tap>", 6: "tap<"}, "P": {0: "waitSlot", 1: "gotSlot", 2: "stored", 3: "arrived"},
       "E": {0: "waitAcc", 1: "gotAcc", 2: "done"}}
lo = int(sys.argv[4]) if len(sys.argv) > 4 else 0
hi = int(sys.argv[5]) if len(sys.argv) > 5 else 400
for t, role, e, idx in ev[lo:hi]:
    rn = names.get(role, "EPI%d" % role)
    k = "MMA" if role == 8 else ("P" if role >= 10 else "E")
    if k == "E" and role not in (0, 4): continue
    print("%9.3f us  %-5s %-9s idx %d" % ((t - t0) / 1e3, rn, evn[k][e], idx))
