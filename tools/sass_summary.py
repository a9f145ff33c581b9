"""SASS evidence table: instruction counts per kernel from `cuobjdump -sass libzerotig.so` (run in the build container)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "zero_tig_b200", "csrc", "libzerotig.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
names = ["UTCHMMA", "LDTM", "UTMALDG", "UBLKCP", "SYNCS", "UTCBAR", "FFMA2", "FFMA"]
rows, cur = collections.OrderedDict(), None
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        rows[cur] = collections.Counter()
        continue
    if cur is None or "/*" not in line:
        continue
    m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if not m:
        continue
    op = m.group(1)
    for n in names:
        if op == n or (n != "FFMA" and op.startswith(n)):
            rows[cur][n] += 1
            break
dem = subprocess.run(["c++filt"] + list(rows), capture_output=True, text=True).stdout.splitlines()
print("# SASS evidence (tools/sass_summary.py: cuobjdump -sass zero_tig_b200/csrc/libzerotig.so, all cubins sm_100a): instruction counts per kernel.")
print("# UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld (TMEM -> registers), UTMALDG = cp.async.bulk.tensor (TMA), UBLKCP = cp.async.bulk,")
print("# SYNCS = mbarrier ops, UTCBAR = tcgen05.commit, FFMA2 = packed fp32x2 FMA.  Sorted by UTCHMMA count.")
print("%-86s %8s %6s %8s %7s %6s %7s %6s %6s" % tuple(["kernel"] + names))
out = []
for (mangled, c), d in zip(rows.items(), dem):
    d = re.sub(r"\(.*", "", d).replace("(int)", "")
    out.append((d, c))
for d, c in sorted(out, key=lambda r: (-r[1]["UTCHMMA"], r[0])):
    if sum(c.values()) == 0:
        continue
    print("%-86s %8d %6d %8d %7d %6d %7d %6d %6d" % tuple([d[:86]] + [c[n] for n in names]))
