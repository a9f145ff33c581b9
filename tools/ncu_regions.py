"""Bucket the warp-stall samples of an ncu report's source page by SASS region."""
import csv, re, subprocess, sys
rep = sys.argv[1]; B = int(sys.argv[2]) if len(sys.argv) > 2 else 60
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
H = rows[1]; R = rows[2:]
si = H.index('# Samples'); src = H.index('Source'); ie = H.index('Instructions Executed')
tot = sum(int(r[si] or 0) for r in R)
print('total samples', tot, 'rows', len(R))
for i in range(0, len(R), B):
    blk = R[i:i + B]
    s = sum(int(r[si] or 0) for r in blk)
    ops = {}
    for r in blk:
        m = re.search(r'\b(LDG|STG|STS|LDS|UTCHMMA|LDTM|SYNCS|UTCBAR|UBLKCP|F2F|MUFU|BAR|EXIT|FENCE|LDC)\b', r[src])
        if m: ops[m.group(1)] = ops.get(m.group(1), 0) + 1
    if s > tot * 0.004: print(i, s, "%.1f%%" % (100.0 * s / tot), ops)
print("-- top rows")
for r in sorted(R, key=lambda r: -int(r[si] or 0))[:14]:
    print(r[si].rjust(7), r[ie].rjust(10), r[src][:110])
