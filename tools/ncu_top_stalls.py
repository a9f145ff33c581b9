"""Top warp-stall instructions from an `ncu --page source --csv` export (one kernel per file)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, data = rows[h], [r for r in rows[h + 1:] if len(r) > 5 and r[0] != "Address"]
isrc, iss, ie = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")


def num(x):
    try:
        return int(float(x))
    except ValueError:
        return 0


tot = sum(num(r[iss]) for r in data)
print(rows[0][1] if len(rows[0]) > 1 else "", "total samples", tot)
for k, r in sorted(enumerate(data), key=lambda kr: -num(kr[1][iss]))[:n]:
    print("%5d %5.1f%% exec %8s  #%04d %s" % (num(r[iss]), 100.0 * num(r[iss]) / max(tot, 1), r[ie], k, r[isrc][:100]))
