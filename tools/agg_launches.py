"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) by kernel name.
    python tools/agg_launches.py gpurun_out/launches.csv [skip_first_n_launches] [n_frames]"""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("=="))]
hdr = rows[0]
ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
nf = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
agg = defaultdict(lambda: [0, 0.0])
tot = 0.0
for r in rows[1 + skip:]:
    name = r[ki].split("(")[0]
    t = float(r[vi].replace(",", "")) / 1e3
    agg[name][0] += 1
    agg[name][1] += t
    tot += t
print("total %.1f us over %d launches (%.1f us / frame)" % (tot, sum(v[0] for v in agg.values()), tot / nf))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-44s n=%5d  %9.1f us  %5.1f%%  avg %7.1f us" % (k, v[0], v[1], 100 * v[1] / tot, v[1] / v[0]))
