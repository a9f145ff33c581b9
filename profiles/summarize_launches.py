"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel (and grid)."""
import collections
import csv
import sys


def main(path, by_grid=False):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
    H, rows = rows[hdr], rows[hdr + 1:]
    ki, vi, gi = H.index("Kernel Name"), H.index("Metric Value"), H.index("Grid Size")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        k = r[ki][:58] + ((" " + r[gi].replace(" ", "")) if by_grid else "")
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-80s n=%5d %9.3f ms %5.1f%% avg %8.1f us" % (k, v[0], v[1] / 1e6, 100 * v[1] / tot, v[1] / v[0] / 1e3))
    print("total %.3f ms over %d launches" % (tot / 1e6, len(rows)))


if __name__ == "__main__":
    main(sys.argv[1], "--grid" in sys.argv)
