"""Summarise an ncu launch list (gpu__time_duration.sum CSV) per kernel.
usage: python scripts/summarize_launches.py gpurun_out/launches.csv [--last-half] > profiles/xxx.txt"""
import collections
import csv
import sys


def main():
    path = sys.argv[1]
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr, rows = rows[0], rows[1:]
    ki, vi, ui, gi = (hdr.index(x) for x in ("Kernel Name", "Metric Value", "Metric Unit", "Grid Size"))
    if "--last-half" in sys.argv:
        rows = rows[len(rows) // 2:]
    agg = collections.OrderedDict()
    for r in rows:
        k = r[ki].split("(")[0].replace("<unnamed>::", "")
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r[ui], 1e-6)
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print("# %s: %d launches, %.3f ms total device time (ncu, cold cache, serialised)" % (path, len(rows), tot))
    print("%-30s %6s %12s %7s" % ("kernel", "n", "ms", "share"))
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-30s %6d %12.3f %6.1f%%" % (k, a[0], a[1], 100 * a[1] / tot))
    if "--list" in sys.argv:
        print("\n# launches in order")
        for r in rows:
            print("%-30s grid=%-16s %10.1f us" % (r[ki].split("(")[0].replace("<unnamed>::", ""), r[gi],
                                                  float(r[vi].replace(",", "")) / 1e3))


if __name__ == "__main__":
    main()
