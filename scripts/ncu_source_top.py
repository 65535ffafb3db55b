"""Top SASS lines by stall samples from `ncu --page source --csv`.
usage: python scripts/ncu_source_top.py src.csv kernel-substring [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2]
N = int(sys.argv[3]) if len(sys.argv) > 3 else 40
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
sec = None
for j, s in enumerate(starts):
    if want in rows[s][1]:
        sec = rows[s: starts[j + 1] if j + 1 < len(starts) else len(rows)]
        break
if sec is None:
    sys.exit("kernel not found: " + ", ".join(rows[s][1].split("(")[0] for s in starts))
hdr = sec[1]
data = [r for r in sec[2:] if len(r) == len(hdr)]
si = hdr.index("# Samples")
ii = hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[si] or 0) for r in data)
print(sec[0][1].split("(")[0], "total samples", tot, "sass lines", len(data))
order = sorted(range(len(data)), key=lambda k: -int(data[k][si] or 0))[:N]
for k in sorted(order):
    r = data[k]
    st = sorted(((int(r[c] or 0), hdr[c][6:]) for c in stall_cols), reverse=True)[:3]
    print("%5d %5.1f%% exec=%-9s %-64s %s" % (k, 100.0 * int(r[si] or 0) / max(1, tot), r[ii], r[1].strip()[:64],
                                            " ".join("%s:%d" % (n, v) for v, n in st if v)))
agg = {}
for r in data:
    for c in stall_cols:
        agg[hdr[c][6:]] = agg.get(hdr[c][6:], 0) + int(r[c] or 0)
print("stall totals:", " ".join("%s:%.1f%%" % (k, 100.0 * v / max(1, tot)) for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v))
