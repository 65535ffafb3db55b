"""DRAM traffic per launch of the main kernels from `ncu --page raw --csv` output.
usage: python scripts/ncu_traffic.py gpurun_out/main_raw.csv > profiles/rNN_traffic.json
k_parse launches are split into the shortened first pass ("k_parse_warm": the shorter
launch of each pair) and the full pass ("k_parse")."""
import csv
import json
import re
import sys


def num(x):
    return float(x.replace(",", "")) if x not in ("", "n/a") else 0.0


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    col = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
    per = {}
    for d in data:
        name = re.sub(r"^void |<unnamed>::|\(.*$", "", d[ki]).strip()
        rd = num(d[col["dram__bytes_read.sum"]]) * scale[units[col["dram__bytes_read.sum"]]]
        wr = num(d[col["dram__bytes_write.sum"]]) * scale[units[col["dram__bytes_write.sum"]]]
        ms = num(d[col["gpu__time_duration.sum"]]) * scale[units[col["gpu__time_duration.sum"]]]
        per.setdefault(name, []).append((rd + wr, ms))
    if "k_parse" in per and len(per["k_parse"]) >= 2:
        ls = per.pop("k_parse")
        med = sorted(x[1] for x in ls)[len(ls) // 2 - 1] * 1.5
        per["k_parse_warm"] = [x for x in ls if x[1] <= med] or ls[:1]
        per["k_parse"] = [x for x in ls if x[1] > med] or ls[1:]
    out = {}
    for name, ls in per.items():
        b = sum(x[0] for x in ls) / len(ls)
        ms = sum(x[1] for x in ls) / len(ls)
        out[name] = {"launches": len(ls), "dram_bytes_per_launch": b, "ms_per_launch_under_ncu": ms,
                     "dram_GBps": b / ms / 1e6 if ms else 0.0}
    json.dump(out, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
