"""Print selected metrics from `ncu -i X.ncu-rep --page raw --csv` output.
usage: python scripts/ncu_metrics.py raw.csv [kernel-substring]"""
import csv
import re
import sys

PAT = re.compile(r"gpu__time_duration.sum|dram__bytes_(read|write).sum$|gpu__dram_throughput.avg.pct|"
                 r"sm__warps_active.avg.pct|registers_per_thread|sm__throughput.avg.pct|sector_hit_rate.pct|"
                 r"issue_active.avg.pct|occupancy_limit|warps_issue_stalled.*_per_warp_active|"
                 r"sm__inst_executed_pipe_.*pct_of_peak_sustained_active|lts__t_bytes.sum$|"
                 r"l1tex__data_bank_conflicts_pipe_lsu.sum$|smsp__inst_executed.sum$|launch__grid_size|launch__block_size|"
                 r"shared_mem_per_block|achieved_occupancy|l1tex__t_bytes.sum$")


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    sub = sys.argv[2] if len(sys.argv) > 2 else ""
    ki = hdr.index("Kernel Name")
    seen = set()
    for d in data:
        name = d[ki].split("(")[0]
        if sub not in name or name in seen:
            continue
        seen.add(name)
        print("=== " + name)
        for i, h in enumerate(hdr):
            if PAT.search(h) and d[i] not in ("", "0", "n/a"):
                print("  %-88s %s %s" % (h, d[i], units[i]))


if __name__ == "__main__":
    main()
