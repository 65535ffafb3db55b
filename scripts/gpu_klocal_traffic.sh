#!/bin/bash
# per-launch time and DRAM bytes of k_local on 256 MiB of the DNA workload, with and without the bag reorder
mkdir -p gpurun_out
export SALZ_BENCH_CACHE=/tmp/c
CMD="python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-extras --no-e2e --workload dna1g --mib 256"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
for r in 0 3; do
  SALZ_BAG_REORDER=$r ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:k_local -c 40 --csv --log-file gpurun_out/klocal_$r.csv $CMD > gpurun_out/ncu_klocal_$r.log 2>&1
  echo "ncu $r exit $?"
done
python - <<'PY'
import csv
for r in (0, 3):
    rows = [x for x in csv.reader(open('gpurun_out/klocal_%d.csv' % r)) if len(x) > 10]
    hdr, rows = rows[0], rows[1:]
    ki, mi, vi, ui, gi, ii = (hdr.index(x) for x in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit", "Grid Size", "ID"))
    per = {}
    for x in rows:
        d = per.setdefault(x[ii], {"grid": x[gi], "name": x[ki][:24]})
        d[x[mi]] = (float(x[vi].replace(",", "")), x[ui])
    print("SALZ_BAG_REORDER=%d" % r)
    for i, d in per.items():
        def val(m):
            v, u = d.get(m, (0, ""))
            return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1, "msecond": 1, "usecond": 1e-3, "nsecond": 1e-6}.get(u, 1)
        print("  %-24s grid %-14s %7.3f ms  read %6.2f GB  write %6.2f GB" % (d["name"], d["grid"], val("gpu__time_duration.sum"), val("dram__bytes_read.sum") / 1e9, val("dram__bytes_write.sum") / 1e9))
PY
