#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file X`).

    python profiles/aggregate_launches.py gpurun_out/launches.csv > profiles/r1_launches_by_kernel.csv

ncu serialises launches and runs them cold, so absolute times are not bench numbers; the SHARE column is what is compared
with bench.py's step time."""
import collections
import csv
import io
import sys

txt = open(sys.argv[1]).read()
rows = list(csv.DictReader(io.StringIO(txt[txt.index('"ID"'):])))
scale = {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}
agg = collections.OrderedDict()
for r in rows:
    key = (r["Kernel Name"], r["Grid Size"], r["Block Size"])
    t = float(r["Metric Value"].replace(",", "")) * scale.get(r["Metric Unit"], 1.0)
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += t
total = sum(a[1] for a in agg.values())
w = csv.writer(sys.stdout)
w.writerow(["kernel", "grid", "block", "launches", "avg_ns", "total_ns", "share_of_listed_gpu_time"])
for (k, g, b), (n, t) in agg.items():
    w.writerow([k, g, b, n, round(t / n, 1), round(t, 1), round(t / total, 4)])
