#!/usr/bin/env python
"""Condense `ncu -i X.ncu-rep --page raw --csv` pages into profiles/r1_ncu_summary.json (or `--out NAME.json`).

    python profiles/make_summary.py mandelbrot=profiles/r1_ncu_mandelbrot_full_raw.csv saxpy=... nbody=...
    python profiles/make_summary.py --out r2_ncu_summary.json blurStrips=profiles/r2_ncu_image_raw.csv@blurStrips ...

`name=path@text` keeps only the launches whose kernel name contains `text` (a report with several kernels).

bench.py reads `traffic_B` (dram__bytes_read.sum + dram__bytes_write.sum of ONE launch) from the summary for
`roofline.traffic`; the other fields are the ones DESIGN.md quotes."""
import csv
import json
import sys
from pathlib import Path

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "second": 1.0}


def summarize(path, only=None):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    if only:
        kcol = hdr.index("Kernel Name")
        rows = rows[:2] + [r for r in rows[2:] if only in r[kcol]]
    # several launches in one report: keep the longest (the resident, whole-range launch)
    tcol = hdr.index("gpu__time_duration.sum")
    row = max(rows[2:], key=lambda r: float(r[tcol].replace(",", "")) * UNIT.get(units[tcol], 1.0))

    def num(name, default=None):
        if name not in hdr:
            return default
        i = hdr.index(name)
        v = row[i].replace(",", "")
        return float(v) * UNIT.get(units[i], 1.0) if v else default

    def text(name):
        return row[hdr.index(name)] if name in hdr else None

    t = num("gpu__time_duration.sum")
    rd, wr = num("dram__bytes_read.sum", 0.0), num("dram__bytes_write.sum", 0.0)
    return {
        "kernel": text("Kernel Name"), "grid": text("Grid Size"), "block": text("Block Size"), "launches_in_report": len(rows) - 2,
        "time_s": t, "dram_read_B": rd, "dram_write_B": wr, "traffic_B": rd + wr, "dram_GBps": (rd + wr) / t / 1e9,
        "regs": text("launch__registers_per_thread"),
        "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "pipe_fma_cycles_pct": num("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
        "sm_throughput_pct": num("sm__throughput.avg.pct_of_peak_sustained_elapsed"),
        "dram_throughput_pct": num("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        "warps_active_pct": num("sm__warps_active.avg.pct_of_peak_sustained_active"),
        "lanes_per_inst": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
    }


if __name__ == "__main__":
    argv = sys.argv[1:]
    out_name = "r1_ncu_summary.json"
    if argv and argv[0] == "--out":
        out_name, argv = argv[1], argv[2:]
    out_path = Path(__file__).resolve().parent / out_name
    out = json.loads(out_path.read_text()) if out_path.exists() else {}
    for arg in argv:
        name, path = arg.split("=", 1)
        path, _, only = path.partition("@")
        out[name] = summarize(path, only or None)
        out[name]["source"] = str(Path(path).name)
    out_path.write_text(json.dumps(out, indent=1) + "\n")
    print(json.dumps(out, indent=1))
