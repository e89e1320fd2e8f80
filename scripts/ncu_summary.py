#!/usr/bin/env python
"""Markdown summary of an .ncu-rep (`ncu --set full`): one row per captured launch.

usage: ncu_summary.py <report.ncu-rep> [title]   (runs `ncu -i ... --page raw --csv`)
"""
import csv, io, subprocess, sys

COLS = [
    ("time ms", "gpu__time_duration.sum", 1e-6, "{:.3f}"),          # ns -> ms (raw page prints ns or ms; handled below)
    ("dram rd GB", "dram__bytes_read.sum", None, "{:.3f}"),
    ("dram wr GB", "dram__bytes_write.sum", None, "{:.3f}"),
    ("warp inst", "smsp__inst_executed.sum", None, "{:.3e}"),
    ("thr/inst", "smsp__thread_inst_executed_per_inst_executed.ratio", None, "{:.1f}"),
    ("issue act %", "smsp__issue_active.avg.pct_of_peak_sustained_active", None, "{:.1f}"),
    ("fma pipe %", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", None, "{:.1f}"),
    ("regs", "launch__registers_per_thread", None, "{:.0f}"),
    ("warps act %", "sm__warps_active.avg.pct_of_peak_sustained_active", None, "{:.1f}"),
]
UNIT_SCALE = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "byte": 1e-9, "Kbyte": 1e-6, "Mbyte": 1e-3, "Gbyte": 1.0}

def main():
    rep = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else rep
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    print(f"# {title}\n")
    print("| kernel | " + " | ".join(c[0] for c in COLS) + " |")
    print("|---|" + "---|" * len(COLS))
    for r in data:
        name = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "")
        cells = []
        for label, metric, _, fmt in COLS:
            if metric not in hdr:
                cells.append("n/a"); continue
            i = hdr.index(metric)
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                cells.append("n/a"); continue
            v *= UNIT_SCALE.get(units[i], 1.0)
            cells.append(fmt.format(v))
        print(f"| {name} | " + " | ".join(cells) + " |")

main()
