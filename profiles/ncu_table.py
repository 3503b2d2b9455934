#!/usr/bin/env python
"""Print a compact per-launch table from `ncu -i X.ncu-rep --page raw --csv` output (stdin or file argument)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin))
h, units, data = rows[0], rows[1], rows[2:]
want = [("Kernel Name", "kernel"), ("Grid Size", "grid"), ("gpu__time_duration.sum", "time"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"), ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"), ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("l1tex__t_sector_hit_rate.pct", "l1hit%"), ("lts__t_sector_hit_rate.pct", "l2hit%"),
        ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"), ("launch__registers_per_thread", "regs"),
        ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64pipe%"), ("sm__inst_executed_pipe_fp64.sum", "fp64inst"),
        ("launch__waves_per_multiprocessor", "waves")]
idx = [(h.index(k), n) for k, n in want if k in h]
print("| " + " | ".join(n for _, n in idx) + " |")
print("|" + "---|" * len(idx))
for r in data:
    cells = []
    for i, n in idx:
        v = r[i]
        if n == "kernel": v = v.split("(")[0].replace("void ", "")
        elif n not in ("grid",):
            try: v = f"{float(v.replace(',', '')):.4g} {units[i]}" if n in ("time", "dram_rd", "dram_wr") else f"{float(v.replace(',', '')):.4g}"
            except ValueError: pass
        cells.append(v)
    print("| " + " | ".join(cells) + " |")
