#!/usr/bin/env python
"""Per-launch table of the FP64 tensor-pipe (DMMA) and memory counters from `ncu -i X.ncu-rep --page raw --csv` (file argument)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h, units, data = rows[0], rows[1], rows[2:]
want = [("Kernel Name", "kernel"), ("Grid Size", "grid"), ("Block Size", "block"), ("gpu__time_duration.sum", "time"), ("launch__registers_per_thread", "regs"),
        ("sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active", "dmma inst % of peak (active)"),
        ("SM_C.TriageCompute.smsp__pipe_tensor_subpipe_dmma_cycles_active.avg", "dmma cycles active / SMSP"),
        ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor pipe % (elapsed)"),
        ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64 pipe inst %"),
        ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex %"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2 %"),
        ("lts__t_sector_hit_rate.pct", "l2 hit %"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("sm__cycles_elapsed.max", "cycles"), ("launch__waves_per_multiprocessor", "waves")]
idx = [(h.index(k), n) for k, n in want if k in h]
print("| " + " | ".join(n for _, n in idx) + " |")
print("|" + "---|" * len(idx))
for r in data:
    cells = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = v.split("(")[0].replace("void ", "").split("::")[-1]
        elif n not in ("grid", "block"):
            try:
                v = f"{float(v.replace(',', '')):.4g} {units[i]}" if n in ("time", "dram rd", "dram wr") else f"{float(v.replace(',', '')):.4g}"
            except ValueError:
                pass
        cells.append(v)
    print("| " + " | ".join(cells) + " |")
