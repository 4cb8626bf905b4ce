"""Summarise an `ncu --page raw --csv` export as a metric x kernel table (first launch of each kernel).
Usage: ncu -i X.ncu-rep --page raw --csv > raw.csv; python tools/ncu_metrics.py raw.csv [title]"""
import csv
import sys

METRICS = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
           "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct"]
rows = list(csv.reader(open(sys.argv[1])))
h, units = rows[0], rows[1]
seen, cols = set(), []
for r in rows[2:]:
    name = r[h.index("Kernel Name")].split("(")[0].replace("<unnamed>::", "")
    if name not in seen:
        seen.add(name); cols.append((name, r))
if len(sys.argv) > 2:
    print(sys.argv[2])
print("metric | unit | " + " | ".join(n for n, _ in cols))
for m in METRICS:
    if m in h:
        i = h.index(m)
        print(f"{m} | {units[i]} | " + " | ".join(r[i] for _, r in cols))
