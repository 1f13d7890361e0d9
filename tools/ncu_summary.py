"""Cut an `ncu -i X.ncu-rep --page raw --csv` dump down to the counters profiles/r*_ncu_full_*_summary.csv keeps.
    python tools/ncu_summary.py raw.csv > profiles/rNN_ncu_full_<tag>_summary.csv
"""
import csv
import sys

KEEP = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__cluster_dim_x", "lts__t_sector_hit_rate.pct",
        "smsp__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_tensor.sum"]

rows = [r for r in csv.reader(open(sys.argv[1])) if r]
start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[start]
# ncu prefixes some counters with their section ("TPC.TriageCompute.sm__pipe_tensor_..."): match by suffix
def find(k):
    for i, h in enumerate(hdr):
        if h == k or h.endswith("." + k):
            return i
    return -1


idx = [find(k) for k in KEEP]
w = csv.writer(sys.stdout)
w.writerow(KEEP)
for r in rows[start + 1:]:
    w.writerow([r[i] if 0 <= i < len(r) else "" for i in idx])
