"""Trim an `ncu --page raw --csv` dump to the metrics profiles/ keeps.

    ncu -i X.ncu-rep --page raw --csv > raw.csv
    python scripts/ncu_summary.py raw.csv > profiles/rN_ncu_full_<what>.csv
"""
import csv
import sys

KEEP = [
    "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "lts__t_bytes.sum", "l1tex__t_bytes.sum",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    head, units = rows[0], rows[1]
    cols = [head.index(k) for k in KEEP if k in head]
    w = csv.writer(sys.stdout)
    w.writerow([head[c] for c in cols])
    w.writerow([units[c] for c in cols])
    for r in rows[2:]:
        w.writerow([r[c] for c in cols])


if __name__ == "__main__":
    main()
