#!/usr/bin/env python
"""tools/ncu_summary.py <report.ncu-rep> -- key rows of an `ncu --set full` report as CSV (stdout)."""
import csv
import subprocess
import sys

KEEP = ['Kernel Name', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'dram__cycles_active.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_op_read_hit_rate.pct',
        'lts__t_sector_op_write_hit_rate.pct', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'dram__cycles_elapsed.avg.per_second']
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
w = csv.writer(sys.stdout)
keep = [k for k in KEEP if k in hdr]
w.writerow(keep)
w.writerow([units[hdr.index(k)] for k in keep])
for r in rows[2:]:
    w.writerow([r[hdr.index(k)] for k in keep])
