"""Print the headline metrics of an .ncu-rep (ncu -i rep --page raw --csv | python ncu_raw.py)."""
import csv, sys
rows = list(csv.reader(sys.stdin)); hdr = rows[0]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_shared_mem',
        'launch__occupancy_limit_registers', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed']
for r in rows[2:]:
    for h, u, v in zip(hdr, rows[1], r):
        if h in want or ('issue_stalled' in h and h.endswith('per_issue_active.ratio') and float(v or 0) > 0.3):
            print(h, u, v)
