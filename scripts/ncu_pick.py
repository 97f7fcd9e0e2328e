"""Prints selected metrics from `ncu -i X.ncu-rep --page raw --csv` output (file argument)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = sys.argv[2:] or ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'pipe_fp64', 'smsp__issue_active.avg.pct', 'smsp__inst_executed.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'lts__t_sector_hit_rate', 'l1tex__t_sector_hit_rate',
        'sm__cycles_elapsed.max', 'launch__shared_mem', 'pipe_lsu', 'pipe_fma', 'achieved_occupancy', 'launch__waves', 'warp_issue_stalled', 'l1tex__data_pipe', 'lsu_mem_shared']
for vals in rows[2:]:
    print("---- kernel:", vals[hdr.index('Kernel Name')] if 'Kernel Name' in hdr else '')
    for h, u, v in zip(hdr, units, vals):
        if any(w in h for w in want):
            print(f"{h:100s} {u:14s} {v}")
