#!/usr/bin/env python
"""Print the metrics that matter from `ncu -i X.ncu-rep --page raw --csv` (stdin or file)."""
import csv
import sys

KEYS = ['Kernel Name', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__time_duration.sum',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__occupancy_limit', 'launch__grid_size', 'launch__waves_per_multiprocessor',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tensor', 'lts__t_sector_hit_rate.pct',
        'l1tex__data_bank_conflicts_pipe_lsu', 'smsp__average_warp', 'smsp__inst_executed.sum',
        'sm__inst_executed_pipe_alu', 'sm__inst_executed_pipe_fma', 'sm__inst_executed_pipe_xu',
        'sm__inst_executed_pipe_lsu', 'sm__inst_executed_pipe_uniform', 'sm__cycles_active.avg',
        'smsp__average_warps_issue_stalled', 'launch__shared_mem', 'lts__t_bytes.sum',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'dram__cycles_active']


def main():
    f = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin
    rows = list(csv.reader(f))
    hdr, units, data = rows[0], rows[1], rows[2:]
    for i, h in enumerate(hdr):
        if any(k in h for k in KEYS):
            print(f"{h} [{units[i]}]: " + " | ".join(d[i][:70] for d in data))


if __name__ == "__main__":
    main()
