#!/usr/bin/env python
"""Prints the roofline-relevant counters of an `ncu --page raw --csv` export (one block per profiled launch)."""
import csv
import sys

KEYS = [
    'Kernel Name', 'Block Size', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
    'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_registers',
    'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
    'smsp__issue_active.avg.pct', 'sm__inst_executed_pipe_fp64.sum', 'smsp__inst_executed.sum',
    'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_lsu.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
    'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'lts__t_sector_hit_rate.pct',
    'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
]


def main(path, extra):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        for i, h in enumerate(hdr):
            if h in KEYS or any(e in h for e in extra):
                print("%-75s %s %s" % (h, r[i], units[i]))
        print('-' * 100)


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2:])
