#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box with `ncu -i`) into one JSON line per captured launch."""
import csv
import json
import subprocess
import sys

WANT = ['Kernel Name', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sectors.sum', 'lts__t_sectors_op_read.sum',
        'lts__t_sectors_op_atom.sum', 'lts__t_sectors_op_red.sum', 'lts__t_sectors_op_write.sum',
        'lts__t_sector_hit_rate.pct', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'smsp__inst_executed.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio']


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader([l for l in out.splitlines() if not l.startswith('==')]))
    head, units = rows[0], rows[1]
    idx = [(w, head.index(w)) for w in WANT if w in head]
    for row in rows[2:]:
        d = {}
        for w, i in idx:
            v = row[i].replace(',', '')
            try:
                v = float(v)
            except ValueError:
                pass
            d[w + (' [%s]' % units[i] if units[i] else '')] = v
        print(json.dumps(d))


if __name__ == '__main__':
    main(sys.argv[1])
