"""Condense an ncu report (--page raw --csv) to the metrics DESIGN.md / bench.py cite.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r01_x.csv"""
import csv
import subprocess
import sys

KEEP = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__pipe_tensor_cycles_active',
        'sm__inst_executed_pipe_tensor', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__occupancy_limit', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__cycles_active.avg',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_active',
        'launch__shared_mem_per_block', 'smsp__cycles_active.avg', 'sm__pipe_tensor_subpipe_hmma_cycles_active_realtime',
        'dram__cycles_active.avg.pct', 'smsp__inst_executed.sum', 'l1tex__t_sector_hit_rate.pct', 'smsp__issue_active.avg.pct_of_peak_sustained_active']
raw = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, body = rows[0], rows[1], rows[2:]
cols = [i for i, h in enumerate(hdr) if any(k in h for k in KEEP) and not h.endswith('.per_second')]
with open(sys.argv[2], 'w', newline='') as f:
    w = csv.writer(f)
    w.writerow(['metric', 'unit'] + ['launch%d' % i for i in range(len(body))])
    for i in cols:
        w.writerow([hdr[i], units[i]] + [r[i] for r in body])
print('wrote', sys.argv[2], len(cols), 'metrics x', len(body), 'launches')
