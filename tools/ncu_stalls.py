"""Top warp-stall reasons (pc sampling) and a few throughput metrics per launch of an ncu report.  usage: ncu_stalls.py rep [n]"""
import csv
import subprocess
import sys

raw = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, body = rows[0], rows[1], rows[2:]
n = int(sys.argv[2]) if len(sys.argv) > 2 else len(body)
idx = {h: i for i, h in enumerate(hdr)}
stall = [h for h in hdr if 'pcsamp_warps_issue_stalled' in h and not h.endswith('_not_issued')]
keep = ['gpu__time_duration.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active']
for li, r in enumerate(body[:n]):
    print('launch', li, r[idx['Kernel Name']][:40], 'grid', r[idx['Grid Size']])
    for k in keep:
        if k in idx:
            print('   %-70s %s %s' % (k, r[idx[k]], units[idx[k]]))
    tot = sum(float(r[idx[h]].replace(',', '') or 0) for h in stall) or 1.0
    top = sorted(stall, key=lambda h: -float(r[idx[h]].replace(',', '') or 0))[:6]
    for h in top:
        print('   stall %-50s %5.1f%%' % (h.replace('smsp__pcsamp_warps_issue_stalled_', ''), 100 * float(r[idx[h]].replace(',', '') or 0) / tot))
