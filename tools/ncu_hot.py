"""Hot spots of an ncu report's SASS page: python tools/ncu_hot.py rep.ncu-rep [launch] [n_launches] [top]
Prints, for one captured launch, total samples / instructions and the instructions with the most stall samples."""
import csv
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
nl = int(sys.argv[3]) if len(sys.argv) > 3 else 1
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h = rows[1]
data = [r for r in rows[2:] if len(r) == len(h) and r[0].startswith('0x')]
n = len(data) // nl
d = data[which * n:(which + 1) * n]
tot = sum(int(r[4]) for r in d)
ex = sum(int(r[5]) for r in d)
print('launch %d: %d SASS instrs, %d samples, %d warp-instructions executed' % (which, n, tot, ex))
order = sorted(range(n), key=lambda i: -int(d[i][4]))[:top]
for i in sorted(order):
    r = d[i]
    st = {h[k].replace('stall_', ''): r[k] for k in range(29, 46) if r[k] not in ('0', '')}
    print('%5d %-60s smp %5s (%4.1f%%) exec %8s %s' % (i, r[1].strip()[:60], r[4], 100.0 * int(r[4]) / max(tot, 1), r[5], st))
