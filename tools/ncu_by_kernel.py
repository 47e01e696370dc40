"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name."""
import csv
import re
import sys
from collections import defaultdict

rows = list(csv.reader(open(sys.argv[1])))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == 'ID')
H = rows[hdr]
ik, iv, iu = H.index('Kernel Name'), H.index('Metric Value'), H.index('Metric Unit')
agg, cnt = defaultdict(float), defaultdict(int)
for r in rows[hdr + 1:]:
    if len(r) <= iv:
        continue
    v = float(r[iv].replace(',', ''))
    v *= {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 'nsecond': 1e-3, 'usecond': 1.0, 'msecond': 1e3}.get(r[iu], 1e-3)
    name = re.sub(r'\(.*', '', r[ik])
    agg[name] += v
    cnt[name] += 1
tot = sum(agg.values())
for k in sorted(agg, key=agg.get, reverse=True):
    print('%10.1f us %5.1f%% %4d x  %s' % (agg[k], 100 * agg[k] / tot, cnt[k], k))
print('%10.1f us total, %d launches' % (tot, sum(cnt.values())))
