"""Print gpurun_out/per_launch_ms.json compactly (TFLOP/s per conv launch)."""
import json, re, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fcn_object_detector_b200.netspec import conv_shapes
f = sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/per_launch_ms.json'
d = json.load(open(f))
shapes = {n: (co, ci, k) for n, co, ci, k, _b in conv_shapes()}
tot = sum(x['ms'] for x in d); tc = 0
for x in d:
    s = x['desc']
    m = re.match(r'\d+: conv_tc (.+?) \d+x\d+/\d+ ', s)
    if m:
        names = m.group(1).split(' + ')
        dims = s.split(' -> ')[1].split('(')[1].split(')')[0].split(',')
        hw = int(dims[1]) * int(dims[2])
        fl = sum(2 * shapes[n][0] * shapes[n][1] * shapes[n][2] ** 2 * hw for n in names) * 64
        cfg = re.search(r'\[(.*?) tiles', s).group(1)
        tc += x['ms']
        print('%7.1f us %5.0f TF/s %-34s %s' % (x['ms'] * 1e3, fl / x['ms'] / 1e9, (names[0] + ('+%d' % (len(names) - 1) if len(names) > 1 else ''))[:34], cfg))
    else:
        print('%7.1f us            %s' % (x['ms'] * 1e3, s[:80]))
print('sum %.3f ms, conv_tc %.3f ms' % (tot, tc))
