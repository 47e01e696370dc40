#!/bin/bash
# One short GPU visit: all parity tests, then the full bench line + per-launch table.  usage: tools/gpu_visit.sh <tag> [pytest args]
tag=${1:-r02x}; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 1500 python -m pytest tests -q -m gpu -x --no-header -p no:cacheprovider "$@" > gpurun_out/${tag}_pytest.log 2>&1
echo "pytest rc=$?"; tail -15 gpurun_out/${tag}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err
python - <<PY
import json
try:
    d = json.load(open('gpurun_out/${tag}_bench.json'))
    print('fps', round(d['value']), 'ms/step', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']), 'one-call', round(d['e2e']['one_call_per_batch']['value']), 'dets', d['e2e']['detections_last_step'])
    r = d['roofline']; print('roof frac', round(r['frac'], 3), 'sustained', round(r['frac_vs_sustained'], 3), 'whole', round(r['whole_step_frac'], 3), 'sum', round(r['sum_of_launches_ms'], 3))
    print('cpu', d['cpu_baseline']); print('latency', d['latency_batch1']); print('layer', {k: v for k, v in d['layer'].items() if k != 'workload'})
    for k, v in d['bandwidth_kernels'].items(): print(k, v)
except Exception as e:
    print('no bench line', e)
PY
cp gpurun_out/per_launch_ms.json gpurun_out/${tag}_per_launch_ms.json 2>/dev/null
python tools/show_launches.py gpurun_out/per_launch_ms.json
