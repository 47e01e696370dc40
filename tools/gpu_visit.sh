#!/bin/bash
# One short GPU visit: all parity tests, then the bench line + per-launch table.  usage: tools/gpu_visit.sh <tag> [pytest args]
tag=${1:-r02x}; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 1500 python -m pytest tests -q -m gpu -x --no-header -p no:cacheprovider "$@" > gpurun_out/${tag}_pytest.log 2>&1
echo "pytest rc=$?"; tail -15 gpurun_out/${tag}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
echo "bench rc=$?"; tail -3 gpurun_out/${tag}_bench.err
python - <<PY
import json
try:
    d = json.load(open('gpurun_out/${tag}_bench.json'))
    print('fps', round(d['value']), 'ms/step', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']), 'roof', d['roofline'].get('frac'))
except Exception as e:
    print('no bench line', e)
PY
cp gpurun_out/per_launch_ms.json gpurun_out/${tag}_per_launch_ms.json 2>/dev/null
python tools/show_launches.py gpurun_out/per_launch_ms.json
