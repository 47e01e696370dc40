#!/bin/bash
# kernel-iteration visit: selected parity tests, bench line, per-launch table, optional timeline steps.  usage: gpu_quick2.sh <tag> "<pytest -k expr>" [timeline steps...]
tag=$1; kexpr=$2; shift; shift
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_net_gpu.py tests/test_plan_gpu.py -q -m gpu -x --no-header -p no:cacheprovider -k "$kexpr" > gpurun_out/${tag}_pytest.log 2>&1
echo "pytest rc=$?"; tail -4 gpurun_out/${tag}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
echo "bench rc=$?"; tail -2 gpurun_out/${tag}_bench.err
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
if [ -n "$1" ]; then python tools/tc_timeline.py 64 "$@" > gpurun_out/${tag}_timeline.txt 2>&1; fi
