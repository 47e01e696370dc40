#!/bin/bash
# quick GPU check after a kernel change: conv/net parity tests, then the bench line and per-launch table
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_net_gpu.py tests/test_detect_gpu.py -q -m gpu -x --no-header -p no:cacheprovider 2>&1 | tail -4
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/q_bench.json')); print('fps', round(d['value']), 'ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'roof', round(d['roofline']['frac'],3))"
python tools/show_launches.py gpurun_out/per_launch_ms.json
