#!/bin/bash
# Quick GPU iteration: selected tests (each group in its own process), then the bench.
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { local name=$1; shift; local to=$1; shift
  timeout $to python -m pytest "$@" -q -m gpu -x --no-header -p no:cacheprovider > gpurun_out/$name.log 2>&1
  echo "$name rc=$?" | tee -a gpurun_out/summary.txt; tail -4 gpurun_out/$name.log; }
run stem 300 tests/test_net_gpu.py -k "s2d"
run net_bf16 600 tests/test_net_gpu.py -k "bf16 or mini"
run detect 300 tests/test_detect_gpu.py
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
