#!/bin/bash
# First-contact GPU run: every test group in its own process (a trapping kernel poisons only its own context).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
run() { # name timeout args...
  local name=$1; shift; local to=$1; shift
  timeout $to python -m pytest "$@" -q -m gpu -x --no-header -p no:cacheprovider > gpurun_out/$name.log 2>&1
  echo "$name rc=$?" | tee -a gpurun_out/summary.txt
  tail -3 gpurun_out/$name.log
}
: > gpurun_out/summary.txt
run postnet 300 tests/test_postnet_gpu.py
run layer 300 tests/test_layer_gpu.py
run detect_pre 200 tests/test_detect_gpu.py -k preprocess
run net_fp32 600 tests/test_net_gpu.py -k fp32
run detect_fp32 300 tests/test_detect_gpu.py -k config1
for c in k1_c64 k3_c64 k5_c16 k3_c96_sw64 k5_c24 k1_c480 k3_c160_n320 k3_w112 mix_3a odd_hw cin48_k5 big_w; do
  run tc_$c 120 "tests/test_net_gpu.py::test_conv_tc_mini_nets[$c]"
done
run stem 300 tests/test_net_gpu.py -k "s2d"
run net_bf16 600 tests/test_net_gpu.py -k "bf16"
run detect_bf16 300 tests/test_detect_gpu.py -k bf16
cat gpurun_out/summary.txt
