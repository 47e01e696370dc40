#!/bin/bash
# one debugging iteration: k3/k5 mini nets, then an ncu source-level capture of two conv_tc2 launches
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { local name=$1; shift; local to=$1; shift
  timeout $to python -m pytest "$@" -q -m gpu -x --no-header -p no:cacheprovider > gpurun_out/$name.log 2>&1
  echo "$name rc=$?" | tee -a gpurun_out/summary.txt; tail -3 gpurun_out/$name.log; }
run k3 120 "tests/test_net_gpu.py::test_conv_tc_mini_nets[k3_c64]"
run k5 120 "tests/test_net_gpu.py::test_conv_tc_mini_nets[k5_c16]"
FCNOD_HALO_BASEOFF=1 run k3_baseoff 120 "tests/test_net_gpu.py::test_conv_tc_mini_nets[k3_c64]"
run minis 600 tests/test_net_gpu.py -k "mini or s2d"
python tools/profile_forward.py 16 1 > gpurun_out/prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_tc2 -s 1 -c 2 -o gpurun_out/prof_tc2 -f python tools/profile_forward.py 16 1 > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_full.log
