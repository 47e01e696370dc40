#!/bin/bash
# ncu --set full + source of one POOL-mode conv launch (the 8th: inception_5a/pool_proj)
mkdir -p gpurun_out
python tools/profile_forward.py 64 2 > gpurun_out/prof_plain.log 2>&1 || { tail -5 gpurun_out/prof_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off --kernel-name-base demangled \
  -k regex:"conv_tc2_kernel<.int.4, .bool.0, .bool.0, .int.0, .bool.1>" -s ${1:-7} -c 1 -o gpurun_out/pool_full -f python tools/profile_forward.py 64 2 > gpurun_out/ncu_pool.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_pool.log
python tools/ncu_hot.py gpurun_out/pool_full.ncu-rep 0 1 60 > gpurun_out/pool_hot.txt 2>&1
ncu -i gpurun_out/pool_full.ncu-rep --page raw --csv > gpurun_out/pool_raw.csv 2>/dev/null
head -c 3000 gpurun_out/pool_hot.txt
