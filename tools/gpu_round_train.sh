#!/bin/bash
# One GPU visit for the training path: tests, bench lines, ncu launch list of one step, ncu --set full of GEMM launches.
# usage: tools/gpu_round_train.sh <tag>
tag=${1:-rXX}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --no-header -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest.log
timeout 600 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
echo "bench rc=$?"; tail -2 gpurun_out/${tag}_bench.err
cp gpurun_out/per_launch_ms.json gpurun_out/${tag}_per_launch_ms.json 2>/dev/null
timeout 300 python bench.py --workload train --steps 10 --warmup 3 > gpurun_out/${tag}_train_bench_n1.json 2> gpurun_out/${tag}_train_bench.err
echo "train bench rc=$?"; tail -2 gpurun_out/${tag}_train_bench.err
timeout 300 python bench.py --workload train --train-precision fp32 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_train_bench_fp32_n1.json 2>> gpurun_out/${tag}_train_bench.err
echo "train fp32 bench rc=$?"
timeout 300 python bench.py --workload train --impl reference --steps 2 > gpurun_out/${tag}_train_bench_ref.json 2>> gpurun_out/${tag}_train_bench.err
echo "train ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${tag}_train_step_bf16_launches.csv \
  python tools/profile_train.py bf16 > gpurun_out/${tag}_ncu_train.log 2>&1
echo "ncu launches rc=$?"
python tools/ncu_by_kernel.py gpurun_out/${tag}_train_step_bf16_launches.csv > gpurun_out/${tag}_train_step_bf16_by_kernel.txt
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"gemm_tc|kgemm_group|wgrad_group" -c 10 -o gpurun_out/${tag}_gemm_fwd -f \
  python tools/profile_train.py bf16 > gpurun_out/${tag}_ncu_gemm_fwd.log 2>&1
echo "ncu gemm fwd rc=$?"
python tools/ncu_summary.py gpurun_out/${tag}_gemm_fwd.ncu-rep gpurun_out/${tag}_gemm_tc_forward_ncu_full.csv
python tools/ncu_stalls.py gpurun_out/${tag}_gemm_fwd.ncu-rep > gpurun_out/${tag}_gemm_tc_forward_stalls.txt 2>&1
rm -f gpurun_out/${tag}_gemm_fwd.ncu-rep
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"gemm_tc|kgemm_group|wgrad_group" --launch-skip 21 -c 14 -o gpurun_out/${tag}_gemm_bwd -f \
  python tools/profile_train.py bf16 > gpurun_out/${tag}_ncu_gemm_bwd.log 2>&1
echo "ncu gemm bwd rc=$?"
python tools/ncu_summary.py gpurun_out/${tag}_gemm_bwd.ncu-rep gpurun_out/${tag}_gemm_tc_backward_ncu_full.csv
python tools/ncu_stalls.py gpurun_out/${tag}_gemm_bwd.ncu-rep > gpurun_out/${tag}_gemm_tc_backward_stalls.txt 2>&1
rm -f gpurun_out/${tag}_gemm_bwd.ncu-rep
timeout 900 ncu --set full --clock-control none --profile-from-start off -k regex:"pool_|lrn_|im2col|col2im|relu_bias|adam|narrow" -c 16 -o gpurun_out/${tag}_train_bw -f \
  python tools/profile_train.py bf16 > gpurun_out/${tag}_ncu_train_bw.log 2>&1
echo "ncu train bw rc=$?"
python tools/ncu_summary.py gpurun_out/${tag}_train_bw.ncu-rep gpurun_out/${tag}_train_bandwidth_kernels_ncu_full.csv && rm -f gpurun_out/${tag}_train_bw.ncu-rep
ls gpurun_out | grep ${tag} | head -40
