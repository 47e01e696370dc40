#!/bin/bash
# One GPU visit: parity tests, bench line, ncu launch list, ncu --set full captures of every kernel family.
# usage: tools/gpu_round.sh <tag> [skip_tests]
tag=${1:-rXX}
mkdir -p gpurun_out
if [ -z "$2" ]; then
  timeout 1500 python -m pytest tests -q -m gpu -x --no-header -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1
  echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest.log
fi
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
echo "bench rc=$?"; cat gpurun_out/${tag}_bench.json
cp gpurun_out/per_launch_ms.json gpurun_out/${tag}_per_launch_ms.json 2>/dev/null
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2>> gpurun_out/${tag}_bench.err
echo "bench ref rc=$?"; cat gpurun_out/${tag}_bench_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
  python tools/profile_forward.py 64 2 > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
# all tensor-core convolutions of the second pass
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:conv_tc2 -c 44 -o gpurun_out/${tag}_conv_full -f \
  python tools/profile_forward.py 64 2 > gpurun_out/${tag}_ncu_full.log 2>&1
echo "ncu conv rc=$?"
python tools/ncu_hot.py gpurun_out/${tag}_conv_full.ncu-rep 0 44 45 > gpurun_out/${tag}_conv1_hot_sass.txt 2>&1
python tools/ncu_hot.py gpurun_out/${tag}_conv_full.ncu-rep 2 44 45 > gpurun_out/${tag}_conv2_3x3_hot_sass.txt 2>&1
python tools/ncu_summary.py gpurun_out/${tag}_conv_full.ncu-rep gpurun_out/${tag}_conv_tc2_ncu_full.csv && rm -f gpurun_out/${tag}_conv_full.ncu-rep
# bandwidth kernels of the network (s2d, pool+lrn, lrn+pool, pools) and the post-net, second pass
timeout 900 ncu --set full --clock-control none --profile-from-start off -k regex:"pool3|lrn|s2d|decode_group|assemble" -c 8 -o gpurun_out/${tag}_bw_full -f \
  python tools/profile_forward.py 64 2 > gpurun_out/${tag}_ncu_bw.log 2>&1
echo "ncu bw rc=$?"
python tools/ncu_summary.py gpurun_out/${tag}_bw_full.ncu-rep gpurun_out/${tag}_bandwidth_kernels_ncu_full.csv && rm -f gpurun_out/${tag}_bw_full.ncu-rep
# layer forward (config 2), prelude, post-net on planted maps
timeout 900 ncu --set full --clock-control none -k regex:"aug_|label_raster|pre_|decode_|assemble" -c 30 -o gpurun_out/${tag}_aux_full -f \
  python tools/profile_aux.py > gpurun_out/${tag}_ncu_aux.log 2>&1
echo "ncu aux rc=$?"; tail -3 gpurun_out/${tag}_ncu_aux.log
python tools/ncu_summary.py gpurun_out/${tag}_aux_full.ncu-rep gpurun_out/${tag}_layer_prelude_postnet_ncu_full.csv && rm -f gpurun_out/${tag}_aux_full.ncu-rep
ls -la gpurun_out | head -30
