#!/bin/bash
# Short GPU visit: trainer tests, the training bench (default, and the ReLU / bias-gradient pass with smaller blocks), and one
# ncu --set full capture of the backward pass's bandwidth kernels.
tag=${1:-r02zzz}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_trainer_gpu.py -q -x -m gpu --no-header -p no:cacheprovider 2>&1 | tail -4
for v in default 64 128; do
  if [ $v = default ]; then unset FCNOD_TRAIN_RB_CHUNK; else export FCNOD_TRAIN_RB_CHUNK=$v; fi
  timeout 150 python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_train_rb_$v.json 2> gpurun_out/${tag}_err_$v.txt
  echo "rb chunk $v rc=$?"
done
unset FCNOD_TRAIN_RB_CHUNK
python - <<PY
import json
for k in ('default', '64', '128'):
    try:
        d = json.loads(open('gpurun_out/${tag}_train_rb_%s.json' % k).read().strip().splitlines()[-1])
        print(k, 'ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'])
    except Exception as e:
        print(k, 'failed', e)
PY
timeout 200 ncu --set full --clock-control none --profile-from-start off -k regex:"relu_bias8|pool_bwd8|lrn_bwd8|col2im|im2col_c3" -c 100 -o gpurun_out/${tag}_bwd_bw -f \
  python tools/profile_train.py bf16 > gpurun_out/${tag}_ncu_bwd_bw.log 2>&1
echo "ncu rc=$?"
python tools/ncu_summary.py gpurun_out/${tag}_bwd_bw.ncu-rep gpurun_out/${tag}_train_backward_bandwidth_kernels_ncu_full.csv && rm -f gpurun_out/${tag}_bwd_bw.ncu-rep
