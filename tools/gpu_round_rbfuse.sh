#!/bin/bash
# Short GPU visit: the per-Concat ReLU / bias-gradient pass against the per-convolution one (tests, then the training bench).
mkdir -p gpurun_out
FCNOD_TRAIN_RB_FUSE=1 timeout 100 python -m pytest tests/test_trainer_gpu.py -q -x -m gpu --no-header -p no:cacheprovider -k "per_concat or bf16_tensor" 2>&1 | tail -4
for v in 0 1; do
  FCNOD_TRAIN_RB_FUSE=$v timeout 40 python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02zzz_train_rbfuse_$v.json 2> gpurun_out/r02zzz_err_rbfuse_$v.txt
  echo "fuse $v rc=$?"
done
python - <<PY
import json
for k in ('0', '1'):
    try:
        d = json.loads(open('gpurun_out/r02zzz_train_rbfuse_%s.json' % k).read().strip().splitlines()[-1])
        print(k, 'ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])
    except Exception as e:
        print(k, 'failed', e)
PY
