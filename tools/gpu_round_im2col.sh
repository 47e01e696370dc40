#!/bin/bash
# Short GPU visit: trainer / GEMM / no-alloc tests, then the training bench with the generic and the staged conv1 im2col.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_trainer_gpu.py tests/test_gemm_tc_gpu.py tests/test_zz_no_alloc.py -q -x -m gpu --no-header -p no:cacheprovider 2>&1 | tail -6
FCNOD_TRAIN_IM2COL_C3=0 timeout 150 python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02zzz_train_generic_im2col.json 2> gpurun_out/r02zzz_err0.txt
echo "generic rc=$?"
timeout 150 python bench.py --workload train --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02zzz_train_staged_im2col.json 2> gpurun_out/r02zzz_err1.txt
echo "staged rc=$?"
python - <<'PY'
import json
for k in ('generic', 'staged'):
    try:
        d = json.loads(open('gpurun_out/r02zzz_train_%s_im2col.json' % k).read().strip().splitlines()[-1])
        print(k, 'ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'])
    except Exception as e:
        print(k, 'failed', e)
PY
tail -2 gpurun_out/r02zzz_err1.txt
