#!/bin/bash
# multi-GPU visit on N GPUs of one box: one-process MultiGpuDetector test, weak- and strong-scaling bench lines.  usage: gpu_multi.sh N tag
N=$1; tag=$2
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m pytest tests/test_detect_gpu.py -q -m gpu -x --no-header -p no:cacheprovider -k "multi_gpu" 2>&1 | tail -3
run() { name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 "$@" > gpurun_out/${tag}_${name}_n${N}.json 2> gpurun_out/${tag}_${name}_n${N}.err
  echo "$name rc=$?"; python - <<PY
import json
try:
    d = [json.loads(l) for l in open("gpurun_out/${tag}_${name}_n${N}.json") if l.startswith("{")][-1]
    print('$name N=$N', 'fps', round(d['value']), 'ms/step', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['value']), 'per-gpu batch', d['config']['per_gpu_batch'], d['scaling'])
except Exception as e:
    print('no line', e)
PY
}
run weak
run strong --global-batch 256
