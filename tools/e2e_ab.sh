#!/bin/bash
# A/B of the streaming e2e path: default vs FCNOD_NO_SLOT_PRELUDE=1, three runs each
for i in 1 2 3; do
  for v in "" "FCNOD_NO_SLOT_PRELUDE=1"; do
    env $v timeout 300 python bench.py --steps 20 --warmup 3 --no-extras --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=[json.loads(l) for l in sys.stdin if l.startswith('{')][-1]; print('${v:-default}', 'fps', round(d['value']), 'e2e', round(d['e2e']['value']), 'one-call', round(d['e2e']['one_call_per_batch']['value']))"
  done
done
