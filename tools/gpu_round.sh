#!/bin/bash
# One GPU visit: parity tests, bench line, ncu launch list, one ncu --set full capture of the heaviest conv launches.
# usage: tools/gpu_round.sh <tag> [skip_tests]
tag=${1:-rXX}
mkdir -p gpurun_out
if [ -z "$2" ]; then
  timeout 1500 python -m pytest tests -q -m gpu -x --no-header -p no:cacheprovider > gpurun_out/${tag}_pytest.log 2>&1
  echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest.log
fi
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
echo "bench rc=$?"; cat gpurun_out/${tag}_bench.json
cp gpurun_out/per_launch_ms.json gpurun_out/${tag}_per_launch_ms.json 2>/dev/null
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
  python tools/profile_forward.py 64 2 > gpurun_out/${tag}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_tc2 -s 40 -c 8 -o gpurun_out/${tag}_conv_full -f \
  python tools/profile_forward.py 64 2 > gpurun_out/${tag}_ncu_full.log 2>&1
echo "ncu full rc=$?"
