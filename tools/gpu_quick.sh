#!/bin/bash
# quick GPU check: MLP + render parity tests, then the bench line (no CPU baseline)
mkdir -p gpurun_out
rc=0
for f in "$@"; do
  name=$(basename $f .py)
  timeout 300 python -m pytest $f -q -m gpu --tb=short -x > gpurun_out/$name.log 2>&1
  r=$?
  echo "$name exit $r"
  [ $r -ne 0 ] && { tail -40 gpurun_out/$name.log; rc=$r; }
done
[ $rc -ne 0 ] && exit $rc
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?"; tail -5 gpurun_out/bench.err
python tools/show_bench.py gpurun_out/bench.json
