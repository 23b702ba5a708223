#!/bin/bash
# ncu --set full capture of selected kernels of one training step of the bench command.
#   usage: tools/gpu_profile.sh <kernel regex> <launch-skip> <launch-count> <out name>
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$1" -s $2 -c $3 \
    -o gpurun_out/$4 -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
tail -3 gpurun_out/ncu_full.log
