#!/bin/bash
# ncu --set full capture (with source) of a few launches of the kernels matching a regex, bench command
#   usage: tools/gpu_ncu_kernel.sh <kernel regex> <launch-skip> <launch-count> <out name>
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k "regex:$1" -s $2 -c $3 \
    -o gpurun_out/$4 -f $CMD > gpurun_out/ncu_$4.log 2>&1
echo "capture rc=$?"; tail -3 gpurun_out/ncu_$4.log
