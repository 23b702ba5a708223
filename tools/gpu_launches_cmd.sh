#!/bin/bash
# launch list (device time per launch, every kernel incl. torch's) of an arbitrary command; the plain run first
#   usage: tools/gpu_launches_cmd.sh <out name> <command ...>
mkdir -p gpurun_out
NAME=$1; shift
"$@" > gpurun_out/plain_$NAME.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$NAME.log; exit 1; }
tail -1 gpurun_out/plain_$NAME.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv \
    --log-file gpurun_out/${NAME}_launches.csv "$@" > gpurun_out/ncu_$NAME.log 2>&1
echo "launch list rc=$?"
python tools/launch_table.py gpurun_out/${NAME}_launches.csv | tee gpurun_out/${NAME}_launches_summary.txt
