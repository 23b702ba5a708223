#!/bin/bash
# ncu --set full over the launches of ONE training step of the bench command (our kernels only); the raw
# page is exported to CSV on the box (the .ncu-rep of ~47 launches is too large to travel) and summarised.
#   usage: [BENCH_FLAGS=--lowres-features] tools/gpu_profile_step.sh <out name>
# The plain run comes first (exit 0 without ncu); its gpu_launches / steps gives the launches per step.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph $BENCH_FLAGS"
RE='^(bin_|composite_|feature_grad|mlp_|pack_weights|project_|readout_|sample_|tile_rows|adam_|images_|lowres_|sweep_|grasp_)'
$CMD > gpurun_out/plain_$1.json 2> gpurun_out/plain_$1.err || { echo "plain run failed"; tail -5 gpurun_out/plain_$1.err; exit 1; }
N=$(python -c "import json;print(json.load(open('gpurun_out/plain_$1.json'))['gpu_launches'])")
echo "launches per step: $N"
ncu --set full --clock-control none -k "regex:$RE" -s $((3 * N)) -c $N -o /tmp/$1 -f $CMD > gpurun_out/ncu_$1.log 2>&1
echo "capture rc=$?"
ncu -i /tmp/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/$1_raw.csv > gpurun_out/$1_summary.txt
python tools/ncu_traffic.py gpurun_out/$1_raw.csv gpurun_out/$1_traffic.json > /dev/null
gzip -f gpurun_out/$1_raw.csv
ls -la gpurun_out/
