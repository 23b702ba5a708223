#!/bin/bash
# ncu --set full over the launches of ONE training step of the bench command (our kernels only); the raw
# page is exported to CSV on the box (the .ncu-rep of 45 launches is too large to travel) and summarised.
#   usage: tools/gpu_profile_step.sh <out name> [launches per step]
mkdir -p gpurun_out
N=${2:-45}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
RE='^(bin_|composite_|feature_grad|mlp_|pack_weights|project_|readout_|sample_|tile_rows)'
$CMD > gpurun_out/plain_$1.json 2> gpurun_out/plain_$1.err || { echo "plain run failed"; exit 1; }
ncu --set full --clock-control none -k "regex:$RE" -s $((3 * N)) -c $N -o /tmp/$1 -f $CMD > gpurun_out/ncu_$1.log 2>&1
echo "capture rc=$?"
ncu -i /tmp/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/$1_raw.csv > gpurun_out/$1_summary.txt
gzip -f gpurun_out/$1_raw.csv
ls -la gpurun_out/
