#!/bin/bash
# ncu --set full over the launches of ONE DNGFOptimizer.optimize_pose step (our kernels only) of tools/bench_query.py;
# raw page exported on the box and summarised.   usage: tools/gpu_profile_query.sh <out name> [launches per step]
mkdir -p gpurun_out
N=${2:-26}
CMD="python tools/bench_query.py --case optimize_pose --steps 1 --warmup 3"
RE='^(bin_|composite_|feature_grad|mlp_|pack_weights|project_|readout_|sample_|tile_rows|tile_add_rows|tile_add_tiles|grasp_|pose_points|query_tangent|generate_rays|quantize|depth_minmax)'
$CMD > gpurun_out/plain_$1.json 2> gpurun_out/plain_$1.err || { echo "plain run failed"; exit 1; }
ncu --set full --clock-control none -k "regex:$RE" -s $((3 * N)) -c $N -o /tmp/$1 -f $CMD > gpurun_out/ncu_$1.log 2>&1
echo "capture rc=$?"
ncu -i /tmp/$1.ncu-rep --page raw --csv > gpurun_out/$1_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/$1_raw.csv > gpurun_out/$1_summary.txt
gzip -f gpurun_out/$1_raw.csv
cat gpurun_out/$1_summary.txt | cut -c1-170
