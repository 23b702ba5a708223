#!/bin/bash
# ncu evidence for the bench command: (1) launch list with device time per launch, (2) one
# --set full capture of every kernel of one training step.  Plain run first (must exit 0).
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:^(composite|mean_bwd|mlp_|pack_weights|project_gather|readout_bwd|sample_|tile_colsum)" -c 400 --csv \
    --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:^(composite|mean_bwd|mlp_|pack_weights|project_gather|readout_bwd|sample_|tile_colsum)" -s 120 -c 40 \
    -o gpurun_out/prof_step $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la gpurun_out/ | head -30
tail -3 gpurun_out/ncu_full.log
