#!/bin/bash
# Runs every GPU test file in its own process (a device-side trap poisons the CUDA context of the
# process it happens in) with a hard timeout; logs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu_info.txt 2>&1
rc=0
for f in tests/test_gpu_probe.py tests/test_gpu_kernels.py tests/test_gpu_mlp.py tests/test_gpu_render.py tests/test_gpu_interface.py tests/test_gpu_query.py tests/test_golden.py tests/test_gpu_fullsize.py tests/test_gpu_edges.py; do
  name=$(basename $f .py)
  timeout 300 python -m pytest $f -q -m gpu --tb=short -rA -s > gpurun_out/$name.log 2>&1
  r=$?
  echo "$name exit $r" | tee -a gpurun_out/summary.txt
  tail -60 gpurun_out/$name.log
  [ $r -ne 0 ] && rc=$r
done
exit $rc
