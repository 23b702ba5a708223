#!/bin/bash
# Where the gather tile sweeps (tile_rows_kernel<0/1>) spend their time: builds the library three times -- as shipped,
# without the tile loads (CN_DIAG=1: the rows read stale shared memory) and without the row work (CN_DIAG=2: loads only)
# and without the dX loads of the backward sweep (CN_DIAG=3) -- into build_variants/ (git-ignored, travels with gpurun), then on the GPU box:
#     for d in 0 1 2 3; do cp build_variants/lib_diag$d.so thesis-clip-nerf_b200/lib/libclipnerf_b200.so; \
#         python bench.py --steps 10 --no-cpu-baseline | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); \
#         print({k: v['ms_per_step'] for k, v in d['stages'].items() if 'gather' in k})"; done
# The diagnostic builds compute WRONG results; the script restores the shipped build at the end.
set -e
cd "$(dirname "$0")/../thesis-clip-nerf_b200/csrc"
mkdir -p ../../build_variants
for d in 1 2 3; do
  make clean > /dev/null
  make EXTRA=-DCN_DIAG=$d > /dev/null
  cp ../lib/libclipnerf_b200.so ../../build_variants/lib_diag$d.so
done
make clean > /dev/null
make > /dev/null
cp ../lib/libclipnerf_b200.so ../../build_variants/lib_diag0.so
ls -la ../../build_variants
