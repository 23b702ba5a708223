#!/bin/bash
# pose-query benches (optimize_pose graph replay / eager, train_step) + the query GPU tests
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_head.py tests/test_gpu_query.py -m gpu -q --tb=short > gpurun_out/query_tests.log 2>&1
echo "query tests rc=$?"; tail -5 gpurun_out/query_tests.log
timeout 600 python tools/bench_query.py > gpurun_out/bench_query.json 2> gpurun_out/bench_query.err
echo "bench_query rc=$?"; tail -3 gpurun_out/bench_query.err; cut -c1-400 gpurun_out/bench_query.json
