#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tree.py tests/test_gpu_multi.py tests/test_gpu_fullsize.py tests/test_gpu_delta.py -x -q > gpurun_out/r2_pytest15.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_pytest15.log
timeout 300 python scripts/time_tree.py 1e8 0.05 3 > gpurun_out/r2_time15.log 2>&1
CF_TT_X=random timeout 300 python scripts/time_tree.py 1e8 0.05 3 >> gpurun_out/r2_time15.log 2>&1
timeout 300 python scripts/time_tree.py 1e6 0.05 3 1 >> gpurun_out/r2_time15.log 2>&1
timeout 300 python scripts/time_tree.py 2e6 0.8 5 >> gpurun_out/r2_time15.log 2>&1
cat gpurun_out/r2_time15.log
