#!/bin/bash
mkdir -p gpurun_out
echo "== dev lib (committed)"; CF_DEV_LIB=1 timeout 300 python scripts/time_tree.py 1e8 0.05 3
echo "== release lib (working tree)"; timeout 300 python scripts/time_tree.py 1e8 0.05 3
timeout 300 python scripts/time_tree.py 1e6 0.05 3 1
timeout 300 python scripts/time_tree.py 2e6 0.8 5
timeout 900 python -m pytest tests/test_gpu_tree.py tests/test_gpu_multi.py tests/test_gpu_delta.py -x -q 2>&1 | tail -4
