#!/bin/bash
mkdir -p gpurun_out
for v in 73 74 75; do
  echo "== CF_TREE_EVAL=$v"
  CF_DEV_LIB=1 CF_TREE_EVAL=$v timeout 300 python scripts/time_tree.py 1e8 0.05 3
done > gpurun_out/r2_tune14.log 2>&1
cat gpurun_out/r2_tune14.log
timeout 600 python -m pytest tests/test_gpu_tree.py -x -q 2>&1 | tail -3
