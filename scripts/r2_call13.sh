#!/bin/bash
mkdir -p gpurun_out
for v in 72 73; do
  echo "== CF_TREE_EVAL=$v"
  CF_DEV_LIB=1 CF_TREE_EVAL=$v timeout 300 python scripts/time_tree.py 1e8 0.05 3
  CF_DEV_LIB=1 CF_TREE_EVAL=$v CF_TREE_BITS=8,8 timeout 300 python scripts/time_tree.py 1e8 0.05 3
  CF_DEV_LIB=1 CF_TREE_EVAL=$v timeout 120 python scripts/time_tree.py 1e6 0.05 3
done > gpurun_out/r2_tune13.log 2>&1
cat gpurun_out/r2_tune13.log
