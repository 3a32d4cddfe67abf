#!/bin/bash
# GPU call 2: parity suite, evaluator variants (DEV lib), launch list of the n=1e8 tree smooth
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest2.log
tail -4 gpurun_out/r2_pytest2.log
for v in 72 62 52 71; do
  echo "== CF_TREE_EVAL=$v"
  CF_DEV_LIB=1 CF_TREE_EVAL=$v timeout 300 python scripts/time_tree.py 1e8 0.05 3
  CF_DEV_LIB=1 CF_TREE_EVAL=$v timeout 120 python scripts/time_tree.py 1e6 0.05 3
done > gpurun_out/r2_tune_eval.log 2>&1
cat gpurun_out/r2_tune_eval.log
CF_TT_X=random timeout 300 python scripts/time_tree.py 1e8 0.05 3 > gpurun_out/r2_tree_random.log 2>&1; cat gpurun_out/r2_tree_random.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_c4_tree_a.csv python scripts/time_tree.py 1e8 0.05 3 > gpurun_out/r2_ncu_a.log 2>&1
echo "ncu rc=$?"
