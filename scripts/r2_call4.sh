#!/bin/bash
# GPU call: evaluator variants and chain geometries (DEV lib), launch list + full capture of the n=1e8 tree smooth
mkdir -p gpurun_out
for v in 72 71; do
  echo "== CF_TREE_EVAL=$v"
  CF_DEV_LIB=1 CF_TREE_EVAL=$v timeout 300 python scripts/time_tree.py 1e8 0.05 3
done > gpurun_out/r2_tune_eval.log 2>&1
for b in 8,8 16 15 14 12,4 13,3; do
  echo "== CF_TREE_BITS=$b"
  CF_DEV_LIB=1 CF_TREE_BITS=$b timeout 300 python scripts/time_tree.py 1e8 0.05 3
done >> gpurun_out/r2_tune_eval.log 2>&1
cat gpurun_out/r2_tune_eval.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_c4_tree_a.csv python scripts/time_tree.py 1e8 0.05 3 > gpurun_out/r2_ncu_a.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tree_eval -s 5 -c 1 -o gpurun_out/r02_tree_eval_a python scripts/time_tree.py 1e8 0.05 3 > gpurun_out/r2_ncu_b.log 2>&1
echo "ncu full rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:residual_hist0 -s 2 -c 1 -o gpurun_out/r02_hist0_a python scripts/time_tree.py 1e8 0.05 3 > gpurun_out/r2_ncu_c.log 2>&1
echo "ncu full2 rc=$?"
