#!/bin/bash
# development: tree evaluator variants (CF_TREE_EVAL) at the headline size and at C2
for v in ${VARIANTS:-0 1 4 5 6 7}; do
  echo "== CF_TREE_EVAL=$v"
  CF_TREE_EVAL=$v timeout 300 python scripts/time_tree.py 1e8 0.05 3 | grep -v "max |"
  CF_TREE_EVAL=$v timeout 120 python scripts/time_tree.py 1e6 0.05 3 | grep -v "max |"
done
