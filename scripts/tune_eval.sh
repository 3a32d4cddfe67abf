#!/bin/bash
# development: tree evaluator variants at the headline size and at C2
# (CF_TREE_EVAL: 5, 6 = CTAs per SM; 70 = 7 CTAs without the L1 prefetch; 7 = the default)
for v in ${VARIANTS:-5 6 70 7}; do
  echo "== CF_TREE_EVAL=$v"
  CF_TREE_EVAL=$v timeout 300 python scripts/time_tree.py 1e8 0.05 3 | grep -v "max |"
  CF_TREE_EVAL=$v timeout 120 python scripts/time_tree.py 1e6 0.05 3 | grep -v "max |"
done
