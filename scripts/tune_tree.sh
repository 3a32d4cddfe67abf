#!/bin/bash
# times the tree evaluator's launch variants (env CF_FIT_VARIANT, csrc/tree.cu) on C2 and C4
for v in ${VARIANTS:-0 1 2 3 4 5 6}; do
  echo "== variant $v"
  CF_FIT_VARIANT=$v python scripts/time_tree.py 1e6 0.05 3 | tail -1
  CF_FIT_VARIANT=$v python scripts/time_tree.py 1e8 0.05 3 | tail -1
done
