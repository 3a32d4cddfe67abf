#!/bin/bash
mkdir -p gpurun_out
echo "== release lib (committed)"; timeout 300 python scripts/time_tree.py 1e8 0.05 3
echo "== dev lib (working tree)"; CF_DEV_LIB=1 timeout 300 python scripts/time_tree.py 1e8 0.05 3
echo "== release lib (committed)"; CF_TT_X=random timeout 300 python scripts/time_tree.py 1e8 0.05 3
echo "== dev lib (working tree)"; CF_TT_X=random CF_DEV_LIB=1 timeout 300 python scripts/time_tree.py 1e8 0.05 3
