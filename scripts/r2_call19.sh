#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest19.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_pytest19.log
timeout 300 python scripts/time_tree.py 1e8 0.05 3
timeout 300 python scripts/time_tree.py 1e6 0.05 3
