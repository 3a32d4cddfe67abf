#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -4
CF_DEV_LIB=0 timeout 300 python scripts/time_multi.py 1e8 0.05 3 2 2>&1 | tail -3
