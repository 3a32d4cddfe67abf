#!/bin/bash
CF_DEV_LIB=1 CF_MULTI_TIMES=1 timeout 300 python scripts/time_multi.py 1e8 0.05 3 8 2>&1 | tail -12
nvidia-smi topo -m 2>&1 | head -12
