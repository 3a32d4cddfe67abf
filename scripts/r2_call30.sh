#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "overlapped" 2>&1 | tail -5
