#!/bin/bash
# 2-GPU check: the whole gpu suite (multi-GPU tests included) and smoke
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke()"
