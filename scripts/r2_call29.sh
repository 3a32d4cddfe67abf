#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_tree.py tests/test_abi.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python bench.py --workload c4 --method tree --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r2_bench29.json 2> gpurun_out/r2_bench29.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench29.json').read().strip().splitlines()[-1])
print("c4", d["ms_per_step"], d["stages_ms_per_step"], d["e2e"])
PY
