#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_tree.py tests/test_gpu_fullsize.py tests/test_gpu_delta.py tests/test_gpu_parity.py -x -q > gpurun_out/r2_pytest10.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_pytest10.log
timeout 300 python scripts/time_tree.py 1e8 0.05 3 > gpurun_out/r2_time10.log 2>&1
timeout 300 python scripts/time_tree.py 1e6 0.05 3 >> gpurun_out/r2_time10.log 2>&1
timeout 300 python scripts/time_tree.py 2e6 0.8 5 >> gpurun_out/r2_time10.log 2>&1
timeout 300 python scripts/time_tree.py 1e7 0.001 3 >> gpurun_out/r2_time10.log 2>&1
cat gpurun_out/r2_time10.log
timeout 600 python bench.py --workload c4 --method tree --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r2_bench10.json 2> gpurun_out/r2_bench10.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench10.json').read().strip().splitlines()[-1])
print("c4", d["ms_per_step"], d["stages_ms_per_step"], d["e2e"])
PY
