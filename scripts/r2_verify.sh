#!/bin/bash
# last check of the round on one GPU: what the driver runs (gpu tests, smoke, both bench arms)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_verify_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_verify_pytest.log
python -c "import __graft_entry__ as g; g.smoke()"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_verify_ref.json 2> gpurun_out/r2_verify_ref.err; echo "ref rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_verify_bench.json 2> gpurun_out/r2_verify_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_verify_bench.json').read().strip().splitlines()[-1])
print("c2", d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"], d["clocks"])
h=d["headline_n1e8"]; print("c4", h["ms_per_step"], h["stages_ms_per_step"], h["e2e"]["ms_per_step"])
r=json.loads(open('gpurun_out/r2_verify_ref.json').read().strip().splitlines()[-1])
print("ref", r["value"], r["ms_per_step"], r["timed_region"])
PY
