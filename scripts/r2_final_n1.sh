#!/bin/bash
# what the driver runs at round end on one GPU: both arms, default flags; then the final launch lists / captures
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference > gpurun_out/r2_final_ref_n1.json 2> gpurun_out/r2_final_ref_n1.err; echo "ref rc=$?"
timeout 900 python bench.py > gpurun_out/r2_final_n1.json 2> gpurun_out/r2_final_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_final_n1.json').read().strip().splitlines()[-1])
print("c2", d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"], d["cpu_baseline"]["value"], d["cpu_baseline"]["cores"])
h=d["headline_n1e8"]; print("c4", h["ms_per_step"], h["stages_ms_per_step"], h["e2e"]["ms_per_step"])
print("sample", d["c4_direct_sample"]["roofline"]["frac"], d["c4_direct_sample"]["oracle_check"])
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_c4_tree.csv python bench.py --workload c4 --method tree --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r2_ncu20b.log 2>&1; echo "list c4 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tree_eval -s 5 -c 1 -o gpurun_out/r02_tree_eval_c4 python bench.py --workload c4 --method tree --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r2_ncu20c.log 2>&1; echo "full eval rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tree_block -s 5 -c 1 -o gpurun_out/r02_tree_block python bench.py --workload c4 --method tree --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r2_ncu20d.log 2>&1; echo "full block rc=$?"
