#!/bin/bash
# final ncu captures of the round (each after the same command ran clean without ncu)
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 3 --no-headline --no-cpu-baseline --no-e2e > gpurun_out/r2_b11a.json 2>gpurun_out/r2_b11a.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_c2.csv python bench.py --steps 1 --warmup 3 --no-headline --no-cpu-baseline --no-e2e > gpurun_out/r2_ncu11a.log 2>&1; echo "list c2 rc=$?"
python bench.py --workload c4 --method tree --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r2_b11b.json 2>gpurun_out/r2_b11b.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_c4_tree.csv python bench.py --workload c4 --method tree --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r2_ncu11b.log 2>&1; echo "list c4 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tree_eval -s 5 -c 1 -o gpurun_out/r02_tree_eval_c4 python bench.py --workload c4 --method tree --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r2_ncu11c.log 2>&1; echo "full eval rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:residual_hist0 -s 2 -c 1 -o gpurun_out/r02_hist0_c4 python bench.py --workload c4 --method tree --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r2_ncu11d.log 2>&1; echo "full hist0 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:windows_kernel -s 1 -c 1 -o gpurun_out/r02_windows_c4 python bench.py --workload c4 --method tree --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r2_ncu11e.log 2>&1; echo "full windows rc=$?"
cat gpurun_out/r2_b11a.json | cut -c1-300
