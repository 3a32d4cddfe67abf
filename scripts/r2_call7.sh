#!/bin/bash
# 1-GPU call: full gpu suite, default bench, sanitizer runs, ncu captures of the round
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest7.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest7.log
tail -12 gpurun_out/r2_pytest7.log
timeout 900 python bench.py > gpurun_out/r2_bench7.json 2> gpurun_out/r2_bench7.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench7.json').read().strip().splitlines()[-1])
print("c2", d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"])
h=d["headline_n1e8"]; print("c4", h["ms_per_step"], h["stages_ms_per_step"], h["e2e"]["ms_per_step"])
PY
tail -3 gpurun_out/r2_bench7.err
# sanitizers on a small end-to-end case (both evaluators, robustness iterations, sort, median)
cat > /tmp/san.py <<'PY'
import numpy as np, sys
sys.path.insert(0, '.')
from curvefit_b200 import datagen, lowess_smoothing, statistics
n = 6000
x = datagen.random_x(n, 1); y = datagen.noisy_y(x, 1)
for m in ("direct", "tree"):
    fit = lowess_smoothing(); fit.set_method(m); fit.initialize(x, y)
    ys = fit.smooth(0.2, nsteps=2)
    print(m, float(ys.sum()))
print("median", statistics.median(y))
PY
for tool in memcheck racecheck initcheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python /tmp/san.py > gpurun_out/r02_sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"; tail -3 gpurun_out/r02_sanitizer_$tool.log
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fit_kernel -s 4 -c 1 -o gpurun_out/r02_fit_kernel_c2 python bench.py --steps 1 --warmup 3 --no-headline --no-cpu-baseline --no-e2e > gpurun_out/r2_ncu7a.log 2>&1; echo "ncu fit rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tree_block -s 5 -c 1 -o gpurun_out/r02_tree_block python scripts/time_tree.py 1e8 0.05 3 > gpurun_out/r2_ncu7b.log 2>&1; echo "ncu block rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:compact_kernel -s 2 -c 1 -o gpurun_out/r02_compact python scripts/time_tree.py 1e8 0.05 3 > gpurun_out/r2_ncu7c.log 2>&1; echo "ncu compact rc=$?"
