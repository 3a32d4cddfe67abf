#!/bin/bash
# N-GPU call (N = $1): default bench under torchrun + tree parity + timing
N=${1:-8}
mkdir -p gpurun_out
timeout 300 python scripts/time_tree.py 1e8 0.05 3 > gpurun_out/r2_time8.log 2>&1; cat gpurun_out/r2_time8.log
timeout 600 python -m pytest tests/test_gpu_tree.py tests/test_gpu_multi.py -x -q > gpurun_out/r2_pytest8.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_pytest8.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_bench8_n$N.json 2> gpurun_out/r2_bench8_n$N.err; echo "bench rc=$?"
tail -5 gpurun_out/r2_bench8_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_bench8_n$N.json').read().strip().splitlines()[-1])
print("c2", d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"], d.get("vs_single_gpu"))
h=d["headline_n1e8"]; print("c4", h["ms_per_step"], h["stages_ms_per_step"], h["e2e"], h.get("e2e_c_abi_one_process"), h.get("vs_single_gpu"))
PY
