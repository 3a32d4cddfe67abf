#!/bin/bash
# default bench under torchrun at N = $1 (what the driver runs at round end), both arms
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_final_n$N.json 2> gpurun_out/r2_final_n$N.err; echo "bench rc=$?"
tail -3 gpurun_out/r2_final_n$N.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r2_final_n$N.json').read().strip().splitlines()[-1])
print("c2", d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"], d.get("vs_single_gpu",{}).get("speedup_vs_single_gpu_same_box"))
h=d["headline_n1e8"]; print("c4", h["ms_per_step"], h["stages_ms_per_step"], h["e2e"]["ms_per_step"], h["e2e"]["h2d_bytes_per_step"], h.get("e2e_c_abi_one_process",{}).get("ms_per_step"), h.get("vs_single_gpu",{}).get("speedup_vs_single_gpu_same_box"))
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29545 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/r2_final_ref_n$N.json 2> gpurun_out/r2_final_ref_n$N.err; echo "ref rc=$?"
python -c "
import json
d=json.loads(open('gpurun_out/r2_final_ref_n$N.json').read().strip().splitlines()[-1]); print('ref', d['value'], d['cpu_baseline']['cores'], d['single_thread']['value'])"
