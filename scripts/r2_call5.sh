#!/bin/bash
# 2-GPU call: multi-GPU tests, default bench under torchrun, reference arm under torchrun
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/r2_pytest5.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest5.log
tail -30 gpurun_out/r2_pytest5.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_bench5_n2.json 2> gpurun_out/r2_bench5_n2.err; echo "bench rc=$?"
cat gpurun_out/r2_bench5_n2.json; tail -5 gpurun_out/r2_bench5_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29545 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2_bench5_ref_n2.json 2> gpurun_out/r2_bench5_ref_n2.err; echo "ref rc=$?"
cut -c1-600 gpurun_out/r2_bench5_ref_n2.json
