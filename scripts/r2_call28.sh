#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/check_sharded.py > gpurun_out/r2_check_sharded.log 2>&1; echo "check rc=$?"
grep -c OK gpurun_out/r2_check_sharded.log; grep -v " OK" gpurun_out/r2_check_sharded.log | tail -20; grep "delta=" gpurun_out/r2_check_sharded.log | grep -v "delta=0:" | head -12 | cut -c1-220
timeout 600 python -m pytest tests/test_gpu_delta.py -x -q 2>&1 | tail -3
