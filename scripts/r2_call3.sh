#!/bin/bash
# GPU call: parity suite, default bench line, reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest3.log
tail -6 gpurun_out/r2_pytest3.log
timeout 900 python bench.py > gpurun_out/r2_bench3.json 2> gpurun_out/r2_bench3.err; echo "bench rc=$?"
cat gpurun_out/r2_bench3.json; tail -5 gpurun_out/r2_bench3.err
timeout 600 python bench.py --impl reference > gpurun_out/r2_bench3_ref.json 2> gpurun_out/r2_bench3_ref.err; echo "ref rc=$?"
cat gpurun_out/r2_bench3_ref.json
which gfortran flang nvfortran ifx ifort f2c 2>&1 | head; echo "fortran probe rc=$?"
nproc
