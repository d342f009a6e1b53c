#!/bin/bash
# round 2: -m gpu suite + the default bench line (config A, N=1) + reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2b_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b_tests.log
tail -3 gpurun_out/r2b_tests.log
python bench.py > gpurun_out/r2b_bench_A.json 2> gpurun_out/r2b_bench_A.err
echo "bench rc=$?"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2b_bench_ref.json 2> gpurun_out/r2b_bench_ref.err
echo "ref rc=$?"
