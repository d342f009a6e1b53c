#!/bin/bash
# round 2: full -m gpu suite with the measured parity figures printed (-rA), the avg_pool2d probe, smoke()
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2_tests_gpu.txt 2>&1
python -c "import os; print('cpus', os.cpu_count())" >> gpurun_out/r2_tests_gpu.txt
python profiles/avgpool_probe.py > gpurun_out/r2_avgpool.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -rA --durations=25 > gpurun_out/r2_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.txt 2>&1
tail -5 gpurun_out/r2_tests.log
