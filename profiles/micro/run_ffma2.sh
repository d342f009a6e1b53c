#!/bin/bash
# Builds and runs the packed-FP32 issue-rate probe; output -> gpurun_out/ffma2_probe.txt
set -e
cd "$(dirname "$0")"
nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o /tmp/ffma2_probe ffma2_probe.cu
mkdir -p ../../gpurun_out
/tmp/ffma2_probe | tee ../../gpurun_out/ffma2_probe.txt
