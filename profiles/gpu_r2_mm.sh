#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 profiles/multimem_check.py > gpurun_out/r2_multimem_n$N.json 2> gpurun_out/r2_multimem_n$N.err; echo "rc=$?"
grep '^{' gpurun_out/r2_multimem_n$N.json; tail -5 gpurun_out/r2_multimem_n$N.err
