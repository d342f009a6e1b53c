#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 profiles/multimem_adam_check.py > gpurun_out/r2_mmadam_n$N.json 2> gpurun_out/r2_mmadam_n$N.err; echo "rc=$?"
grep '^{' gpurun_out/r2_mmadam_n$N.json; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/r2_mmadam_n$N.err | tail -15
