#!/bin/bash
# 2 GPUs: second-device test, bench with / without the flat gradient buffer
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_streams.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/r2i_tests.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2i_tests.log
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$RUN --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2i_bench_A_n2.json 2> gpurun_out/r2i_bench_A_n2.err; echo "bench rc=$?"
$RUN --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 --no-flat-grads > gpurun_out/r2i_bench_A_n2_five.json 2> gpurun_out/r2i_bench_A_n2_five.err; echo "bench five rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r2i_bench_A_n2.json", "gpurun_out/r2i_bench_A_n2_five.json"):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 1), round(d["ms_per_step"], 3), d["multi_gpu"], "store", round(d["f1_store_path"]["value"], 1), "train", round(d["train_step"]["it_per_s"], 1), "e2e", round(d["e2e"]["value"], 1))
    except Exception as e:
        print(f, "failed", e)
PY
