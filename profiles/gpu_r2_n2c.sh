#!/bin/bash
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$RUN --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 5 --flat-grads sym > gpurun_out/r2j_bench_A_n2_sym.json 2> gpurun_out/r2j_bench_A_n2_sym.err; echo "bench rc=$?"
$RUN --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2j_bench_A_n2.json 2> gpurun_out/r2j_bench_A_n2.err; echo "bench default rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r2j_bench_A_n2_sym.json", "gpurun_out/r2j_bench_A_n2.json"):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 1), round(d["ms_per_step"], 3), d["multi_gpu"]["allreduce_ms"], d["gradient_collective"][:40], "store", round(d["f1_store_path"]["value"], 1), "train", round(d["train_step"]["it_per_s"], 1), d["train_step"]["optimizer"][:60], "e2e", round(d["e2e"]["value"], 1))
    except Exception as e:
        print(f, "failed", e)
PY
grep -n "Error\|Traceback" -A5 gpurun_out/r2j_bench_A_n2_sym.err | head -20
