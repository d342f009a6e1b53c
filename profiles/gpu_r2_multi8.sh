#!/bin/bash
# round 2, 8 GPUs: hardware check, bench A weak scaling (balanced / plain), config B strong scaling
N=${1:-8}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$RUN --master-port 29511 profiles/multi_gpu_check.py > gpurun_out/r2_multi_check_n$N.json 2> gpurun_out/r2_multi_check_n$N.err; echo "check rc=$?"
$RUN --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench_A_n$N.json 2> gpurun_out/r2_bench_A_n$N.err; echo "bench rc=$?"
$RUN --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 --no-balance > gpurun_out/r2_bench_A_n${N}_plain.json 2> gpurun_out/r2_bench_A_n${N}_plain.err; echo "bench plain rc=$?"
$RUN --master-port 29514 bench.py --gpus $N --config B --total-views 64 --steps 5 --warmup 3 > gpurun_out/r2_bench_B_n$N.json 2> gpurun_out/r2_bench_B_n$N.err; echo "bench B rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r2_multi_check_n$N.json", "gpurun_out/r2_bench_A_n$N.json", "gpurun_out/r2_bench_A_n${N}_plain.json", "gpurun_out/r2_bench_B_n$N.json"):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        if "all_ranks" in d:
            print(f, d["all_ranks"], d["ok"])
        else:
            print(f, round(d["value"], 1), round(d["ms_per_step"], 3), d["multi_gpu"], "store", round(d["f1_store_path"]["value"], 1), "train", round(d["train_step"]["it_per_s"], 1), "e2e", round(d["e2e"]["value"], 1))
    except Exception as e:
        print(f, "failed", e)
PY
