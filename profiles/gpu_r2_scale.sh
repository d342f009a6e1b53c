#!/bin/bash
# final-code multi-GPU bench lines: config A weak scaling (driver's launch form) and, at 8 ranks, config B strong scaling
N=${1:-8}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$RUN --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2s_bench_A_n$N.json 2> gpurun_out/r2s_bench_A_n$N.err; echo "bench A rc=$?"
if [ "$N" = "8" ]; then
$RUN --master-port 29514 bench.py --gpus $N --config B --total-views 64 --steps 5 --warmup 3 > gpurun_out/r2s_bench_B_n$N.json 2> gpurun_out/r2s_bench_B_n$N.err; echo "bench B rc=$?"
fi
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/r2s_bench_*_n$N.json")):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 1), round(d["ms_per_step"], 3), d["multi_gpu"]["render_ms_max"], d["multi_gpu"]["rank_spread_ms"], d["multi_gpu"]["allreduce_ms"], "e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_step"], 3), "store", round(d["f1_store_path"]["value"], 1), round(d["f1_store_path"]["ms_per_step"], 3), "train", round(d["train_step"]["it_per_s"], 1), round(d["train_step"]["ms_per_step"], 3), d["gradient_collective"])
    except Exception as e:
        print(f, "failed", e)
PY
