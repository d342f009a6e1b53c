#!/bin/bash
# round 2, 8 GPUs, final code: fused-optimiser check, bench A default (sym flat grads + fused optimiser) and with NCCL only, config B
N=${1:-8}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$RUN --master-port 29531 profiles/multimem_adam_check.py > gpurun_out/r2_mmadam_n$N.json 2> gpurun_out/r2_mmadam_n$N.err; echo "mm check rc=$?"
$RUN --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2k_bench_A_n$N.json 2> gpurun_out/r2k_bench_A_n$N.err; echo "bench rc=$?"
$RUN --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 --flat-grads off --no-fused-optimizer > gpurun_out/r2k_bench_A_n${N}_nccl.json 2> gpurun_out/r2k_bench_A_n${N}_nccl.err; echo "bench nccl rc=$?"
$RUN --master-port 29514 bench.py --gpus $N --config B --total-views 64 --steps 5 --warmup 3 > gpurun_out/r2k_bench_B_n$N.json 2> gpurun_out/r2k_bench_B_n$N.err; echo "bench B rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r2_mmadam_n$N.json", "gpurun_out/r2k_bench_A_n$N.json", "gpurun_out/r2k_bench_A_n${N}_nccl.json", "gpurun_out/r2k_bench_B_n$N.json"):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        if "fused_multimem_step_ms" in d:
            print(f, d)
        else:
            print(f, round(d["value"], 1), round(d["ms_per_step"], 3), d["multi_gpu"]["render_ms_max"], d["multi_gpu"]["rank_spread_ms"], d["multi_gpu"]["allreduce_ms"], d["gradient_collective"][:30], "store", round(d["f1_store_path"]["value"], 1), "train", round(d["train_step"]["it_per_s"], 1), round(d["train_step"]["ms_per_step"], 3), d["train_step"]["optimizer"][:30], "e2e", round(d["e2e"]["value"], 1))
    except Exception as e:
        print(f, "failed", e)
PY
