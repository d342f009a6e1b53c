#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$RUN --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 --flat-grads arena > gpurun_out/r2w_bench_A_n${N}_arena.json 2> gpurun_out/r2w_bench_A_n${N}_arena.err; echo "bench arena rc=$?"
$RUN --master-port 29513 bench.py --gpus $N --steps 20 --warmup 5 --flat-grads off > gpurun_out/r2w_bench_A_n${N}_off.json 2> gpurun_out/r2w_bench_A_n${N}_off.err; echo "bench off rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r2w_bench_A_n${N}_arena.json", "gpurun_out/r2w_bench_A_n${N}_off.json"):
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, round(d["value"], 1), round(d["ms_per_step"], 3), "one", round(d["ms_per_step_one_stream"], 3), d["multi_gpu"]["render_ms_max"], d["multi_gpu"]["allreduce_ms"], "e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_step"], 3), d.get("gradient_arena_hits"))
    except Exception as e:
        print(f, "failed", e)
PY
grep -n "Error\|Traceback" -A6 gpurun_out/r2w_bench_A_n${N}_arena.err | head -20
