#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 --no-reference-trainer > gpurun_out/r2q_bench_A.json 2> gpurun_out/r2q_bench_A.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads([l for l in open("gpurun_out/r2q_bench_A.json") if l.startswith("{")][-1])
print("value", round(d["value"], 1), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_step"], 3), "e2e async", round(d["e2e_async_readback"]["value"], 1), round(d["e2e_async_readback"]["ms_per_step"], 3), "train", round(d["train_step"]["ms_per_step"], 3))
PY
tail -3 gpurun_out/r2q_bench_A.err
