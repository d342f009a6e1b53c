#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r2y_tests.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2y_tests.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2y_bench_A.json 2> gpurun_out/r2y_bench_A.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads([l for l in open("gpurun_out/r2y_bench_A.json") if l.startswith("{")][-1])
print(round(d["value"], 1), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), "ref trainer", d["reference_trainer_it_per_s"], d["reference_trainer"].get("unavailable"))
PY
