#!/bin/bash
# generic round-2 iteration: GPU tests (fail fast), bench A + C short, launch list A
TAG=${1:-it}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/${TAG}_tests.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/${TAG}_tests.log
tail -3 gpurun_out/${TAG}_tests.log
python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_A.json 2> gpurun_out/${TAG}_bench_A.err; echo "bench A rc=$?"
python bench.py --config C --views-per-rank 1 --steps 6 --warmup 3 --no-reference-trainer > gpurun_out/${TAG}_bench_C.json 2> gpurun_out/${TAG}_bench_C.err; echo "bench C rc=$?"
python profiles/config_d.py > gpurun_out/${TAG}_config_d.json 2> gpurun_out/${TAG}_config_d.err; echo "config D rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches_A.csv python profiles/one_view.py 3 > gpurun_out/ncu_l.log 2>&1
python profiles/summarize.py launches gpurun_out/${TAG}_launches_A.csv | head -30
python - <<PY
import json
for cfg in ("A", "C"):
    try:
        d = json.loads([l for l in open(f"gpurun_out/${TAG}_bench_{cfg}.json") if l.startswith("{")][-1])
        st = d["roofline"].get("stages_one_stream_ms") or {k: v["ms"] for k, v in d["roofline"]["stages"].items()}
        print(cfg, "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), "one-stream", round(d["ms_per_step_one_stream"], 3),
              "e2e", round(d["e2e"]["value"], 1), "store", round(d["f1_store_path"]["value"], 1), "train it/s", round(d["train_step"]["it_per_s"], 1),
              "ref trainer", d.get("reference_trainer"))
        print("   ", st)
    except Exception as e:
        print(cfg, "failed", e)
PY
tail -5 gpurun_out/${TAG}_config_d.json
