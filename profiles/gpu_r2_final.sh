#!/bin/bash
# round 2 evidence on ONE B200, final code: GPU tests with the measured parity figures, smoke, the default bench line and
# the reference arm exactly as the driver launches them, config C, config D, ncu launch lists + one --set full capture
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,driver_version --format=csv > gpurun_out/r2z_gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -rA --durations=15 > gpurun_out/r2z_gpu_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2z_gpu_tests.log
tail -3 gpurun_out/r2z_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2z_smoke.txt 2>&1; echo "smoke rc=$?"
python bench.py --impl reference > gpurun_out/r2z_bench_reference_arm.json 2> gpurun_out/r2z_bench_reference_arm.err; echo "reference arm rc=$?"
python bench.py > gpurun_out/r2z_bench_A.json 2> gpurun_out/r2z_bench_A.err; echo "bench A rc=$?"
python bench.py --config C --views-per-rank 1 --steps 8 --warmup 3 --no-reference-trainer > gpurun_out/r2z_bench_C.json 2> gpurun_out/r2z_bench_C.err; echo "bench C rc=$?"
python profiles/config_d.py > gpurun_out/r2z_config_d.json 2> gpurun_out/r2z_config_d.err; echo "config D rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2z_launches_A.csv python profiles/one_view.py 3 > gpurun_out/ncu_l.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'raster_|project_sh|onesweep|coarse|expand|rs_hist|tile_|rs_scan|cam_pack' -c 24 -o gpurun_out/r2z_full_A python profiles/one_view.py 1 > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2z_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-reference-trainer > gpurun_out/ncu_lb.log 2>&1; echo "ncu bench launch list rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2z_launches_C.csv python profiles/one_view.py 2 C > gpurun_out/ncu_lc.log 2>&1
python - <<PY
import json
for cfg in ("A", "C"):
    try:
        d = json.loads([l for l in open(f"gpurun_out/r2z_bench_{cfg}.json") if l.startswith("{")][-1])
        st = d["roofline"].get("stages_one_stream_ms") or {k: v["ms"] for k, v in d["roofline"]["stages"].items()}
        print(cfg, "value", round(d["value"], 1), "ms/step", round(d["ms_per_step"], 3), "one-stream", round(d["ms_per_step_one_stream"], 3),
              "e2e", round(d["e2e"]["value"], 1), "store", round(d["f1_store_path"]["value"], 1), "train it/s", round(d["train_step"]["it_per_s"], 1),
              "ref trainer it/s", d.get("reference_trainer_it_per_s"), "launches", d["gpu_launches"], d["clocks"])
        print("   ", st)
    except Exception as e:
        print(cfg, "failed", e)
r = json.loads([l for l in open("gpurun_out/r2z_bench_reference_arm.json") if l.startswith("{")][-1])
print("reference arm", round(r["value"], 2), r["steps"], r["warmup"], round(r["ms_per_step"], 1), r["cpu_baseline"]["cores"])
PY
