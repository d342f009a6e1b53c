#!/bin/bash
mkdir -p gpurun_out
for v in "" items8; do
  if [ -n "$v" ]; then export B200_LIB=$PWD/3d-gaussian-splatting_b200/libb200splat_$v.so; else unset B200_LIB; fi
  timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "radix or binning or two_level" > gpurun_out/r2v_tests_$v.log 2>&1; echo "tests[$v] rc=$?"; tail -1 gpurun_out/r2v_tests_$v.log
  python bench.py --steps 10 --warmup 3 --no-reference-trainer > gpurun_out/r2v_bench_A_$v.json 2> gpurun_out/r2v_bench_A_$v.err
  python bench.py --config C --views-per-rank 1 --steps 6 --warmup 3 --no-reference-trainer > gpurun_out/r2v_bench_C_$v.json 2> gpurun_out/r2v_bench_C_$v.err
done
python - <<PY
import json
for v in ("", "items8"):
    for cfg in ("A", "C"):
        try:
            d = json.loads([l for l in open(f"gpurun_out/r2v_bench_{cfg}_{v}.json") if l.startswith("{")][-1])
            st = d["roofline"].get("stages_one_stream_ms") or {k: x["ms"] for k, x in d["roofline"]["stages"].items()}
            print(v or "items16", cfg, round(d["value"], 1), round(d["ms_per_step"], 3), "sort_depth", st["sort_depth"], "bin_coarse", st["bin_coarse"])
        except Exception as e:
            print(v, cfg, "failed", e)
PY
