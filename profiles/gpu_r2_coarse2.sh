#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_streams.py -m gpu -q -x > gpurun_out/r2g_tests.log 2>&1
echo "pytest rc=$?" | tee -a gpurun_out/r2g_tests.log
tail -3 gpurun_out/r2g_tests.log
python bench.py --steps 10 --warmup 3 --no-reference-trainer > gpurun_out/r2g_bench_A.json 2> gpurun_out/r2g_bench_A.err; echo "bench A rc=$?"
python bench.py --config C --views-per-rank 1 --steps 6 --warmup 3 --no-reference-trainer > gpurun_out/r2g_bench_C.json 2> gpurun_out/r2g_bench_C.err; echo "bench C rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2g_launches_A.csv python profiles/one_view.py 3 > gpurun_out/ncu_l.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2g_launches_C.csv python profiles/one_view.py 2 C > gpurun_out/ncu_lc.log 2>&1
python profiles/summarize.py launches gpurun_out/r2g_launches_A.csv | head -24
python profiles/summarize.py launches gpurun_out/r2g_launches_C.csv | head -24
ncu --set full --clock-control none --import-source on --kernel-name regex:"emit_coarse|expand_kernel|coarse_prep" -c 3 -o gpurun_out/r2g_coarse_A python profiles/one_view.py 1 > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out/*.ncu-rep
