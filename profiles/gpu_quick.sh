python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py > gpurun_out/bench_A2.json 2> gpurun_out/bench_A2.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_A2.err
python profiles/show_bench.py gpurun_out/bench_A2.json
