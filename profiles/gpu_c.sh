python bench.py --config C --steps 5 --warmup 3 --views-per-rank 1 > gpurun_out/bench_C.json 2> gpurun_out/bench_C.err; tail -2 gpurun_out/bench_C.err
python profiles/show_bench.py gpurun_out/bench_C.json
