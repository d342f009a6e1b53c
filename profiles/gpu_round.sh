set -x
nproc; free -g | head -2
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_A.json 2> gpurun_out/bench_A.err; echo "bench rc=$?"
cat gpurun_out/bench_A.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>&1; cat gpurun_out/bench_ref.json
python profiles/one_view.py 2 > gpurun_out/one_view.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_A.csv python profiles/one_view.py 2 > gpurun_out/ncu_l.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'raster_|project_sh|onesweep|emit|rs_hist' -c 16 -o gpurun_out/full_A python profiles/one_view.py 1 > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
python bench.py --config C --steps 5 --warmup 3 --views-per-rank 1 > gpurun_out/bench_C.json 2> gpurun_out/bench_C.err; cat gpurun_out/bench_C.json
ls -la gpurun_out
