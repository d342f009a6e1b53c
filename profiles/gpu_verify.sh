# Round-end verification on one B200: every GPU test, the smoke entry point, the default bench line, the reference arm.
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/bench_A.json 2> gpurun_out/bench_A.err; echo "bench rc=$?"; python profiles/show_bench.py gpurun_out/bench_A.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2>gpurun_out/bench_ref.err; cut -c1-400 gpurun_out/bench_ref.json
