# full ncu capture of one config-A view (all library kernels) + launch list of three views
python profiles/one_view.py 3 > gpurun_out/one_view.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_A.csv python profiles/one_view.py 3 > gpurun_out/ncu_l.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'raster_|project_sh|onesweep|emit|rs_hist|tile_prep|bin_|rs_scan|cam_pack' -c 20 -o gpurun_out/full_A python profiles/one_view.py 1 > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
# launch list of the bench command itself (first 400 launches = the device leg's warm-up and timed steps)
python bench.py --steps 2 --warmup 1 > gpurun_out/bench_s2.json 2> gpurun_out/bench_s2.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_lb.log 2>&1; echo "ncu bench launch list rc=$?"
python bench.py --config C --steps 5 --warmup 3 --views-per-rank 1 > gpurun_out/bench_C.json 2> gpurun_out/bench_C.err; python profiles/show_bench.py gpurun_out/bench_C.json
