#!/bin/bash
# round 2 evidence: ncu --set full of one config-A view (every library kernel, source-level for the blends), launch lists
# of three views and of the bench command, one config-C view's launch list
TAG=${1:-r2n}
mkdir -p gpurun_out
python profiles/one_view.py 3 > gpurun_out/one_view.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches_A.csv python profiles/one_view.py 3 > gpurun_out/ncu_l.log 2>&1
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'raster_|project_sh|onesweep|coarse|expand|rs_hist|tile_|rs_scan|cam_pack' -c 24 -o gpurun_out/${TAG}_full_A python profiles/one_view.py 1 > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
python bench.py --steps 2 --warmup 3 --no-reference-trainer > gpurun_out/${TAG}_bench_s2.json 2> gpurun_out/${TAG}_bench_s2.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-reference-trainer > gpurun_out/ncu_lb.log 2>&1; echo "ncu bench launch list rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/${TAG}_launches_C.csv python profiles/one_view.py 2 C > gpurun_out/ncu_lc.log 2>&1
ls -la gpurun_out/*.ncu-rep
