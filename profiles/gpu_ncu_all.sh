# full ncu capture of one config-A view (all library kernels) + launch list of three views
python profiles/one_view.py 3 > gpurun_out/one_view.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_A.csv python profiles/one_view.py 3 > gpurun_out/ncu_l.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'raster_|project_sh|onesweep|emit|rs_hist|tile_prep|bin_|rs_scan|cam_pack' -c 20 -o gpurun_out/full_A python profiles/one_view.py 1 > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
