python profiles/one_view.py 2 > gpurun_out/one_view.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'raster_' -c 2 -o gpurun_out/full_raster python profiles/one_view.py 1 > gpurun_out/ncu_f.log 2>&1; echo "ncu full rc=$?"
