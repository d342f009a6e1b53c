python profiles/fwd_only.py 2 > gpurun_out/fwd_only.log 2>&1 || { tail -5 gpurun_out/fwd_only.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'onesweep|bin_emit|rs_hist|project_sh_fwd' -c 9 -o gpurun_out/full_sort python profiles/fwd_only.py 1 > gpurun_out/ncu_s.log 2>&1; echo "ncu rc=$?"
