python profiles/one_view.py 3 > gpurun_out/one_view.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_A.csv python profiles/one_view.py 3 > gpurun_out/ncu_l.log 2>&1
python profiles/summarize.py launches gpurun_out/launches_A.csv
