python scripts/_lat.py && ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_lat.csv python scripts/_lat.py > gpurun_out/ncu_lat.log 2>&1
tail -2 gpurun_out/ncu_lat.log
