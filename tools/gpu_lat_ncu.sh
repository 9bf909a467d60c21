ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/lat_launches.csv python tools/small_prof.py 1 > gpurun_out/lat_ncu.log 2>&1
tail -2 gpurun_out/lat_ncu.log
