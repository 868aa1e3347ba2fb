set -x
mkdir -p gpurun_out
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_odd24.csv python tools/time_config.py odd24 > gpurun_out/ncu_odd24.log 2>&1
