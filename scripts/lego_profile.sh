set -x
python scripts/time_lego_render.py 640000 gpurun_out/lego_now.json > gpurun_out/lego_now.log 2>&1
tail -3 gpurun_out/lego_now.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/lego_launches.csv python scripts/time_lego_render.py 8192 > gpurun_out/lego_ncu.log 2>&1
python scripts/launch_summary.py gpurun_out/lego_launches.csv 40
